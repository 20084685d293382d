"""Transport part of correctThermo (mu, alpha, rhoD) and the rates-of-progress entry (calculateRR).

CPU tests: the Cantera-made coefficient file the reference ships (tests/golden/reference_thermo_es80.npz,
see make_transport_fixture.py) pins the mechanism compiler's molecular weights and NASA-7 tables
bit for bit, and its fits drive the oracle (oracle/transport.py) through physical anchors.
GPU tests: transport_kernel and rop_kernel against the oracle through the C ABI.
"""
import os

import numpy as np
import pytest
from conftest import ROOT, golden

from dfchem_host import synth
from oracle import kinetics as K
from oracle import mech as OM
from oracle import transport as OT

MECHS = ["es80", "h2_li", "drm19", "gri30", "c7_onestep", "ethanol66"]


def ref_thermo():
    return np.load(os.path.join(ROOT, "tests", "golden", "reference_thermo_es80.npz"))


synthetic_fits = synth.synthetic_transport_fits


def test_mechanism_tables_match_the_cantera_made_file(built):
    """W, T_mid and both NASA-7 ranges of the ES80 blob equal, bit for bit, what Cantera wrote into the
    reference's thermo_ES80_H2-7-16.txt (a8, a16, a17 of SURVEY section 8)."""
    d = ref_thermo()
    m = OM.deserialize(synth.load_blob("es80"))
    assert np.array_equal(d["W"], np.asarray(m.W))
    assert np.array_equal(d["nasa"][:, 0], np.asarray(m.Tmid))
    assert np.array_equal(d["nasa"][:, 1:8], np.asarray(m.nasa_hi).reshape(-1, 7))
    assert np.array_equal(d["nasa"][:, 8:15], np.asarray(m.nasa_lo).reshape(-1, 7))


def test_thermo_file_round_trip(built, tmp_path):
    import dfchem_host

    d = ref_thermo()
    path = tmp_path / "thermo_es80.txt"
    dfchem_host.write_thermo_file(path, d["W"], d["nasa"], d["visc"], d["cond"], d["diff"])
    assert os.path.getsize(path) == 3420  # size of the reference's file: 4 + 8*(7 + 105 + 35 + 35 + 245)
    back = dfchem_host.read_thermo_file(path)
    for k in ("W", "nasa", "visc", "cond", "diff"):
        assert np.array_equal(back[k], d[k])


def test_transport_oracle_physical_anchors(built):
    """Air-like O2/N2 at 300 K, 1 atm with Cantera's fits: textbook values mu = 1.85e-5 Pa s,
    lambda = 0.026 W/m/K; D(H2 trace in air) about 0.78 cm^2/s.  Loose (5 %) anchors: they catch a wrong
    basis (log vs log10), a swapped index or a missing sqrt(T), not the last digits."""
    d = ref_thermo()
    m = OM.deserialize(synth.load_blob("es80"))
    sp = list(m.species)
    Y = np.zeros((m.N, 1))
    Y[sp.index("O2")], Y[sp.index("N2")], Y[sp.index("H2")] = 0.233, 0.767 - 1e-6, 1e-6
    T, p = np.array([300.0]), np.array([101325.0])
    W = np.asarray(m.W)
    rho = p / (8314.46261815324 * T) / (Y[:, 0] / W).sum()
    cp = K.cp_mass(m, T, Y.T)
    mu, alpha, rhoD = OT.transport(T, p, rho, cp, Y, W, d["visc"], d["cond"], d["diff"])
    assert abs(mu[0] / 1.85e-5 - 1) < 0.05
    assert abs(alpha[0] * cp[0] / 0.026 - 1) < 0.05
    assert abs(rhoD[sp.index("H2"), 0] / rho[0] / 0.78e-4 - 1) < 0.1
    # UnityLewis (dfChemistryModel.C:521-527): every rhoD_i equals alpha
    _, a2, rD2 = OT.transport(T, p, rho, cp, Y, W, d["visc"], d["cond"], d["diff"], unity_lewis=True)
    assert np.array_equal(rD2, np.broadcast_to(a2, rD2.shape))
    # the binary fits are symmetric, as Cantera's are
    assert np.allclose(d["diff"], d["diff"].transpose(1, 0, 2), rtol=1e-13, atol=0)


# ---- GPU -----------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def handles(built):
    import dfchem_host

    cache = {}

    def get(name):
        if name not in cache:
            cache[name] = dfchem_host.Handle(synth.load_blob(name), 0)
        return cache[name]

    yield get
    for h in cache.values():
        h.close()


def batch_for(h, name, n, seed=99):
    b = synth.make_batch(name, h.species, h.W, n, seed=seed)
    m = OM.deserialize(synth.load_blob(name))
    Yn = b["Y"] / b["Y"].sum(0, keepdims=True)
    b["cp"] = K.cp_mass(m, b["T"], np.ascontiguousarray(Yn.T))
    return b


def rel(a, b):
    return np.abs(a - b).max() / np.abs(b).max()


@pytest.mark.gpu
def test_transport_es80_reference_fits(handles, tmp_path):
    import dfchem_host

    d = ref_thermo()
    h = handles("es80")
    path = tmp_path / "thermo_ES80_H2-7-16.txt"
    dfchem_host.write_thermo_file(path, d["W"], d["nasa"], d["visc"], d["cond"], d["diff"])
    h.load_transport_file(str(path))
    b = batch_for(h, "es80", 4099)
    # one pure-species cell: X_i = 1 -> rhoD_i = 0 (dfThermo.cu:232-235)
    b["Y"][:, 0] = 0.0
    b["Y"][h.species.index("N2"), 0] = 1.0
    mu, alpha, rhoD = h.transport(b["T"], b["p"], b["rho"], b["cp"], b["Y"])
    omu, oal, orD = OT.transport(b["T"], b["p"], b["rho"], b["cp"], b["Y"], h.W, d["visc"], d["cond"], d["diff"])
    assert np.abs(mu / omu - 1).max() < 1e-12
    assert np.abs(alpha / oal - 1).max() < 1e-12
    assert rhoD[h.species.index("N2"), 0] == 0.0
    ok = orD > 0
    assert np.abs(rhoD[ok] / orD[ok] - 1).max() < 1e-11
    mu2, al2, rD2 = h.transport(b["T"], b["p"], b["rho"], b["cp"], b["Y"], unity_lewis=True)
    assert np.array_equal(mu2, mu) and np.array_equal(al2, alpha)
    assert np.array_equal(rD2, np.broadcast_to(alpha, rD2.shape))


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["drm19", "gri30"])
def test_transport_synthetic_fits(handles, name):
    h = handles(name)
    visc, cond, diff = synthetic_fits(h.W)
    h.set_transport_fits(visc, cond, diff)
    b = batch_for(h, name, 2500)
    mu, alpha, rhoD = h.transport(b["T"], b["p"], b["rho"], b["cp"], b["Y"])
    omu, oal, orD = OT.transport(b["T"], b["p"], b["rho"], b["cp"], b["Y"], h.W, visc, cond, diff)
    assert np.isfinite(mu).all() and np.isfinite(rhoD).all()
    assert np.abs(mu / omu - 1).max() < 1e-12
    assert np.abs(alpha / oal - 1).max() < 1e-12
    assert np.abs(rhoD / orD - 1).max() < 1e-11


@pytest.mark.gpu
def test_transport_errors(handles):
    import dfchem_host

    h = dfchem_host.Handle(synth.load_blob("h2_li"), 0)
    b = batch_for(h, "h2_li", 16)
    with pytest.raises(dfchem_host.DfchemError):  # no fits loaded
        h.transport(b["T"], b["p"], b["rho"], b["cp"], b["Y"])
    with pytest.raises(dfchem_host.DfchemError):  # missing file
        h.load_transport_file("/nonexistent/thermo_x.txt")
    h.close()


@pytest.mark.gpu
def test_wrong_species_count_in_transport_file(handles, tmp_path):
    import dfchem_host

    d = ref_thermo()
    path = tmp_path / "thermo.txt"
    dfchem_host.write_thermo_file(path, d["W"], d["nasa"], d["visc"], d["cond"], d["diff"])
    with pytest.raises(dfchem_host.DfchemError):
        handles("drm19").load_transport_file(str(path))


@pytest.mark.gpu
@pytest.mark.parametrize("name", MECHS)
def test_rates_of_progress_match_oracle(handles, name):
    """dfchem_rop (getNetRatesOfProgress of calculateRR) against the oracle, and the identity
    sum_r nu_ir q_r W_i = wrate_i with dfchem_wdot."""
    g = golden(name)
    h = handles(name)
    m = OM.deserialize(synth.load_blob(name))
    T, p, Y = (np.ascontiguousarray(g[k]) for k in ("T", "p", "Y"))
    q = h.rop(T, p, Y)
    Yn = Y / Y.sum(0, keepdims=True)
    rho = K.density(m, T, p, np.ascontiguousarray(Yn.T))
    C = rho[:, None] * Yn.T / np.asarray(m.W)[None]
    qf, qr = K.rates_of_progress(m, K.RateConstants(m, T), C)
    qo = (qf - qr).T
    scale = np.maximum(np.abs(qf), np.abs(qr)).max(1)[None, :] + 1e-300  # per cell: largest one-way rate
    assert (np.abs(q - qo) / scale).max() < 1e-11
    S = np.stack([h.reaction_stoich(r) for r in range(h.R)], axis=1)  # [N, R]
    assert np.array_equal(S, K.stoich_matrix(m))
    wr = h.wdot(T, p, Y)
    ws = (S @ q) * h.W[:, None]
    assert (np.abs(ws - wr) / (np.abs(wr).max(0, keepdims=True) + 1e-300)).max() < 1e-10


@pytest.mark.gpu
def test_model_correct_thermo_with_transport_and_calculate_rr(built, tmp_path):
    import dfchem_host
    from dfchem_host import DfChemistryModel

    d = ref_thermo()
    path = tmp_path / "thermo_ES80_H2-7-16.txt"
    dfchem_host.write_thermo_file(path, d["W"], d["nasa"], d["visc"], d["cond"], d["diff"])
    g = golden("es80")
    n = g["T"].shape[0]
    for model in ("Mix", "UnityLewis"):
        chem = DfChemistryModel(synth.load_blob("es80"), n, transport_file=str(path), transport_model=model)
        chem.T[:] = g["T"]
        chem.p[:] = g["p"]
        chem.Y()[:] = g["Y"]
        chem.updateEnergy()
        chem.T[:] = g["T"] + 25.0
        chem.correctThermo()
        assert np.abs(chem.T - g["T"]).max() < 1e-7
        m = OM.deserialize(synth.load_blob("es80"))
        Yn = g["Y"] / g["Y"].sum(0, keepdims=True)
        cp = K.cp_mass(m, chem.T, np.ascontiguousarray(Yn.T))
        omu, oal, orD = OT.transport(chem.T, chem.p, chem.rho, cp, g["Y"], np.asarray(m.W), d["visc"], d["cond"],
                                     d["diff"], unity_lewis=model == "UnityLewis")
        assert np.abs(chem.mu / omu - 1).max() < 1e-11
        assert np.abs(chem.alpha / oal - 1).max() < 1e-11
        rD = np.stack([chem.rhoD(i) for i in range(chem.nSpecies())])
        assert np.abs(rD / orD - 1).max() < 1e-10
        # hai_i = h_i(T)/W_i (dfChemistryModel.C:533-540)
        hai = np.stack([chem.hai(i) for i in range(chem.nSpecies())])
        ho = (K.h_RT(m, chem.T) * 8314.46261815324 * chem.T[:, None] / np.asarray(m.W)[None]).T
        assert np.abs(hai - ho).max() <= 1e-12 * np.abs(ho).max()
    # calculateRR(reaction, species) = nu_net * q * W
    sp = chem.species().index("H2O")
    tot = sum(chem.calculateRR(r, sp) for r in range(chem.nReactions()))
    chem.calculateW()
    assert np.abs(tot - chem.wrate(sp)).max() <= 1e-10 * np.abs(chem.wrate(sp)).max()


@pytest.mark.gpu
def test_patch_thermo_and_internal_energy_form(built, tmp_path):
    """Boundary-face loop of correctThermo (dfChemistryModel.C:546-727): a fixed-T patch returns he = h(T) and the
    properties at T; a free patch solves T from he; the "ea" form round-trips through e = h - p/rho."""
    import dfchem_host
    from dfchem_host import DfChemistryModel

    d = ref_thermo()
    path = tmp_path / "thermo_ES80_H2-7-16.txt"
    dfchem_host.write_thermo_file(path, d["W"], d["nasa"], d["visc"], d["cond"], d["diff"])
    g = golden("es80")
    n = g["T"].shape[0]
    m = OM.deserialize(synth.load_blob("es80"))
    Yn = np.ascontiguousarray((g["Y"] / g["Y"].sum(0, keepdims=True)).T)
    h_ref = K.h_mass(m, g["T"], Yn)
    rho_ref = K.density(m, g["T"], g["p"], Yn)
    for he_name in ("ha", "ea"):
        chem = DfChemistryModel(synth.load_blob("es80"), n, transport_file=str(path), he_name=he_name)

        def fresh(T):
            N = chem.nSpecies()
            return dict(T=np.array(T, float), p=np.array(g["p"], float), rho=np.array(rho_ref), he=np.zeros(n), psi=np.zeros(n),
                        Y=np.ascontiguousarray(g["Y"]), mu=np.zeros(n), alpha=np.zeros(n), rhoD=np.zeros((N, n)),
                        hai=np.zeros((N, n)))

        fixed = fresh(g["T"])
        chem.correctThermoPatch(fixed, fixes_value=True)
        he_ref = h_ref if he_name == "ha" else h_ref - g["p"] / rho_ref
        assert np.abs(fixed["he"] - he_ref).max() <= 1e-9 * np.abs(h_ref).max() + 1e-6
        assert np.abs(fixed["T"] - g["T"]).max() < 1e-7
        assert np.abs(fixed["rho"] / rho_ref - 1).max() < 1e-12
        free = fresh(g["T"] + 40.0)
        free["he"][:] = fixed["he"]
        chem.correctThermoPatch(free, fixes_value=False)
        assert np.abs(free["T"] - g["T"]).max() < 1e-6
        assert np.abs(free["mu"] / fixed["mu"] - 1).max() < 1e-9
        # cell form in the same energy variable
        chem.T[:] = g["T"]
        chem.p[:] = g["p"]
        chem.rho[:] = rho_ref
        chem.Y()[:] = g["Y"]
        chem.updateEnergy()
        assert np.abs(chem.he - he_ref).max() <= 1e-9 * np.abs(h_ref).max() + 1e-6
        chem.T[:] = g["T"] - 35.0
        chem.correctThermo()
        assert np.abs(chem.T - g["T"]).max() < 1e-6
