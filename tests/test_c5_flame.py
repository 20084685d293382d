"""Config C5 (BASELINE.json configs[4]): the real CH4 flame state the reference ships for dfLowMachFoam
(examples/dfLowMachFoam/notorch/twoD_reactingTGV/CH4/0/*.gz: 262,144 cells, drm19, p = 101325 Pa, odeCoeffs 1e-6 / 1e-10,
deltaT 1e-6) through the chemistry step and correctThermo back to back, against the oracle.

Fixtures (tests/golden/make_c5_fixture.py): the whole field set and a 1,024-state sample = the 512 distinct states of
the field (kind 0: every one sits at chemical equilibrium, RR ~ 0) + the same columns blended with their neighbours at
constant enthalpy (kind 1: off equilibrium, the chemistry step has work to do), with converged oracle RR / Qdot."""
import gzip
import os
import re

import numpy as np
import pytest
from conftest import HAVE_REFERENCE, ROOT, PKG, rr_parity_error
from parity_common import port_solve, stats

FIELDS = os.path.join(PKG, "data", "c5_drm19_fields.npz")
SAMPLE = os.path.join(ROOT, "tests", "golden", "c5_drm19_sample.npz")
R_GAS = 8314.46261815324


def test_fixture_is_the_shipped_field():
    from dfchem_host import synth
    from oracle import mech as OM

    f = np.load(FIELDS)
    s = np.load(SAMPLE)
    m = OM.deserialize(synth.load_blob("drm19"))
    assert [str(x) for x in f["species"]] == list(m.species) == [str(x) for x in s["species"]]
    assert f["T"].shape == (262144,) and f["Y"].shape == (21, 262144) and float(f["p"]) == 101325.0
    assert abs(f["Y"].sum(0) - 1.0).max() < 1e-12 and 499.0 < f["T"].min() < 501.0 and 2200.0 < f["T"].max() < 2210.0
    T2 = f["T"].reshape(512, 512)
    assert np.abs(np.diff(T2, axis=0)).max() == 0.0          # a 1-D profile repeated on every row
    raw = s["kind"] == 0
    assert raw.sum() == 512 and np.array_equal(s["T"][raw], T2[0]) and np.array_equal(s["Y"][:, raw], f["Y"][:, :512])
    # the shipped states are equilibria: the oracle's converged RR is round-off; the blended ones react
    assert np.abs(s["RR"][:, raw]).max() < 1e-6 and np.abs(s["RR"][:, ~raw]).max() > 1.0
    if HAVE_REFERENCE:
        case = "/root/reference/examples/dfLowMachFoam/notorch/twoD_reactingTGV/CH4/0"
        txt = gzip.open(os.path.join(case, "OH.gz"), "rt").read()
        mm = re.search(r"nonuniform\s+List<scalar>\s*(\d+)\s*\(", txt)
        vals = np.array(txt[mm.end():txt.index(")", mm.end())].split(), dtype=np.float64)
        assert np.array_equal(vals, f["Y"][list(m.species).index("OH")])


@pytest.fixture(scope="module")
def handle(built):
    import dfchem_host
    from dfchem_host import synth

    h = dfchem_host.Handle(synth.load_blob("drm19"), 0)
    yield h
    h.close()


def _solve(h, s, rtol, atol):
    n = s["T"].shape[0]
    h.set_tolerances(rtol, atol)
    cost, st = np.zeros(n), np.zeros(n, np.int32)
    RR, Q = h.solve(np.ascontiguousarray(s["T"]), np.ascontiguousarray(s["p"]), np.ascontiguousarray(s["rho"]),
                    np.ascontiguousarray(s["dt"]), np.ascontiguousarray(s["Y"]), cost=cost, status=st)
    assert (st == 0).all()
    return RR, Q, cost


@pytest.mark.gpu
def test_chemistry_parity_on_real_flame_states(handle):
    s = np.load(SAMPLE)
    RR, Q, cost = _solve(handle, s, 1e-11, 1e-20)
    err = rr_parity_error(RR, s["RR"], s["Y"], s["rho"], s["dt"])
    assert err.max() <= 1.0, (float(err.max()), np.unravel_index(err.argmax(), err.shape))
    qscale = (np.abs(handle.hc)[:, None] * (1e-6 * np.abs(s["RR"]) + (s["rho"] / s["dt"])[None] * (1e-9 * np.abs(s["Y"]) + 1e-15))).sum(0)
    assert (np.abs(Q - s["Qdot"]) <= qscale).all()
    # at the case's own tolerances (odeCoeffs 1e-6 / 1e-10): no worse than the CVODE-equivalent port, on the reacting states
    RRb, _, costb = _solve(handle, s, 1e-6, 1e-10)
    g = {k: s[k] for k in s.files}
    RRp, _, _ = port_solve("drm19", g, 1e-6, 1e-10)
    react = s["kind"] == 1
    dev, port = stats(RRb, g, 1e-6, 1e-10, sel=react), stats(RRp, g, 1e-6, 1e-10, sel=react)
    for key in dev:
        assert dev[key] <= max(port[key], 1.0), (key, dev, port)
    assert costb[s["kind"] == 0].max() <= 4        # equilibrium states: nothing to integrate


@pytest.mark.gpu
def test_chemistry_then_correct_thermo_back_to_back(handle):
    """One dfLowMachFoam chemistry + thermo update without transport: Y += RR dt / rho at constant ha, then
    correctThermo (T from (ha, p, Y), psi, rho, cp) -- device vs oracle on every sample state."""
    from dfchem_host import synth
    from oracle import integrate as OI
    from oracle import kinetics as K
    from oracle import mech as OM

    s = np.load(SAMPLE)
    m = OM.deserialize(synth.load_blob("drm19"))
    RR, Q, _ = _solve(handle, s, 1e-11, 1e-20)
    Ynew = np.ascontiguousarray(s["Y"] + RR * (s["dt"] / s["rho"])[None])
    he = handle.h_from_T(np.ascontiguousarray(s["T"]), np.ascontiguousarray(s["Y"]))
    assert np.allclose(he, s["he"], rtol=1e-13, atol=1e-6)
    T2, psi, rho, cp, _ = handle.correct_thermo(he, np.ascontiguousarray(s["p"]), Ynew, np.ascontiguousarray(s["T"]), want_cp=True)
    Yo = s["Y"] + s["RR"] * (s["dt"] / s["rho"])[None]               # the oracle's own updated composition
    sel = np.concatenate([np.arange(0, 512, 16), 512 + np.arange(0, 512, 8)])   # every blended flame state region + a few raw
    To = np.array([OI.T_from_h(m, s["he"][c], Yo[:, c] / Yo[:, c].sum(), s["T"][c]) for c in sel])
    assert np.abs(T2[sel] - To).max() < 1e-6, float(np.abs(T2[sel] - To).max())
    Yn = (Yo[:, sel] / Yo[:, sel].sum(0, keepdims=True)).T
    assert np.allclose(psi[sel], K.mean_W(m, Yn) / (R_GAS * To), rtol=1e-9)
    assert np.allclose(rho[sel], s["p"][sel] * psi[sel], rtol=1e-14)
    assert np.allclose(cp[sel], K.cp_mass(m, To, Yn), rtol=1e-8)
    # heat release shows up as a temperature rise of the right sign and size: cp dT ~ -sum hc RR dt / rho ... via ha = const
    react = np.nonzero(s["kind"] == 1)[0]
    hot = react[np.abs(Q[react]).argmax()]
    assert np.sign(T2[hot] - s["T"][hot]) == np.sign(Q[hot])


@pytest.mark.gpu
def test_whole_field_chemistry_and_thermo(handle):
    """All 262,144 shipped cells at the case's tolerances: conservation, equilibrium states stay put, thermo round trip."""
    f = np.load(FIELDS)
    n = f["T"].shape[0]
    T, Y = np.ascontiguousarray(f["T"]), np.ascontiguousarray(f["Y"])
    p = np.full(n, float(f["p"]))
    he = handle.h_from_T(T, Y)
    T0, psi, rho, _, _ = handle.correct_thermo(he, p, Y, T + 25.0)
    assert np.abs(T0 - T).max() < 1e-7
    handle.set_tolerances(1e-6, 1e-10)
    cost, st = np.zeros(n), np.zeros(n, np.int32)
    RR, Q = handle.solve(T, p, rho, 1e-6, Y, cost=cost, status=st)
    assert (st == 0).all() and np.isfinite(RR).all()
    assert (np.abs(RR.sum(0)) <= 1e-8 * np.abs(RR).sum(0) + 1e-6).all()
    assert cost.max() <= 4 and np.abs(RR * 1e-6 / rho[None]).max() < 1e-9   # |dY| over the step: equilibria do not move
    assert np.array_equal(RR.reshape(21, 512, 512)[:, 0], RR.reshape(21, 512, 512)[:, 511])   # identical rows, identical results
