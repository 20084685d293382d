"""Oracle self-consistency: the reference publishes no known-answer vectors at the
solve() boundary and Cantera is absent, so the restated arithmetic is pinned by invariants that
do not need Cantera, by the two published ignition anchors, and by the committed golden vectors."""
import numpy as np
import pytest
from conftest import golden

from dfchem_host import synth
from oracle import integrate as OI
from oracle import kinetics as K
from oracle import mech as OM

MECHS = ["es80", "h2_li", "drm19", "gri30", "c7_onestep", "ethanol66"]


def load(name):
    return OM.deserialize(synth.load_blob(name))


def random_state(m, name, n, seed):
    b = synth.make_batch(name, m.species, m.W, n, seed=seed, T_range=(700.0, 2500.0))
    Y = b["Y"].T.copy()
    rng = np.random.default_rng(seed)
    Y += rng.uniform(0, 1e-3, Y.shape)
    Y /= Y.sum(1, keepdims=True)
    rho = K.density(m, b["T"], b["p"], Y)
    return b["T"], b["p"], Y, rho[:, None] * Y / m.W[None]


@pytest.mark.parametrize("name", MECHS)
def test_element_and_mass_conservation(name):
    m = load(name)
    T, p, Y, C = random_state(m, name, 16, 1)
    w = K.wdot(m, K.RateConstants(m, T), C)
    scale = np.abs(w).max(1, keepdims=True) * m.W.max()
    assert np.abs((w * m.W[None]).sum(1, keepdims=True) / scale).max() < 1e-12       # sum W_k wdot_k = 0
    assert np.abs(w @ m.comp / np.abs(w).max(1, keepdims=True)).max() < 1e-11         # every element


@pytest.mark.parametrize("name", MECHS)
def test_nasa7_continuity_and_formation_enthalpies(name):
    m = load(name)
    for k in range(m.N):
        if m.Thi[k] == m.Tmid[k]:
            continue
        T = np.array([m.Tmid[k]])
        for fn, tol in ((K.cp_R, 5e-2), (K.h_RT, 5e-2), (K.s_R, 5e-2)):  # file data; a swapped range gives O(1)
            lo = fn(m, T)[0, k]
            hi = fn(m, T + 1e-9)[0, k]
            assert abs(lo - hi) <= tol * max(1.0, abs(lo)), (m.species[k], fn.__name__)
    known = {"H2O": -241.8e6, "CO2": -393.5e6, "CH4": -74.6e6, "H2": 0.0, "O2": 0.0, "N2": 0.0}  # J/kmol
    if name == "c7_onestep":  # that file's O2 record carries a non-zero reference enthalpy (its own data)
        known = {"H2O": -241.8e6, "CO2": -393.5e6}
    for sp, hf in known.items():
        if sp in m.species:
            k = m.species.index(sp)
            assert abs(m.hc[k] * m.W[k] - hf) < 0.6e6, sp


@pytest.mark.parametrize("name", ["h2_li", "drm19", "ethanol66", "c7_onestep"])
def test_analytic_jacobian_matches_finite_differences(name):
    m = load(name)
    T, p, Y, C = random_state(m, name, 2, 3)
    rc = K.RateConstants(m, T)
    J = K.jacobian_C(m, rc, C)
    for c in range(2):
        rc1 = K.RateConstants(m, T[c:c + 1])
        Jfd = np.zeros((m.N, m.N))
        for j in range(m.N):
            d = 1e-6 * max(C[c, j], 1e-8 * C[c].max())
            Cp, Cm = C[c].copy(), C[c].copy()
            Cp[j] += d
            Cm[j] -= d
            Jfd[:, j] = (K.wdot(m, rc1, Cp[None])[0] - K.wdot(m, rc1, Cm[None])[0]) / (2 * d)
        assert np.abs(J[c] - Jfd).max() <= 2e-6 * np.abs(Jfd).max()


def test_detailed_balance_at_equilibrium():
    """With Kc from the same NASA-7 data, net rates of progress vanish at the composition that a
    long constant-(T, rho) integration converges to."""
    m = load("h2_li")
    T, p = 2200.0, 101325.0
    Y0 = synth.fuel_air(m.species, m.W, "H2", 0.5, [1.0])[0]
    rho = K.density(m, np.array([T]), np.array([p]), Y0[None])[0]
    _, _, Yeq = OI.solve_single(m, T, p, rho, Y0, 1.0, rtol=1e-12, atol=1e-22)
    C = (rho * Yeq / m.W)[None]
    qf, qr = K.rates_of_progress(m, K.RateConstants(m, np.array([T])), C)
    rev = m.reversible == 1
    assert np.abs(qf - qr)[0, rev].max() <= 1e-6 * np.abs(qf)[0, rev].max()


def test_h2_ignition_anchors():
    """docs/source/qs/0Dcvode.jpg and test/corrtest.cpp:117 (C1 case: ES80, T0 = 1000 K, 1 atm,
    Y from examples/df0DFoam/zeroD_cubicReactor/H2/cvodeIntegrator/0): ignition inflection
    ~2.45e-4 s and T(1 ms) ~2590 K.  Read off a screenshot, hence the loose bands."""
    from scipy.integrate import solve_ivp

    m = load("es80")
    Y0 = np.zeros(m.N)
    Y0[m.species.index("H2")] = 0.02555366261652
    Y0[m.species.index("O2")] = 0.227143668370244
    Y0[m.species.index("N2")] = 0.7473026689382
    p = 101325.0

    def f(t, z):
        T, Y = z[0], z[1:][None]
        Tn = np.array([T])
        rho = p * K.mean_W(m, Y) / (OM.GAS_CONSTANT * Tn)
        w = K.wdot(m, K.RateConstants(m, Tn), rho[:, None] * Y / m.W[None])[0]
        h = K.h_RT(m, Tn)[0] * OM.GAS_CONSTANT * T
        return np.concatenate([[-(h * w).sum() / (rho[0] * K.cp_mass(m, Tn, Y)[0])], w * m.W / rho[0]])

    # atol must sit far below the initial radical pool (~1e-20): the example case itself runs CVODE
    # with relTol 1e-15 / absTol 1e-24 (its constant/CanteraTorchProperties) for this reason
    sol = solve_ivp(f, (0, 1e-3), np.concatenate([[1000.0], Y0]), method="BDF", rtol=1e-8, atol=1e-26,
                    dense_output=True)
    t = np.linspace(0, 1e-3, 4001)
    T = sol.sol(t)[0]
    t_ign = t[np.argmax(np.gradient(T, t))]
    assert 2.2e-4 < t_ign < 2.7e-4, t_ign
    assert 2540.0 < T[-1] < 2640.0, T[-1]


@pytest.mark.parametrize("name", MECHS)
def test_golden_instantaneous_rates_and_thermo(name):
    g = golden(name)
    m = load(name)
    assert [str(s) for s in g["species"]] == m.species
    Yn = (g["Y"] / g["Y"].sum(0, keepdims=True)).T
    rho_ct = K.density(m, g["T"], g["p"], Yn)
    wr = (K.wdot(m, K.RateConstants(m, g["T"]), rho_ct[:, None] * Yn / m.W[None]) * m.W[None]).T
    assert np.allclose(wr, g["wrate"], rtol=1e-12, atol=1e-12 * np.abs(g["wrate"]).max())
    assert np.allclose(K.h_mass(m, g["T"], Yn), g["he"], rtol=1e-13)
    for c in range(3):
        assert abs(OI.T_from_h(m, g["he"][c], Yn[c], g["T"][c] + 80.0) - g["T"][c]) < 1e-8


@pytest.mark.parametrize("name", ["es80", "h2_li", "c7_onestep"])
def test_golden_solve_is_reproduced_and_method_independent(name):
    """The goldens are Radau at rtol 1e-12; BDF at 1e-11 must land on the same converged answer."""
    g = golden(name)
    m = load(name)
    for c in (0, 3):
        RR, Q, _ = OI.solve_single(m, g["T"][c], g["p"][c], g["rho"][c], g["Y"][:, c], g["dt"][c],
                                   rtol=1e-12, atol=1e-22)
        assert np.allclose(RR, g["RR"][:, c], rtol=1e-9, atol=1e-12 * np.abs(g["RR"][:, c]).max())
        RRb, _, _ = OI.solve_single(m, g["T"][c], g["p"][c], g["rho"][c], g["Y"][:, c], g["dt"][c],
                                    rtol=1e-11, atol=1e-22, method="BDF")
        big = np.abs(g["RR"][:, c]) > 1e-6 * np.abs(g["RR"][:, c]).max()
        assert np.allclose(RRb[big], g["RR"][big, c], rtol=2e-6)
