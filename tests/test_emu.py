"""Kernel logic on the CPU: tests/emu compiles the very device source (chem_core.cuh,
integrator.cuh, cell_driver.cuh) as a sequential emulation of the 32-lane team, so the device
tables, gather schedules, Jacobian, Gauss-Jordan and the RODAS4 controller are checked against
the oracle and the golden vectors without a GPU.  (The GPU parity tests are in test_gpu_parity.)"""
import ctypes

import numpy as np
import pytest
from conftest import golden, rr_parity_error

from dfchem_host import synth
from oracle import kinetics as K
from oracle import mech as OM

P = ctypes.POINTER(ctypes.c_double)
MECHS = ["es80", "h2_li", "drm19", "gri30", "c7_onestep", "ethanol66"]


def dp(a):
    return a.ctypes.data_as(P)


@pytest.fixture(scope="module")
def emu(built):
    lib = ctypes.CDLL(built.EMU)
    lib.emu_create.restype = ctypes.c_void_p
    lib.emu_create.argtypes = [ctypes.c_char_p, ctypes.c_size_t]
    return lib


def create(lib, name):
    blob = synth.load_blob(name)
    h = ctypes.c_void_p(lib.emu_create(blob, len(blob)))
    assert h.value
    info = (ctypes.c_int * 15)()
    lib.emu_info(h, info)
    keys = "N NA NP R msets fall pow ndq npart celldbl itA itB itJ jacent wdotent".split()
    return h, dict(zip(keys, info))


def emu_solve(lib, h, g, rtol, atol, N):
    n = g["T"].shape[0]
    RR, Q, cost = np.zeros((N, n)), np.zeros(n), np.zeros(n)
    st = np.zeros(n, np.int32)
    cnt = (ctypes.c_ulonglong * 4)()
    Y = np.ascontiguousarray(g["Y"])
    dt = np.ascontiguousarray(g["dt"])
    rc = lib.emu_solve(h, n, dp(np.ascontiguousarray(g["T"])), dp(np.ascontiguousarray(g["p"])),
                  dp(np.ascontiguousarray(g["rho"])), dp(dt), ctypes.c_double(0.0), dp(Y), n, dp(RR), n, dp(Q),
                  dp(cost), st.ctypes.data_as(ctypes.POINTER(ctypes.c_int)), ctypes.c_double(rtol),
                  ctypes.c_double(atol), 200000, cnt)
    assert rc == 0          # -1: a table-driven address left the emulated cell block
    return RR, Q, cost, st


def test_address_check_is_armed(emu):
    """The emulator checks every table-driven address against the cell block (DFC_CHECK_ADDR)."""
    assert emu.emu_bounds_selftest() == 1


def test_table_shapes(emu):
    sizes = {"es80": (7, 6), "h2_li": (9, 8), "drm19": (21, 19), "gri30": (36, 34), "c7_onestep": (5, 4),
             "ethanol66": (34, 31)}  # (species, species with non-zero net stoichiometry)
    for name, (N, NA) in sizes.items():
        h, info = create(emu, name)
        assert (info["N"], info["NA"]) == (N, NA) and info["NP"] >= NA
        if name in ("gri30", "drm19"):  # large schedules are balanced to within the 4-entry padding
            assert info["itJ"] <= (info["jacent"] + 31) // 32 + 4
        emu.emu_destroy(h)
    h, info = create(emu, "c7_onestep")
    assert info["pow"] == 1                                  # fractional orders take the pow path
    emu.emu_destroy(h)


@pytest.mark.parametrize("name", MECHS)
def test_wdot_matches_golden(emu, name):
    g = golden(name)
    h, info = create(emu, name)
    n = g["T"].shape[0]
    wr = np.zeros((info["N"], n))
    rc = emu.emu_wdot(h, n, dp(np.ascontiguousarray(g["T"])), dp(np.ascontiguousarray(g["p"])),
                 dp(np.ascontiguousarray(g["Y"])), n, dp(wr), n)
    assert rc == 0
    emu.emu_destroy(h)
    scale = np.abs(g["wrate"]).max(0, keepdims=True) + 1e-300
    assert (np.abs(wr - g["wrate"]) / scale).max() < 1e-12


@pytest.mark.parametrize("name", ["h2_li", "drm19", "gri30", "ethanol66", "c7_onestep"])
def test_jacobian_and_inverse_match_oracle(emu, name):
    g = golden(name)
    m = OM.deserialize(synth.load_blob(name))
    h, info = create(emu, name)
    N, NA = info["N"], info["NA"]
    perm = (ctypes.c_int * N)()
    emu.emu_perm(h, perm)
    act = np.array(list(perm))[:NA]
    for c in (0, 2):
        Yn = g["Y"][:, c] / g["Y"][:, c].sum()
        T, p = g["T"][c:c + 1], g["p"][c:c + 1]
        rho = K.density(m, T, p, Yn[None])
        C = rho[:, None] * Yn[None] / m.W[None]
        Jref = K.jacobian_C(m, K.RateConstants(m, T), C)[0]
        J, Ainv = np.zeros((N, N)), np.zeros((NA, NA))
        ihg = 1.0 / (2e-8 * 0.25)
        rc = emu.emu_jacobian(h, ctypes.c_double(T[0]), ctypes.c_double(p[0]), dp(np.ascontiguousarray(Yn)),
                         ctypes.c_double(ihg), dp(J), dp(Ainv))
        assert rc == 0
        assert np.abs(J - Jref)[:, act].max() <= 1e-12 * np.abs(Jref).max()
        A = np.eye(NA) * ihg - Jref[np.ix_(act, act)]
        assert np.abs(Ainv @ A - np.eye(NA)).max() < 1e-9
    emu.emu_destroy(h)


@pytest.mark.parametrize("name", MECHS)
def test_solve_parity_with_golden(emu, name):
    """RR and Qdot of the emulated kernel at tight tolerance vs the converged oracle; includes
    per-cell dt, an un-normalised cell (RR against the raw Y) and cold cells that take one step."""
    g = golden(name)
    h, info = create(emu, name)
    RR, Q, cost, st = emu_solve(emu, h, g, 1e-11, 1e-20, info["N"])
    emu.emu_destroy(h)
    assert (st == 0).all()
    err = rr_parity_error(RR, g["RR"], g["Y"], g["rho"], g["dt"])
    assert err.max() <= 1.0, (name, float(err.max()), np.unravel_index(err.argmax(), err.shape))
    hc = OM.deserialize(synth.load_blob(name)).hc
    qscale = (np.abs(hc)[:, None] * (1e-6 * np.abs(g["RR"]) +
                                     (g["rho"] / g["dt"])[None] * (1e-9 * np.abs(g["Y"]) + 1e-15))).sum(0)
    assert (np.abs(Q - g["Qdot"]) <= qscale).all()


def test_reference_tolerances_stay_close(emu):
    """At the tolerances of 35 of the reference's 38 cases (rtol 1e-6 / atol 1e-10) the result is an
    rtol-level approximation, as CVODE's is: every species within 50*(rtol*|Y| + atol)*rho/dt."""
    g = golden("drm19")
    h, info = create(emu, "drm19")
    RR, Q, cost, st = emu_solve(emu, h, g, 1e-6, 1e-10, info["N"])
    emu.emu_destroy(h)
    assert (st == 0).all()
    bound = 50.0 * (g["rho"] / g["dt"])[None] * (1e-6 * np.abs(g["Y"]) + 1e-10)
    assert (np.abs(RR - g["RR"]) <= bound).all()
    assert cost.min() >= 1 and cost.max() < 400


def emu_solve_with(lib, fn, h, g, rtol, atol, N):
    n = g["T"].shape[0]
    RR, Q, cost = np.zeros((N, n)), np.zeros(n), np.zeros(n)
    st = np.zeros(n, np.int32)
    cnt = (ctypes.c_ulonglong * 4)()
    keep = [np.ascontiguousarray(g[k]) for k in ("T", "p", "rho", "dt", "Y")]
    rc = getattr(lib, fn)(h, n, dp(keep[0]), dp(keep[1]), dp(keep[2]), dp(keep[3]), ctypes.c_double(0.0), dp(keep[4]), n,
                          dp(RR), n, dp(Q), dp(cost), st.ctypes.data_as(ctypes.POINTER(ctypes.c_int)),
                          ctypes.c_double(rtol), ctypes.c_double(atol), 200000, cnt)
    assert rc == 0
    return RR, Q, cost, st, [int(c) for c in cnt[:3]]


@pytest.mark.parametrize("name", ["es80", "h2_li", "c7_onestep"])
@pytest.mark.parametrize("rtol,atol", [(1e-11, 1e-20), (1e-6, 1e-10)])
def test_lane_per_cell_kernel_equals_warp_per_cell_kernel(emu, name, rtol, atol):
    """lane_cell.cuh (lane stride 1 on the host) against the warp-per-cell integrator on the golden states: the
    same step sequence (attempt counts may differ in a few cells where an error norm sits on the accept
    threshold, the sums are taken in a different order), RR equal far below the tolerance, golden parity."""
    g = golden(name)
    h, info = create(emu, name)
    RRw, Qw, costw, stw, cw = emu_solve_with(emu, "emu_solve", h, g, rtol, atol, info["N"])
    RRl, Ql, costl, stl, cl = emu_solve_with(emu, "emu_solve_lpc", h, g, rtol, atol, info["N"])
    emu.emu_destroy(h)
    assert (stw == 0).all() and (stl == 0).all()
    assert (costw != costl).mean() <= 0.05 and abs(cw[0] + cw[1] - cl[0] - cl[1]) <= 0.01 * (cw[0] + cw[1])
    same = costw == costl
    floor = (g["rho"] / g["dt"])[None] * (rtol * np.abs(g["Y"]) + atol)
    assert (np.abs(RRl - RRw)[:, same] <= 1e-3 * floor[:, same] + 1e-9 * np.abs(RRw)[:, same]).all()
    if rtol < 1e-10:
        err = rr_parity_error(RRl, g["RR"], g["Y"], g["rho"], g["dt"])
        assert err.max() <= 1.0, (name, float(err.max()))
