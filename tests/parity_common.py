"""Shared pieces of the equal-tolerance accuracy tests (CPU: kernel emulation; GPU: the library).

Question the tests answer: at the tolerances the benchmark is timed at -- rtol 1e-6 /
atol 1e-10, 35 of the reference's 38 cases -- and at the code default 1e-9 / 1e-15 (dfChemistryModel.C:57-58), is the
device integrator (RODAS4) at least as close to the CONVERGED solution as the CVODE-equivalent BDF the reference runs
(restated in oracle/c) at the same (rtol, atol), on the same cells?  Two error measures, both against the converged
oracle RR of the golden file (SciPy Radau, rtol 1e-12):

  tolerance units   |RR - RR_conv| / (rho/dt * (rtol*|Y| + atol))     the error in units of what the tolerance in use admits
  strict parity     conftest.rr_parity_error: |RR - RR_conv| / (1e-6*|RR_conv| + rho/dt*(1e-9*|Y| + 1e-15))

and two statistics over the per-cell maxima: the maximum and the 99th percentile.
"""
import ctypes
import os

import numpy as np
from conftest import ROOT, rr_parity_error

TOLS = [(1e-6, 1e-10), (1e-9, 1e-15)]
P = ctypes.POINTER(ctypes.c_double)


def golden256(name):
    return np.load(os.path.join(ROOT, "tests", "golden", f"solve256_{name}.npz"))


def tol_units(RR, RR_ref, Y, rho, dt, rtol, atol):
    return np.abs(RR - RR_ref) / ((rho / dt)[None, :] * (rtol * np.abs(Y) + atol))


def stats(RR, g, rtol, atol, sel=None):
    """(max, p99) of the per-cell maxima of both error measures."""
    tu = tol_units(RR, g["RR"], g["Y"], g["rho"], g["dt"], rtol, atol).max(0)
    sp = rr_parity_error(RR, g["RR"], g["Y"], g["rho"], g["dt"]).max(0)
    if sel is not None:
        tu, sp = tu[sel], sp[sel]
    return dict(tol_units_max=float(tu.max()), tol_units_p99=float(np.percentile(tu, 99)),
                parity_max=float(sp.max()), parity_p99=float(np.percentile(sp, 99)))


_oc = None


def port_solve(name, g, rtol, atol, threads=None):
    """The CVODE-equivalent CPU port (oracle/c) on the cells of g.  Returns RR, steps per cell."""
    global _oc
    from dfchem_host import synth

    if _oc is None:
        _oc = ctypes.CDLL(os.path.join(ROOT, "oracle", "c", "liboracle_c.so"))
        _oc.oc_load.restype = ctypes.c_void_p
        _oc.oc_load.argtypes = [ctypes.c_char_p, ctypes.c_size_t]
    blob = synth.load_blob(name)
    m = ctypes.c_void_p(_oc.oc_load(blob, len(blob)))
    n = g["T"].shape[0]
    T, p, rho, dt = (np.ascontiguousarray(g[k], dtype=np.float64) for k in ("T", "p", "rho", "dt"))
    Y = np.ascontiguousarray(g["Y"], dtype=np.float64)
    N = Y.shape[0]
    RR, Q, cost = np.zeros((N, n)), np.zeros(n), np.zeros(n)
    st = np.zeros(n, np.int32)
    stats_ = (ctypes.c_long * 3)()
    rc = _oc.oc_solve_batch(m, n, T.ctypes.data_as(P), p.ctypes.data_as(P), rho.ctypes.data_as(P), dt.ctypes.data_as(P),
                            ctypes.c_double(0.0), Y.ctypes.data_as(P), n, ctypes.c_double(rtol), ctypes.c_double(atol), 200000,
                            RR.ctypes.data_as(P), n, Q.ctypes.data_as(P), cost.ctypes.data_as(P),
                            st.ctypes.data_as(ctypes.POINTER(ctypes.c_int)), threads or (os.cpu_count() or 1), stats_)
    assert rc == 0
    _oc.oc_free(m)
    return RR, cost, st


def subset(g, idx):
    out = {}
    for k in g.files if hasattr(g, "files") else g:
        v = g[k]
        if k == "species":
            out[k] = v
        elif v.ndim == 2:
            out[k] = np.ascontiguousarray(v[:, idx])
        else:
            out[k] = np.ascontiguousarray(v[idx])
    return out
