"""Host-side mirror of the reference's chemistry interface, bound to libdfchem_b200.so via ctypes.

`DfChemistryModel` keeps the OpenFOAM-facing surface of the reference class
`Foam::dfChemistryModel<ThermoType>` (src/dfChemistryModel/dfChemistryModel.H:281-354) for the
hot path: `solve(deltaT)`, `RR(i)`, `Qdot()`, `correctThermo()`, `updateEnergy()`,
`calculateW()`/`wrate(i)`, `species()`, `Y()`.  Fields are plain NumPy arrays in the
reference's layout (one contiguous array per species = species-major SoA).  All arithmetic runs
in the CUDA library; there is no NumPy/CPU fallback: if the library or a GPU is missing the
constructor raises.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, c_char_p, c_double, c_int, c_size_t, c_ulonglong, c_void_p

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), "libdfchem_b200.so")   # the in-tree build, nothing else
GREAT = 1.0e15  # OpenFOAM's `great` (double precision), returned by updateReactionRates (.C:871,894)

_dp = POINTER(c_double)
_ip = POINTER(c_int)

# every symbol include/dfchem.h declares: (restype, argtypes)
SIGNATURES = {
    "dfchem_create": (c_int, [c_char_p, c_int, POINTER(c_void_p)]),
    "dfchem_create_from_blob": (c_int, [c_void_p, c_size_t, c_int, POINTER(c_void_p)]),
    "dfchem_destroy": (None, [c_void_p]),
    "dfchem_last_error": (c_char_p, [c_void_p]),
    "dfchem_species_count": (c_int, [c_void_p]),
    "dfchem_reaction_count": (c_int, [c_void_p]),
    "dfchem_species_name": (c_char_p, [c_void_p, c_int]),
    "dfchem_species_index": (c_int, [c_void_p, c_char_p]),
    "dfchem_species_constants": (c_int, [c_void_p, c_int, _dp]),
    "dfchem_set_tolerances": (c_int, [c_void_p, c_double, c_double]),
    "dfchem_set_max_steps": (c_int, [c_void_p, c_int]),
    "dfchem_solve": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_double, c_void_p,
                             c_int, c_void_p, c_int, c_void_p, c_void_p, c_void_p]),
    "dfchem_solve_dev": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_double, c_void_p,
                                 c_int, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    "dfchem_solve_ptrs": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_double, c_void_p, c_void_p,
                                  c_void_p, c_void_p, c_void_p]),
    "dfchem_bind_deltaTChem": (c_int, [c_void_p, c_void_p]),
    "dfchem_rop_fwd_rev": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_int]),
    "dfchem_rop_fwd_rev_dev": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_int, c_void_p]),
    "dfchem_tc": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p]),
    "dfchem_tc_dev": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    "dfchem_wdot": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_int]),
    "dfchem_wdot_static": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_int]),
    "dfchem_wdot_dev": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_int, c_void_p]),
    "dfchem_correct_thermo": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p,
                                      c_void_p, c_void_p, c_void_p, c_int]),
    "dfchem_correct_thermo_dev": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p,
                                          c_void_p, c_void_p, c_void_p, c_int, c_void_p]),
    "dfchem_h_from_T": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_int, c_void_p]),
    "dfchem_h_from_T_dev": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    "dfchem_rop": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_int]),
    "dfchem_rop_dev": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_int, c_void_p]),
    "dfchem_reaction_stoich": (c_int, [c_void_p, c_int, _dp]),
    "dfchem_set_transport_fits": (c_int, [c_void_p, _dp, _dp, _dp]),
    "dfchem_load_transport_file": (c_int, [c_void_p, c_char_p]),
    "dfchem_generate_transport_fits": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p]),
    "dfchem_transport": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int,
                                 c_void_p, c_void_p, c_void_p, c_int, c_int]),
    "dfchem_transport_dev": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int,
                                     c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p]),
    "dfchem_correct_thermo_patches_dev": (c_int, [c_void_p, c_int, _ip, _ip, c_int] + [c_void_p] * 10 + [c_int, c_int] + [c_void_p] * 9),
    "dfchem_last_counters": (c_int, [c_void_p, POINTER(c_ulonglong)]),
    "dfchem_balance_plan": (c_int, [_dp, c_int, _dp]),
    "dfchem_balance_plan2": (c_int, [_dp, c_int, c_int, _dp]),
    "dfchem_balance_counts": (c_int, [_dp, c_int, c_int, c_int, _dp, c_int, _ip]),
    "dfchem_comm_unique_id": (c_int, [c_void_p]),
    "dfchem_comm_init": (c_int, [c_void_p, c_int, c_int, c_void_p]),
    "dfchem_comm_attach": (c_int, [c_void_p, c_void_p, c_int, c_int]),
    "dfchem_solve_balanced_dev": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_double, c_void_p,
                                          c_int, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p]),
    "dfchem_balance_stats": (c_int, [c_void_p, _dp]),
    "dfchem_bench_fp64_peak": (c_int, [c_int, _dp]),
    "dfchem_version": (c_char_p, []),
    "dfchem_solver_name": (c_char_p, [c_void_p]),
}

_lib = None


def load_library(path: str = LIB_PATH):
    """dlopen libdfchem_b200.so and type every export.  Raises if the library is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(path):
            raise RuntimeError(f"{path} not found: build it with `python __graft_entry__.py` "
                               "(there is no CPU fallback)")
        lib = ctypes.CDLL(path)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


class DfchemError(RuntimeError):
    code = 0   # the DFCHEM_E* value


ECELLFAIL = -6  # DFCHEM_ECELLFAIL


def _ptr(a):
    if a is None:
        return None
    if isinstance(a, int):
        return c_void_p(a)
    return c_void_p(a.ctypes.data)


class Handle:
    """Thin RAII wrapper over dfchem_handle."""

    def __init__(self, mechanism: str | bytes, device: int = 0):
        self.lib = load_library()
        h = c_void_p()
        if isinstance(mechanism, (bytes, bytearray)):
            buf = ctypes.create_string_buffer(bytes(mechanism), len(mechanism))
            rc = self.lib.dfchem_create_from_blob(ctypes.cast(buf, c_void_p), len(mechanism), device, ctypes.byref(h))
        else:
            rc = self.lib.dfchem_create(os.fsencode(mechanism), device, ctypes.byref(h))
        if rc != 0:
            raise DfchemError(f"dfchem_create failed ({rc}): {self.lib.dfchem_last_error(None).decode()}")
        self.h = h
        self.N = self.lib.dfchem_species_count(h)
        self.R = self.lib.dfchem_reaction_count(h)
        self.species = [self.lib.dfchem_species_name(h, i).decode() for i in range(self.N)]
        self.W = np.zeros(self.N)
        self.hc = np.zeros(self.N)
        self.check(self.lib.dfchem_species_constants(h, 0, self.W.ctypes.data_as(_dp)))
        self.check(self.lib.dfchem_species_constants(h, 1, self.hc.ctypes.data_as(_dp)))

    def check(self, rc):
        if rc != 0:
            err = DfchemError(f"dfchem error {rc}: {self.lib.dfchem_last_error(self.h).decode()}")
            err.code = rc
            raise err

    def close(self):
        if getattr(self, "h", None):
            self.lib.dfchem_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_tolerances(self, rtol, atol):
        self.check(self.lib.dfchem_set_tolerances(self.h, rtol, atol))

    def counters(self):
        out = (c_ulonglong * 4)()
        self.check(self.lib.dfchem_last_counters(self.h, out))
        return dict(accepted=out[0], rejected=out[1], failed=out[2], launches=out[3])

    # ---- host-pointer entry points (NumPy arrays; Y, RR are [N, ld] C-contiguous) --------
    def solve(self, T, p, rho, dt, Y, RR=None, Qdot=None, cost=None, status=None):
        nC = T.shape[0]
        assert Y.shape[0] == self.N and Y.flags.c_contiguous and Y.dtype == np.float64
        RR = np.empty((self.N, nC)) if RR is None else RR
        Qdot = np.empty(nC) if Qdot is None else Qdot
        dt_arr = None if np.isscalar(dt) else np.ascontiguousarray(dt, dtype=np.float64)
        self.check(self.lib.dfchem_solve(self.h, nC, _ptr(T), _ptr(p), _ptr(rho), _ptr(dt_arr),
                                         float(dt) if dt_arr is None else 0.0, _ptr(Y), Y.shape[1],
                                         _ptr(RR), RR.shape[1], _ptr(Qdot), _ptr(cost), _ptr(status)))
        return RR, Qdot

    def solve_ptrs(self, T, p, rho, dt, Ys, RRs=None, Qdot=None, cost=None, status=None):
        """dfchem_solve_ptrs: Ys / RRs are lists of N separately allocated 1-D arrays (the PtrList<volScalarField>
        Y_ / RR_ of dfChemistryModel.H:105,121 as they are)."""
        nC = T.shape[0]
        assert len(Ys) == self.N and all(y.dtype == np.float64 and y.flags.c_contiguous and y.shape == (nC,) for y in Ys)
        RRs = [np.empty(nC) for _ in range(self.N)] if RRs is None else RRs
        Qdot = np.empty(nC) if Qdot is None else Qdot
        yp = (c_void_p * self.N)(*[y.ctypes.data for y in Ys])
        rp = (c_void_p * self.N)(*[r.ctypes.data for r in RRs])
        dt_arr = None if np.isscalar(dt) else np.ascontiguousarray(dt, dtype=np.float64)
        self.check(self.lib.dfchem_solve_ptrs(self.h, nC, _ptr(T), _ptr(p), _ptr(rho), _ptr(dt_arr),
                                              float(dt) if dt_arr is None else 0.0, ctypes.cast(yp, c_void_p),
                                              ctypes.cast(rp, c_void_p), _ptr(Qdot), _ptr(cost), _ptr(status)))
        return RRs, Qdot

    def bind_deltaTChem(self, arr):
        """deltaTChem in/out array (host NumPy array for solve / solve_ptrs, a device address for the _dev form);
        None unbinds.  The array must stay alive while bound."""
        self._dtchem = arr
        self.check(self.lib.dfchem_bind_deltaTChem(self.h, _ptr(arr)))

    def rop_fwd_rev(self, T, p, Y):
        """Forward and reverse rates of progress [R, nC] (PaSR.C:349-350)."""
        nC = T.shape[0]
        qf, qr = np.empty((self.R, nC)), np.empty((self.R, nC))
        self.check(self.lib.dfchem_rop_fwd_rev(self.h, nC, _ptr(T), _ptr(p), _ptr(Y), Y.shape[1], _ptr(qf), _ptr(qr), nC))
        return qf, qr

    def tc(self, T, p, rho, Y):
        """PaSR 'reactionRate' chemical time scale [nC] (PaSR.C:326-380)."""
        nC = T.shape[0]
        out = np.empty(nC)
        self.check(self.lib.dfchem_tc(self.h, nC, _ptr(T), _ptr(p), _ptr(rho), _ptr(Y), Y.shape[1], _ptr(out)))
        return out

    def wdot(self, T, p, Y):
        nC = T.shape[0]
        out = np.empty((self.N, nC))
        self.check(self.lib.dfchem_wdot(self.h, nC, _ptr(T), _ptr(p), _ptr(Y), Y.shape[1], _ptr(out), nC))
        return out

    def solver_name(self):
        """Kernel that solve() launches on large batches (dfchem_solver_name)."""
        return self.lib.dfchem_solver_name(self.h).decode()

    def wdot_static(self, T, p, Y):
        """dfchem_wdot through the allocation-free entry point (static device scratch, include/dfchem.h)."""
        nC = T.shape[0]
        out = np.empty((self.N, nC))
        self.check(self.lib.dfchem_wdot_static(self.h, nC, _ptr(T), _ptr(p), _ptr(Y), Y.shape[1], _ptr(out), nC))
        return out

    def correct_thermo(self, he, p, Y, T, want_cp=False, want_hai=False):
        n = he.shape[0]
        T = np.array(T, dtype=np.float64, copy=True)
        psi = np.empty(n)
        rho = np.empty(n)
        cp = np.empty(n) if want_cp else None
        hai = np.empty((self.N, n)) if want_hai else None
        self.check(self.lib.dfchem_correct_thermo(self.h, n, _ptr(he), _ptr(p), _ptr(Y), Y.shape[1], _ptr(T),
                                                  _ptr(psi), _ptr(rho), _ptr(cp), _ptr(hai), n))
        return T, psi, rho, cp, hai

    def h_from_T(self, T, Y):
        n = T.shape[0]
        he = np.empty(n)
        self.check(self.lib.dfchem_h_from_T(self.h, n, _ptr(T), _ptr(Y), Y.shape[1], _ptr(he)))
        return he


    def rop(self, T, p, Y):
        """Net rates of progress [R, nC] (kmol/m^3/s), reaction order of the mechanism file."""
        nC = T.shape[0]
        out = np.empty((self.R, nC))
        self.check(self.lib.dfchem_rop(self.h, nC, _ptr(T), _ptr(p), _ptr(Y), Y.shape[1], _ptr(out), nC))
        return out

    def reaction_stoich(self, reaction):
        nu = np.zeros(self.N)
        self.check(self.lib.dfchem_reaction_stoich(self.h, int(reaction), nu.ctypes.data_as(_dp)))
        return nu

    def set_transport_fits(self, visc, cond, diff):
        visc = np.ascontiguousarray(visc, dtype=np.float64).reshape(self.N, 5)
        cond = np.ascontiguousarray(cond, dtype=np.float64).reshape(self.N, 5)
        diff = np.ascontiguousarray(diff, dtype=np.float64).reshape(self.N, self.N, 5)
        self.check(self.lib.dfchem_set_transport_fits(self.h, visc.ctypes.data_as(_dp), cond.ctypes.data_as(_dp),
                                                      diff.ctypes.data_as(_dp)))

    def generate_transport_fits(self):
        """Fits from the mechanism's own transport data (dfchem_generate_transport_fits); returns (visc, cond, diff)."""
        visc, cond, diff = np.zeros((self.N, 5)), np.zeros((self.N, 5)), np.zeros((self.N, self.N, 5))
        self.check(self.lib.dfchem_generate_transport_fits(self.h, _ptr(visc), _ptr(cond), _ptr(diff)))
        return visc, cond, diff

    def load_transport_file(self, path):
        self.check(self.lib.dfchem_load_transport_file(self.h, os.fsencode(path)))

    def transport(self, T, p, rho, cp, Y, unity_lewis=False, want_mu=True, want_alpha=True, want_rhoD=True):
        n = T.shape[0]
        mu = np.empty(n) if want_mu else None
        alpha = np.empty(n) if want_alpha else None
        rhoD = np.empty((self.N, n)) if want_rhoD else None
        self.check(self.lib.dfchem_transport(self.h, n, _ptr(T), _ptr(p), _ptr(rho), _ptr(cp), _ptr(Y), Y.shape[1],
                                             _ptr(mu), _ptr(alpha), _ptr(rhoD), n, int(bool(unity_lewis))))
        return mu, alpha, rhoD


def read_thermo_file(path):
    """Parse the reference's thermo_<mech>.txt (src_gpu/dfThermo.cu:361-435): int N; W[N];
    N x 15 NASA [Tmid, high7, low7]; N x 5 viscosity; N x 5 conductivity; N x N x 5 diffusion."""
    raw = open(path, "rb").read()
    N = int(np.frombuffer(raw[:4], dtype=np.int32)[0])
    a = np.frombuffer(raw[4:], dtype=np.float64)
    if a.size != N + 15 * N + 10 * N + 5 * N * N:
        raise DfchemError(f"{path}: size does not match N = {N}")
    o = 0
    out = {}
    for key, shape in (("W", (N,)), ("nasa", (N, 15)), ("visc", (N, 5)), ("cond", (N, 5)), ("diff", (N, N, 5))):
        cnt = int(np.prod(shape))
        out[key] = a[o:o + cnt].reshape(shape).copy()
        o += cnt
    return out


def write_thermo_file(path, W, nasa, visc, cond, diff):
    with open(path, "wb") as f:
        f.write(np.int32(len(W)).tobytes())
        for a in (W, nasa, visc, cond, diff):
            f.write(np.ascontiguousarray(a, dtype=np.float64).tobytes())


def balance_plan(loads):
    lib = load_library()
    loads = np.ascontiguousarray(loads, dtype=np.float64)
    n = loads.shape[0]
    send = np.zeros((n, n))
    rc = lib.dfchem_balance_plan(loads.ctypes.data_as(_dp), n, send.ctypes.data_as(_dp))
    if rc != 0:
        raise DfchemError(f"dfchem_balance_plan failed ({rc})")
    return send


from .model import DfChemistryModel  # noqa: E402,F401
