"""Our fused thermo + transport kernels against the REFERENCE's own device kernels (oracle/_ref/libdfthermo_ref.so, built
by oracle/ref_thermo/build_ref.sh from /root/reference/src_gpu/dfThermo.cu:8-357) on a synthetic ES80 batch: maximum
differences and CUDA-event times of one correctThermo pass.  Prints one JSON line.  usage: python scripts/ref_thermo_compare.py [cells]"""
import ctypes
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "deepflame-dev_b200"))


def load_ref():
    path = os.path.join(ROOT, "oracle", "_ref", "libdfthermo_ref.so")
    lib = ctypes.CDLL(path)
    P = ctypes.POINTER(ctypes.c_double)
    lib.ref_thermo_init.argtypes = [ctypes.c_int, P, P, P, P, P]
    lib.ref_correct_thermo.argtypes = [ctypes.c_int] + [P] * 9 + [ctypes.c_int, ctypes.POINTER(ctypes.c_float)]
    return lib


def dp(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_double))


def compare(n=1 << 20, reps=5):
    import torch

    import dfchem_host
    from dfchem_host import synth

    d = np.load(os.path.join(ROOT, "tests", "golden", "reference_thermo_es80.npz"))
    h = dfchem_host.Handle(synth.load_blob("es80"), 0)
    h.set_transport_fits(d["visc"], d["cond"], d["diff"])
    b = synth.make_batch("es80", h.species, h.W, n, seed=4242)
    he = h.h_from_T(b["T"], b["Y"])
    Tg = b["T"] + 30.0
    N = h.N
    # reference device kernels
    ref = load_ref()
    assert ref.ref_thermo_init(N, dp(np.ascontiguousarray(d["W"])), dp(np.ascontiguousarray(d["nasa"])), dp(np.ascontiguousarray(d["visc"])),
                               dp(np.ascontiguousarray(d["cond"])), dp(np.ascontiguousarray(d["diff"]))) == 0
    rT = Tg.copy()
    rpsi, rrho, rmu, ral = (np.empty(n) for _ in range(4))
    rrD = np.empty((N, n))
    ms_ref = ctypes.c_float()
    rc = ref.ref_correct_thermo(n, dp(rT), dp(he), dp(b["p"]), dp(b["Y"]), dp(rpsi), dp(rrho), dp(rmu), dp(ral), dp(rrD), reps,
                                ctypes.byref(ms_ref))
    assert rc == 0, rc
    # ours, through the C ABI (host form for the values, device form for the timing)
    T, psi, rho, cp, hai = h.correct_thermo(he, b["p"], b["Y"], Tg, want_cp=True, want_hai=True)
    mu, alpha, rhoD = h.transport(T, b["p"], rho, cp, b["Y"])
    dev = torch.device("cuda:0")
    t = {k: torch.from_numpy(np.ascontiguousarray(v)).to(dev) for k, v in dict(he=he, p=b["p"], Y=b["Y"], Tg=Tg).items()}
    o = {k: torch.empty(n, dtype=torch.float64, device=dev) for k in ("T", "psi", "rho", "cp", "mu", "al")}
    o["hai"] = torch.empty((N, n), dtype=torch.float64, device=dev)
    o["rD"] = torch.empty((N, n), dtype=torch.float64, device=dev)
    lib = h.lib
    st = torch.cuda.current_stream().cuda_stream
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    tot = 0.0
    for rep in range(reps + 1):
        o["T"].copy_(t["Tg"])
        e0.record()
        h.check(lib.dfchem_correct_thermo_dev(h.h, n, t["he"].data_ptr(), t["p"].data_ptr(), t["Y"].data_ptr(), n, o["T"].data_ptr(),
                                              o["psi"].data_ptr(), o["rho"].data_ptr(), o["cp"].data_ptr(), None, n, st))
        h.check(lib.dfchem_transport_dev(h.h, n, o["T"].data_ptr(), t["p"].data_ptr(), o["rho"].data_ptr(), o["cp"].data_ptr(),
                                         t["Y"].data_ptr(), n, o["mu"].data_ptr(), o["al"].data_ptr(), o["rD"].data_ptr(), n, 0, st))
        e1.record()
        torch.cuda.synchronize()
        if rep > 0:
            tot += e0.elapsed_time(e1)
    ok = rrD > 0
    out = {"mechanism": "es80 (ES80_H2-7-16, 7 species)", "cells": n,
           "max_abs_dT_K": float(np.abs(T - rT).max()),
           "max_rel_psi": float(np.abs(psi / rpsi - 1).max()), "max_rel_rho": float(np.abs(rho / rrho - 1).max()),
           "max_rel_mu": float(np.abs(mu / rmu - 1).max()), "max_rel_alpha": float(np.abs(alpha / ral - 1).max()),
           "max_rel_rhoD": float(np.abs(rhoD[ok] / rrD[ok] - 1).max()),
           "ms_reference_dfThermo_8_launches": float(ms_ref.value), "ms_ours_2_launches": tot / reps,
           "reference": "src_gpu/dfThermo.cu:8-357 kernels compiled for sm_100a with -fmad=false, launch sequence of dfThermo::correctThermo (:588-601)"}
    h.close()
    return out


if __name__ == "__main__":
    print(json.dumps(compare(int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20)))
