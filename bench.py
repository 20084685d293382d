#!/usr/bin/env python
"""Benchmark of the hot path: cell-chemistry integrations/s (BASELINE.json metric).

A "step" is one chemistry sub-step (dfChemistryModel::solve(deltaT), deltaT = 1e-6 s) over one
synthetic batch: by default the CH4 / GRI-3.0-class mechanism (mechanisms/CH4/gri30.yaml, 36
species, 219 reactions), 4,194,304 cells PER GPU (weak scaling: each rank owns its own batch, the
way each MPI rank owns its mesh partition in the reference; no data-path collective), at the
tolerances of 35 of the reference's 38 cases (rtol 1e-6, atol 1e-10).

  python bench.py --gpus N --steps K --warmup W            our arm  (torchrun for N > 1)
  python bench.py --impl reference ...                     the reference's CPU algorithm
                                                            (oracle/c port; rank 0 only)
One JSON line on stdout (rank 0).  See DESIGN.md 'Measurement' for every field.
"""
from __future__ import annotations

import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "deepflame-dev_b200"))

METRIC = "cell-chemistry integrations/s"
UNIT = "cells/s"

# Algorithmic FP64 work per unit (flops: FMA = 2, exp/log/pow = 1), dense N x N
# Jacobian / LU model with N = species that the ODE actually moves.  Keys: per-cell set-up,
# per RHS evaluation, per analytic Jacobian, per (I/(h g) - J) formation, per LU (2/3 N^3),
# per triangular solve pair (2 N^2).
FLOP_MODEL = {
    "gri30": dict(setup=3612 + 524, rhs=3807 + 50, jac=15453, form=2 * 36 * 36, lu=2 * 36 ** 3 // 3, solve=2 * 36 * 36),
    "drm19": dict(setup=1598 + 201, rhs=1478 + 16, jac=4924, form=882, lu=6174, solve=882),
    "es80": dict(setup=306 + 17, rhs=173, jac=552, form=98, lu=229, solve=98),
    "h2_li": dict(setup=512 + 51, rhs=354 + 4, jac=1018, form=162, lu=486, solve=162),
    "c7_onestep": dict(setup=156 + 2, rhs=36, jac=46, form=50, lu=83, solve=50),
}
STAGES = 6  # RODAS4: per step attempt 1 LU + 6 RHS + 6 solves; per accepted step 1 Jacobian


def algorithmic_flops(mech, n_cells, accepted, rejected):
    f = FLOP_MODEL[mech]
    attempts = accepted + rejected
    return (n_cells * f["setup"] + attempts * (STAGES * f["rhs"] + f["form"] + f["lu"] + STAGES * f["solve"])
            + accepted * f["jac"])


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        busy = [s for s in sm if s > 0]
        return {"sm_mhz": float(np.median(busy)) if busy else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def load_oracle_c():
    """The CPU arm is built where it runs, with -O3 -march=native (oracle/c/Makefile target `native`): the portable
    liboracle_c.so that travels with the snapshot is only the fall-back when no compiler is on the box."""
    path = os.path.join(ROOT, "oracle", "c", "liboracle_c.so")
    native = False
    try:
        import hashlib

        flags = next((ln for ln in open("/proc/cpuinfo") if ln.startswith("flags")), "")
        tag = hashlib.sha1(flags.encode()).hexdigest()[:10]          # one build per distinct CPU feature set
        subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle", "c"), "native", f"TAG={tag}"], check=True,
                       stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, timeout=120)
        cand = os.path.join(ROOT, "oracle", "c", f"liboracle_c_native_{tag}.so")
        if os.path.exists(cand):
            path, native = cand, True
    except (OSError, subprocess.SubprocessError):
        pass
    lib = ctypes.CDLL(path)
    lib.built_native = native
    lib.oc_load.restype = ctypes.c_void_p
    lib.oc_load.argtypes = [ctypes.c_char_p, ctypes.c_size_t]
    return lib


def cpu_reference_rate(mech, batch, n_sample, rtol, atol, dt, threads):
    """Times the reference's CPU algorithm (oracle/c: per-cell variable-order BDF + Newton + dense
    LU + finite-difference Jacobian, fresh per cell, one cell per task on `threads` host threads)
    on the first n_sample cells of the batch.  Returns (cells/s, seconds, mean BDF steps/cell)."""
    from dfchem_host import synth

    lib = load_oracle_c()
    blob = synth.load_blob(mech)
    m = ctypes.c_void_p(lib.oc_load(blob, len(blob)))
    P = ctypes.POINTER(ctypes.c_double)
    n = n_sample
    T, p, rho = (np.ascontiguousarray(batch[k][:n]) for k in ("T", "p", "rho"))
    Y = np.ascontiguousarray(batch["Y"][:, :n])
    N = Y.shape[0]
    RR, Q = np.empty((N, n)), np.empty(n)
    stats = (ctypes.c_long * 3)()
    t0 = time.perf_counter()
    lib.oc_solve_batch(m, n, T.ctypes.data_as(P), p.ctypes.data_as(P), rho.ctypes.data_as(P), None,
                       ctypes.c_double(dt), Y.ctypes.data_as(P), n, ctypes.c_double(rtol), ctypes.c_double(atol),
                       20000, RR.ctypes.data_as(P), n, Q.ctypes.data_as(P), None, None, threads, stats)
    el = time.perf_counter() - t0
    lib.oc_free(m)
    return n / el, el, stats[0] / n


def run_reference(args, rank):
    """--impl reference: the reference's own CPU algorithm for the path on the box's host cores
    (Cantera/CVODE are absent, so this is the oracle/c port; DESIGN.md 'CPU baseline')."""
    if rank != 0:
        return
    from dfchem_host import synth
    from oracle import mech as OM

    m = OM.deserialize(synth.load_blob(args.mech))
    threads = os.cpu_count() or 1
    sample = args.ref_cells or 2048 * threads     # ~6 s of CPU work per step on 16 cores
    batch = synth.make_batch(args.mech, m.species, m.W, sample, seed=12345)
    times = []
    for i in range(args.warmup + args.steps):
        rate, el, steps = cpu_reference_rate(args.mech, batch, sample, args.rtol, args.atol, args.dt, threads)
        if i >= args.warmup:
            times.append(el)
    ms = 1e3 * float(np.mean(times))
    value = sample / (ms * 1e-3)
    desc = f"first {sample} cells of the seed-12345 batch per step ({steps:.1f} BDF steps/cell)"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args, args.cells, impl="reference"),   # same workload; each step times a bounded sample of it
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": desc},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }), flush=True)


def workload_config(args, cells, impl="ours"):
    integ = ("RODAS4 (Rosenbrock 4(3), 6 stages, analytic Jacobian)" if impl == "ours" else
             "variable-order BDF 1-5 + Newton + dense LU + finite-difference Jacobian (oracle/c, CVODE-equivalent port)")
    return {"workload": f"{args.mech} ({'mechanisms/CH4/gri30.yaml, 36 species / 219 reactions' if args.mech == 'gri30' else args.mech}), "
                        f"{cells} synthetic cells per GPU, T~U[300,2500] K, p~U[1,10] atm, 50% fuel/air phi~U[0.5,2] + 50% "
                        f"ignition-trajectory states, deltaT={args.dt:g} s",
            "cells_per_gpu": cells, "rtol": args.rtol, "atol": args.atol, "integrator": integ,
            "l2": "inputs (T,p,rho,Y = 1.3 GB at 4M cells) exceed the 126 MB L2; no explicit flush"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--mech", default="gri30")
    ap.add_argument("--cells", type=int, default=4194304)
    ap.add_argument("--rtol", type=float, default=1e-6)
    ap.add_argument("--atol", type=float, default=1e-10)
    ap.add_argument("--dt", type=float, default=1e-6)
    ap.add_argument("--ref-cells", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-thermo", action="store_true", help="skip the correctThermo (HBM-bound) leg")
    ap.add_argument("--thermo-only", action="store_true", help="run only the correctThermo leg (ncu captures)")
    ap.add_argument("--skew", type=float, default=0.0,
                    help="make rank loads unequal: rank r draws T from [300 + skew*r*200, 2500] K (N > 1)")
    ap.add_argument("--rebalance", action="store_true",
                    help="headline step through dfchem_solve_balanced_dev (in-library NCCL rebalancer, N > 1)")
    ap.add_argument("--workload", default="synthetic", choices=["synthetic", "c5"],
                    help="c5: the reference's real 262,144-cell drm19 flame field (config C5), chemistry + correctThermo back to back")
    ap.add_argument("--no-configs", action="store_true", help="skip the other BASELINE.json configurations (configs object)")
    ap.add_argument("--no-strong", action="store_true", help="skip the strong-scaled, skewed, rebalanced leg (N > 1)")
    ap.add_argument("--no-parity", action="store_true", help="skip the accuracy statement (parity object)")
    ap.add_argument("--no-flow", action="store_true", help="skip the consecutive-flow-steps leg (flow_steps object, N = 1)")
    ap.add_argument("--flow-cells", type=int, default=1 << 20, help="cells of the consecutive-flow-steps leg")
    ap.add_argument("--strong-cells", type=int, default=4194304, help="total cells of the strong-scaled leg")
    ap.add_argument("--balance-algorithm", default="allAverage", choices=["allAverage", "headTail"])
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return

    # stdout carries exactly one JSON line: anything libraries print there meanwhile (NCCL's version
    # banner, ...) is sent to stderr instead
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)

    import torch
    import torch.distributed as dist

    import dfchem_host
    from dfchem_host import synth

    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib = dfchem_host.load_library()   # raises if the CUDA library is missing: no fallback
    stream = torch.cuda.current_stream()
    alg = 0 if args.balance_algorithm == "allAverage" else 1

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    class DeviceBatch:
        """A batch resident in HBM plus its output arrays."""

        def __init__(self, h, batch):
            self.h, self.n, self.N = h, batch["T"].shape[0], h.N
            self.d = {k: torch.from_numpy(np.ascontiguousarray(v)).to(dev) for k, v in batch.items()}
            self.RR = torch.empty((self.N, self.n), dtype=torch.float64, device=dev)
            self.Q = torch.empty(self.n, dtype=torch.float64, device=dev)
            self.cost = torch.empty(self.n, dtype=torch.float64, device=dev)
            self.stat = torch.empty(self.n, dtype=torch.int32, device=dev)

        def solve(self, dt):
            d = self.d
            self.h.check(lib.dfchem_solve_dev(self.h.h, self.n, d["T"].data_ptr(), d["p"].data_ptr(), d["rho"].data_ptr(), None, dt,
                                              d["Y"].data_ptr(), self.n, self.RR.data_ptr(), self.n, self.Q.data_ptr(),
                                              self.cost.data_ptr(), self.stat.data_ptr(), stream.cuda_stream))

        def solve_balanced(self, dt, prev_cost):
            d = self.d
            self.h.check(lib.dfchem_solve_balanced_dev(self.h.h, self.n, d["T"].data_ptr(), d["p"].data_ptr(), d["rho"].data_ptr(),
                                                       None, dt, d["Y"].data_ptr(), self.n, self.RR.data_ptr(), self.n,
                                                       self.Q.data_ptr(), prev_cost.data_ptr(), self.cost.data_ptr(),
                                                       self.stat.data_ptr(), alg, stream.cuda_stream))

    def time_steps(fn, steps, warmup, sampler=None):
        """W untimed + exactly K timed steps, barrier + synchronize on both sides, CUDA events, max over ranks."""
        for _ in range(warmup):
            fn()
        barrier()
        if sampler:
            sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            fn()
        e1.record(stream)
        barrier()
        ms_local = e0.elapsed_time(e1) / steps
        return ms_local, max_over_ranks(ms_local)

    def solve_stats(h, db):
        cnt = h.counters()
        failed = int((db.stat != 0).sum().item())
        assert failed == 0, f"{failed} cells failed to integrate"      # a failed cell invalidates the number
        att = float(db.cost.mean().item())
        # one (I/(h g) - J) factorisation per step attempt, one analytic Jacobian per accepted step (RODAS4, no reuse)
        return {"accepted_steps": cnt["accepted"], "rejected_steps": cnt["rejected"], "mean_attempts_per_cell": att,
                "max_attempts_per_cell": float(db.cost.max().item()), "factorisations_per_cell": att,
                "jacobians_per_cell": cnt["accepted"] / db.n, "failed_cells": failed}, cnt

    h = dfchem_host.Handle(synth.load_blob(args.mech if args.workload == "synthetic" else "drm19"), local)
    h.set_tolerances(args.rtol, args.atol)
    N = h.N

    if args.workload == "c5":
        out = c5_leg(args, h, lib, dev, stream, time_steps, world, torch)
        sys.stdout.flush()
        os.dup2(saved_stdout, 1)
        if rank == 0:
            print(json.dumps(out), flush=True)
        if world > 1:
            dist.destroy_process_group()
        return

    n = args.cells
    T_range = (300.0 + args.skew * rank * 200.0, 2500.0) if args.skew else (300.0, 2500.0)
    batch = synth.make_batch(args.mech, h.species, h.W, n, seed=12345 + rank, T_range=T_range)
    db = DeviceBatch(h, batch)
    d, dRR, dQ, dcost, dstat = db.d, db.RR, db.Q, db.cost, db.stat

    if world > 1:
        from dfchem_host import balance

        balance.init_library_comm(h, dev)

    step_dev = lambda: db.solve(args.dt)   # noqa: E731
    if args.rebalance and world > 1:
        db.solve(args.dt)                  # cost of the previous flow step = the balancer's input
        torch.cuda.synchronize()
        prev_cost = dcost.clone()
        step_dev = lambda: db.solve_balanced(args.dt, prev_cost)   # noqa: E731

    def thermo_leg():
    # ---- correctThermo leg (HBM-bound): T from (h, p, Y), psi, rho [+ cp, hai] ------------------
        thermo = None
        if not args.no_thermo:
            he = torch.empty(n, dtype=torch.float64, device=dev)
            h.check(lib.dfchem_h_from_T_dev(h.h, n, d["T"].data_ptr(), d["Y"].data_ptr(), n, he.data_ptr(), stream.cuda_stream))
            Tg = (d["T"] + 30.0).contiguous()                      # a flow step moves T by tens of K at most
            Tw = torch.empty_like(Tg)
            psi, rho2, cpo = (torch.empty(n, dtype=torch.float64, device=dev) for _ in range(3))
            hai = torch.empty((N, n), dtype=torch.float64, device=dev)
            thermo = {}
            for label, cp_ptr, hai_ptr, nbytes in (("T_psi_rho", None, None, (N + 3) * 8 + 3 * 8),
                                                   ("T_psi_rho_cp_hai", cpo.data_ptr(), hai.data_ptr(), (N + 3) * 8 + (4 + N) * 8)):
                def step_thermo():
                    Tw.copy_(Tg)                                   # in/out argument: restore the guess (counted in the bytes)
                    h.check(lib.dfchem_correct_thermo_dev(h.h, n, he.data_ptr(), d["p"].data_ptr(), d["Y"].data_ptr(), n,
                                                          Tw.data_ptr(), psi.data_ptr(), rho2.data_ptr(), cp_ptr, hai_ptr, n,
                                                          stream.cuda_stream))
                for _ in range(3):
                    step_thermo()
                barrier()
                reps = 10
                t0e, t1e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                tc0, tc1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                tc0.record(stream)
                for _ in range(reps):
                    Tw.copy_(Tg)
                tc1.record(stream)
                t0e.record(stream)
                for _ in range(reps):
                    step_thermo()
                t1e.record(stream)
                barrier()
                ms_t = (t0e.elapsed_time(t1e) - tc0.elapsed_time(tc1)) / reps   # kernel alone (copy of the guess subtracted)
                ms_t = max_over_ranks(ms_t)
                # Round trip T -> h -> T.  h(T) of NASA-7 data jumps by ~1e-4 K worth of enthalpy at a species' switch temperature,
                # so a cell whose T lies within that distance BELOW a switch temperature has a second valid root just above it
                # (reached from the guess above): a handful of cells per 4M may differ by ~1e-4 K, every other cell by round-off.
                dT = (Tw - d["T"]).abs()
                at_switch = dT >= 1e-6
                n_sw = int(at_switch.sum().item())
                err = float(dT[~at_switch].max().item())
                err_sw = float(dT.max().item())
                assert n_sw <= 16 and err_sw < 1e-2, f"thermo round trip: {n_sw} cells off by up to {err_sw} K"
                thermo[label] = {"cells_per_s": world * n / (ms_t * 1e-3), "ms": ms_t, "algorithmic_bytes_per_cell": nbytes,
                                 "achieved_GBps": nbytes * n / (ms_t * 1e-3) / 1e9, "max_abs_T_error_K": err,
                                 "cells_on_a_nasa_switch_temperature": n_sw, "max_abs_T_error_on_switch_K": err_sw}
            # ---- transport part of correctThermo (FP64-bound: N^2 pair terms per cell) -------------
            fits_kind = load_transport(h, args.mech, synth)
            mu, al = (torch.empty(n, dtype=torch.float64, device=dev) for _ in range(2))

            def step_tr():
                h.check(lib.dfchem_transport_dev(h.h, n, Tw.data_ptr(), d["p"].data_ptr(), rho2.data_ptr(), cpo.data_ptr(),
                                                 d["Y"].data_ptr(), n, mu.data_ptr(), al.data_ptr(), hai.data_ptr(), n, 0,
                                                 stream.cuda_stream))
            for _ in range(2):
                step_tr()
            barrier()
            t0e, t1e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0e.record(stream)
            for _ in range(3):
                step_tr()
            t1e.record(stream)
            barrier()
            ms_t = max_over_ranks(t0e.elapsed_time(t1e) / 3)
            # per cell: N poly (4 fma) x2, Wilke N^2 x 5, diffusion N(N-1) x (4 fma poly + mul + rcp + 3), flops with fma = 2
            flops = N * 2 * 9 + N * N * 9 + N * (N - 1) * 16 + 8 * N
            thermo["mu_alpha_rhoD"] = {"cells_per_s": world * n / (ms_t * 1e-3), "ms": ms_t, "bound": "fp64",
                                       "algorithmic_flops_per_cell": flops,
                                       "achieved_TFLOPs": flops * n / (ms_t * 1e-3) / 1e12,
                                       "algorithmic_bytes_per_cell": (N + 4) * 8 + (N + 2) * 8,
                                       "fits": fits_kind,
                                       "finite": bool(torch.isfinite(mu).all().item() and torch.isfinite(hai).all().item())}
            del hai
        return thermo

    if args.thermo_only:
        th = thermo_leg()
        sys.stdout.flush()
        os.dup2(saved_stdout, 1)
        if rank == 0:
            print(json.dumps({"thermo": th, "n_gpus": world, "cells_per_gpu": n}), flush=True)
        if world > 1:
            dist.destroy_process_group()
        return

    sampler = ClockSampler(local) if rank == 0 else None
    ms_local, ms = time_steps(step_dev, args.steps, args.warmup, sampler)
    clocks = sampler.stop() if rank == 0 else None
    istats, cnt = solve_stats(h, db)
    value = world * n / (ms * 1e-3)

    # ---- end-to-end arm: host buffers through the C ABI, copies inside the timed region --------
    pin = {k: torch.from_numpy(v).pin_memory() for k, v in batch.items()}
    hRR = torch.empty((N, n), dtype=torch.float64).pin_memory()
    hQ = torch.empty(n, dtype=torch.float64).pin_memory()

    def step_host():
        h.check(lib.dfchem_solve(h.h, n, pin["T"].data_ptr(), pin["p"].data_ptr(), pin["rho"].data_ptr(), None, args.dt,
                                 pin["Y"].data_ptr(), n, hRR.data_ptr(), n, hQ.data_ptr(), None, None))

    step_host()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_host()
    torch.cuda.synchronize()
    e2e_ms = max_over_ranks(1e3 * (time.perf_counter() - t0) / args.steps)
    barrier()
    same = bool(torch.equal(hRR, dRR.cpu()))
    e2e = {"value": world * n / (e2e_ms * 1e-3), "unit": UNIT, "ms_per_step": e2e_ms,
           "h2d_bytes_per_step": (N + 3) * n * 8, "d2h_bytes_per_step": (N + 1) * n * 8,
           "matches_device_arm": same}
    del pin, hRR, hQ

    thermo = thermo_leg()

    # ---- strong-scaled, skewed batch with and without the in-library rebalancer (config C3) ------
    strong = None
    if world > 1 and not args.no_strong:
        strong = strong_leg(args, h, lib, dev, rank, world, DeviceBatch, time_steps, sum_over_ranks, max_over_ranks, synth, torch)

    # ---- the other BASELINE.json configurations, device-timed the same way (rank 0's GPU, N = 1 semantics) ----
    configs = None
    if world == 1 and not args.no_configs and args.mech == "gri30":   # (N = 1 only: the driver's BENCH run)
        del db, d, dRR, dQ, dcost, dstat
        torch.cuda.empty_cache()
        configs = {}
        for name, cells in (("es80", 1 << 20), ("c7_onestep", 1 << 20), ("drm19", 1 << 20)):
            hc = dfchem_host.Handle(synth.load_blob(name), local)
            hc.set_tolerances(args.rtol, args.atol)
            bc = DeviceBatch(hc, synth.make_batch(name, hc.species, hc.W, cells, seed=12345))
            for _ in range(3):
                bc.solve(args.dt)
            torch.cuda.synchronize()
            c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            c0.record(stream)
            for _ in range(3):
                bc.solve(args.dt)
            c1.record(stream)
            torch.cuda.synchronize()
            msc = c0.elapsed_time(c1) / 3
            st_c, cnt_c = solve_stats(hc, bc)
            fl = algorithmic_flops(name, cells, cnt_c["accepted"], cnt_c["rejected"])
            configs[f"{name}_{cells}"] = {"value": cells / (msc * 1e-3), "unit": UNIT, "ms_per_step": msc, "cells": cells,
                                          "mean_attempts_per_cell": st_c["mean_attempts_per_cell"],
                                          "achieved_TFLOPs": fl / (msc * 1e-3) / 1e12}
            del bc
            hc.close()
        c5 = c5_leg(args, None, lib, dev, stream, time_steps, 1, torch, local=local)
        configs["c5_drm19_real_flame_262144"] = c5["c5"]

    # ---- consecutive flow steps: chemistry -> Y update -> correctThermo, deltaTChem carried or not (N = 1) ----
    flow = None
    if world == 1 and not args.no_flow and args.workload == "synthetic" and not args.thermo_only:
        flow = flow_leg(args, lib, dev, stream, torch, local)

    if rank == 0:
        pk = ctypes.c_double()
        rc = lib.dfchem_bench_fp64_peak(local, ctypes.byref(pk))
        peak = pk.value if rc == 0 else None
        flops = algorithmic_flops(args.mech, n, cnt["accepted"], cnt["rejected"]) if args.mech in FLOP_MODEL else None
        achieved = flops / (ms_local * 1e-3) / 1e12 if flops else None
        traffic = None
        tfile = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tfile):
            try:
                rec = json.load(open(tfile)).get(args.mech)
                traffic = rec["dram_bytes_per_cell"] * n if rec else None   # per launch, scaled from the ncu capture
            except (OSError, ValueError):
                traffic = None
        roof = {"bound": "fp64", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                "frac": (achieved / peak) if (achieved and peak) else None, "traffic": traffic,
                "peak_source": "dfchem_bench_fp64_peak (DFMA microbenchmark, measured in this run; MEASURED_PEAKS.json has no FP64 entry)",
                "kernel": lib.dfchem_solver_name(h.h).decode(), "algorithmic_flops_per_launch": flops,
                "flop_model": "SURVEY 8d with N = all species of the mechanism (36 for gri30; the kernel factorises the 34 active ones)",
                "hbm_algorithmic_GBps": (2 * N + 6) * 8 * n / (ms_local * 1e-3) / 1e9,
                "step_attempts_per_s": (cnt["accepted"] + cnt["rejected"]) / (ms_local * 1e-3)}
        if configs and peak:
            for v in configs.values():
                if "achieved_TFLOPs" in v:
                    v["frac_of_fp64_peak"] = v["achieved_TFLOPs"] / peak
        out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
               "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
               "data": "synthetic", "config": workload_config(args, n), "clocks": clocks, "e2e": e2e,
               "gpu_launches": args.steps * 1, "roofline": roof, "rebalance": bool(args.rebalance and world > 1), "skew": args.skew,
               "integrator_stats": istats}
        if strong is not None:
            out["strong"] = strong
        if flow is not None:
            out["flow_steps"] = flow
        if configs is not None:
            out["configs"] = configs
        if thermo is not None:
            hbm = None
            try:
                hbm = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
                src = "MEASURED_PEAKS.json hbm_gbs (measured copy bandwidth)"
            except (OSError, KeyError, ValueError):
                hbm, src = 6650.0, "fallback 6.65 TB/s (B200_PROFILING.md)"
            for v in thermo.values():
                if v.get("bound") == "fp64":   # transport_kernel: N^2 pair terms per cell
                    v["roofline"] = {"bound": "fp64", "achieved": v["achieved_TFLOPs"], "peak": peak, "unit": "TFLOP/s",
                                     "frac": v["achieved_TFLOPs"] / peak if peak else None,
                                     "peak_source": "dfchem_bench_fp64_peak (measured in this run)", "kernel": "transport_kernel"}
                    continue
                v["roofline"] = {"bound": "hbm", "achieved": v["achieved_GBps"], "peak": hbm, "unit": "GB/s",
                                 "frac": v["achieved_GBps"] / hbm, "peak_source": src, "kernel": "thermo_kernel"}
            out["thermo"] = thermo
        if not args.no_parity and args.mech == "gri30":
            out["parity"] = parity_statement(args, dfchem_host, synth, local)
        if world == 1 and not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            probe = min(n, 16 * threads)
            r0, el0, _ = cpu_reference_rate(args.mech, batch, probe, args.rtol, args.atol, args.dt, threads)
            sample = int(min(n, max(probe, r0 * 15.0)))
            r1, el1, bsteps = cpu_reference_rate(args.mech, batch, sample, args.rtol, args.atol, args.dt, threads)
            out["cpu_baseline"] = {"value": r1, "unit": UNIT, "cores": threads, "kind": "port",
                                   "sample": f"first {sample} cells of the same batch, {el1:.1f} s, "
                                             f"{bsteps:.1f} BDF steps/cell (oracle/c: CVODE-equivalent BDF + FD Jacobian + dense LU, "
                                             f"CVODE matrix-reuse policy, -O3 -march=native)"}
        sys.stdout.flush()
        os.dup2(saved_stdout, 1)
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


def load_transport(h, mech, synth):
    """Transport fits for the bench's mu/alpha/rhoD leg: the mechanism compiler's own generator when the library
    has one (dfchem_generate_transport_fits), else synthetic fits of Cantera's functional form."""
    lib = h.lib
    rc = lib.dfchem_generate_transport_fits(h.h, None, None, None)
    if rc == 0:
        return "generated by the mechanism compiler from the species' Lennard-Jones data (dfchem_generate_transport_fits)"
    visc, cond, diff = synth.synthetic_transport_fits(h.W)
    h.set_transport_fits(visc, cond, diff)
    return "synthetic (Cantera's ln T polynomial form)"


def flow_leg(args, lib, dev, stream, torch, local, steps=4):
    """What a running flow solver sees: the same cells through `steps` consecutive flow steps -- chemistry sub-step, Y += RR dt/rho,
    correctThermo at constant enthalpy (T, psi, rho follow) -- once with every sub-step cold-started (the reference's CVODE path:
    a new ReactorNet per cell and call, dfChemistryModel.C:748-764) and once with the deltaTChem field carried from step to step
    (dfchem_bind_deltaTChem: the warm start of OpenFOAM's StandardChemistryModel, part of the boundary).  Reported per flow step:
    chemistry time and step attempts per cell; and how far the two final states are apart in units of the tolerance."""
    import dfchem_host
    from dfchem_host import synth

    h = dfchem_host.Handle(synth.load_blob(args.mech), local)
    h.set_tolerances(args.rtol, args.atol)
    n, N, dt = args.flow_cells, h.N, args.dt
    b = synth.make_batch(args.mech, h.species, h.W, n, seed=12345)
    res, finals = {}, {}
    for mode in ("cold_start", "deltaTChem_carried"):
        T = torch.from_numpy(b["T"]).to(dev)
        p = torch.from_numpy(b["p"]).to(dev)
        Y = torch.from_numpy(np.ascontiguousarray(b["Y"])).to(dev)
        Y /= Y.sum(0, keepdim=True)
        he, psi, rho, Q, cost = (torch.empty(n, dtype=torch.float64, device=dev) for _ in range(5))
        RR = torch.empty((N, n), dtype=torch.float64, device=dev)
        stat = torch.empty(n, dtype=torch.int32, device=dev)
        dtc = torch.zeros(n, dtype=torch.float64, device=dev)
        h.check(lib.dfchem_h_from_T_dev(h.h, n, T.data_ptr(), Y.data_ptr(), n, he.data_ptr(), stream.cuda_stream))

        def thermo():
            h.check(lib.dfchem_correct_thermo_dev(h.h, n, he.data_ptr(), p.data_ptr(), Y.data_ptr(), n, T.data_ptr(), psi.data_ptr(),
                                                  rho.data_ptr(), None, None, n, stream.cuda_stream))

        thermo()
        h.bind_deltaTChem(dtc.data_ptr() if mode == "deltaTChem_carried" else None)
        ms, att, amax = [], [], []
        for _ in range(steps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record(stream)
            h.check(lib.dfchem_solve_dev(h.h, n, T.data_ptr(), p.data_ptr(), rho.data_ptr(), None, dt, Y.data_ptr(), n, RR.data_ptr(), n,
                                         Q.data_ptr(), cost.data_ptr(), stat.data_ptr(), stream.cuda_stream))
            e1.record(stream)
            torch.cuda.synchronize()
            assert int((stat != 0).sum().item()) == 0
            ms.append(e0.elapsed_time(e1))
            att.append(float(cost.mean().item()))
            amax.append(float(cost.max().item()))
            Y += RR * (dt / rho)[None, :]
            Y.clamp_(min=0.0)
            Y /= Y.sum(0, keepdim=True)
            thermo()
        h.bind_deltaTChem(None)
        torch.cuda.synchronize()
        finals[mode] = (Y.clone(), T.clone())
        res[mode] = {"ms_chemistry_per_flow_step": ms, "mean_attempts_per_cell": att, "max_attempts_per_cell": amax,
                     "cells_per_s_last_step": n / (ms[-1] * 1e-3)}
    Yc, Tc = finals["cold_start"]
    Yw, Tw = finals["deltaTChem_carried"]
    per_cell = ((Yw - Yc).abs() / (args.rtol * Yc.abs() + args.atol)).max(0).values
    tol_units = per_cell.max().item()
    tol_p99 = torch.quantile(per_cell[: 1 << 23], 0.99).item()
    h.close()
    return {"workload": f"{args.mech}, {n} synthetic cells (the headline batch's first cells), {steps} consecutive flow steps of deltaT = {dt:g} s: "
                        "chemistry sub-step, Y += RR dt/rho, correctThermo at constant enthalpy; rtol/atol as the headline",
            **res, "final_state_difference": {"max_abs_dY_in_tolerance_units": tol_units, "p99_abs_dY_in_tolerance_units": tol_p99,
                                              "max_abs_dT_K": float((Tw - Tc).abs().max().item()),
                                              "note": "carried vs cold after the last flow step, per-cell maximum over the species; tolerance unit = "
                                                      "rtol |Y| + atol per step; cells in the induction phase of an ignition amplify any "
                                                      "tolerance-level difference exponentially from step to step (the maximum), the bulk does not (p99)"}}


def c5_leg(args, h, lib, dev, stream, time_steps, world, torch, local=0):
    """Config C5: the real 262,144-cell drm19 flame field of examples/dfLowMachFoam/notorch/twoD_reactingTGV/CH4
    (fixture deepflame-dev_b200/data/c5_drm19_fields.npz) through chemistry + correctThermo back to back, as shipped
    (every cell at its chemical equilibrium) and neighbour-blended at constant enthalpy (off equilibrium)."""
    import dfchem_host
    from dfchem_host import synth

    own = h is None
    if own:
        h = dfchem_host.Handle(synth.load_blob("drm19"), local)
    h.set_tolerances(1e-6, 1e-10)     # the case's odeCoeffs (constant/CanteraTorchProperties)
    f = np.load(os.path.join(ROOT, "deepflame-dev_b200", "data", "c5_drm19_fields.npz"))
    n, N = f["T"].shape[0], h.N
    dt = 1e-6
    res = {}
    T0 = torch.from_numpy(f["T"]).to(dev)
    Y0 = torch.from_numpy(np.ascontiguousarray(f["Y"])).to(dev)
    p = torch.full((n,), float(f["p"]), dtype=torch.float64, device=dev)
    he0 = torch.empty(n, dtype=torch.float64, device=dev)
    h.check(lib.dfchem_h_from_T_dev(h.h, n, T0.data_ptr(), Y0.data_ptr(), n, he0.data_ptr(), stream.cuda_stream))
    for label in ("as_shipped", "neighbour_blended"):
        if label == "as_shipped":
            T, Y, he = T0.clone(), Y0, he0
        else:   # blend every cell with the one 8 columns further along x (512-cell rows) at constant enthalpy
            sh = 8
            Y = (0.5 * (Y0.view(N, 512, 512) + torch.roll(Y0.view(N, 512, 512), -sh, dims=2))).reshape(N, n).contiguous()
            he = (0.5 * (he0.view(512, 512) + torch.roll(he0.view(512, 512), -sh, dims=1))).reshape(n).contiguous()
            T = (0.5 * (T0.view(512, 512) + torch.roll(T0.view(512, 512), -sh, dims=1))).reshape(n).contiguous()
        psi, rho, cp = (torch.empty(n, dtype=torch.float64, device=dev) for _ in range(3))
        hai = torch.empty((N, n), dtype=torch.float64, device=dev)
        RR = torch.empty((N, n), dtype=torch.float64, device=dev)
        Q, cost = (torch.empty(n, dtype=torch.float64, device=dev) for _ in range(2))
        stat = torch.empty(n, dtype=torch.int32, device=dev)
        Tw = T.clone()

        def thermo():
            h.check(lib.dfchem_correct_thermo_dev(h.h, n, he.data_ptr(), p.data_ptr(), Y.data_ptr(), n, Tw.data_ptr(), psi.data_ptr(),
                                                  rho.data_ptr(), cp.data_ptr(), hai.data_ptr(), n, stream.cuda_stream))

        def chem():
            h.check(lib.dfchem_solve_dev(h.h, n, Tw.data_ptr(), p.data_ptr(), rho.data_ptr(), None, dt, Y.data_ptr(), n,
                                         RR.data_ptr(), n, Q.data_ptr(), cost.data_ptr(), stat.data_ptr(), stream.cuda_stream))

        thermo()                                   # consistent (T, psi, rho) for the state
        torch.cuda.synchronize()

        def step():                                # one flow step's cell-local work: chemistry, then correctThermo
            chem()
            thermo()

        _, ms = time_steps(step, 5, 3)
        _, ms_chem = time_steps(chem, 5, 1)
        assert int((stat != 0).sum().item()) == 0
        res[label] = {"cells": n, "ms_chem_plus_thermo": ms, "ms_chem": ms_chem, "chem_cells_per_s": n / (ms_chem * 1e-3),
                      "step_cells_per_s": n / (ms * 1e-3), "mean_attempts_per_cell": float(cost.mean().item()),
                      "max_attempts_per_cell": float(cost.max().item()),
                      "T_range_K": [float(Tw.min().item()), float(Tw.max().item())],
                      "max_abs_RR": float(RR.abs().max().item())}
    out = {"c5": {"workload": "examples/dfLowMachFoam/notorch/twoD_reactingTGV/CH4/0 (drm19, 512 x 512 cells, p = 101325 Pa), "
                              "rtol 1e-6 / atol 1e-10, deltaT 1e-6: chemistry + correctThermo (T, psi, rho, cp, hai) back to back; "
                              "data: the reference's own field files (fixture c5_drm19_fields.npz)", **res},
           "metric": METRIC, "unit": UNIT, "value": res["neighbour_blended"]["chem_cells_per_s"], "n_gpus": world}
    if own:
        h.close()
    return out


def strong_leg(args, h, lib, dev, rank, world, DeviceBatch, time_steps, sum_over_ranks, max_over_ranks, synth, torch):
    """Config C3 as written: ONE batch of --strong-cells cells split over the N GPUs, flame-like skew (rank r holds
    the cells with T in the r-th of N equal slices of [300, 2500] K, so the hot ranks carry about twice the work),
    timed without and with the in-library rebalancer (NCCL in the timed region)."""
    n_s = args.strong_cells // world
    lo = 300.0 + 2200.0 * rank / world
    b = synth.make_batch(args.mech, h.species, h.W, n_s, seed=777 + rank, T_range=(lo, lo + 2200.0 / world))
    sb = DeviceBatch(h, b)
    sb.solve(args.dt)
    torch.cuda.synchronize()
    prev_cost = sb.cost.clone()
    ref_RR = sb.RR.clone()
    load = float(prev_cost.sum().item())
    loads = [0.0] * world
    t = torch.zeros(world, dtype=torch.float64, device=dev)
    t[rank] = load
    import torch.distributed as dist
    dist.all_reduce(t)
    loads = [float(x) for x in t.tolist()]
    _, ms_plain = time_steps(lambda: sb.solve(args.dt), 2, 1)
    _, ms_bal = time_steps(lambda: sb.solve_balanced(args.dt, prev_cost), 2, 1)
    torch.cuda.synchronize()
    identical = bool(torch.equal(sb.RR, ref_RR))
    st = (ctypes.c_double * 8)()
    h.check(lib.dfchem_balance_stats(h.h, st))
    sent_cells = sum_over_ranks(st[0])
    sent_bytes = sum_over_ranks(st[2])
    plan_ms, xchg_ms = max_over_ranks(st[3]), max_over_ranks(st[4])
    guest_ms, own_ms, call_ms = max_over_ranks(st[5]), max_over_ranks(st[6]), max_over_ranks(st[7])
    same_all = sum_over_ranks(1.0 if identical else 0.0) == world
    total = n_s * world
    del sb
    return {"cells_total": total, "cells_per_gpu": n_s, "skew": "rank r: T ~ U[300 + 2200 r/N, 300 + 2200 (r+1)/N] K",
            "algorithm": args.balance_algorithm, "rank_loads_step_attempts": loads,
            "load_imbalance_max_over_mean": max(loads) / (sum(loads) / world),
            "without_rebalance": {"value": total / (ms_plain * 1e-3), "unit": UNIT, "ms_per_step": ms_plain},
            "with_rebalance": {"value": total / (ms_bal * 1e-3), "unit": UNIT, "ms_per_step": ms_bal},
            "speedup": ms_plain / ms_bal, "cells_migrated_per_step": sent_cells, "bytes_moved_per_step": sent_bytes,
            "ms_plan": plan_ms, "ms_exchange_out": xchg_ms, "ms_solve_guests": guest_ms, "ms_solve_residents": own_ms,
            "ms_balanced_call": call_ms, "exchange_fraction_of_step": (plan_ms + xchg_ms) / ms_bal,
            "results_bit_identical_to_unbalanced": same_all}


def parity_statement(args, dfchem_host, synth, local):
    """Accuracy of the timed configuration, on the committed 256-cell golden set of the mechanism (converged oracle
    RR, tests/golden/solve256_<mech>.npz): the device at the benchmark's (rtol, atol), the CVODE-equivalent port at
    the same tolerances, and the device tolerance at which the strict 1e-6 RR gate passes on all 256 cells."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from parity_common import golden256, port_solve, stats

    g = golden256(args.mech)
    n = g["T"].shape[0]
    h = dfchem_host.Handle(synth.load_blob(args.mech), local)
    h.check(h.lib.dfchem_set_max_steps(h.h, 400000))
    arrs = [np.ascontiguousarray(g[k]) for k in ("T", "p", "rho", "dt", "Y")]

    def dev(rtol, atol):
        h.set_tolerances(rtol, atol)
        cost = np.zeros(n)
        RR, _ = h.solve(*arrs, cost=cost)
        return stats(RR, g, rtol, atol), float(cost.mean())

    at_bench, att = dev(args.rtol, args.atol)
    RRp, steps, _ = port_solve(args.mech, g, args.rtol, args.atol)
    port = stats(RRp, g, args.rtol, args.atol)
    gate = None
    for rtol, atol in [(1e-6, 1e-10), (1e-7, 1e-12), (1e-8, 1e-14), (1e-9, 1e-15), (1e-10, 1e-18), (1e-11, 1e-20)]:
        st, a = dev(rtol, atol)
        if st["parity_max"] <= 1.0:
            gate = {"rtol": rtol, "atol": atol, "parity_max": st["parity_max"], "attempts_per_cell": a}
            break
    h.close()
    return {"cells": n, "golden": f"tests/golden/solve256_{args.mech}.npz (SciPy Radau, rtol 1e-12)",
            "measures": "per-cell maxima; tol_units = |RR - RR_conv| / (rho/dt (rtol |Y| + atol)); parity = |RR - RR_conv| / "
                        "(1e-6 |RR_conv| + rho/dt (1e-9 |Y| + 1e-15))",
            "err_at_bench_tol": at_bench, "attempts_per_cell_at_bench_tol": att,
            "err_cvode_port_at_bench_tol": port, "port_steps_per_cell": float(steps.mean()),
            "rtol_gate_passes_at": gate}


if __name__ == "__main__":
    main()
