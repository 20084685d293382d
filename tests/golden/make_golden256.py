"""Generates tests/golden/solve256_<mech>.npz in the build container: 256 cells per mechanism -- stratified in T
(8 strata x p strata), half unburnt fuel/air and half ignition-trajectory states like the bench batch, plus 32 edge cells
(slightly negative trace mass fractions, zero-radical hot mixtures, X_i -> 1 compositions, un-normalised sums, per-cell dt
from 1e-7 to 1e-5 s) -- with the converged oracle RR / Qdot of `solveSingle` (oracle/integrate.py: SciPy Radau at
rtol 1e-12, dfChemistryModel.C:738-780 semantics) and the CVODE-equivalent port's result is NOT stored (it is re-run by
the tests).  usage: python tests/golden/make_golden256.py [--workers 6] [mech ...]
"""
import multiprocessing as mp
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "deepflame-dev_b200"))
from dfchem_host import synth  # noqa: E402
from oracle import integrate as OI  # noqa: E402
from oracle import kinetics as K  # noqa: E402
from oracle import mech as OM  # noqa: E402

MECHS = ["es80", "h2_li", "c7_onestep", "drm19", "gri30", "ethanol66"]
N_MAIN, N_EDGE = 224, 32
_cache = {}


def _solve(args):
    name, T, p, rho, Y, dt = args
    if name not in _cache:
        _cache[name] = OM.deserialize(synth.load_blob(name))
    RR, Q, _ = OI.solve_single(_cache[name], T, p, rho, Y, dt, rtol=1e-12, atol=1e-22)
    return RR, Q


def build_inputs(name, m):
    rng = np.random.default_rng(256)
    b = synth.make_batch(name, m.species, m.W, N_MAIN, seed=4242)
    # stratify: T in 8 equal strata of [300, 2500] K, p in [1, 10] atm, preserving each cell's composition class
    strata = np.arange(N_MAIN) % 8
    T = 300.0 + (strata + rng.uniform(0, 1, N_MAIN)) * (2200.0 / 8)
    p = rng.uniform(1.0, 10.0, N_MAIN) * 101325.0
    Y = b["Y"].copy()
    dt = np.full(N_MAIN, 1e-6)
    # edge cells
    _, fuel, nu_o2, _ = synth.MECHANISMS[name]
    sp = list(m.species)
    Ye = synth.fuel_air(sp, m.W, fuel, nu_o2, rng.uniform(0.6, 1.6, N_EDGE)).T.copy()    # [N, 32]
    Te = rng.uniform(900.0, 2200.0, N_EDGE)
    pe = rng.uniform(1.0, 10.0, N_EDGE) * 101325.0
    dte = np.full(N_EDGE, 1e-6)
    inert = sp.index("N2")
    others = [k for k in range(m.N) if k != inert]
    for c in range(N_EDGE):
        kind = c % 8
        if kind == 0:      # slightly negative trace species (what a transport step can leave behind)
            ks = rng.choice(others, size=min(3, len(others)), replace=False)
            Ye[ks, c] = -rng.uniform(1e-12, 1e-9, ks.size)
        elif kind == 1:    # zero radicals, hot and dense: the chain-branching start (unstable-mode guard)
            Te[c] = rng.uniform(1500.0, 1900.0)
            pe[c] = rng.uniform(6.0, 10.0) * 101325.0
        elif kind == 2:    # X_i -> 1: a single species
            Ye[:, c] = 0.0
            Ye[rng.choice(m.N), c] = 1.0
        elif kind == 3:    # almost pure inert with a trace of fuel/air
            Ye[:, c] *= 1e-6
            Ye[inert, c] = 1.0 - Ye[others, c].sum()
        elif kind == 4:    # un-normalised sum (RR is taken against the raw Y)
            Ye[:, c] *= rng.choice([0.9, 1.1])
        elif kind == 5:    # short and long flow steps
            dte[c] = rng.choice([1e-7, 1e-5])
        elif kind == 6:    # burnt state pushed to a cold temperature (radical recombination transient)
            Ye[:, c] = Y[:, rng.integers(0, N_MAIN)]
            Te[c] = rng.uniform(300.0, 600.0)
        else:              # cold unburnt: nothing happens
            Te[c] = rng.uniform(300.0, 500.0)
    T = np.concatenate([T, Te]); p = np.concatenate([p, pe]); dt = np.concatenate([dt, dte])
    Y = np.ascontiguousarray(np.concatenate([Y, Ye], axis=1))
    Yn = (Y / Y.sum(0, keepdims=True)).T
    rho = K.density(m, T, p, Yn)
    return T, p, rho, Y, dt


def main():
    args = [a for a in sys.argv[1:]]
    workers = 6
    if "--workers" in args:
        i = args.index("--workers")
        workers = int(args[i + 1])
        del args[i:i + 2]
    names = args or MECHS
    with mp.Pool(workers) as pool:
        for name in names:
            m = OM.deserialize(synth.load_blob(name))
            T, p, rho, Y, dt = build_inputs(name, m)
            n = T.size
            res = pool.map(_solve, [(name, T[c], p[c], rho[c], Y[:, c], dt[c]) for c in range(n)], chunksize=2)
            RR = np.stack([r[0] for r in res], axis=1)
            Q = np.array([r[1] for r in res])
            Yn = (Y / Y.sum(0, keepdims=True)).T
            C = rho[:, None] * Yn / m.W[None]
            wrate = (K.wdot(m, K.RateConstants(m, T), C) * m.W[None]).T
            np.savez_compressed(os.path.join(ROOT, "tests", "golden", f"solve256_{name}.npz"), T=T, p=p, rho=rho, Y=Y, dt=dt,
                                RR=RR, Qdot=Q, wrate=wrate, he=K.h_mass(m, T, Yn), cp=K.cp_mass(m, T, Yn),
                                species=np.array(m.species))
            print(name, n, "cells, max |RR|", np.abs(RR).max(), flush=True)


if __name__ == "__main__":
    main()
