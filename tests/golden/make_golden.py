"""Generates tests/golden/solve_<mech>.npz in the build container: inputs (T, p, rho, Y, dt) and
the converged oracle outputs (RR, Qdot) of `solveSingle` for a handful of cells per mechanism,
plus instantaneous wdot and the thermo round trip.  The oracle (oracle/*.py, SciPy Radau at
rtol 1e-12) reads the mechanism blobs produced by the mechanism compiler from the reference's
mechanisms/*.yaml.  usage: python tests/golden/make_golden.py [mech ...]
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "deepflame-dev_b200"))
from dfchem_host import synth  # noqa: E402  (pure NumPy part of the host package)
from oracle import integrate as OI  # noqa: E402
from oracle import kinetics as K  # noqa: E402
from oracle import mech as OM  # noqa: E402

CASES = {"es80": 16, "h2_li": 12, "drm19": 10, "gri30": 8, "c7_onestep": 12, "ethanol66": 6}


def main():
    names = sys.argv[1:] or list(CASES)
    for name in names:
        n = CASES[name]
        m = OM.deserialize(synth.load_blob(name))
        b = synth.make_batch(name, m.species, m.W, n, seed=2024)
        # make the batch less trivial: a few per-cell time steps and one un-normalised cell
        dt = np.full(n, 1e-6)
        dt[1::4] = 5e-7
        dt[2::4] = 2e-6
        Y = b["Y"].copy()
        Y[:, 0] *= 1.001
        RR = np.zeros((m.N, n))
        Q = np.zeros(n)
        for c in range(n):
            RR[:, c], Q[c], _ = OI.solve_single(m, b["T"][c], b["p"][c], b["rho"][c], Y[:, c], dt[c],
                                                rtol=1e-12, atol=1e-22)
            print(name, c, b["T"][c], np.abs(RR[:, c]).max(), flush=True)
        Yn = (Y / Y.sum(0, keepdims=True)).T
        rho_ct = K.density(m, b["T"], b["p"], Yn)
        C = rho_ct[:, None] * Yn / m.W[None]
        wrate = (K.wdot(m, K.RateConstants(m, b["T"]), C) * m.W[None]).T
        he = K.h_mass(m, b["T"], Yn)
        cp = K.cp_mass(m, b["T"], Yn)
        np.savez_compressed(os.path.join(ROOT, "tests", "golden", f"solve_{name}.npz"), T=b["T"], p=b["p"],
                            rho=b["rho"], Y=Y, dt=dt, RR=RR, Qdot=Q, wrate=wrate, he=he, cp=cp,
                            species=np.array(m.species))


if __name__ == "__main__":
    main()
