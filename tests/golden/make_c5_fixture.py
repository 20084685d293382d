"""Config C5 (BASELINE.json configs[4]): the real CH4 flame state the reference ships for dfLowMachFoam --
examples/dfLowMachFoam/notorch/twoD_reactingTGV/CH4/0/{T,<21 drm19 species>}.gz, 262,144 cells (512 x 512),
p uniform 101325 Pa, mechanism drm19, odeCoeffs 1e-6 / 1e-10, deltaT 1e-6 (constant/CanteraTorchProperties,
system/controlDict).  Run in the build container (needs /root/reference); writes

  deepflame-dev_b200/data/c5_drm19_fields.npz   the whole field set (T [n], Y [21, n] in the mechanism's species order, p)
  tests/golden/c5_drm19_sample.npz              the 512 distinct states of the field (kind 0) + 512 neighbour-blended,
                                                 off-equilibrium states (kind 1), with the converged oracle RR / Qdot
                                                 (SciPy Radau, rtol 1e-12) and the thermo oracle (he, cp, psi, rho)

usage: python tests/golden/make_c5_fixture.py [--workers 6]
"""
import gzip
import multiprocessing as mp
import os
import re
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "deepflame-dev_b200"))
from dfchem_host import synth  # noqa: E402
from oracle import integrate as OI  # noqa: E402
from oracle import kinetics as K  # noqa: E402
from oracle import mech as OM  # noqa: E402

CASE = "/root/reference/examples/dfLowMachFoam/notorch/twoD_reactingTGV/CH4/0"
DT = 1e-6
BLEND = 8


def read_field(path):
    """internalField of an OpenFOAM ASCII volScalarField (.gz): nonuniform List<scalar> or uniform."""
    txt = gzip.open(path, "rt").read()
    mu = re.search(r"internalField\s+uniform\s+([-+0-9.eE]+)", txt)
    if mu:
        return float(mu.group(1))
    mm = re.search(r"internalField\s+nonuniform\s+List<scalar>\s*(\d+)\s*\(", txt)
    n = int(mm.group(1))
    body = txt[mm.end():]
    vals = np.array(body[:body.index(")")].split(), dtype=np.float64)
    assert vals.size == n, (path, vals.size, n)
    return vals


_m = None


def _solve(args):
    global _m
    if _m is None:
        _m = OM.deserialize(synth.load_blob("drm19"))
    T, p, rho, Y = args
    RR, Q, _ = OI.solve_single(_m, T, p, rho, Y, DT, rtol=1e-12, atol=1e-22)
    return RR, Q


def main():
    workers = int(sys.argv[sys.argv.index("--workers") + 1]) if "--workers" in sys.argv else 6
    m = OM.deserialize(synth.load_blob("drm19"))
    T = read_field(os.path.join(CASE, "T.gz"))
    n = T.size
    p = read_field(os.path.join(CASE, "p.gz"))
    assert isinstance(p, float)
    Y = np.zeros((m.N, n))
    for k, s in enumerate(m.species):
        v = read_field(os.path.join(CASE, f"{s}.gz"))
        Y[k] = v
    print("cells", n, "T", T.min(), T.max(), "sum Y", Y.sum(0).min(), Y.sum(0).max())
    np.savez_compressed(os.path.join(ROOT, "deepflame-dev_b200", "data", "c5_drm19_fields.npz"), T=T, p=np.float64(p), Y=Y,
                        species=np.array(m.species))
    # The shipped field is a 1-D double-flame profile along x repeated on all 512 rows, and every cell sits at the
    # chemical equilibrium of its own (T, p, elements): net rates vanish to round-off (max |RR| ~ 1e-10 kg/m3/s).
    # Parity sample = (a) the 512 distinct states of one row, as shipped, and (b) the same 512 columns each blended
    # 50/50 with the column `BLEND` cells further along x at constant enthalpy (what diffusion between neighbours does
    # over a few flow steps), which moves the states off equilibrium so the chemistry step has work to do.
    T2, Y3 = T.reshape(512, 512), Y.reshape(m.N, 512, 512)
    assert np.abs(np.diff(T2, axis=0)).max() == 0.0
    row_T, row_Y = T2[0], np.ascontiguousarray(Y3[:, 0, :])
    Yn0 = (row_Y / row_Y.sum(0, keepdims=True)).T
    h0 = K.h_mass(m, row_T, Yn0)
    Yb = 0.5 * (row_Y + np.roll(row_Y, -BLEND, axis=1))
    hb = 0.5 * (h0 + np.roll(h0, -BLEND))
    Tb = np.array([OI.T_from_h(m, hb[c], Yb[:, c] / Yb[:, c].sum(), 0.5 * (row_T[c] + np.roll(row_T, -BLEND)[c])) for c in range(512)])
    pick = np.concatenate([np.arange(512), np.arange(512)])
    kind = np.concatenate([np.zeros(512, int), np.ones(512, int)])
    Ts = np.concatenate([row_T, Tb])
    Ys = np.ascontiguousarray(np.concatenate([row_Y, Yb], axis=1))
    ps = np.full(pick.size, p)
    Yn = (Ys / Ys.sum(0, keepdims=True)).T
    rho = K.density(m, Ts, ps, Yn)       # the OpenFOAM rho of a consistent thermo state: p Wbar / (R T)
    with mp.Pool(workers) as pool:
        res = pool.map(_solve, [(Ts[c], ps[c], rho[c], Ys[:, c]) for c in range(pick.size)], chunksize=4)
    RR = np.stack([r[0] for r in res], axis=1)
    Q = np.array([r[1] for r in res])
    he = K.h_mass(m, Ts, Yn)
    cp = K.cp_mass(m, Ts, Yn)
    psi = K.mean_W(m, Yn) / (8314.46261815324 * Ts)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "c5_drm19_sample.npz"), cell=pick, kind=kind, blend=np.int64(BLEND), T=Ts, p=ps, rho=rho, Y=Ys,
                        dt=np.full(pick.size, DT), RR=RR, Qdot=Q, he=he, cp=cp, psi=psi, species=np.array(m.species))
    print("sample written", pick.size, "max |RR|", np.abs(RR).max())


if __name__ == "__main__":
    main()
