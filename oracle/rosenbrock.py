"""TEST INFRASTRUCTURE (oracle) -- never imported by the product path.

NumPy model of the batched Rosenbrock integrator the CUDA path implements
(deepflame-dev_b200/csrc/integrator.cuh).  Used (a) to choose the scheme, (b) by the
CPU tests to pin the step controller's host logic.  All cells advance in lock-step with
per-cell step sizes; finished cells are masked out.

Unknowns are molar concentrations C_k = rho Y_k / W_k of the frozen-(T, rho) reactor of
solveSingle (src/dfChemistryModel/dfChemistryModel.C:748-764); the error weights restate
Cantera/CVODE's rtol*|Y_k| + atol (dfChemistryModel.C:414) in concentration units.
"""
from __future__ import annotations

import numpy as np

from . import kinetics as K

# RODAS4 (Hairer & Wanner, Solving ODEs II, sec. IV.7), 6 stages, order 4(3), stiffly accurate.
RODAS4 = dict(
    s=6, order=4, gamma=0.25,
    a={(1, 0): 0.1544000000000000e+01,
       (2, 0): 0.9466785280815826, (2, 1): 0.2557011698983284,
       (3, 0): 0.3314825187068521e+01, (3, 1): 0.2896124015972201e+01, (3, 2): 0.9986419139977817,
       (4, 0): 0.1221224509226641e+01, (4, 1): 0.6019134481288629e+01, (4, 2): 0.1253708332932087e+02,
       (4, 3): -0.6878860361058950,
       (5, 0): 0.1221224509226641e+01, (5, 1): 0.6019134481288629e+01, (5, 2): 0.1253708332932087e+02,
       (5, 3): -0.6878860361058950, (5, 4): 1.0},
    c={(1, 0): -0.5668800000000000e+01,
       (2, 0): -0.2430093356833875e+01, (2, 1): -0.2063599157091915,
       (3, 0): -0.1073529058151375, (3, 1): -0.9594562251023355e+01, (3, 2): -0.2047028614809616e+02,
       (4, 0): 0.7496443313967647e+01, (4, 1): -0.1024680431464352e+02, (4, 2): -0.3399990352819905e+02,
       (4, 3): 0.1170890893206160e+02,
       (5, 0): 0.8083246795921522e+01, (5, 1): -0.7981132988064893e+01, (5, 2): -0.3152159432874371e+02,
       (5, 3): 0.1631930543123136e+02, (5, 4): -0.6058818238834054e+01},
    m=[0.1221224509226641e+01, 0.6019134481288629e+01, 0.1253708332932087e+02, -0.6878860361058950, 1.0, 1.0],
    e=[0, 0, 0, 0, 0, 1.0],
)

# RODAS3 (Sandu et al. 1997), 4 stages, order 3(2), stiffly accurate.
RODAS3 = dict(
    s=4, order=3, gamma=0.5,
    a={(1, 0): 0.0, (2, 0): 2.0, (2, 1): 0.0, (3, 0): 2.0, (3, 1): 0.0, (3, 2): 1.0},
    c={(1, 0): 4.0, (2, 0): 1.0, (2, 1): -1.0, (3, 0): 1.0, (3, 1): -1.0, (3, 2): -8.0 / 3.0},
    m=[2.0, 0.0, 1.0, 1.0],
    e=[0, 0, 0, 1.0],
)


def integrate(m, T, rho, C0, dt, rtol, atol_Y, scheme=RODAS4, h0=None, max_steps=20000,
              fac_min=0.2, fac_max=6.0, safety=0.9, stats=None, err_on_increment=0.0):
    """Advance C' = wdot(C) over dt for every cell.  T, rho [nc] frozen.  Returns C(dt).

    err_on_increment > 0 adds the weight  err_on_increment * |C - C0|  to the tolerance scale
    replacing rtol*|C| when smaller (increment-relative control used for RR parity)."""
    nc, N = C0.shape
    rc = K.RateConstants(m, T)
    C = C0.copy()
    dt = np.broadcast_to(np.asarray(dt, float), (nc,)).copy()
    t = np.zeros(nc)
    atolC = atol_Y * rho[:, None] / m.W[None]
    s = scheme["s"]; g = scheme["gamma"]; A = scheme["a"]; Cc = scheme["c"]
    p = scheme["order"]
    h = dt.copy() if h0 is None else np.minimum(np.broadcast_to(h0, (nc,)).copy(), dt)
    nsteps = np.zeros(nc, int); nrej = np.zeros(nc, int)
    active = np.ones(nc, bool)
    eye = np.eye(N)[None]
    it = 0
    while active.any():
        it += 1
        if it > max_steps:
            raise RuntimeError("max steps")
        idx = np.nonzero(active)[0]
        Ca = C[idx]; ha = np.minimum(h[idx], dt[idx] - t[idx])
        rca = _subset(rc, idx)
        J = K.jacobian_C(m, rca, Ca)
        Amat = eye / (ha[:, None, None] * g) - J
        f0 = K.wdot(m, rca, Ca)
        ks = []
        ok = np.ones(len(idx), bool)
        try:
            lu = np.linalg.inv(Amat)  # prototype only
        except np.linalg.LinAlgError:
            lu = np.linalg.pinv(Amat)
        for i in range(s):
            if i == 0:
                rhs = f0
            else:
                yi = Ca.copy()
                for j in range(i):
                    aij = A.get((i, j), 0.0)
                    if aij != 0.0:
                        yi = yi + aij * ks[j]
                rhs = K.wdot(m, rca, yi)
                for j in range(i):
                    cij = Cc.get((i, j), 0.0)
                    if cij != 0.0:
                        rhs = rhs + (cij / ha)[:, None] * ks[j]
            ks.append(np.einsum("cij,cj->ci", lu, rhs))
        Cn = Ca.copy(); err = np.zeros_like(Ca)
        for i in range(s):
            if scheme["m"][i] != 0:
                Cn = Cn + scheme["m"][i] * ks[i]
            if scheme["e"][i] != 0:
                err = err + scheme["e"][i] * ks[i]
        sc = atolC[idx] + rtol * np.maximum(np.abs(Ca), np.abs(Cn))
        if err_on_increment > 0:
            inc = np.maximum(np.abs(Cn - C0[idx]), np.abs(Ca - C0[idx]))
            sc = np.minimum(sc, atolC[idx] * err_on_increment_floor(err_on_increment) + err_on_increment * inc)
        en = np.sqrt(np.mean((err / sc) ** 2, axis=1))
        en = np.where(np.isfinite(en), en, 1e10)
        fac = np.clip(safety * np.maximum(en, 1e-10) ** (-1.0 / p), fac_min, fac_max)
        acc = en <= 1.0
        # accepted
        ia = idx[acc]
        C[ia] = Cn[acc]; t[ia] += ha[acc]; nsteps[ia] += 1
        nrej[idx[~acc]] += 1
        h[idx] = ha * fac
        done = t >= dt * (1 - 1e-14)
        active &= ~done
    if stats is not None:
        stats["nsteps"] = nsteps; stats["nrej"] = nrej
    return C


def err_on_increment_floor(_):
    return 1.0


def _subset(rc, idx):
    o = object.__new__(K.RateConstants)
    o.T = rc.T[idx]; o.kf = rc.kf[idx]; o.k0 = rc.k0[idx]; o.log10Fc = rc.log10Fc[idx]; o.rKc = rc.rKc[idx]
    return o
