"""TEST INFRASTRUCTURE (oracle) -- never imported by the product path.

Converged CPU restatement of `dfChemistryModel::solveSingle`
(src/dfChemistryModel/dfChemistryModel.C:738-780) and of the operator-split 0-D loop of
df0DFoam (applications/solvers/df0DFoam/df0DFoam.C:106-111, YEqn.H:15, EEqn.H:5), using
SciPy's implicit integrators at tolerances far tighter than the reference's CVODE settings so
that the result is the *converged* solution of the same ODE.  parity unpinned.
"""
from __future__ import annotations

import numpy as np
from scipy.integrate import solve_ivp

from . import kinetics as K
from .mech import GAS_CONSTANT, Mechanism


def numerical_jacobian(f, y, eps=1e-7):
    f0 = f(y)
    J = np.zeros((len(y), len(y)))
    for j in range(len(y)):
        d = eps * max(abs(y[j]), 1e-12)
        yp = y.copy(); yp[j] += d
        ym = y.copy(); ym[j] -= d
        J[:, j] = (f(yp) - f(ym)) / (2 * d)
    return J


def solve_single(m: Mechanism, T, p, rho_of, Y, dt, rtol=1e-12, atol=1e-20, method="Radau"):
    """One cell of solveSingle.  Returns (RR[N] kg/m3/s, Qdot J/m3/s, Ynew[N]).

    setState_TPY normalises Y and sets rho_ct = p Wbar/(R T) (dfChemistryModel.C:755);
    Reactor with energy off integrates dY/dt = wdot W / rho_ct at frozen T, rho_ct (:748-764);
    RR_i = (Ynew_i - Yraw_i)/dt * rho_OpenFOAM, Qdot = -sum hc_i RR_i (:768-772).
    """
    Y = np.asarray(Y, float)
    Yn = Y / Y.sum()
    rho = float(K.density(m, np.array([T]), np.array([p]), Yn[None])[0])
    rc = K.RateConstants(m, np.array([T]))
    scale = m.W / rho

    def f(t, y):
        C = (y / scale)[None]
        return K.wdot(m, rc, C)[0] * scale

    def jac(t, y):
        return numerical_jacobian(lambda z: f(0.0, z), y)

    sol = solve_ivp(f, (0.0, dt), Yn, method=method, rtol=rtol, atol=atol, jac=jac)
    assert sol.success, sol.message
    Ynew = sol.y[:, -1]
    RR = (Ynew - Y) / dt * rho_of
    Qdot = -(m.hc * RR).sum()
    return RR, Qdot, Ynew


def T_from_h(m: Mechanism, h, Y, T0, tol=1e-12, maxit=100):
    """Newton on T with cp as slope (setState_HP, dfChemistryModel.C:495-503; same iteration as
    src_gpu/dfThermo.cu:296-320 run to 1e-12)."""
    T = float(T0)
    Y = np.asarray(Y)[None]
    for _ in range(maxit):
        hh = K.h_mass(m, np.array([T]), Y)[0]
        cp = K.cp_mass(m, np.array([T]), Y)[0]
        dT = (hh - h) / cp
        dT = max(-100.0, min(100.0, dT))
        T -= dT
        if abs(dT) < tol * T:
            break
    return T


def zero_d_const_pressure(m: Mechanism, T0, p, Y0, dt, nsteps, rtol=1e-10, atol=1e-18):
    """df0DFoam operator split: chemistry at frozen (T, rho) for dt, Y += RR dt / rho,
    h conserved (constant pressure), T from h.  Returns T history [nsteps+1]."""
    Y = np.asarray(Y0, float).copy()
    T = float(T0)
    h = K.h_mass(m, np.array([T]), Y[None])[0]
    Ts = [T]
    for _ in range(nsteps):
        rho = K.density(m, np.array([T]), np.array([p]), Y[None])[0]
        RR, _, _ = solve_single(m, T, p, rho, Y, dt, rtol, atol)
        Y = Y + RR * dt / rho
        T = T_from_h(m, h, Y, T)
        Ts.append(T)
    return np.array(Ts)
