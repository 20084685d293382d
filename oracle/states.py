"""TEST INFRASTRUCTURE (oracle) -- never imported by the product path.

Synthetic cell states for parity tests and benchmarks: unburnt fuel/air
mixtures and partially-burnt states sampled from constant-pressure adiabatic ignition
trajectories integrated with SciPy (the df0DFoam situation, applications/solvers/df0DFoam).
"""
from __future__ import annotations

import numpy as np
from scipy.integrate import solve_ivp

from . import kinetics as K
from .mech import GAS_CONSTANT, Mechanism

FUELS = {"H2": "H2", "CH4": "CH4", "C7H16": "C7H16", "C2H5OH": "C2H5OH"}


def fuel_air(m: Mechanism, fuel: str, phi):
    """Mass fractions of fuel/air at equivalence ratio phi; air = O2:N2 0.233:0.767 by mass."""
    phi = np.atleast_1d(np.asarray(phi, float))
    k = m.species.index(fuel)
    el = {e: m.comp[k, i] for i, e in enumerate(m.elements)}
    nO2 = el.get("C", 0) + el.get("H", 0) / 4.0 - el.get("O", 0) / 2.0  # kmol O2 per kmol fuel
    WO2 = m.W[m.species.index("O2")]
    st = nO2 * WO2 / m.W[k]                      # kg O2 per kg fuel, stoichiometric
    Y = np.zeros((len(phi), m.N))
    # mass of fuel per kg air: phi * 0.233 / st
    mf = phi * 0.233 / st
    tot = 1.0 + mf
    Y[:, k] = mf / tot
    Y[:, m.species.index("O2")] = 0.233 / tot
    Y[:, m.species.index("N2")] = 0.767 / tot
    return Y


def ignition_trajectory(m: Mechanism, T0, p, Y0, t_end, n_samples, rtol=1e-8, atol=1e-14):
    """Constant-pressure adiabatic ignition; returns (T[ns], Y[ns,N]) sampled uniformly in
    temperature progress (so that the thin ignition front is well represented)."""
    N = m.N

    def f(t, z):
        T = z[0]; Y = np.maximum(z[1:], 0.0)[None]
        Tn = np.array([T])
        rho = p * K.mean_W(m, Y) / (GAS_CONSTANT * Tn)
        C = rho[:, None] * Y / m.W[None]
        w = K.wdot(m, K.RateConstants(m, Tn), C)[0]
        cp = K.cp_mass(m, Tn, Y)[0]
        h = K.h_RT(m, Tn)[0] * GAS_CONSTANT * T
        dT = -(h * w).sum() / (rho[0] * cp)
        return np.concatenate([[dT], w * m.W / rho[0]])

    sol = solve_ivp(f, (0, t_end), np.concatenate([[T0], Y0]), method="BDF", rtol=rtol, atol=atol,
                    dense_output=False)
    T = sol.y[0]
    Y = sol.y[1:].T
    # sample uniformly in arclength of T (progress)
    prog = np.concatenate([[0], np.cumsum(np.abs(np.diff(T)) + 1e-3 * (T.max() - T.min()) / len(T))])
    target = np.linspace(0, prog[-1], n_samples)
    ii = np.clip(np.searchsorted(prog, target), 0, len(T) - 1)
    return T[ii], np.maximum(Y[ii], 0.0)
