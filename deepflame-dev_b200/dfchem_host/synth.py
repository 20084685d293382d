"""Synthetic cell batches for benchmarks and tests (BASELINE.json configs 1-3):
per cell T ~ U[300, 2500] K, p ~ U[1, 10] atm; half the cells hold unburnt fuel/air at
phi ~ U[0.5, 2] (air = O2:N2 0.233:0.767 by mass), half hold partially-burnt compositions
sampled from the ignition-trajectory tables in data/states_<mech>.npz (generated once by
oracle/make_states.py in the build container).  rho = p Wbar/(R T).  Pure NumPy.
"""
from __future__ import annotations

import os

import numpy as np

R_GAS = 8314.46261815324
ATM = 101325.0
_DATA = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "data")

# name -> (blob file, fuel species, kmol O2 per kmol fuel at stoichiometry, states table)
MECHANISMS = {
    "es80": ("es80.blob", "H2", 0.5, "states_es80.npz"),
    "h2_li": ("h2_li.blob", "H2", 0.5, None),
    "drm19": ("drm19.blob", "CH4", 2.0, "states_drm19.npz"),
    "gri30": ("gri30.blob", "CH4", 2.0, "states_gri30.npz"),
    "c7_onestep": ("c7_onestep.blob", "C7H16", 11.0, None),
    "ethanol66": ("ethanol66.blob", "C2H5OH", 3.0, None),
}


def blob_path(name):
    return os.path.join(_DATA, "mech", MECHANISMS[name][0])


def load_blob(name):
    with open(blob_path(name), "rb") as f:
        return f.read()


def fuel_air(species, W, fuel, nu_o2, phi):
    phi = np.atleast_1d(np.asarray(phi, float))
    kf, ko, kn = species.index(fuel), species.index("O2"), species.index("N2")
    st = nu_o2 * W[ko] / W[kf]  # kg O2 per kg fuel
    mf = phi * 0.233 / st
    tot = 1.0 + mf
    Y = np.zeros((len(phi), len(species)))
    Y[:, kf] = mf / tot
    Y[:, ko] = 0.233 / tot
    Y[:, kn] = 0.767 / tot
    return Y


def make_batch(name, species, W, n_cells, seed=12345, burning_fraction=0.5, T_range=(300.0, 2500.0),
               p_range_atm=(1.0, 10.0)):
    """Returns dict(T, p, rho [n], Y [N, n] species-major C-contiguous)."""
    _, fuel, nu_o2, table = MECHANISMS[name]
    rng = np.random.Generator(np.random.PCG64(seed))
    n = int(n_cells)
    T = rng.uniform(T_range[0], T_range[1], n)
    p = rng.uniform(p_range_atm[0], p_range_atm[1], n) * ATM
    Y = fuel_air(species, W, fuel, nu_o2, rng.uniform(0.5, 2.0, n))
    if table is not None and burning_fraction > 0:
        tab = np.load(os.path.join(_DATA, table))
        assert [str(s) for s in tab["species"]] == list(species)
        burning = rng.uniform(0, 1, n) < burning_fraction
        pick = rng.integers(0, tab["Y"].shape[0], n)
        Yb = np.maximum(tab["Y"][pick], 0.0)
        Yb /= Yb.sum(1, keepdims=True)
        Y[burning] = Yb[burning]
    wbar = 1.0 / (Y / W[None, :]).sum(1)
    rho = p * wbar / (R_GAS * T)
    return dict(T=T, p=p, rho=rho, Y=np.ascontiguousarray(Y.T))


def synthetic_transport_fits(W, seed=7):
    """Plausible fits for mechanisms without a coefficient file: the hard-sphere forms
    mu ~ sqrt(W T)/sigma^2, lambda ~ mu cv-like / W, D_ij ~ T^1.5 sqrt(1/W_i + 1/W_j)/sigma_ij^2 with a
    small ln T dependence; symmetric in (i, j) as Cantera's are."""
    rng = np.random.default_rng(seed)
    W = np.asarray(W, float)
    N = W.shape[0]
    sig = rng.uniform(2.5, 5.0, N)
    visc = np.zeros((N, 5))
    cond = np.zeros((N, 5))
    diff = np.zeros((N, N, 5))
    visc[:, 0] = np.sqrt(2.67e-6 * np.sqrt(W) / sig**2)
    visc[:, 1] = visc[:, 0] * rng.uniform(0.01, 0.05, N)
    visc[:, 2] = -visc[:, 0] * rng.uniform(0.0, 0.003, N)
    cond[:, 0] = 2.67e-6 * np.sqrt(W) / sig**2 * 3.0e4 / W
    cond[:, 1] = cond[:, 0] * rng.uniform(0.02, 0.1, N)
    cond[:, 3] = cond[:, 0] * rng.uniform(0.0, 1e-4, N)
    sij = 0.5 * (sig[:, None] + sig[None, :])
    d0 = 2.6e-2 * np.sqrt(0.5 * (1.0 / W[:, None] + 1.0 / W[None, :])) / sij**2
    e1 = rng.uniform(0.01, 0.04, (N, N))
    e1 = 0.5 * (e1 + e1.T)
    diff[:, :, 0] = d0
    diff[:, :, 1] = d0 * e1
    diff[:, :, 4] = -d0 * 1e-5
    return visc, cond, diff
