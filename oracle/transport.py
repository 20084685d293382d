"""TEST INFRASTRUCTURE (oracle): mixture transport properties of correctThermo, CPU restatement.

Follows the arithmetic the reference evaluates from Cantera's polynomial fits in ln T
(dfChemistryModel.C:513-541 through MixTransport; restated in-tree on the device in
src_gpu/dfThermo.cu):
  mole fractions, mean molecular weight   dfThermo.cu:54-71
  polynomial basis 1, lnT .. lnT^4        dfThermo.cu:73-86
  Wilke mixture viscosity                 dfThermo.cu:111-151  (constants :40-43)
  mixture conductivity, alpha = lambda/cp dfThermo.cu:172-206
  mixture-averaged diffusion (mass form)  dfThermo.cu:208-255
Fits come in the thermo_<mech>.txt layout (dfThermo.cu:361-435).  Only tests/, smoke() and the
bench's CPU leg may import this module.
"""
import numpy as np

SQRT8 = 2.8284271247461903


def mole_fractions(Y, W):
    """Y [N, n] -> X [N, n], Wbar [n]"""
    s = (Y / W[:, None]).sum(0)
    X = Y / (W[:, None] * s[None])
    return X, (X * W[:, None]).sum(0)


def poly_basis(T):
    l = np.log(T)
    return np.stack([np.ones_like(l), l, l * l, l * l * l, (l * l) * (l * l)])


def viscosity(T, X, W, visc):
    P = poly_basis(T)                      # [5, n]
    sv = visc @ P                          # [N, n]  sqrt(mu_i / sqrt(T))
    c1 = (1.0 + W[:, None] / W[None, :]) ** -0.5
    c2 = (W[None, :] / W[:, None]) ** 0.25
    tmp = 1.0 + (sv[:, None, :] / sv[None, :, :]) * c2[:, :, None]
    den = (X[None, :, :] / SQRT8 * c1[:, :, None] * tmp * tmp).sum(1)
    return (X * sv * sv / den).sum(0) * np.sqrt(T)


def conductivity(T, X, cond):
    lam = (cond @ poly_basis(T)) * np.sqrt(T)[None]
    return 0.5 * ((X * lam).sum(0) + 1.0 / (X / lam).sum(0))


def mix_diff_mass(T, p, rho, X, Wbar, W, diff):
    """rho * D_i^mix [N, n] (kg/m/s)"""
    N = W.shape[0]
    P = poly_basis(T)
    powT = T * np.sqrt(T)
    out = np.zeros_like(X)
    for i in range(N):
        s1 = np.zeros_like(T)
        s2 = np.zeros_like(T)
        for j in range(N):
            if i == j:
                continue
            D = (diff[i, j] @ P) * powT
            s1 += X[j] / D
            s2 += X[j] * W[j] / D
        s2 *= X[i] / (Wbar - X[i] * W[i])
        d = 1.0 / (s1 + s2) * (rho / p)
        out[i] = np.where(X[i] + 1e-10 > 1.0, 0.0, d)
    return out


def transport(T, p, rho, cp, Y, W, visc, cond, diff, unity_lewis=False):
    X, Wbar = mole_fractions(Y, W)
    mu = viscosity(T, X, W, visc)
    alpha = conductivity(T, X, cond) / cp
    if unity_lewis:
        rhoD = np.broadcast_to(alpha, Y.shape).copy()
    else:
        with np.errstate(divide="ignore", invalid="ignore"):
            rhoD = mix_diff_mass(T, p, rho, X, Wbar, W, diff)
    return mu, alpha, rhoD
