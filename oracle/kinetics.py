"""TEST INFRASTRUCTURE (oracle) -- never imported by the product path.

NumPy restatement of the arithmetic the reference obtains from Cantera 2.6 on the hot path
`dfChemistryModel::solveSingle` (src/dfChemistryModel/dfChemistryModel.C:738-780) and
`calculateW` (:897-923): NASA-7 thermo, Arrhenius / three-body / falloff (Lindemann, Troe)
rate constants, equilibrium constants, rates of progress, net production rates, and the
isothermal-isochoric reactor right-hand side dY_k/dt = wdot_k W_k / rho (Reactor with
setEnergy(0), dfChemistryModel.C:748-764).  Cantera is an absent third-party dependency
(libcantera-devel=2.6, .github/workflows/CPU_inferencce_validation.yml:35): the formulas are
its published rate laws => "parity unpinned" against Cantera itself.

All functions are vectorised over cells: T [nc], C [nc, N].
"""
from __future__ import annotations

import numpy as np

from .mech import (ELEMENTARY, FALLOFF_LINDEMANN, FALLOFF_TROE, GAS_CONSTANT, ONE_ATM,
                   THREE_BODY, Mechanism)

SMALL = 1e-300
BIG = 1e300


def _sel(m: Mechanism, T):
    """[nc,N,7] NASA coefficients of the active range (low range for T <= Tmid)."""
    T = np.atleast_1d(np.asarray(T, float))
    use_lo = T[:, None] <= m.Tmid[None, :]
    return np.where(use_lo[:, :, None], m.nasa_lo[None], m.nasa_hi[None]), T


def cp_R(m, T):
    a, T = _sel(m, T)
    t = T[:, None]
    return a[..., 0] + a[..., 1] * t + a[..., 2] * t**2 + a[..., 3] * t**3 + a[..., 4] * t**4


def h_RT(m, T):
    a, T = _sel(m, T)
    t = T[:, None]
    return (a[..., 0] + a[..., 1] * t / 2 + a[..., 2] * t**2 / 3 + a[..., 3] * t**3 / 4
            + a[..., 4] * t**4 / 5 + a[..., 5] / t)


def s_R(m, T):
    a, T = _sel(m, T)
    t = T[:, None]
    return (a[..., 0] * np.log(t) + a[..., 1] * t + a[..., 2] * t**2 / 2 + a[..., 3] * t**3 / 3
            + a[..., 4] * t**4 / 4 + a[..., 6])


def mean_W(m, Y):
    return 1.0 / (Y / m.W[None]).sum(1)


def density(m, T, p, Y):
    """setState_TPY: rho = p Wbar / (R T) with Y normalised."""
    Yn = Y / Y.sum(1, keepdims=True)
    return p * mean_W(m, Yn) / (GAS_CONSTANT * T)


def h_mass(m, T, Y):
    T = np.atleast_1d(T)
    return (Y * h_RT(m, T) * (GAS_CONSTANT * T[:, None]) / m.W[None]).sum(1)


def cp_mass(m, T, Y):
    return (Y * cp_R(m, T) * GAS_CONSTANT / m.W[None]).sum(1)


class RateConstants:
    """Per-cell constants of one chemistry sub-step (T frozen): kf/k_inf, k0, 1/Kc, log10 Fcent."""

    def __init__(self, m: Mechanism, T):
        T = np.atleast_1d(np.asarray(T, float))
        self.T = T
        logT = np.log(T)[:, None]
        rT = 1.0 / T[:, None]
        self.kf = m.A[None] * np.exp(m.b[None] * logT - m.Ta[None] * rT)
        fall = m.kind >= FALLOFF_LINDEMANN
        with np.errstate(divide="ignore", invalid="ignore", over="ignore"):
            self.k0 = np.where(fall[None], m.A0[None] * np.exp(m.b0[None] * logT - m.Ta0[None] * rT), 0.0)
            # Troe centre (Cantera Troe::updateTemp)
            a, t3, t1, t2 = (m.troe[:, j][None] for j in range(4))
            t = T[:, None]
            is_troe = (m.kind == FALLOFF_TROE)[None]
            fc = (1 - a) * np.exp(-t / np.where(is_troe, t3, 1.0)) + a * np.exp(-t / np.where(is_troe, t1, 1.0))
            fc = fc + np.where(np.isfinite(t2), np.exp(-np.where(np.isfinite(t2), t2, 0.0) / t), 0.0)
            self.log10Fc = np.where(is_troe, np.log10(np.maximum(fc, SMALL)), 0.0)
        # equilibrium constants (Cantera GasKinetics::updateKc)
        g_RT = h_RT(m, T) - s_R(m, T)          # [nc,N] at p_ref = 1 atm
        dG = np.zeros((len(T), m.R))
        dn = np.zeros(m.R)
        for j in range(m.r_idx.shape[1]):
            ok = m.r_idx[:, j] >= 0
            dG[:, ok] -= m.r_nu[ok, j][None] * g_RT[:, m.r_idx[ok, j]]
            dn[ok] -= m.r_nu[ok, j]
            ok = m.p_idx[:, j] >= 0
            dG[:, ok] += m.p_nu[ok, j][None] * g_RT[:, m.p_idx[ok, j]]
            dn[ok] += m.p_nu[ok, j]
        log_c0 = np.log(ONE_ATM / (GAS_CONSTANT * T))[:, None]
        with np.errstate(over="ignore"):
            rkc = np.minimum(np.exp(dG - dn[None] * log_c0), BIG)
        self.rKc = np.where(m.reversible[None] == 1, rkc, 0.0)


def _conc_product(C, idx, order):
    """prod_j C[idx_j]^order_j with Cantera's rules (integer orders by multiplication,
    sign-preserving; non-integer order: C <= 0 -> 0)."""
    out = np.ones((C.shape[0], idx.shape[0]))
    for j in range(idx.shape[1]):
        ok = (idx[:, j] >= 0)[None]
        c = C[:, np.where(idx[:, j] >= 0, idx[:, j], 0)]
        o = order[:, j][None]
        frac = ok & (o != 0.0) & (o != 1.0) & (o != 2.0) & (o != 3.0)
        f = np.where(o == 1.0, c, np.where(o == 2.0, c * c, np.where(o == 3.0, c * c * c, 1.0)))
        if frac.any():
            with np.errstate(invalid="ignore", divide="ignore"):
                pw = np.where(c > 0, np.power(np.abs(c), np.where(frac, o, 1.0)), 0.0)
            f = np.where(frac, pw, f)
        out *= np.where(ok, f, 1.0)
    return out


def rates_of_progress(m: Mechanism, rc: RateConstants, C):
    """Forward / reverse rates of progress [nc,R] (kmol/m^3/s)."""
    M = C @ m.eff.T                                   # [nc,R]; rows of ones for M-less reactions
    kf = rc.kf.copy()
    fall = m.kind >= FALLOFF_LINDEMANN
    if fall.any():
        with np.errstate(divide="ignore", invalid="ignore"):
            pr = M * rc.k0 / (rc.kf + SMALL)
            lpr = np.log10(np.maximum(pr, SMALL))
            lf = rc.log10Fc
            cc = -0.4 - 0.67 * lf
            nn = 0.75 - 1.27 * lf
            f1 = (lpr + cc) / (nn - 0.14 * (lpr + cc))
            F = np.power(10.0, lf / (1.0 + f1 * f1))
            F = np.where((m.kind == FALLOFF_TROE)[None], F, 1.0)
            kfall = rc.kf * (pr / (1.0 + pr)) * F
        kf = np.where(fall[None], kfall, kf)
    tb = (m.kind == THREE_BODY)[None]
    kf = np.where(tb, kf * M, kf)
    qf = kf * _conc_product(C, m.r_idx, m.r_ord)
    qr = kf * rc.rKc * _conc_product(C, m.p_idx, m.p_nu)
    return qf, qr


def wdot(m: Mechanism, rc: RateConstants, C):
    """Net molar production rates [nc,N] (kmol/m^3/s)  (GasKinetics::getNetProductionRates)."""
    qf, qr = rates_of_progress(m, rc, C)
    q = qf - qr
    w = np.zeros_like(C)
    for j in range(m.r_idx.shape[1]):
        ok = m.r_idx[:, j] >= 0
        np.add.at(w, (slice(None), m.r_idx[ok, j]), -q[:, ok] * m.r_nu[ok, j][None])
        ok = m.p_idx[:, j] >= 0
        np.add.at(w, (slice(None), m.p_idx[ok, j]), q[:, ok] * m.p_nu[ok, j][None])
    return w


def stoich_matrix(m: Mechanism):
    """S [N,R]: wdot = S @ (qf - qr)."""
    S = np.zeros((m.N, m.R))
    for j in range(m.r_idx.shape[1]):
        for r in range(m.R):
            if m.r_idx[r, j] >= 0:
                S[m.r_idx[r, j], r] -= m.r_nu[r, j]
            if m.p_idx[r, j] >= 0:
                S[m.p_idx[r, j], r] += m.p_nu[r, j]
    return S


def jacobian_C(m: Mechanism, rc: RateConstants, C):
    """Analytic d(wdot)/dC [nc,N,N] in concentration space, including the third-body and
    falloff ([M], Pr, Troe F) derivatives.  Checked against finite differences in
    tests/test_oracle.py.  Integer reaction orders only get the polynomial derivative;
    fractional orders use o*C^(o-1) for C > 0 and 0 otherwise."""
    nc, N = C.shape
    R = m.R
    M = C @ m.eff.T
    fall = (m.kind >= FALLOFF_LINDEMANN)[None]
    tb = (m.kind == THREE_BODY)[None]
    keff = rc.kf.copy()
    dk_dM = np.zeros((nc, R))
    with np.errstate(divide="ignore", invalid="ignore"):
        pr = M * rc.k0 / (rc.kf + SMALL)
        lpr = np.log10(np.maximum(pr, SMALL))
        lf = rc.log10Fc
        cc = -0.4 - 0.67 * lf
        nn = 0.75 - 1.27 * lf
        x = lpr + cc
        den = nn - 0.14 * x
        f1 = x / den
        lgF = lf / (1.0 + f1 * f1)
        is_troe = (m.kind == FALLOFF_TROE)[None]
        F = np.where(is_troe, np.power(10.0, lgF), 1.0)
        # dF/dPr = F ln10 dlgF/dlpr / (Pr ln10)
        dlgF = -lf * 2.0 * f1 / (1.0 + f1 * f1) ** 2 * (nn / den**2)
        dF_dpr = np.where(is_troe & (pr > SMALL), F * dlgF / pr, 0.0)
        kfall = rc.kf * (pr / (1.0 + pr)) * F
        dkfall_dM = rc.kf * (F / (1.0 + pr) ** 2 + pr / (1.0 + pr) * dF_dpr) * (rc.k0 / (rc.kf + SMALL))
    keff = np.where(fall, kfall, keff)
    dk_dM = np.where(fall, dkfall_dM, dk_dM)
    dk_dM = np.where(tb, rc.kf, dk_dM)
    keff = np.where(tb, rc.kf * M, keff)

    def dprod(idx, order):
        """returns P [nc,R] and list over slots j of dP/dC_{idx_j} [nc,R]"""
        fac = []
        dfac = []
        for j in range(idx.shape[1]):
            ok = (idx[:, j] >= 0)[None]
            c = C[:, np.where(idx[:, j] >= 0, idx[:, j], 0)]
            o = order[:, j][None]
            frac = ok & (o != 0.0) & (o != 1.0) & (o != 2.0) & (o != 3.0)
            f = np.where(o == 1.0, c, np.where(o == 2.0, c * c, np.where(o == 3.0, c * c * c, 1.0)))
            d = np.where(o == 1.0, 1.0, np.where(o == 2.0, 2 * c, np.where(o == 3.0, 3 * c * c, 0.0)))
            if frac.any():
                with np.errstate(invalid="ignore", divide="ignore"):
                    oo = np.where(frac, o, 1.0)
                    pw = np.where(c > 0, np.power(np.abs(c), oo), 0.0)
                    dpw = np.where(c > 0, oo * np.power(np.abs(c), oo - 1.0), 0.0)
                f = np.where(frac, pw, f)
                d = np.where(frac, dpw, d)
            fac.append(np.where(ok, f, 1.0))
            dfac.append(np.where(ok, d, 0.0))
        P = fac[0] * fac[1] * fac[2]
        dP = [dfac[0] * fac[1] * fac[2], fac[0] * dfac[1] * fac[2], fac[0] * fac[1] * dfac[2]]
        return P, dP

    Pf, dPf = dprod(m.r_idx, m.r_ord)
    Pr_, dPr = dprod(m.p_idx, m.p_nu)
    # dq/dC [nc,R,N]
    dq = np.zeros((nc, R, N))
    ar = np.arange(R)
    for j in range(m.r_idx.shape[1]):
        ok = m.r_idx[:, j] >= 0
        np.add.at(dq, (slice(None), ar[ok], m.r_idx[ok, j]), (keff * dPf[j])[:, ok])
        ok = m.p_idx[:, j] >= 0
        np.add.at(dq, (slice(None), ar[ok], m.p_idx[ok, j]), -(keff * rc.rKc * dPr[j])[:, ok])
    has_m = (m.kind != ELEMENTARY)
    dq[:, has_m, :] += (dk_dM * (Pf - rc.rKc * Pr_))[:, has_m, None] * m.eff[None, has_m, :]
    S = stoich_matrix(m)
    return np.einsum("kr,crj->ckj", S, dq)
