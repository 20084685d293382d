// LANE-PER-CELL chemistry sub-step: the small-mechanism form of dfChemistryModel::solveSingle
// (src/dfChemistryModel/dfChemistryModel.C:738-780).  One lane owns one cell for its whole
// integration; the lanes of a warp execute the same table entry on 32 different cells, so the
// reaction decode is warp-uniform and every per-cell array lives in shared memory as
// X[i * kLS + lane] (one bank per lane: conflict-free by construction).  GATHER FORM: the reaction
// loop only stores q_r and d q_r/d c_j; wdot_i, J_ij, the LU factors and the stage solves are loops
// over COMPILE-TIME indices (template parameter NAP = padded active species), so their accumulators
// and the matrix live in registers and shared memory sees loads and stores but no dependent
// read-modify-write chain.  Same ODE, same rate laws
// (SURVEY.md Appendix A), same RODAS4 scheme, controller and unstable-mode guard as the
// warp-per-cell integrator (integrator.cuh); only the mapping differs.  No lane ever talks to
// another lane, so the code is plain scalar C++ and compiles unchanged for the host emulation
// (tests/emu, kLS = 1).
#pragma once
#include "cell_driver.cuh"
#include "lpc_tables.hpp"

namespace dfchem {

#ifdef __CUDA_ARCH__
constexpr int kLS = 32;  // lane stride of the per-lane arrays (doubles)
#else
constexpr int kLS = 1;
#endif
constexpr int kLpcLanes = 32;

#define LPC_AT(off, i) L[((off) + (i)) * kLS]

struct LpcState {  // per-lane integrator state (registers)
    double t, h, dt, rho_ct, ysum, h_free;
    int cell, acc, rej, status;
    bool after_reject, need_base, first;
};

// compile-time offsets (doubles per lane) of the fixed region
template <int NAP>
struct LpcFixed {
    static constexpr int J = 0, LU = NAP * NAP, K = (lpc_lu_in_regs(NAP) ? 1 : 2) * NAP * NAP, F0 = K + 6 * NAP, DC = F0 + NAP;
    static_assert(DC + NAP == lpc_fixed_doubles(NAP), "fixed region");
};

// LpcMech whose tables point into a copy of the blob at `nb` (shared memory)
DFC_HD LpcMech lpc_rebase(const LpcMech& m, const uint8_t* nb) {
    LpcMech r = m;
#define LPC_REBASE(field, T) r.field = reinterpret_cast<const T*>(nb + (reinterpret_cast<const uint8_t*>(m.field) - m.base))
    LPC_REBASE(rx, LpcRx);
    LPC_REBASE(powtab, double);
    LPC_REBASE(eff, double);
    LPC_REBASE(fall, double);
    LPC_REBASE(arrh, double);
    LPC_REBASE(fidx, uint32_t);
    LPC_REBASE(fcoef, double);
    LPC_REBASE(jrow, int);
    LPC_REBASE(jidx, uint32_t);
    LPC_REBASE(jcoef, double);
    LPC_REBASE(gptr, int);
    LPC_REBASE(gidx, int);
    LPC_REBASE(gcoef, double);
    LPC_REBASE(W, double);
    LPC_REBASE(Tmid, double);
    LPC_REBASE(nasa, double);
    LPC_REBASE(hc, double);
    LPC_REBASE(perm, int);
#undef LPC_REBASE
    r.base = nb;
    return r;
}

// setState_TPY (dfChemistryModel.C:755): normalise Y, rho_ct = p Wbar/(R T), c_k = rho_ct Y_k / W_k
template <int NAP>
DFC_HD void lpc_load_state(const LpcMech& m, double* L, const double* Y, int ldY, int cell, double T, double p, LpcState& S) {
    double s = 0.0;
    for (int k = 0; k < m.N; ++k) {
        const double y = Y[(size_t)m.perm[k] * ldY + cell];
        LPC_AT(m.off_c, k) = y;
        s += y;
    }
    const double rys = 1.0 / s;
    double rw = 0.0;
    for (int k = 0; k < m.N; ++k) rw += LPC_AT(m.off_c, k) * rys / m.W[k];
    const double rho_ct = p / (rw * kGasConstant * T);
    for (int k = 0; k < m.N; ++k) {
        const double ck = rho_ct * (LPC_AT(m.off_c, k) * rys) / m.W[k];
        LPC_AT(m.off_c, k) = ck;
        LPC_AT(m.off_cs, k) = ck;
    }
    LPC_AT(m.off_c, m.N) = 1.0;   // the "absent factor" slot
    LPC_AT(m.off_cs, m.N) = 1.0;
    LPC_AT(m.off_q, m.q_zero) = 0.0;   // the padding slot of the wdot lists
    DFC_UNROLL
    for (int i = 0; i < NAP; ++i) LPC_AT(LpcFixed<NAP>::DC, i) = 0.0;
    S.ysum = s;
    S.rho_ct = rho_ct;
}

// per-cell constants of the frozen-T sub-step (same arithmetic as setup_rate_constants)
DFC_HD void lpc_setup(const LpcMech& m, double* L, double T) {
    const double logT = log(T), rT = 1.0 / T;
    for (int k = 0; k < m.N; ++k) {
        // (plain loads: the tables sit in shared memory here, the __ldg helpers of chem_core.cuh are global-only)
        const double* a = m.nasa + (size_t)k * 14 + (T <= m.Tmid[k] ? 0 : 7);
        const double h_RT = a[0] + T * (a[1] / 2 + T * (a[2] / 3 + T * (a[3] / 4 + T * a[4] / 5))) + a[5] / T;
        const double s_R = a[0] * logT + T * (a[1] + T * (a[2] / 2 + T * (a[3] / 3 + T * a[4] / 4))) + a[6];
        LPC_AT(m.off_q, k) = h_RT - s_R;
    }
    const double log_c0 = log(kOneAtm / (kGasConstant * T));
    for (int r = 0; r < m.R; ++r) {
        const LpcRx x = m.rx[r];
        const double* ar = m.arrh + (size_t)r * 3;
        const double lnk = ar[0] + ar[1] * logT - ar[2] * rT;
        double rk = 0.0;
        if (x.x & 4u) {
            double dG = 0.0, dn = 0.0;
            for (int e = m.gptr[r]; e < m.gptr[r + 1]; ++e) {
                dG = fma(m.gcoef[e], LPC_AT(m.off_q, m.gidx[e]), dG);
                dn += m.gcoef[e];
            }
            rk = fmin(exp(dG - dn * log_c0), kBig);
        }
        LPC_AT(m.off_kf, r) = exp(lnk);
        LPC_AT(m.off_rkc, r) = rk;
        if ((x.x & 3u) >= (uint32_t)kFalloffLindemann) {
            const int fid = (int)(x.z & 0xFFu);
            const double* f = m.fall + (size_t)fid * 8;
            LPC_AT(m.off_k0, fid) = fmin(exp(f[0] + f[1] * logT - f[2] * rT - lnk), 1e200);
            double lf = 0.0;
            if (f[7] != 0.0) {
                double fc = (1.0 - f[3]) * exp(-T / f[4]) + f[3] * exp(-T / f[5]);
                if (f[6] < 1e299) fc += exp(-f[6] / T);
                lf = log10(fmax(fc, kSmall));
            }
            LPC_AT(m.off_lfc, fid) = lf;
        }
    }
}

// Rate of progress of reaction r at the concentrations in column off_cx, all in registers: q, the forward and
// reverse parts, and (if JAC) the partial derivatives in slot order (present reactant factors, present product
// factors, d q/d[M]); `mask` bit a = derivative a is present.  Loads and arithmetic only; lpc_rate_store writes.
struct LpcRate {
    double q, qf, qr, d[7];
    uint32_t mask;
};
template <bool JAC>
DFC_HD void lpc_rate(const LpcMech& m, const double* L, int off_cx, int r, const LpcRx& x, LpcRate& o) {
    const uint32_t none = (uint32_t)m.N;
    const uint32_t kind = x.x & 3u;
    const uint32_t fo[3] = {(x.x >> 8) & 63u, (x.x >> 14) & 63u, (x.x >> 20) & 63u};
    const uint32_t po[3] = {x.y & 63u, (x.y >> 6) & 63u, (x.y >> 12) & 63u};
    double fv[3], fd[3], pv[3], pd[3];
    if (!(x.x & 8u)) {
        DFC_UNROLL
        for (int a = 0; a < 3; ++a) {
            fv[a] = LPC_AT(off_cx, fo[a]);
            pv[a] = LPC_AT(off_cx, po[a]);
            fd[a] = 1.0;
            pd[a] = 1.0;
        }
    } else {  // fractional / large orders: c^o with Cantera's rule (non-positive c -> 0)
        const double* pw = m.powtab + (size_t)((x.z >> 8) & 0xFFu) * 9;
        for (int a = 0; a < 3; ++a) {
            fv[a] = pv[a] = 1.0;
            fd[a] = pd[a] = 0.0;
            if (fo[a] != none) {
                const D2 t = pow_factor(pw[a], LPC_AT(off_cx, fo[a]));
                fv[a] = t.x;
                fd[a] = t.y;
            }
            if (po[a] != none) {
                const D2 t = pow_factor(pw[6 + a], LPC_AT(off_cx, po[a]));
                pv[a] = t.x;
                pd[a] = t.y;
            }
        }
    }
    double kfe = LPC_AT(m.off_kf, r);
    const double rkc = LPC_AT(m.off_rkc, r);
    double dkdM = 0.0;
    if (kind != (uint32_t)kElementary) {
        const double M = LPC_AT(m.off_M, x.y >> 20);  // third-body concentration of this reaction's efficiency set
        if (kind == (uint32_t)kThreeBody) {
            dkdM = kfe;
            kfe *= M;
        } else {
            const int fid = (int)(x.z & 0xFFu);
            const double rk = LPC_AT(m.off_k0, fid);  // k0 / kinf
            const double pr = M * rk;
            double F = 1.0, dF_dpr = 0.0;
            if (kind == (uint32_t)kFalloffTroe) {
                const double lf = LPC_AT(m.off_lfc, fid);
                const double lpr = log10(fmax(pr, kSmall));
                const double cc = -0.4 - 0.67 * lf;
                const double nn = 0.75 - 1.27 * lf;
                const double xx = lpr + cc;
                const double rden = 1.0 / (nn - 0.14 * xx);
                const double f1 = xx * rden;
                const double g = 1.0 / (1.0 + f1 * f1);
                F = exp10(lf * g);
                if (JAC && pr > kSmall) dF_dpr = F * (-lf * 2.0 * f1 * g * g * (nn * rden * rden)) / pr;
            }
            const double ip = 1.0 / (1.0 + pr);
            if (JAC) dkdM = kfe * (F * ip * ip + pr * ip * dF_dpr) * rk;
            kfe = kfe * (pr * ip) * F;
        }
    }
    const double Pf = fv[0] * fv[1] * fv[2];
    const double Pp = pv[0] * pv[1] * pv[2];
    const double kre = kfe * rkc;
    o.qf = kfe * Pf;
    o.qr = kre * Pp;
    o.q = kfe * Pf - kre * Pp;
    if (JAC) {
        o.mask = (fo[0] != none ? 1u : 0u) | (fo[1] != none ? 2u : 0u) | (fo[2] != none ? 4u : 0u) | (po[0] != none ? 8u : 0u) |
                 (po[1] != none ? 16u : 0u) | (po[2] != none ? 32u : 0u) | (kind != (uint32_t)kElementary ? 64u : 0u);
        o.d[0] = kfe * fd[0] * fv[1] * fv[2];
        o.d[1] = kfe * fv[0] * fd[1] * fv[2];
        o.d[2] = kfe * fv[0] * fv[1] * fd[2];
        o.d[3] = -kre * pd[0] * pv[1] * pv[2];
        o.d[4] = -kre * pv[0] * pd[1] * pv[2];
        o.d[5] = -kre * pv[0] * pv[1] * pd[2];
        o.d[6] = dkdM * (Pf - rkc * Pp);
    }
}
template <bool JAC>
DFC_HD void lpc_rate_store(const LpcMech& m, double* L, int r, const LpcRx& x, const LpcRate& o, double* qf_out, double* qr_out,
                           size_t ldq) {
    LPC_AT(m.off_q, r) = o.q;
    if (qf_out) {
        const size_t p = (size_t)(x.z >> 16) * ldq;
        qf_out[p] = o.qf;
        qr_out[p] = o.qr;
    }
    if (JAC) {
        int slot = m.off_dq + (int)x.w;
        DFC_UNROLL
        for (int a = 0; a < 7; ++a)
            if (o.mask & (1u << a)) LPC_AT(slot++, 0) = o.d[a];
    }
}

// Reaction loop: q_r -> column off_q (and, if JAC, the partial derivatives -> column off_dq).
template <bool JAC>
DFC_HD void lpc_reactions(const LpcMech& m, double* L, int off_cx, double* qf_out = nullptr, double* qr_out = nullptr,
                          size_t ldq = 0) {
    // third-body concentrations, once per efficiency set instead of once per reaction
    for (int sset = 0; sset < m.n_msets; ++sset) {
        const double* ef = m.eff + (size_t)sset * m.N;
        double M0 = 0.0, M1 = 0.0;
        int k = 0;
        for (; k + 1 < m.N; k += 2) {
            M0 = fma(ef[k], LPC_AT(off_cx, k), M0);
            M1 = fma(ef[k + 1], LPC_AT(off_cx, k + 1), M1);
        }
        if (k < m.N) M0 = fma(ef[k], LPC_AT(off_cx, k), M0);
        LPC_AT(m.off_M, sset) = M0 + M1;
    }
    int r = 0;
    DFC_NOUNROLL
    for (; r < m.R; ++r) {
        const LpcRx x = m.rx[r];
        LpcRate o;
        lpc_rate<JAC>(m, L, off_cx, r, x, o);
        lpc_rate_store<JAC>(m, L, r, x, o, qf_out, qr_out, ldq);
    }
}

// acc[t] = sum_k coef[k][t] * column[idx[k][t]]: one ELL row (lpc_tables.hpp) feeds all NAP targets, i.e. NAP
// independent load -> multiply-add chains per iteration, accumulators in registers
template <int NAP>
DFC_HD void lpc_gather(const uint32_t* idx, const double* coef, int rows, const double* L, int off_col, double (&acc)[NAP]) {
    DFC_UNROLL
    for (int t = 0; t < NAP; ++t) acc[t] = 0.0;
    for (int k = 0; k < rows; ++k, idx += NAP / 2, coef += NAP) {
        DFC_UNROLL
        for (int w = 0; w < NAP / 2; ++w) {
            const uint32_t u = idx[w];
            const D2 c = ld2(coef + 2 * w);
            acc[2 * w] = fma(c.x, LPC_AT(off_col, u & 0xFFFFu), acc[2 * w]);
            acc[2 * w + 1] = fma(c.y, LPC_AT(off_col, u >> 16), acc[2 * w + 1]);
        }
    }
}

// f[i] = wdot_i(cx) in registers; if JAC also J (dense NAP x NAP, row-major, fixed region) from the derivative column
template <int NAP, bool JAC>
DFC_HD void lpc_eval(const LpcMech& m, double* L, int off_cx, double (&f)[NAP]) {
    lpc_reactions<JAC>(m, L, off_cx);
    lpc_gather<NAP>(m.fidx, m.fcoef, m.f_rows, L, m.off_q, f);
    if (JAC) {
        LPC_AT(m.off_dq, m.n_dq) = 0.0;  // the padding slot of the lists (the column aliases the LU / stage region)
        DFC_NOUNROLL
        for (int i = 0; i < NAP; ++i) {
            const int k0 = m.jrow[i];
            double a[NAP];
            lpc_gather<NAP>(m.jidx + (size_t)k0 * (NAP / 2), m.jcoef + (size_t)k0 * NAP, m.jrow[i + 1] - k0, L, m.off_dq, a);
            DFC_UNROLL
            for (int j = 0; j < NAP; ++j) LPC_AT(LpcFixed<NAP>::J, i * NAP + j) = a[j];
        }
    }
}

// A = I/(h g) - J, LU in registers without pivoting (1/pivot stored on the diagonal).  The factors stay in the
// caller's registers for the six stage solves when the matrix is small (lpc_lu_in_regs), else they go to the
// fixed region.  Returns the number of negative pivots (unstable-mode guard, integrator.cuh).
template <int NAP>
constexpr bool lpc_lu_in_regs() { return lpc_lu_in_regs(NAP); }

template <int NAP>
DFC_HD int lpc_factor_matrix(double* L, double inv_hg, double (&A)[NAP][NAP]) {
    DFC_UNROLL
    for (int i = 0; i < NAP; ++i) {
        DFC_UNROLL
        for (int j = 0; j < NAP; ++j) A[i][j] = (i == j ? inv_hg : 0.0) - LPC_AT(LpcFixed<NAP>::J, i * NAP + j);
    }
    int neg = 0;
    DFC_UNROLL
    for (int k = 0; k < NAP; ++k) {
        const double p = 1.0 / A[k][k];
        A[k][k] = p;
        neg += p < 0.0 ? 1 : 0;
        DFC_UNROLL
        for (int i = k + 1; i < NAP; ++i) {
            const double l = A[i][k] * p;
            A[i][k] = l;
            DFC_UNROLL
            for (int j = k + 1; j < NAP; ++j) A[i][j] = fma(-l, A[k][j], A[i][j]);
        }
    }
    if (!lpc_lu_in_regs<NAP>()) {
        DFC_UNROLL
        for (int i = 0; i < NAP; ++i) {
            DFC_UNROLL
            for (int j = 0; j < NAP; ++j) LPC_AT(LpcFixed<NAP>::LU, i * NAP + j) = A[i][j];
        }
    }
    return neg;
}
// x <- A^-1 x, x in registers; factors from registers or from the fixed region (compile-time offsets)
template <int NAP>
DFC_HD void lpc_lu_solve(const double* L, const double (&A)[NAP][NAP], double (&x)[NAP]) {
#define LPC_LU(i, j) (lpc_lu_in_regs<NAP>() ? A[i][j] : LPC_AT(LpcFixed<NAP>::LU, (i) * NAP + (j)))
    DFC_UNROLL
    for (int i = 1; i < NAP; ++i) {
        DFC_UNROLL
        for (int j = 0; j < i; ++j) x[i] = fma(-LPC_LU(i, j), x[j], x[i]);
    }
    DFC_UNROLL
    for (int i = NAP - 1; i >= 0; --i) {
        DFC_UNROLL
        for (int j = i + 1; j < NAP; ++j) x[i] = fma(-LPC_LU(i, j), x[j], x[i]);
        x[i] *= LPC_LU(i, i);
    }
#undef LPC_LU
}

template <int NAP>
DFC_HD void lpc_begin_cell(const LpcMech& m, double* L, const SolveArgs& a, int cell, LpcState& S) {
    const double T = a.T[cell];
    S.cell = cell;
    S.dt = a.dt ? a.dt[cell] : a.dt_uniform;
    lpc_load_state<NAP>(m, L, a.Y, a.ldY, cell, T, a.p[cell], S);
    lpc_setup(m, L, T);
    S.t = 0.0;
    S.h = S.dt;
    S.h_free = S.dt;
    S.acc = S.rej = 0;
    S.status = kCellOk;
    S.after_reject = false;
    S.need_base = true;
    S.first = true;
    if (a.dtchem && a.dtchem[cell] > 0.0) {  // deltaTChem warm start (integrator.cuh)
        S.h = fmin(a.dtchem[cell], S.dt);
        S.first = false;
    }
}

// stage i of RODAS4: stage point, rate evaluation, right-hand side, solve.  i is a run-time value and
// the six stages are one loop body (the coefficient of a (stage, j) pair is looked up once and applied to all
// rows; same operation order per row as the unrolled form, so the results are identical).
// RODAS4 coefficients for a run-time (stage, j) pair: constant memory on the device
#define LPC_TAB_ROW(f, i) f(i, 0), f(i, 1), f(i, 2), f(i, 3), f(i, 4), f(i, 5)
#define LPC_TAB(f) {LPC_TAB_ROW(f, 0), LPC_TAB_ROW(f, 1), LPC_TAB_ROW(f, 2), LPC_TAB_ROW(f, 3), LPC_TAB_ROW(f, 4), LPC_TAB_ROW(f, 5)}
#if defined(__CUDACC__)
__constant__ double lpc_tab_a[36] = LPC_TAB(rodas_a);
__constant__ double lpc_tab_c[36] = LPC_TAB(rodas_c);
#endif
static const double lpc_tab_a_host[36] = LPC_TAB(rodas_a);
static const double lpc_tab_c_host[36] = LPC_TAB(rodas_c);
DFC_HD double lpc_a(int i, int j) {
#ifdef __CUDA_ARCH__
    return lpc_tab_a[i * 6 + j];
#else
    return lpc_tab_a_host[i * 6 + j];
#endif
}
DFC_HD double lpc_c(int i, int j) {
#ifdef __CUDA_ARCH__
    return lpc_tab_c[i * 6 + j];
#else
    return lpc_tab_c_host[i * 6 + j];
#endif
}
template <int NAP>
DFC_HD void lpc_stage(const LpcMech& m, double* L, const double (&A)[NAP][NAP], double ih, int i) {
    using F = LpcFixed<NAP>;
    double x[NAP];
    if (i == 0) {
        DFC_UNROLL
        for (int row = 0; row < NAP; ++row) x[row] = LPC_AT(F::F0, row);
    } else {
        double v[NAP];
        DFC_UNROLL
        for (int row = 0; row < NAP; ++row) v[row] = LPC_AT(m.off_c, row);
        for (int j = 0; j < i; ++j) {
            const double aj = lpc_a(i, j);
            DFC_UNROLL
            for (int row = 0; row < NAP; ++row) v[row] = fma(aj, LPC_AT(F::K, j * NAP + row), v[row]);
        }
        DFC_UNROLL
        for (int row = 0; row < NAP; ++row)
            if (row < m.NA) LPC_AT(m.off_cs, row) = v[row];
        lpc_eval<NAP, false>(m, L, m.off_cs, x);
        for (int j = 0; j < i; ++j) {
            const double cj = lpc_c(i, j) * ih;
            DFC_UNROLL
            for (int row = 0; row < NAP; ++row) x[row] = fma(cj, LPC_AT(F::K, j * NAP + row), x[row]);
        }
    }
    lpc_lu_solve<NAP>(L, A, x);
    DFC_UNROLL
    for (int row = 0; row < NAP; ++row) LPC_AT(F::K, i * NAP + row) = x[row];
}

// One step attempt of the cell on this lane (the body of integrate_cell's loops).  Returns true
// when the sub-step is complete (or has failed: S.status).
template <int NAP>
DFC_HD bool lpc_attempt(const LpcMech& m, double* L, LpcState& S, double rtol, double atol, int max_steps) {
    using F = LpcFixed<NAP>;
    const double atol_rho = atol * S.rho_ct;
    if (!(S.dt > 0.0)) return true;
    if (S.need_base) {
        double f0[NAP];
        lpc_eval<NAP, true>(m, L, m.off_c, f0);
        DFC_UNROLL
        for (int i = 0; i < NAP; ++i) LPC_AT(F::F0, i) = f0[i];
        S.need_base = false;
        if (S.first) {
            double e2 = 0.0;
            DFC_UNROLL
            for (int i = 0; i < NAP; ++i)
                if (i < m.NA) {
                    const double sc = atol_rho / m.W[i] + rtol * fabs(LPC_AT(m.off_c, i));
                    const double e = f0[i] / sc;
                    e2 = fma(e, e, e2);
                }
            const double d1 = sqrt(e2 / (double)m.N);
            if (d1 * S.dt > DFC_H0K) S.h = fmax(DFC_H0K / d1, 1e-3 * S.dt);
            S.first = false;
        }
    }
    if (S.acc + S.rej >= max_steps) {
        S.status = kCellMaxSteps;
        return true;
    }
    const double hs = fmin(S.h, S.dt - S.t);
    if (!(hs > 1e-14 * S.dt)) {
        S.status = kCellUnderflow;
        return true;
    }
    const bool cut = hs < S.h;
    const double ih = 1.0 / hs;
    double A[NAP][NAP];
    if (lpc_factor_matrix<NAP>(L, ih / kGamma, A) & 1) {  // unstable-mode guard
        ++S.rej;
        S.after_reject = true;
        S.h = hs * kFacUnstable;
        return false;
    }
    DFC_NOUNROLL
    for (int i = 0; i < kStages; ++i) lpc_stage<NAP>(m, L, A, ih, i);
    // new solution = argument of the last stage + k_6; error estimate = k_6
    double inc[NAP];
    double e2 = 0.0;
    DFC_UNROLL
    for (int row = 0; row < NAP; ++row) {
        const double k6 = LPC_AT(F::K, 5 * NAP + row);
        double v = k6;
        DFC_UNROLL
        for (int j = 0; j < 5; ++j) v = fma(rodas_a(5, j), LPC_AT(F::K, j * NAP + row), v);
        inc[row] = v;
        if (row < m.NA) {
            const double c0 = LPC_AT(m.off_c, row);
            const double sc = atol_rho / m.W[row] + rtol * fmax(fabs(c0), fabs(c0 + v));
            const double e = k6 / sc;
            e2 = fma(e, e, e2);
        }
    }
    double en = sqrt(e2 / (double)m.N);
    if (!(en < 1e30)) en = 1e30;
    double fac = kFacSafe / sqrt(sqrt(fmax(en, 1e-12)));
    const bool accepted = en <= 1.0;
    if (accepted) {
        fac = fmin(fmax(fac, kFacMin), S.after_reject ? 1.0 : kFacMax);
        S.after_reject = false;
        ++S.acc;
        S.t += hs;
        DFC_UNROLL
        for (int row = 0; row < NAP; ++row)
            if (row < m.NA) {
                LPC_AT(F::DC, row) += inc[row];
                LPC_AT(m.off_c, row) += inc[row];
            }
        S.need_base = true;
    } else {
        fac = fmin(fmax(fac, kFacMinReject), 1.0);
        ++S.rej;
        S.after_reject = true;
    }
    S.h_free = (accepted && cut) ? fmax(S.h, hs * fac) : hs * fac;
    S.h = hs * fac;
    return S.t >= S.dt * (1.0 - 1e-14);
}

// RR_i = (Ynew_i - Yraw_i)/dt * rho_OF against the RAW Y, Qdot = -sum hc_i RR_i (dfChemistryModel.C:768-772)
template <int NAP>
DFC_HD void lpc_finish_cell(const LpcMech& m, double* L, const SolveArgs& a, const LpcState& S) {
    const int cell = S.cell;
    const double scale = S.dt > 0.0 ? a.rho[cell] / S.dt : 0.0;
    const double rys = 1.0 / S.ysum;
    double qd = 0.0;
    for (int k = 0; k < m.N; ++k) {
        const int ko = m.perm[k];
        const double yraw = a.Y[(size_t)ko * a.ldY + cell];
        double dY = yraw * rys - yraw;
        if (k < m.NA) dY += LPC_AT(LpcFixed<NAP>::DC, k) * m.W[k] / S.rho_ct;
        const double rr = dY * scale;
        a.RR[(size_t)ko * a.ldRR + cell] = rr;
        qd = fma(-m.hc[k], rr, qd);
    }
    a.Qdot[cell] = qd;
    if (a.cost) a.cost[cell] = (double)(S.acc + S.rej);
    if (a.status) a.status[cell] = S.status;
    if (a.dtchem && S.dt > 0.0 && S.status == kCellOk) a.dtchem[cell] = S.h_free;
}

// instantiation list of the padded sizes (lpc_pad)
#define DFC_DISPATCH_LPC(NAP_VALUE, CALL) \
    switch (NAP_VALUE) {                  \
        case 4: CALL(4); break;           \
        case 6: CALL(6); break;           \
        case 8: CALL(8); break;           \
        case 12: CALL(12); break;         \
        default: break;                   \
    }

}  // namespace dfchem
