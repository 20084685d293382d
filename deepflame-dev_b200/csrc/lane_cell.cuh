// LANE-PER-CELL chemistry sub-step: the small-mechanism form of dfChemistryModel::solveSingle
// (src/dfChemistryModel/dfChemistryModel.C:738-780).  One lane owns one cell for its whole
// integration; the lanes of a warp execute the same table entry on 32 different cells, so the
// reaction decode is warp-uniform and every per-cell array lives in shared memory as
// X[i * kLS + lane] (one bank per lane: conflict-free by construction).  Same ODE, same rate laws
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

// RODAS4 coefficient tables as run-time indexable arrays (the stage loop is a real loop): constant memory on
// the device (warp-uniform index), a local table in the host emulation
#define LPC_RODAS_A                                                                                               \
    {{0, 0, 0, 0, 0},                                                                                             \
     {0.1544000000000000e+01, 0, 0, 0, 0},                                                                        \
     {0.9466785280815826, 0.2557011698983284, 0, 0, 0},                                                           \
     {0.3314825187068521e+01, 0.2896124015972201e+01, 0.9986419139977817, 0, 0},                                  \
     {0.1221224509226641e+01, 0.6019134481288629e+01, 0.1253708332932087e+02, -0.6878860361058950, 0},            \
     {0.1221224509226641e+01, 0.6019134481288629e+01, 0.1253708332932087e+02, -0.6878860361058950, 1.0}}
#define LPC_RODAS_C                                                                                               \
    {{0, 0, 0, 0, 0},                                                                                             \
     {-0.5668800000000000e+01, 0, 0, 0, 0},                                                                       \
     {-0.2430093356833875e+01, -0.2063599157091915, 0, 0, 0},                                                     \
     {-0.1073529058151375, -0.9594562251023355e+01, -0.2047028614809616e+02, 0, 0},                               \
     {0.7496443313967647e+01, -0.1024680431464352e+02, -0.3399990352819905e+02, 0.1170890893206160e+02, 0},       \
     {0.8083246795921522e+01, -0.7981132988064893e+01, -0.3152159432874371e+02, 0.1631930543123136e+02,           \
      -0.6058818238834054e+01}}
#if defined(__CUDACC__)
__constant__ double lpc_tab_a[6][5] = LPC_RODAS_A;
__constant__ double lpc_tab_c[6][5] = LPC_RODAS_C;
#endif
DFC_HD double lpc_a(int i, int j) {
#ifdef __CUDA_ARCH__
    return lpc_tab_a[i][j];
#else
    constexpr double t[6][5] = LPC_RODAS_A;
    return t[i][j];
#endif
}
DFC_HD double lpc_c(int i, int j) {
#ifdef __CUDA_ARCH__
    return lpc_tab_c[i][j];
#else
    constexpr double t[6][5] = LPC_RODAS_C;
    return t[i][j];
#endif
}

// LpcMech whose tables point into a copy of the blob at `nb` (shared memory)
DFC_HD LpcMech lpc_rebase(const LpcMech& m, const uint8_t* nb) {
    LpcMech r = m;
#define LPC_REBASE(field, T) r.field = reinterpret_cast<const T*>(nb + (reinterpret_cast<const uint8_t*>(m.field) - m.base))
    LPC_REBASE(reac, LpcReaction);
    LPC_REBASE(eff, double);
    LPC_REBASE(fall, double);
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
DFC_HD void lpc_load_state(const LpcMech& m, double* L, const double* Y, int ldY, int cell, double T, double p,
                           LpcState& S) {
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
    for (int i = 0; i < m.NA; ++i) LPC_AT(m.off_dc, i) = 0.0;
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
        LPC_AT(m.off_x, k) = h_RT - s_R;
    }
    const double log_c0 = log(kOneAtm / (kGasConstant * T));
    for (int r = 0; r < m.R; ++r) {
        const LpcReaction& x = m.reac[r];
        const double lnk = x.lnA + x.b * logT - x.Ta * rT;
        double rk = 0.0;
        if (x.rev) {
            double dG = 0.0, dn = 0.0;
            for (int n = 0; n < x.nnet; ++n) {
                dG = fma(x.nnu[n], LPC_AT(m.off_x, x.nsp[n]), dG);
                dn += x.nnu[n];
            }
            rk = fmin(exp(dG - dn * log_c0), kBig);
        }
        LPC_AT(m.off_kf, r) = exp(lnk);
        LPC_AT(m.off_rkc, r) = rk;
        if (x.fall >= 0) {
            const double* f = m.fall + (size_t)x.fall * 8;
            LPC_AT(m.off_k0, x.fall) = fmin(exp(f[0] + f[1] * logT - f[2] * rT - lnk), 1e200);
            double lf = 0.0;
            if (f[7] != 0.0) {
                double fc = (1.0 - f[3]) * exp(-T / f[4]) + f[3] * exp(-T / f[5]);
                if (f[6] < 1e299) fc += exp(-f[6] / T);
                lf = log10(fmax(fc, kSmall));
            }
            LPC_AT(m.off_lfc, x.fall) = lf;
        }
    }
}

DFC_HD D2 lpc_factor(bool simple, double o, double c) {
    if (simple) {
        if (o == 1.0) return D2{c, 1.0};
        if (o == 2.0) return D2{c * c, 2.0 * c};
        return D2{c * c * c, 3.0 * c * c};
    }
    return pow_factor(o, c);
}

// f = wdot(cx) [and J = d wdot / d c, dense NA x NA row-major]; q_out (optional): net rates of progress
template <bool JAC>
DFC_HD void lpc_eval(const LpcMech& m, double* L, int off_cx, double* q_out = nullptr, size_t ldq = 0) {
    const int NA = m.NA, N = m.N;
    for (int i = 0; i < NA; ++i) LPC_AT(m.off_f, i) = 0.0;
    if (JAC)
        for (int e = 0; e < NA * NA; ++e) LPC_AT(m.off_J, e) = 0.0;
    DFC_NOUNROLL
    for (int r = 0; r < m.R; ++r) {
        const LpcReaction& x = m.reac[r];
        double fv[3] = {1.0, 1.0, 1.0}, fd[3] = {0.0, 0.0, 0.0}, pv[3] = {1.0, 1.0, 1.0}, pd[3] = {0.0, 0.0, 0.0};
        const bool simple = x.simple != 0;
        DFC_UNROLL
        for (int a = 0; a < 3; ++a) {
            if (a < x.nr) {
                const D2 t = lpc_factor(simple, x.rord[a], LPC_AT(off_cx, x.rsp[a]));
                fv[a] = t.x;
                fd[a] = t.y;
            }
            if (a < x.np) {
                const D2 t = lpc_factor(simple, x.pord[a], LPC_AT(off_cx, x.psp[a]));
                pv[a] = t.x;
                pd[a] = t.y;
            }
        }
        double kfe = LPC_AT(m.off_kf, r);
        const double rkc = LPC_AT(m.off_rkc, r);
        double dkdM = 0.0;
        if (x.kind != kElementary) {
            const double* ef = m.eff + (size_t)x.mset * N;
            double M = 0.0;
            for (int k = 0; k < N; ++k) M = fma(ef[k], LPC_AT(off_cx, k), M);
            if (x.kind == kThreeBody) {
                dkdM = kfe;
                kfe *= M;
            } else {
                const double rk = LPC_AT(m.off_k0, x.fall);  // k0 / kinf
                const double pr = M * rk;
                double F = 1.0, dF_dpr = 0.0;
                if (x.kind == kFalloffTroe) {
                    const double lf = LPC_AT(m.off_lfc, x.fall);
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
        const double q = kfe * Pf - kre * Pp;
        if (q_out) q_out[(size_t)x.caller_index * ldq] = q;
        DFC_UNROLL
        for (int n = 0; n < kLpcMaxNet; ++n)
            if (n < x.nnet) LPC_AT(m.off_f, x.nsp[n]) = fma(x.nnu[n], q, LPC_AT(m.off_f, x.nsp[n]));
        if (JAC) {
            DFC_UNROLL
            for (int a = 0; a < 3; ++a) {
                if (a < x.nr && x.rsp[a] < NA) {
                    const double d = kfe * fd[a] * (a == 0 ? fv[1] * fv[2] : a == 1 ? fv[0] * fv[2] : fv[0] * fv[1]);
                    for (int n = 0; n < x.nnet; ++n) {
                        const int e = x.nsp[n] * NA + x.rsp[a];
                        LPC_AT(m.off_J, e) = fma(x.nnu[n], d, LPC_AT(m.off_J, e));
                    }
                }
                if (a < x.np && x.psp[a] < NA) {
                    const double d = -kre * pd[a] * (a == 0 ? pv[1] * pv[2] : a == 1 ? pv[0] * pv[2] : pv[0] * pv[1]);
                    for (int n = 0; n < x.nnet; ++n) {
                        const int e = x.nsp[n] * NA + x.psp[a];
                        LPC_AT(m.off_J, e) = fma(x.nnu[n], d, LPC_AT(m.off_J, e));
                    }
                }
            }
            if (x.kind != kElementary) {
                const double dM = dkdM * (Pf - rkc * Pp);
                const double* ef = m.eff + (size_t)x.mset * N;
                for (int j = 0; j < NA; ++j) {
                    const double w = dM * ef[j];
                    if (ef[j] != 0.0)
                        for (int n = 0; n < x.nnet; ++n) {
                            const int e = x.nsp[n] * NA + j;
                            LPC_AT(m.off_J, e) = fma(x.nnu[n], w, LPC_AT(m.off_J, e));
                        }
                }
            }
        }
    }
}

// A = I/(h g) - J, LU in place without pivoting (1/pivot stored on the diagonal).  Returns the number
// of negative pivots (unstable-mode guard, integrator.cuh).
DFC_HD int lpc_factor_matrix(const LpcMech& m, double* L, double inv_hg) {
    const int NA = m.NA;
    for (int i = 0; i < NA; ++i)
        for (int j = 0; j < NA; ++j) LPC_AT(m.off_A, i * NA + j) = (i == j ? inv_hg : 0.0) - LPC_AT(m.off_J, i * NA + j);
    int neg = 0;
    for (int k = 0; k < NA; ++k) {
        const double p = 1.0 / LPC_AT(m.off_A, k * NA + k);
        LPC_AT(m.off_A, k * NA + k) = p;
        neg += p < 0.0 ? 1 : 0;
        for (int i = k + 1; i < NA; ++i) {
            const double l = LPC_AT(m.off_A, i * NA + k) * p;
            LPC_AT(m.off_A, i * NA + k) = l;
            for (int j = k + 1; j < NA; ++j)
                LPC_AT(m.off_A, i * NA + j) = fma(-l, LPC_AT(m.off_A, k * NA + j), LPC_AT(m.off_A, i * NA + j));
        }
    }
    return neg;
}
// x <- A^-1 x (x = the off_x array)
DFC_HD void lpc_lu_solve(const LpcMech& m, double* L) {
    const int NA = m.NA;
    for (int i = 1; i < NA; ++i) {
        double s = LPC_AT(m.off_x, i);
        for (int j = 0; j < i; ++j) s = fma(-LPC_AT(m.off_A, i * NA + j), LPC_AT(m.off_x, j), s);
        LPC_AT(m.off_x, i) = s;
    }
    for (int i = NA - 1; i >= 0; --i) {
        double s = LPC_AT(m.off_x, i);
        for (int j = i + 1; j < NA; ++j) s = fma(-LPC_AT(m.off_A, i * NA + j), LPC_AT(m.off_x, j), s);
        LPC_AT(m.off_x, i) = s * LPC_AT(m.off_A, i * NA + i);
    }
}

DFC_HD void lpc_begin_cell(const LpcMech& m, double* L, const SolveArgs& a, int cell, LpcState& S) {
    const double T = a.T[cell];
    S.cell = cell;
    S.dt = a.dt ? a.dt[cell] : a.dt_uniform;
    lpc_load_state(m, L, a.Y, a.ldY, cell, T, a.p[cell], S);
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

// One step attempt of the cell on this lane (the body of integrate_cell's loops).  Returns true
// when the sub-step is complete (or has failed: S.status).
DFC_HD bool lpc_attempt(const LpcMech& m, double* L, LpcState& S, double rtol, double atol, int max_steps) {
    const int NA = m.NA;
    const double atol_rho = atol * S.rho_ct;
    if (!(S.dt > 0.0)) return true;
    if (S.need_base) {
        lpc_eval<true>(m, L, m.off_c);
        for (int i = 0; i < NA; ++i) LPC_AT(m.off_f0, i) = LPC_AT(m.off_f, i);
        S.need_base = false;
        if (S.first) {
            double e2 = 0.0;
            for (int i = 0; i < NA; ++i) {
                const double sc = atol_rho / m.W[i] + rtol * fabs(LPC_AT(m.off_c, i));
                const double e = LPC_AT(m.off_f0, i) / sc;
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
    if (lpc_factor_matrix(m, L, ih / kGamma) & 1) {  // unstable-mode guard
        ++S.rej;
        S.after_reject = true;
        S.h = hs * kFacUnstable;
        return false;
    }
    DFC_NOUNROLL
    for (int i = 0; i < kStages; ++i) {
        if (i > 0) {
            for (int row = 0; row < NA; ++row) {
                double v = LPC_AT(m.off_c, row);
                for (int j = 0; j < i; ++j) v = fma(lpc_a(i, j), LPC_AT(m.off_k, j * NA + row), v);
                LPC_AT(m.off_cs, row) = v;
            }
            lpc_eval<false>(m, L, m.off_cs);
        }
        for (int row = 0; row < NA; ++row) {
            double v = i == 0 ? LPC_AT(m.off_f0, row) : LPC_AT(m.off_f, row);
            for (int j = 0; j < i; ++j) v = fma(lpc_c(i, j) * ih, LPC_AT(m.off_k, j * NA + row), v);
            LPC_AT(m.off_x, row) = v;
        }
        lpc_lu_solve(m, L);
        for (int row = 0; row < NA; ++row) LPC_AT(m.off_k, i * NA + row) = LPC_AT(m.off_x, row);
    }
    // new solution = argument of the last stage + k_6; error estimate = k_6.  The increment is left in off_x.
    double e2 = 0.0;
    for (int row = 0; row < NA; ++row) {
        double inc = LPC_AT(m.off_k, 5 * NA + row);
        for (int j = 0; j < 5; ++j) inc = fma(lpc_a(5, j), LPC_AT(m.off_k, j * NA + row), inc);
        LPC_AT(m.off_x, row) = inc;
        const double c0 = LPC_AT(m.off_c, row);
        const double sc = atol_rho / m.W[row] + rtol * fmax(fabs(c0), fabs(c0 + inc));
        const double e = LPC_AT(m.off_k, 5 * NA + row) / sc;
        e2 = fma(e, e, e2);
    }
    double en = sqrt(e2 / (double)m.N);
    if (!(en < 1e30)) en = 1e30;
    double fac = kFacSafe / sqrt(sqrt(fmax(en, 1e-12)));
    if (en <= 1.0) {
        fac = fmin(fmax(fac, kFacMin), S.after_reject ? 1.0 : kFacMax);
        S.after_reject = false;
        ++S.acc;
        S.t += hs;
        for (int row = 0; row < NA; ++row) {
            const double inc = LPC_AT(m.off_x, row);
            LPC_AT(m.off_dc, row) += inc;
            LPC_AT(m.off_c, row) += inc;
        }
        S.need_base = true;
    } else {
        fac = fmin(fmax(fac, kFacMinReject), 1.0);
        ++S.rej;
        S.after_reject = true;
    }
    S.h_free = (S.need_base && cut) ? fmax(S.h, hs * fac) : hs * fac;
    S.h = hs * fac;
    return S.t >= S.dt * (1.0 - 1e-14);
}

// RR_i = (Ynew_i - Yraw_i)/dt * rho_OF against the RAW Y, Qdot = -sum hc_i RR_i (dfChemistryModel.C:768-772)
DFC_HD void lpc_finish_cell(const LpcMech& m, double* L, const SolveArgs& a, const LpcState& S) {
    const int cell = S.cell;
    const double scale = S.dt > 0.0 ? a.rho[cell] / S.dt : 0.0;
    const double rys = 1.0 / S.ysum;
    double qd = 0.0;
    for (int k = 0; k < m.N; ++k) {
        const int ko = m.perm[k];
        const double yraw = a.Y[(size_t)ko * a.ldY + cell];
        double dY = yraw * rys - yraw;
        if (k < m.NA) dY += LPC_AT(m.off_dc, k) * m.W[k] / S.rho_ct;
        const double rr = dY * scale;
        a.RR[(size_t)ko * a.ldRR + cell] = rr;
        qd = fma(-m.hc[k], rr, qd);
    }
    a.Qdot[cell] = qd;
    if (a.cost) a.cost[cell] = (double)(S.acc + S.rej);
    if (a.status) a.status[cell] = S.status;
    if (a.dtchem && S.dt > 0.0 && S.status == kCellOk) a.dtchem[cell] = S.h_free;
}

}  // namespace dfchem
