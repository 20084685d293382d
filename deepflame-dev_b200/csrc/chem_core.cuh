// Warp-per-cell chemistry core: rate constants, rates of progress, wdot, analytic Jacobian,
// register-resident Gauss-Jordan inverse, and the adaptive RODAS4 Rosenbrock integrator.
//
// This is the B200 replacement for what the reference does per cell in
// dfChemistryModel::solveSingle (src/dfChemistryModel/dfChemistryModel.C:738-780): Cantera
// Reactor (energy off) + ReactorNet/CVODE advance(deltaT) at frozen (T, rho).  The rate laws
// are Cantera 2.6's; the integrator is a different (one-step, L-stable)
// scheme solving the same ODE  d c_k/dt = wdot_k(c)  to the same rtol/atol semantics.
//
// Execution model: one team of kTeam = 32 lanes (a warp) owns one cell.  Per-cell data that all
// lanes read lives in the cell's shared-memory block (layout in DevMech); per-lane data (the
// lane's rows of the inverse, stage vectors) lives in registers (LaneRegs).  Lanes communicate
// ONLY through shared memory + team barriers, never shuffles, so the same source also compiles
// as a sequential host emulation (tests/emu) in which a "lane block" is a for-loop over lanes.
// Code between DFC_LANES_BEGIN/END runs once per lane (`tl` = lane in team); code outside runs
// team-uniformly (every lane computes the same value from shared memory).
#pragma once
#include <cmath>

#include "dev_tables.hpp"

#if defined(__CUDACC__)
#define DFC_HD __host__ __device__ __forceinline__
#else
#define DFC_HD inline
#endif

#ifdef __CUDA_ARCH__
#define DFC_LANES_BEGIN \
    {                   \
        const int tl = dfc_lane;
#define DFC_LANES_END \
    }                 \
    __syncwarp();
#define DFC_LANES_END_NOSYNC }
#define DFC_LR (*regs)
#define DFC_LDG(p) __ldg(p)
#define DFC_UNROLL _Pragma("unroll")
#define DFC_NOINLINE __noinline__
#else
#define DFC_LANES_BEGIN for (int tl = 0; tl < ::dfchem::kTeam; ++tl) {
#define DFC_LANES_END }
#define DFC_LANES_END_NOSYNC }
#define DFC_LR (regs[tl])
#define DFC_LDG(p) (*(p))
#define DFC_UNROLL
#define DFC_NOINLINE
#endif

namespace dfchem {

constexpr double kSmall = 1e-300;
constexpr double kBig = 1e300;
constexpr int kCellOk = 0, kCellMaxSteps = 1, kCellUnderflow = 2;  // = DFCHEM_CELL_* of dfchem.h

// RODAS4 (Hairer & Wanner, Solving ODEs II, IV.7): 6 stages, order 4(3), stiffly accurate,
// in the transformed form  (I/(h g) - J) k_i = f(y + sum_j a_ij k_j) + sum_j (c_ij/h) k_j.
constexpr int kStages = 6;
constexpr double kGamma = 0.25;
constexpr int kOrder = 4;
DFC_HD double rodas_a(int i, int j) {
    switch (i * 8 + j) {
        case 1 * 8 + 0: return 0.1544000000000000e+01;
        case 2 * 8 + 0: return 0.9466785280815826;
        case 2 * 8 + 1: return 0.2557011698983284;
        case 3 * 8 + 0: return 0.3314825187068521e+01;
        case 3 * 8 + 1: return 0.2896124015972201e+01;
        case 3 * 8 + 2: return 0.9986419139977817;
        case 4 * 8 + 0: case 5 * 8 + 0: return 0.1221224509226641e+01;
        case 4 * 8 + 1: case 5 * 8 + 1: return 0.6019134481288629e+01;
        case 4 * 8 + 2: case 5 * 8 + 2: return 0.1253708332932087e+02;
        case 4 * 8 + 3: case 5 * 8 + 3: return -0.6878860361058950;
        case 5 * 8 + 4: return 1.0;
        default: return 0.0;
    }
}
DFC_HD double rodas_c(int i, int j) {
    switch (i * 8 + j) {
        case 1 * 8 + 0: return -0.5668800000000000e+01;
        case 2 * 8 + 0: return -0.2430093356833875e+01;
        case 2 * 8 + 1: return -0.2063599157091915;
        case 3 * 8 + 0: return -0.1073529058151375;
        case 3 * 8 + 1: return -0.9594562251023355e+01;
        case 3 * 8 + 2: return -0.2047028614809616e+02;
        case 4 * 8 + 0: return 0.7496443313967647e+01;
        case 4 * 8 + 1: return -0.1024680431464352e+02;
        case 4 * 8 + 2: return -0.3399990352819905e+02;
        case 4 * 8 + 3: return 0.1170890893206160e+02;
        case 5 * 8 + 0: return 0.8083246795921522e+01;
        case 5 * 8 + 1: return -0.7981132988064893e+01;
        case 5 * 8 + 2: return -0.3152159432874371e+02;
        case 5 * 8 + 3: return 0.1631930543123136e+02;
        case 5 * 8 + 4: return -0.6058818238834054e+01;
        default: return 0.0;
    }
}

template <int NP, int NSLOT>
struct LaneRegs {
    double a[NSLOT][NP];      // this lane's rows of (I/(h g) - J)^-1, row-unscaled (see gj_invert)
    double invd[NSLOT];       // 1/pivot of each row: x_row = invd * sum_j a[row][j] b_j
    double k[kStages][NSLOT]; // stage increments
    double c0[NSLOT];         // concentrations at t = 0
    double dc[NSLOT];         // accumulated increment c(t) - c0 (kept separately: RR needs c - c0)
    double f0[NSLOT];         // wdot at the step's base point
    double inc[NSLOT];        // increment of the step being tested
};

struct D2 {
    double x, y;
};
DFC_HD D2 ld2(const double* p) {
#ifdef __CUDA_ARCH__
    double2 v = *reinterpret_cast<const double2*>(p);
    return D2{v.x, v.y};
#else
    return D2{p[0], p[1]};
#endif
}
DFC_HD void st2(double* p, double a, double b) {
#ifdef __CUDA_ARCH__
    *reinterpret_cast<double2*>(p) = make_double2(a, b);
#else
    p[0] = a;
    p[1] = b;
#endif
}

// NASA-7 (same arithmetic as src_gpu/dfThermo.cu:153-170,257-274)
DFC_HD const double* nasa_sel(const DevMech& m, int k, double T) {
    return m.nasa + (size_t)k * 14 + (T <= DFC_LDG(m.Tmid + k) ? 0 : 7);
}
DFC_HD double nasa_cp_R(const double* a, double T) {
    return DFC_LDG(a) + T * (DFC_LDG(a + 1) + T * (DFC_LDG(a + 2) + T * (DFC_LDG(a + 3) + T * DFC_LDG(a + 4))));
}
DFC_HD double nasa_h_RT(const double* a, double T) {
    return DFC_LDG(a) + T * (DFC_LDG(a + 1) / 2 + T * (DFC_LDG(a + 2) / 3 + T * (DFC_LDG(a + 3) / 4 + T * DFC_LDG(a + 4) / 5))) +
           DFC_LDG(a + 5) / T;
}
DFC_HD double nasa_s_R(const double* a, double T, double logT) {
    return DFC_LDG(a) * logT + T * (DFC_LDG(a + 1) + T * (DFC_LDG(a + 2) / 2 + T * (DFC_LDG(a + 3) / 3 + T * DFC_LDG(a + 4) / 4))) +
           DFC_LDG(a + 6);
}

// DevMech whose hot tables point into a copy of the hot prefix at `nb` (shared memory)
DFC_HD DevMech rebase_hot(const DevMech& m, const uint8_t* nb) {
    DevMech r = m;
    const uint8_t* ob = m.hot_base;
    r.hot_base = nb;
    r.rdesc = reinterpret_cast<const U4*>(nb + (reinterpret_cast<const uint8_t*>(m.rdesc) - ob));
    r.coef = reinterpret_cast<const double*>(nb + (reinterpret_cast<const uint8_t*>(m.coef) - ob));
    r.ms_dflt = reinterpret_cast<const double*>(nb + (reinterpret_cast<const uint8_t*>(m.ms_dflt) - ob));
    r.ms_val = reinterpret_cast<const double*>(nb + (reinterpret_cast<const uint8_t*>(m.ms_val) - ob));
    r.ms_ptr = reinterpret_cast<const int*>(nb + (reinterpret_cast<const uint8_t*>(m.ms_ptr) - ob));
    r.ms_idx = reinterpret_cast<const int*>(nb + (reinterpret_cast<const uint8_t*>(m.ms_idx) - ob));
    r.wdotA.ent = reinterpret_cast<const uint32_t*>(nb + (reinterpret_cast<const uint8_t*>(m.wdotA.ent) - ob));
    r.wdotB.ent = reinterpret_cast<const uint32_t*>(nb + (reinterpret_cast<const uint8_t*>(m.wdotB.ent) - ob));
    r.jac.ent = reinterpret_cast<const uint32_t*>(nb + (reinterpret_cast<const uint8_t*>(m.jac.ent) - ob));
    return r;
}

// team-uniform sum of the 32 partials in red[]
DFC_HD double team_sum(const double* red) {
    double s = 0.0;
    DFC_UNROLL
    for (int i = 0; i < kTeam; i += 2) {
        D2 v = ld2(red + i);
        s += v.x;
        s += v.y;
    }
#ifdef __CUDA_ARCH__
    __syncwarp();  // red[] may be rewritten by the next lane block
#endif
    return s;
}

// ---- third-body concentrations M_s for the efficiency sets -----------------------------------
DFC_HD void compute_msets(const DevMech& m, double* sm, const double* cx, int dfc_lane) {
    (void)dfc_lane;
    if (m.n_msets == 0) return;
    double* red = sm + m.off_red;
    DFC_LANES_BEGIN
    double s = 0.0;
    for (int k = tl; k < m.N; k += kTeam) s += cx[k];
    red[tl] = s;
    DFC_LANES_END
    const double ctot = team_sum(red);
    DFC_LANES_BEGIN
    for (int s = tl; s < m.n_msets; s += kTeam) {
        double M = m.ms_dflt[s] * ctot;  // hot tables may live in shared memory: plain loads
        for (int q = m.ms_ptr[s]; q < m.ms_ptr[s + 1]; ++q) M += m.ms_val[q] * cx[m.ms_idx[q]];
        sm[m.off_M + s] = M;
    }
    DFC_LANES_END
}

// ---- rates of progress q_r (and, if JAC, the partial derivatives) ----------------------------
// Lane-per-reaction.  Rate laws: SURVEY Appendix A.2-A.6 (Cantera GasKinetics semantics).
template <bool JAC>
DFC_HD void eval_reactions(const DevMech& m, double* sm, const double* cx, int dfc_lane) {
    (void)dfc_lane;
    const double* kf = sm + m.off_kf;
    const double* rkc = sm + m.off_rkc;
    double* q = sm + m.off_q;
    double* dq = sm + m.off_dq;
    DFC_LANES_BEGIN
    for (int r = tl; r < m.R; r += kTeam) {
#ifdef __CUDA_ARCH__
        const uint4 dd = *(reinterpret_cast<const uint4*>(m.rdesc) + r);
        const U4 d{dd.x, dd.y, dd.z, dd.w};
#else
        const U4 d = m.rdesc[r];
#endif
        const uint32_t flags = (d.y >> 16) & 0xFFu;
        const uint32_t kind = flags & 3u;
        int fi[3] = {(int)(d.x & 0xFF), (int)((d.x >> 8) & 0xFF), (int)((d.x >> 16) & 0xFF)};
        int pi[3] = {(int)(d.x >> 24), (int)(d.y & 0xFF), (int)((d.y >> 8) & 0xFF)};
        double fv[3], fd[3], pv[3], pd[3];
        if (!(flags & kFlagPow)) {
            DFC_UNROLL
            for (int a = 0; a < 3; ++a) {
                fv[a] = fi[a] != 255 ? cx[fi[a]] : 1.0;
                pv[a] = pi[a] != 255 ? cx[pi[a]] : 1.0;
                fd[a] = 1.0;
                pd[a] = 1.0;
            }
        } else {
            // fractional / large orders: c^o with Cantera's rule (non-positive c -> 0)
            const double* pw = m.powtab + (size_t)(d.z >> 16) * 9;
            for (int a = 0; a < 3; ++a) {
                fv[a] = pv[a] = 1.0;
                fd[a] = pd[a] = 0.0;
                if (fi[a] != 255) {
                    const double o = DFC_LDG(pw + a), c = cx[fi[a]];
                    if (o == 1.0) { fv[a] = c; fd[a] = 1.0; }
                    else if (o == 2.0) { fv[a] = c * c; fd[a] = 2.0 * c; }
                    else if (o == 3.0) { fv[a] = c * c * c; fd[a] = 3.0 * c * c; }
                    else if (c > 0.0) { fv[a] = pow(c, o); fd[a] = o * pow(c, o - 1.0); }
                    else { fv[a] = 0.0; fd[a] = 0.0; }
                }
                if (pi[a] != 255) {
                    const double o = DFC_LDG(pw + 6 + a), c = cx[pi[a]];
                    if (o == 1.0) { pv[a] = c; pd[a] = 1.0; }
                    else if (o == 2.0) { pv[a] = c * c; pd[a] = 2.0 * c; }
                    else if (o == 3.0) { pv[a] = c * c * c; pd[a] = 3.0 * c * c; }
                    else if (c > 0.0) { pv[a] = pow(c, o); pd[a] = o * pow(c, o - 1.0); }
                    else { pv[a] = 0.0; pd[a] = 0.0; }
                }
            }
        }
        double kfe = kf[r];
        double dkdM = 0.0;
        if (kind != (uint32_t)kElementary) {
            const double M = sm[m.off_M + (d.y >> 24)];
            if (kind == (uint32_t)kThreeBody) {
                dkdM = kfe;
                kfe *= M;
            } else {
                const int fid = (int)(d.z & 0xFFFF);
                const double k0 = sm[m.off_k0 + fid];
                const double rk = k0 / (kfe + kSmall);
                const double pr = M * rk;
                double F = 1.0, dF_dpr = 0.0;
                if (kind == (uint32_t)kFalloffTroe) {
                    const double lf = sm[m.off_lfc + fid];
                    const double lpr = log10(fmax(pr, kSmall));
                    const double cc = -0.4 - 0.67 * lf;
                    const double nn = 0.75 - 1.27 * lf;
                    const double x = lpr + cc;
                    const double den = nn - 0.14 * x;
                    const double f1 = x / den;
                    const double g = 1.0 / (1.0 + f1 * f1);
                    F = exp10(lf * g);
                    if (JAC && pr > kSmall) dF_dpr = F * (-lf * 2.0 * f1 * g * g * (nn / (den * den))) / pr;
                }
                const double ip = 1.0 / (1.0 + pr);
                if (JAC) dkdM = kfe * (F * ip * ip + pr * ip * dF_dpr) * rk;
                kfe = kfe * (pr * ip) * F;
            }
        }
        const double Pf = fv[0] * fv[1] * fv[2];
        const double Pp = pv[0] * pv[1] * pv[2];
        const double kre = kfe * rkc[r];
        q[r] = kfe * Pf - kre * Pp;
        if (JAC) {
            int slot = (int)d.w;
            if (fi[0] != 255) dq[slot++] = kfe * fd[0] * fv[1] * fv[2];
            if (fi[1] != 255) dq[slot++] = kfe * fv[0] * fd[1] * fv[2];
            if (fi[2] != 255) dq[slot++] = kfe * fv[0] * fv[1] * fd[2];
            if (pi[0] != 255) dq[slot++] = -kre * pd[0] * pv[1] * pv[2];
            if (pi[1] != 255) dq[slot++] = -kre * pv[0] * pd[1] * pv[2];
            if (pi[2] != 255) dq[slot++] = -kre * pv[0] * pv[1] * pd[2];
            if (kind != (uint32_t)kElementary) dq[slot] = dkdM * (Pf - rkc[r] * Pp);
        }
    }
    DFC_LANES_END
}

// ---- segmented gather (see GatherProg) --------------------------------------------------------
DFC_HD void run_gather(const GatherProg& g, const double* coef, const double* src, double* dst,
                       int dfc_lane) {
    (void)dfc_lane;
    DFC_LANES_BEGIN
    double acc = 0.0;
    const uint32_t* e = g.ent + tl * 4;
    for (int it = 0; it < g.n_iter; it += 4, e += 4 * kTeam) {
        // four entries per 16-byte load; all loads of the batch are issued before the FMA chain
#ifdef __CUDA_ARCH__
        const uint4 uu = *reinterpret_cast<const uint4*>(e);
        const uint32_t u[4] = {uu.x, uu.y, uu.z, uu.w};
#else
        const uint32_t u[4] = {e[0], e[1], e[2], e[3]};
#endif
        double cf[4], sv[4];
        DFC_UNROLL
        for (int b = 0; b < 4; ++b) {
            cf[b] = coef[(u[b] >> 12) & 0x7Fu];
            sv[b] = src[u[b] & 0xFFFu];
        }
        DFC_UNROLL
        for (int b = 0; b < 4; ++b) {
            acc = fma(cf[b], sv[b], acc);
            if (u[b] & (1u << 19)) {
                dst[u[b] >> 20] = acc;
                acc = 0.0;
            }
        }
    }
    DFC_LANES_END
}

// wdot_i (i < NA) at concentrations cx -> out[i].  Deliberately NOT inlined on the device: it is
// called from seven sites per Rosenbrock step and the kernel must stay instruction-cache resident.
#if defined(__CUDACC__)
__host__ __device__ DFC_NOINLINE
#else
inline
#endif
void eval_wdot(const DevMech& m, double* sm, const double* cx, double* out, int dfc_lane,
                      bool jac) {
    compute_msets(m, sm, cx, dfc_lane);
    if (jac)
        eval_reactions<true>(m, sm, cx, dfc_lane);
    else
        eval_reactions<false>(m, sm, cx, dfc_lane);
    run_gather(m.wdotA, m.coef, sm + m.off_q, sm + m.off_part, dfc_lane);
    run_gather(m.wdotB, m.coef, sm + m.off_part, out, dfc_lane);
}

// ---- per-cell constants of the frozen-T sub-step: kf, 1/Kc, falloff k0 and log10 Fcent --------
DFC_HD void setup_rate_constants(const DevMech& m, double* sm, double T, int dfc_lane) {
    (void)dfc_lane;
    const double logT = log(T), rT = 1.0 / T;
    double* g = sm + m.off_dq;  // scratch: g_k/RT at 1 atm
    DFC_LANES_BEGIN
    for (int k = tl; k < m.N; k += kTeam) {
        const double* a = nasa_sel(m, k, T);
        g[k] = nasa_h_RT(a, T) - nasa_s_R(a, T, logT);
    }
    if (tl == 0) sm[m.off_part] = 0.0;
    DFC_LANES_END
    const double log_c0 = log(kOneAtm / (kGasConstant * T));
    DFC_LANES_BEGIN
    for (int r = tl; r < m.R; r += kTeam) {
        const U4 d = m.rdesc[r];
        const uint32_t flags = (d.y >> 16) & 0xFFu;
        const double* ar = m.arrh + (size_t)r * 3;
        sm[m.off_kf + r] = exp(DFC_LDG(ar) + DFC_LDG(ar + 1) * logT - DFC_LDG(ar + 2) * rT);
        double rk = 0.0;
        if (flags & kFlagRev) {
            int fi[3] = {(int)(d.x & 0xFF), (int)((d.x >> 8) & 0xFF), (int)((d.x >> 16) & 0xFF)};
            int pi[3] = {(int)(d.x >> 24), (int)(d.y & 0xFF), (int)((d.y >> 8) & 0xFF)};
            double dG = 0.0, dn = 0.0;
            if (!(flags & kFlagPow)) {
                for (int a = 0; a < 3; ++a) {
                    if (fi[a] != 255) { dG -= g[fi[a]]; dn -= 1.0; }
                    if (pi[a] != 255) { dG += g[pi[a]]; dn += 1.0; }
                }
            } else {
                const double* pw = m.powtab + (size_t)(d.z >> 16) * 9;
                for (int a = 0; a < 3; ++a) {
                    if (fi[a] != 255) { dG -= pw[3 + a] * g[fi[a]]; dn -= pw[3 + a]; }
                    if (pi[a] != 255) { dG += pw[6 + a] * g[pi[a]]; dn += pw[6 + a]; }
                }
            }
            rk = fmin(exp(dG - dn * log_c0), kBig);
        }
        sm[m.off_rkc + r] = rk;
        if ((flags & 3u) >= (uint32_t)kFalloffLindemann) {
            const int fid = (int)(d.z & 0xFFFF);
            const double* f = m.fall + (size_t)fid * 8;
            sm[m.off_k0 + fid] = exp(f[0] + f[1] * logT - f[2] * rT);
            double lf = 0.0;
            if (f[7] != 0.0) {
                double fc = (1.0 - f[3]) * exp(-T / f[4]) + f[3] * exp(-T / f[5]);
                if (f[6] < 1e299) fc += exp(-f[6] / T);
                lf = log10(fmax(fc, kSmall));
            }
            sm[m.off_lfc + fid] = lf;
        }
    }
    DFC_LANES_END
}

}  // namespace dfchem
