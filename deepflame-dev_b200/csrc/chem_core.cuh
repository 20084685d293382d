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
// through shared memory + team barriers; the only shuffles are the team reductions (team_sum), which
// have a host form, so the same source also compiles
// as a sequential host emulation (tests/emu) in which a "lane block" is a for-loop over lanes.
// Code between DFC_LANES_BEGIN/END runs once per lane (`tl` = lane in team); code outside runs
// team-uniformly (every lane computes the same value from shared memory).
#pragma once
#include <cmath>
#include <cstring>

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
#define DFC_NOUNROLL _Pragma("unroll 1")
#define DFC_NOINLINE __noinline__
#else
#define DFC_LANES_BEGIN for (int tl = 0; tl < ::dfchem::kTeam; ++tl) {
#define DFC_LANES_END }
#define DFC_LANES_END_NOSYNC }
#define DFC_LR (regs[tl])
#define DFC_LDG(p) (*(p))
#define DFC_UNROLL
#define DFC_NOUNROLL
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
constexpr DFC_HD double rodas_a(int i, int j) {
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
constexpr DFC_HD double rodas_c(int i, int j) {
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

// Matrices with 33..36 rows use the bordered layout (integrator.cuh): 32 core rows, one per lane,
// and NB = NP - 32 border rows handled through the Schur complement.
constexpr int border_rows(int NP) { return (NP > 32 && NP <= 36) ? NP - 32 : 0; }
// stage-A rows of the wdot schedule a lane can keep in registers (WdotProg), by kernel size class
constexpr int wdot_rows_cap(int NP) { return NP <= 16 ? 4 : NP <= 24 ? 6 : NP <= 36 ? 10 : kWdotRowsMax; }

template <int NP, int NSLOT>
struct LaneRegs {
    static constexpr int NB = border_rows(NP);
    static constexpr int NBX = NB ? NB : 1;
    static constexpr int NCOL = NB ? 32 : NP;
    static constexpr int NROW = NB ? 1 : NSLOT;
    double a[NROW][NCOL];     // this lane's rows of (I/(h g) - J)^-1, row-unscaled (see gj_invert)
    double invd[NROW];        // 1/pivot of each row: x_row = invd * sum_j a[row][j] b_j
    double g[NBX];            // bordered: this row's entries in the border columns -> A11^-1 A12
    double v[NBX];            // bordered: column `lane` of the border rows -> -A21 A11^-1
    double S[NBX * NBX];      // bordered: Schur complement, then its inverse (same on every lane)
    double redv[NBX];         // bordered: this lane's terms of the border dot products
    double tmp;
    double k[kStages][NSLOT]; // stage increments
    double dc[NSLOT];         // accumulated increment c(t) - c(0) (kept separately: RR needs c - c0)
    double f0[NSLOT];         // wdot at the step's base point
    double atolc[NSLOT];      // atol * rho / W_row: the absolute part of the error weights
    double red;               // this lane's term of a team reduction (team_sum)
    // this lane's rows of the wdot schedule: byte offsets of s0..s3 and dst (load_wdot_regs)
    static constexpr int RW = wdot_rows_cap(NP);
    uint32_t wa[RW][5];       // [4]: dst offset | bit 31 = the row starts a new partial sum
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
#define DFC_REBASE(field, T) r.field = reinterpret_cast<const T*>(nb + (reinterpret_cast<const uint8_t*>(m.field) - ob))
    r.hot_base = nb;
    DFC_REBASE(rdesc, U4);
    DFC_REBASE(coef, double);
    DFC_REBASE(ms_ent, U4);
    DFC_REBASE(ms_rng, uint32_t);
    DFC_REBASE(jac.ent, U4);
    DFC_REBASE(jac.dst, uint32_t);
#undef DFC_REBASE
    return r;
}

// Host emulation only (tests/emu, -DDFC_EMU_BOUNDS): every table-driven address is checked
// against the cell block the emulator registered; on the device the macro is empty.
#if !defined(__CUDA_ARCH__) && defined(DFC_EMU_BOUNDS)
struct EmuWindow {
    const char* lo;
    const char* hi;
    void (*fail)(const void*);
};
extern thread_local EmuWindow dfc_emu_window;
inline void dfc_check_addr(const void* p) {
    const char* q = static_cast<const char*>(p);
    if (dfc_emu_window.lo && (q < dfc_emu_window.lo || q + sizeof(double) > dfc_emu_window.hi))
        dfc_emu_window.fail(p);
}
#define DFC_CHECK_ADDR(p) dfc_check_addr(p)
#else
#define DFC_CHECK_ADDR(p) ((void)0)
#endif

// byte-offset addressing of the per-cell arrays (offsets come pre-scaled from the tables)
DFC_HD double ldb(const double* base, uint32_t byte_off) {
    DFC_CHECK_ADDR(reinterpret_cast<const char*>(base) + byte_off);
    return *reinterpret_cast<const double*>(reinterpret_cast<const char*>(base) + byte_off);
}
DFC_HD double* ptrb(double* base, uint32_t byte_off) {
    DFC_CHECK_ADDR(reinterpret_cast<char*>(base) + byte_off);
    return reinterpret_cast<double*>(reinterpret_cast<char*>(base) + byte_off);
}
DFC_HD U4 ld_u4(const U4* p) {
#ifdef __CUDA_ARCH__
    const uint4 v = *reinterpret_cast<const uint4*>(p);
    return U4{v.x, v.y, v.z, v.w};
#else
    return *p;
#endif
}
DFC_HD double u2d(uint32_t lo, uint32_t hi) {
#ifdef __CUDA_ARCH__
    return __hiloint2double((int)hi, (int)lo);
#else
    const uint64_t b = ((uint64_t)hi << 32) | lo;
    double v;
    memcpy(&v, &b, 8);
    return v;
#endif
}
// +/- v selected by bit 31 of the entry word (one LOP3 on the high half of the double)
DFC_HD double signed_by(double v, uint32_t e) {
#ifdef __CUDA_ARCH__
    return __hiloint2double(__double2hiint(v) ^ (int)(e & 0x80000000u), __double2loint(v));
#else
    return (e & 0x80000000u) ? -v : v;
#endif
}

// Sum over the team of one value per lane (the lane's value is DFC_LR.red).  Device: xor
// butterfly through warp shuffles, every lane ends with the same bits.  Host emulation: the same
// butterfly evaluated for lane 0.
template <class Regs>
DFC_HD double team_sum(Regs* regs) {
#ifdef __CUDA_ARCH__
    double v = regs->red;
    DFC_UNROLL
    for (int d = 16; d >= 1; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    return v;
#else
    double t[kTeam];
    for (int i = 0; i < kTeam; ++i) t[i] = regs[i].red;
    for (int d = 16; d >= 1; d >>= 1)
        for (int i = 0; i < d; ++i) t[i] += t[i + d];
    return t[0];
#endif
}

// the same for NB values per lane (DFC_LR.redv[b]) -> z[b]
template <int NB, class Regs>
DFC_HD void team_sum_n(Regs* regs, double* z) {
#ifdef __CUDA_ARCH__
    DFC_UNROLL
    for (int b = 0; b < NB; ++b) z[b] = regs->redv[b];
    DFC_UNROLL
    for (int d = 16; d >= 1; d >>= 1) {
        DFC_UNROLL
        for (int b = 0; b < NB; ++b) z[b] += __shfl_xor_sync(0xffffffffu, z[b], d);
    }
#else
    for (int b = 0; b < NB; ++b) {
        double t[kTeam];
        for (int i = 0; i < kTeam; ++i) t[i] = regs[i].redv[b];
        for (int d = 16; d >= 1; d >>= 1)
            for (int i = 0; i < d; ++i) t[i] += t[i + d];
        z[b] = t[0];
    }
#endif
}

// ---- register-resident wdot schedule (see WdotProg) ------------------------------------------
// Offsets are bytes from `DFC_SMEM_BASE`: the CTA's dynamic shared memory on the device (so the
// compiler emits LDS/STS with the offset register as the whole address), the cell block in the
// host emulation.
#ifdef __CUDA_ARCH__
extern __shared__ double2 dfc_smem[];
#define DFC_SMEM_BASE(sm) reinterpret_cast<char*>(dfc_smem)
#else
#define DFC_SMEM_BASE(sm) reinterpret_cast<char*>(sm)
#endif
DFC_HD double ldo(const char* base, uint32_t off) {
    DFC_CHECK_ADDR(base + off);
    return *reinterpret_cast<const double*>(base + off);
}
DFC_HD void sto(char* base, uint32_t off, double v) {
    DFC_CHECK_ADDR(base + off);
    *reinterpret_cast<double*>(base + off) = v;
}

// cell_off: byte offset of this team's cell block from DFC_SMEM_BASE (0 in the emulation)
template <class Regs>
DFC_HD bool load_wdot_regs(const DevMech& m, Regs* regs, int dfc_lane, uint32_t cell_off) {
    (void)dfc_lane;
    if (m.wdot.nA > Regs::RW) return false;
    DFC_LANES_BEGIN
    DFC_UNROLL
    for (int i = 0; i < Regs::RW; ++i) {
        DFC_UNROLL
        for (int k = 0; k < 5; ++k) DFC_LR.wa[i][k] = 0;
        if (i < m.wdot.nA) {
            const uint16_t* t = m.wdot.tab + ((size_t)i * kTeam + tl) * 6;
            DFC_UNROLL
            for (int k = 0; k < 5; ++k) DFC_LR.wa[i][k] = cell_off + 8u * DFC_LDG(t + k);
            DFC_LR.wa[i][4] |= (uint32_t)(DFC_LDG(t + 5) & 1u) << 31;
        }
    }
    DFC_LANES_END
    return true;
}

// w = S q: q[R+1] -> part[4 NA] -> vec[0..NA)
template <class Regs>
DFC_HD void run_wdot(const DevMech& m, double* sm, Regs* regs, int dfc_lane) {
    (void)dfc_lane;
    char* base = DFC_SMEM_BASE(sm);
    DFC_LANES_BEGIN
    double acc = 0.0;
    DFC_UNROLL
    for (int i = 0; i < Regs::RW; ++i) {
        if (i < m.wdot.nA) {
            const double s = (ldo(base, DFC_LR.wa[i][0]) + ldo(base, DFC_LR.wa[i][1])) -
                             (ldo(base, DFC_LR.wa[i][2]) + ldo(base, DFC_LR.wa[i][3]));
            const uint32_t d = DFC_LR.wa[i][4];
            acc = ((int32_t)d < 0) ? s : acc + s;
            sto(base, d & 0x7FFFFFFFu, acc);
        }
    }
    DFC_LANES_END
    const double* part = sm + m.off_part;
    double* vec = sm + m.off_vec;
    DFC_LANES_BEGIN
    for (int i = tl; i < m.NA; i += kTeam) {
        const D2 p0 = ld2(part + 4 * i), p1 = ld2(part + 4 * i + 2);
        vec[i] = (p0.x + p0.y) + (p1.x + p1.y);
    }
    DFC_LANES_END
}

// ---- row programs (see RowProg) ---------------------------------------------------------------
DFC_HD void run_rows(const RowProg& g, const double* coef, const double* src, double* dst, int dfc_lane) {
    (void)dfc_lane;
    DFC_LANES_BEGIN
    const U4* e = g.ent + tl;
    const uint32_t* dp = g.dst + tl;
    DFC_NOUNROLL
    for (int i = 0; i < g.nA; ++i, e += kTeam, dp += kTeam) {
        const U4 u = ld_u4(e);
        const uint32_t d = *dp;
        const double v0 = signed_by(ldb(src, u.x & 0xFFFFu), u.x);
        const double v1 = signed_by(ldb(src, u.y & 0xFFFFu), u.y);
        const double v2 = signed_by(ldb(src, u.z & 0xFFFFu), u.z);
        const double v3 = signed_by(ldb(src, u.w & 0xFFFFu), u.w);
        double sum = (v0 + v1) + (v2 + v3);
        double* out = ptrb(dst, d & 0xFFFFFFu);
        if (d & kRowAcc) sum += *out;
        *out = sum;
    }
    DFC_NOUNROLL
    for (int i = 0; i < g.nB; ++i, e += kTeam, dp += kTeam) {
        const U4 u = ld_u4(e);
        const uint32_t d = *dp;
        double s0 = coef[(u.x >> 16) & 0x7Fu] * ldb(src, u.x & 0xFFFFu);
        double s1 = coef[(u.y >> 16) & 0x7Fu] * ldb(src, u.y & 0xFFFFu);
        s0 = fma(coef[(u.z >> 16) & 0x7Fu], ldb(src, u.z & 0xFFFFu), s0);
        s1 = fma(coef[(u.w >> 16) & 0x7Fu], ldb(src, u.w & 0xFFFFu), s1);
        double sum = s0 + s1;
        double* out = ptrb(dst, d & 0xFFFFFFu);
        if (d & kRowAcc) sum += *out;
        *out = sum;
    }
    DFC_LANES_END
}

// c^o and its derivative with Cantera's rule for non-integer orders (non-positive c -> 0).
// Rare (global one-step mechanisms): kept out of line so the hot loop stays small.
#if defined(__CUDACC__)
__host__ __device__ DFC_NOINLINE
#else
inline
#endif
D2 pow_factor(double o, double c) {
    if (o == 1.0) return D2{c, 1.0};
    if (o == 0.0) return D2{1.0, 0.0};  // Cantera's C_AnyN: order 0 contributes the factor 1 even at c <= 0
    if (o == 2.0) return D2{c * c, 2.0 * c};
    if (o == 3.0) return D2{c * c * c, 3.0 * c * c};
    if (c > 0.0) return D2{pow(c, o), o * pow(c, o - 1.0)};
    return D2{0.0, 0.0};
}

// ---- rates of progress q_r (and, if JAC, the partial derivatives) ----------------------------
// Lane-per-reaction.  Rate laws: SURVEY Appendix A.2-A.6 (Cantera GasKinetics semantics).
// cx: [N + 2] concentrations with cx[N] = 1 (absent factor); M_s already evaluated.
// FR: also store the forward and reverse rates of progress (getFwdRatesOfProgress / getRevRatesOfProgress,
// PaSR.C:349-350) in dq[2r], dq[2r + 1] (the derivative scratch, unused without JAC).
template <bool JAC, bool FR = false>
DFC_HD void eval_reactions(const DevMech& m, double* sm, const double* cx, int dfc_lane) {
    (void)dfc_lane;
    const double* kfk = sm + m.off_kf;
    double* q = sm + m.off_q;
    double* dq = sm + m.off_dq;
    const uint32_t none = (uint32_t)m.N << 3;
    DFC_LANES_BEGIN
    for (int r = tl; r < m.R; r += kTeam) {
        const U4 d = ld_u4(m.rdesc + r);
        const uint32_t kind = d.x & 3u;
        const uint32_t fo[3] = {d.x & 0x7F8u, (d.x >> 8) & 0x7F8u, (d.x >> 16) & 0x7F8u};
        const uint32_t po[3] = {d.y & 0x7F8u, (d.y >> 8) & 0x7F8u, (d.y >> 16) & 0x7F8u};
        double fv[3], fd[3], pv[3], pd[3];
        if (!(d.x & kFlagPow)) {
            DFC_UNROLL
            for (int a = 0; a < 3; ++a) {
                fv[a] = ldb(cx, fo[a]);
                pv[a] = ldb(cx, po[a]);
                fd[a] = 1.0;
                pd[a] = 1.0;
            }
        } else {
            // fractional / large orders: c^o with Cantera's rule (non-positive c -> 0)
            const double* pw = m.powtab + (size_t)(d.z >> 20) * 9;
            for (int a = 0; a < 3; ++a) {
                fv[a] = pv[a] = 1.0;
                fd[a] = pd[a] = 0.0;
                if (fo[a] != none) {
                    const D2 t = pow_factor(DFC_LDG(pw + a), ldb(cx, fo[a]));
                    fv[a] = t.x;
                    fd[a] = t.y;
                }
                if (po[a] != none) {
                    const D2 t = pow_factor(DFC_LDG(pw + 6 + a), ldb(cx, po[a]));
                    pv[a] = t.x;
                    pd[a] = t.y;
                }
            }
        }
        const D2 kk = ld2(kfk + 2 * r);  // kf, 1/Kc
        double kfe = kk.x;
        double dkdM = 0.0;
        if (kind != (uint32_t)kElementary) {
            // third-body concentration of this reaction's efficiency set
            const uint32_t rg = m.ms_rng[d.z & 0xFFu];
            const U4* me = m.ms_ent + (rg & 0xFFFFu);
            double M = 0.0;
            for (int t = 0, nt = (int)(rg >> 16); t < nt; ++t) {
                const U4 u = ld_u4(me + t);
                M = fma(u2d(u.x, u.y), ldb(cx, u.z), M);
            }
            if (kind == (uint32_t)kThreeBody) {
                dkdM = kfe;
                kfe *= M;
            } else {
                const int fid = (int)((d.z >> 8) & 0xFFFu);
                const double rk = sm[m.off_k0 + fid];  // k0 / kinf
                const double pr = M * rk;
                double F = 1.0, dF_dpr = 0.0;
                if (kind == (uint32_t)kFalloffTroe) {
                    const double lf = sm[m.off_lfc + fid];
                    const double lpr = log10(fmax(pr, kSmall));
                    const double cc = -0.4 - 0.67 * lf;
                    const double nn = 0.75 - 1.27 * lf;
                    const double x = lpr + cc;
                    const double den = nn - 0.14 * x;
                    const double rden = 1.0 / den;
                    const double f1 = x * rden;
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
        const double kre = kfe * kk.y;
        q[r] = kfe * Pf - kre * Pp;
        if (FR) st2(dq + 2 * r, kfe * Pf, kre * Pp);
        if (JAC) {
            int slot = (int)d.w;
            if (fo[0] != none) dq[slot++] = kfe * fd[0] * fv[1] * fv[2];
            if (fo[1] != none) dq[slot++] = kfe * fv[0] * fd[1] * fv[2];
            if (fo[2] != none) dq[slot++] = kfe * fv[0] * fv[1] * fd[2];
            if (po[0] != none) dq[slot++] = -kre * pd[0] * pv[1] * pv[2];
            if (po[1] != none) dq[slot++] = -kre * pv[0] * pd[1] * pv[2];
            if (po[2] != none) dq[slot++] = -kre * pv[0] * pv[1] * pd[2];
            if (kind != (uint32_t)kElementary) dq[slot] = dkdM * (Pf - kk.y * Pp);
        }
    }
    DFC_LANES_END
}

// wdot_i (i < NA) at concentrations cx -> vec[i] (sm + off_vec).  cx is one of the cell's c / cs
// arrays ([N + 2]: slot N = 1.0, slot N+1 receives the total concentration).
template <bool JAC, class Regs>
DFC_HD void eval_wdot(const DevMech& m, double* sm, double* cx, Regs* regs, int dfc_lane) {
    (void)dfc_lane;
    if (m.n_msets > 0) {
        DFC_LANES_BEGIN
        double s = 0.0;
        for (int k = tl; k < m.N; k += kTeam) s += cx[k];
        DFC_LR.red = s;
        DFC_LANES_END_NOSYNC
        const double ctot = team_sum(regs);
        DFC_LANES_BEGIN
        if (tl == 0) cx[m.N + 1] = ctot;
        DFC_LANES_END
    }
    if (JAC) {
        // the dq scratch is shared with part / vec / row (dead by now): restore its zero slot
        DFC_LANES_BEGIN
        if (tl == 0) sm[m.off_dq + m.n_dq] = 0.0;
        DFC_LANES_END_NOSYNC
    }
    eval_reactions<JAC>(m, sm, cx, dfc_lane);
    // J = d wdot / d c from the reaction derivatives; must run before part/out overwrite dq
    if (JAC) run_rows(m.jac, m.coef, sm + m.off_dq, sm + m.off_J, dfc_lane);
    run_wdot(m, sm, regs, dfc_lane);
}

// ---- per-cell constants of the frozen-T sub-step: kf, 1/Kc, falloff k0/kinf and log10 Fcent ---
DFC_HD void setup_rate_constants(const DevMech& m, double* sm, double T, int dfc_lane) {
    (void)dfc_lane;
    const double logT = log(T), rT = 1.0 / T;
    double* g = sm + m.off_dq;  // scratch: g_k/RT at 1 atm
    DFC_LANES_BEGIN
    for (int k = tl; k < m.N; k += kTeam) {
        const double* a = nasa_sel(m, k, T);
        g[k] = nasa_h_RT(a, T) - nasa_s_R(a, T, logT);
    }
    DFC_LANES_END
    const double log_c0 = log(kOneAtm / (kGasConstant * T));
    const uint32_t none = (uint32_t)m.N << 3;
    DFC_LANES_BEGIN
    for (int r = tl; r < m.R; r += kTeam) {
        const U4 d = m.rdesc[r];
        const double* ar = m.arrh + (size_t)r * 3;
        const double lnk = DFC_LDG(ar) + DFC_LDG(ar + 1) * logT - DFC_LDG(ar + 2) * rT;
        double rk = 0.0;
        if (d.x & kFlagRev) {
            const uint32_t fo[3] = {d.x & 0x7F8u, (d.x >> 8) & 0x7F8u, (d.x >> 16) & 0x7F8u};
            const uint32_t po[3] = {d.y & 0x7F8u, (d.y >> 8) & 0x7F8u, (d.y >> 16) & 0x7F8u};
            double dG = 0.0, dn = 0.0;
            if (!(d.x & kFlagPow)) {
                for (int a = 0; a < 3; ++a) {
                    if (fo[a] != none) { dG -= ldb(g, fo[a]); dn -= 1.0; }
                    if (po[a] != none) { dG += ldb(g, po[a]); dn += 1.0; }
                }
            } else {
                const double* pw = m.powtab + (size_t)(d.z >> 20) * 9;
                for (int a = 0; a < 3; ++a) {
                    if (fo[a] != none) { dG -= pw[3 + a] * ldb(g, fo[a]); dn -= pw[3 + a]; }
                    if (po[a] != none) { dG += pw[6 + a] * ldb(g, po[a]); dn += pw[6 + a]; }
                }
            }
            rk = fmin(exp(dG - dn * log_c0), kBig);
        }
        st2(sm + m.off_kf + 2 * r, exp(lnk), rk);
        if ((d.x & 3u) >= (uint32_t)kFalloffLindemann) {
            const int fid = (int)((d.z >> 8) & 0xFFFu);
            const double* f = m.fall + (size_t)fid * 8;
            // k0 / kinf as one exponential (kinf = 0 cannot occur: ln A is finite for falloff)
            sm[m.off_k0 + fid] = fmin(exp(f[0] + f[1] * logT - f[2] * rT - lnk), 1e200);
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
