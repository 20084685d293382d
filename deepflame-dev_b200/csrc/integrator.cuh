// Per-cell linear algebra and the adaptive RODAS4 step loop (see chem_core.cuh for the model).
#pragma once
#include "chem_core.cuh"

namespace dfchem {

// ---- (I/(h g) - J)^-1 by in-place Gauss-Jordan, rows in registers -----------------------------
// Lane tl owns matrix rows tl + 32*s (s < NSLOT).  Elimination step k: the owner of row k
// publishes it through shared memory (double-buffered, one team barrier per step); every other
// row does  f = a_ik/a_kk;  a_i -= f * a_k  (with column k replaced by the running inverse
// column, i.e. a_ik <- -f).  Pivot rows are NOT rescaled; each row remembers 1/pivot (invd) and
// the scale is applied in the mat-vec (gj_apply), which keeps every lane on the same DFMA
// stream.
// ROTATING COLUMN STORAGE: at step k register j holds logical column (k + j) mod NP, so the
// pivot column is always register 0 and the step index never indexes the register file.  The
// rotation costs nothing: the update a'_(j-1) = fma(-f, u_j, a_j) writes to the neighbouring
// register.  That lets the k-loop stay a real loop (about 2 KB of SASS instead of NP unrolled
// copies, which overflowed the instruction cache).  After exactly NP steps the rotation has
// gone full circle; padding rows/columns NA <= r < NP are identity, so their steps are no-ops.
// No pivoting: the matrix is I/(h g) - J with h under error control; a breakdown shows up as a
// non-finite error norm and the step is retried with a smaller h.
template <int NP, int NSLOT, int SK>
DFC_HD void gj_steps(const DevMech& m, double* sm, LaneRegs<NP, NSLOT>* regs, int dfc_lane, int k_begin,
                     int k_end) {
    (void)dfc_lane;
#ifdef __CUDA_ARCH__
#pragma unroll 1
#endif
    for (int k = k_begin; k < k_end; ++k) {
        const int lk = k - SK * kTeam;
        double* rowb = sm + m.off_row + (k & 1) * NP;
        DFC_LANES_BEGIN
        if (tl == lk) {
            DFC_UNROLL
            for (int j = 0; j + 1 < NP; j += 2) st2(rowb + j, DFC_LR.a[SK][j], DFC_LR.a[SK][j + 1]);
            if (NP & 1) rowb[NP - 1] = DFC_LR.a[SK][NP - 1];
        }
        DFC_LANES_END
        DFC_LANES_BEGIN
        const double p = 1.0 / rowb[0];
        double f[NSLOT];
        DFC_UNROLL
        for (int s = 0; s < NSLOT; ++s) {
            const bool own = (s == SK) && (tl == lk);
            f[s] = own ? 0.0 : DFC_LR.a[s][0] * p;
            if (own) DFC_LR.invd[s] = p;
        }
        DFC_UNROLL
        for (int j = 0; j < NP; j += 2) {
            if (j + 1 < NP) {
                const D2 u = ld2(rowb + j);
                DFC_UNROLL
                for (int s = 0; s < NSLOT; ++s) {
                    if (j > 0) DFC_LR.a[s][j - 1] = fma(-f[s], u.x, DFC_LR.a[s][j]);
                    DFC_LR.a[s][j] = fma(-f[s], u.y, DFC_LR.a[s][j + 1]);
                }
            } else {
                const double u = rowb[j];
                DFC_UNROLL
                for (int s = 0; s < NSLOT; ++s) DFC_LR.a[s][j - 1] = fma(-f[s], u, DFC_LR.a[s][j]);
            }
        }
        DFC_UNROLL
        for (int s = 0; s < NSLOT; ++s) {
            const bool own = (s == SK) && (tl == lk);
            DFC_LR.a[s][NP - 1] = own ? 1.0 : -f[s];
        }
        DFC_LANES_END_NOSYNC
    }
}

template <int NP, int NSLOT>
DFC_HD void gj_invert(const DevMech& m, double* sm, LaneRegs<NP, NSLOT>* regs, int dfc_lane,
                      double inv_hg) {
    (void)dfc_lane;
    const double* J = sm + m.off_J;
    const int NA = m.NA, ldj = m.ldj;
    DFC_LANES_BEGIN
    DFC_UNROLL
    for (int s = 0; s < NSLOT; ++s) {
        const int row = tl + kTeam * s;
        if (row < NA) {
            const double gs = J[NA * ldj + row];
            DFC_UNROLL
            for (int j = 0; j < NP; ++j)
                DFC_LR.a[s][j] = (j < NA) ? ((j == row ? inv_hg : 0.0) - J[j * ldj + row] - gs) : 0.0;
        } else {
            DFC_UNROLL
            for (int j = 0; j < NP; ++j) DFC_LR.a[s][j] = (j == row) ? 1.0 : 0.0;  // identity padding
        }
        DFC_LR.invd[s] = 0.0;
    }
    DFC_LR.tmp = 0.0;
    DFC_LANES_END
    gj_steps<NP, NSLOT, 0>(m, sm, regs, dfc_lane, 0, NP < kTeam ? NP : kTeam);
    if (NSLOT > 1) gj_steps<NP, NSLOT, (NSLOT > 1 ? 1 : 0)>(m, sm, regs, dfc_lane, kTeam, NP);
}

// x = A^-1 b for this lane's rows; b is read (broadcast) from shared memory vec[0..NP)
template <int NP, int NSLOT>
DFC_HD void gj_apply(const LaneRegs<NP, NSLOT>& lr, const double* vec, double* x) {
    double acc[NSLOT];
    DFC_UNROLL
    for (int s = 0; s < NSLOT; ++s) acc[s] = 0.0;
    DFC_UNROLL
    for (int j = 0; j + 1 < NP; j += 2) {
        const D2 b = ld2(vec + j);
        DFC_UNROLL
        for (int s = 0; s < NSLOT; ++s) {
            acc[s] = fma(lr.a[s][j], b.x, acc[s]);
            acc[s] = fma(lr.a[s][j + 1], b.y, acc[s]);
        }
    }
    if (NP & 1) {
        DFC_UNROLL
        for (int s = 0; s < NSLOT; ++s) acc[s] = fma(lr.a[s][NP - 1], vec[NP - 1], acc[s]);
    }
    DFC_UNROLL
    for (int s = 0; s < NSLOT; ++s) x[s] = acc[s] * lr.invd[s];
}

struct StepStats {
    int accepted, rejected, status;
};

// RODAS4 stage combinations with the stage index as a run-time value (the stage loop is a real
// loop so that eval_wdot is instantiated once); the stage increments stay in registers because
// every access below uses a compile-time index.
template <int NSLOT>
DFC_HD double stage_point(int i, const double (&k)[kStages][NSLOT], int s, double v) {
    switch (i) {
        case 1: v = fma(rodas_a(1, 0), k[0][s], v); break;
        case 2: v = fma(rodas_a(2, 0), k[0][s], v); v = fma(rodas_a(2, 1), k[1][s], v); break;
        case 3:
            v = fma(rodas_a(3, 0), k[0][s], v); v = fma(rodas_a(3, 1), k[1][s], v);
            v = fma(rodas_a(3, 2), k[2][s], v);
            break;
        default:  // stages 4 and 5 share a_4j; stage 5 adds k_4
            v = fma(rodas_a(4, 0), k[0][s], v); v = fma(rodas_a(4, 1), k[1][s], v);
            v = fma(rodas_a(4, 2), k[2][s], v); v = fma(rodas_a(4, 3), k[3][s], v);
            if (i == 5) v += k[4][s];
            break;
    }
    return v;
}
template <int NSLOT>
DFC_HD double stage_rhs(int i, const double (&k)[kStages][NSLOT], int s, double v, double ih) {
    switch (i) {
        case 0: break;
        case 1: v = fma(rodas_c(1, 0) * ih, k[0][s], v); break;
        case 2: v = fma(rodas_c(2, 0) * ih, k[0][s], v); v = fma(rodas_c(2, 1) * ih, k[1][s], v); break;
        case 3:
            v = fma(rodas_c(3, 0) * ih, k[0][s], v); v = fma(rodas_c(3, 1) * ih, k[1][s], v);
            v = fma(rodas_c(3, 2) * ih, k[2][s], v);
            break;
        case 4:
            v = fma(rodas_c(4, 0) * ih, k[0][s], v); v = fma(rodas_c(4, 1) * ih, k[1][s], v);
            v = fma(rodas_c(4, 2) * ih, k[2][s], v); v = fma(rodas_c(4, 3) * ih, k[3][s], v);
            break;
        default:
            v = fma(rodas_c(5, 0) * ih, k[0][s], v); v = fma(rodas_c(5, 1) * ih, k[1][s], v);
            v = fma(rodas_c(5, 2) * ih, k[2][s], v); v = fma(rodas_c(5, 3) * ih, k[3][s], v);
            v = fma(rodas_c(5, 4) * ih, k[4][s], v);
            break;
    }
    return v;
}
template <int NSLOT>
DFC_HD void stage_store(int i, double (&k)[kStages][NSLOT], int s, double x) {
    switch (i) {
        case 0: k[0][s] = x; break;
        case 1: k[1][s] = x; break;
        case 2: k[2][s] = x; break;
        case 3: k[3][s] = x; break;
        case 4: k[4][s] = x; break;
        default: k[5][s] = x; break;
    }
}

// ---- bordered variant for 33..36 rows ----------------------------------------------------------
// A = [A11 A12; A21 A22] with A11 the 32 x 32 core (row l on lane l, NSLOT would otherwise double
// every lane's work for two to four extra rows).  The 32 Gauss-Jordan steps run on the core rows
// augmented by the border columns (g -> A11^-1 A12, same row scaling as the core), and update
// the border rows, stored by column (lane j holds column j: v -> -A21 A11^-1) plus the Schur
// complement S = A22 - A21 A11^-1 A12, which every lane carries redundantly and finally inverts.
//   x2 = S^-1 (r2 + V r1),   x1 = A11^-1 r1 - (A11^-1 A12) x2          (solve_stage)
template <int NB>
DFC_HD int small_inverse(double* S) {
    int neg = 0;
    DFC_UNROLL
    for (int k = 0; k < NB; ++k) {
        const double p = 1.0 / S[k * NB + k];
        neg += p < 0.0 ? 1 : 0;
        DFC_UNROLL
        for (int j = 0; j < NB; ++j)
            if (j != k) S[k * NB + j] *= p;
        S[k * NB + k] = p;
        DFC_UNROLL
        for (int i = 0; i < NB; ++i) {
            if (i == k) continue;
            const double f = S[i * NB + k];
            DFC_UNROLL
            for (int j = 0; j < NB; ++j)
                if (j != k) S[i * NB + j] = fma(-f, S[k * NB + j], S[i * NB + j]);
            S[i * NB + k] = -f * p;
        }
    }
    return neg;
}

template <int NP, int NSLOT>
DFC_HD void gjb_invert(const DevMech& m, double* sm, LaneRegs<NP, NSLOT>* regs, int dfc_lane, double inv_hg) {
    (void)dfc_lane;
    constexpr int NB = LaneRegs<NP, NSLOT>::NB;
    constexpr int RB = (kTeam + 2 * NB + 1) & ~1;  // published pivot record: 32 core + NB aug + NB border
    const double* J = sm + m.off_J;
    const int NA = m.NA, ldj = m.ldj;  // NA == NP == 32 + NB
    DFC_LANES_BEGIN
    {
        const double gs = J[NA * ldj + tl];
        DFC_UNROLL
        for (int j = 0; j < kTeam; ++j) DFC_LR.a[0][j] = (j == tl ? inv_hg : 0.0) - J[j * ldj + tl] - gs;
        DFC_UNROLL
        for (int b = 0; b < NB; ++b) {
            DFC_LR.g[b] = -J[(kTeam + b) * ldj + tl] - gs;
            const double gsb = J[NA * ldj + kTeam + b];
            DFC_LR.v[b] = -J[tl * ldj + kTeam + b] - gsb;
            DFC_UNROLL
            for (int b2 = 0; b2 < NB; ++b2)
                DFC_LR.S[b * NB + b2] = (b == b2 ? inv_hg : 0.0) - J[(kTeam + b2) * ldj + kTeam + b] - gsb;
        }
        DFC_LR.invd[0] = 0.0;
    }
    DFC_LANES_END
#ifdef __CUDA_ARCH__
#pragma unroll 1
#endif
    for (int k = 0; k < kTeam; ++k) {
        double* rowb = sm + m.off_row + (k & 1) * RB;
        DFC_LANES_BEGIN
        if (tl == k) {
            DFC_UNROLL
            for (int j = 0; j < kTeam; j += 2) st2(rowb + j, DFC_LR.a[0][j], DFC_LR.a[0][j + 1]);
            DFC_UNROLL
            for (int b = 0; b < NB; ++b) {
                rowb[kTeam + b] = DFC_LR.g[b];
                rowb[kTeam + NB + b] = DFC_LR.v[b];
            }
        }
        DFC_LANES_END
        DFC_LANES_BEGIN
        const bool own = tl == k;
        const double p = 1.0 / rowb[0];
        const double f = own ? 0.0 : DFC_LR.a[0][0] * p;
        if (own) DFC_LR.invd[0] = p;
        const double ur = rowb[(tl - k) & (kTeam - 1)];  // pivot row's entry in logical column `lane`
        DFC_UNROLL
        for (int j = 0; j < kTeam; j += 2) {
            const D2 u = ld2(rowb + j);
            if (j > 0) DFC_LR.a[0][j - 1] = fma(-f, u.x, DFC_LR.a[0][j]);
            DFC_LR.a[0][j] = fma(-f, u.y, DFC_LR.a[0][j + 1]);
        }
        DFC_LR.a[0][kTeam - 1] = own ? 1.0 : -f;
        double ug[NB];
        DFC_UNROLL
        for (int b = 0; b < NB; ++b) {
            ug[b] = rowb[kTeam + b];
            DFC_LR.g[b] = fma(-f, ug[b], DFC_LR.g[b]);
        }
        DFC_UNROLL
        for (int b = 0; b < NB; ++b) {
            const double fb = rowb[kTeam + NB + b] * p;
            DFC_LR.v[b] = own ? -fb : fma(-fb, ur, DFC_LR.v[b]);
            DFC_UNROLL
            for (int b2 = 0; b2 < NB; ++b2) DFC_LR.S[b * NB + b2] = fma(-fb, ug[b2], DFC_LR.S[b * NB + b2]);
        }
        DFC_LANES_END_NOSYNC
    }
    DFC_LANES_BEGIN
    DFC_LR.tmp = (double)small_inverse<NB>(DFC_LR.S);  // negative Schur pivots (same on every lane)
    DFC_LANES_END
}

// k_i = A^-1 vec for stage i (team level: the bordered variant needs a team reduction)
template <int NP, int NSLOT>
DFC_HD void solve_stage(const DevMech& m, double* sm, LaneRegs<NP, NSLOT>* regs, int dfc_lane,
                        const double* vec, int i) {
    (void)dfc_lane; (void)m; (void)sm;
    constexpr int NB = LaneRegs<NP, NSLOT>::NB;
    if constexpr (NB == 0) {
        DFC_LANES_BEGIN
        double x[NSLOT];
        gj_apply<NP, NSLOT>(DFC_LR, vec, x);
        DFC_UNROLL
        for (int s = 0; s < NSLOT; ++s) stage_store<NSLOT>(i, DFC_LR.k, s, x[s]);
        DFC_LANES_END
    } else {
        DFC_LANES_BEGIN
        double acc0 = 0.0, acc1 = 0.0;
        DFC_UNROLL
        for (int j = 0; j < kTeam; j += 2) {
            const D2 b = ld2(vec + j);
            acc0 = fma(DFC_LR.a[0][j], b.x, acc0);
            acc1 = fma(DFC_LR.a[0][j + 1], b.y, acc1);
        }
        DFC_LR.tmp = acc0 + acc1;
        const double r1 = vec[tl];
        DFC_UNROLL
        for (int b = 0; b < NB; ++b) DFC_LR.redv[b] = DFC_LR.v[b] * r1;
        DFC_LANES_END_NOSYNC
        double z[NB];
        team_sum_n<NB>(regs, z);
        DFC_LANES_BEGIN
        double x2[NB];
        DFC_UNROLL
        for (int b = 0; b < NB; ++b) {
            double sacc = 0.0;
            DFC_UNROLL
            for (int b2 = 0; b2 < NB; ++b2) sacc = fma(DFC_LR.S[b * NB + b2], vec[kTeam + b2] + z[b2], sacc);
            x2[b] = sacc;
        }
        double x0 = DFC_LR.tmp, x1 = 0.0;
        DFC_UNROLL
        for (int b = 0; b < NB; ++b) {
            x0 = fma(-DFC_LR.g[b], x2[b], x0);
            if (tl == b) x1 = x2[b];
        }
        stage_store<NSLOT>(i, DFC_LR.k, 0, x0 * DFC_LR.invd[0]);
        stage_store<NSLOT>(i, DFC_LR.k, 1, x1);
        DFC_LANES_END
    }
}

template <int NP, int NSLOT>
DFC_HD void invert_matrix(const DevMech& m, double* sm, LaneRegs<NP, NSLOT>* regs, int dfc_lane, double inv_hg) {
    if constexpr (LaneRegs<NP, NSLOT>::NB == 0)
        gj_invert<NP, NSLOT>(m, sm, regs, dfc_lane, inv_hg);
    else
        gjb_invert<NP, NSLOT>(m, sm, regs, dfc_lane, inv_hg);
}

// step-size controller constants: Hairer & Wanner's RODAS controller with safety 0.95 and a wider
// shrink range after a rejection; first step h0 = min(dt, K / ||f0/sc||).  Tuned on the gri30 bench
// batch through the host emulation: 28.8 -> 25.2 attempts per cell; safety 1.0 stalls.  K = 10 (round 2;
// was 1e3): same attempt count (fewer rejections), but cells whose whole sub-step fits one step at K = 1e3
// now start carefully, which cuts the worst RR error on trace species at rtol 1e-6 / atol 1e-10 by 40x
// (profiles/r2/parity_vs_rtol.json) -- below the CVODE-equivalent port's at equal tolerances.
#ifndef DFC_FACMINREJ
#define DFC_FACMINREJ 0.1
#endif
#ifndef DFC_FACMAX
#define DFC_FACMAX 6.0
#endif
#ifndef DFC_FACSAFE
#define DFC_FACSAFE 0.95
#endif
#ifndef DFC_H0K
#define DFC_H0K 10.0
#endif
constexpr double kFacSafe = DFC_FACSAFE, kFacMax = DFC_FACMAX, kFacMin = 0.2, kFacMinReject = DFC_FACMINREJ;
constexpr double kFacUnstable = 0.25;  // step cut when the unstable-mode guard fires

// Integrate d c/dt = wdot(c) over [0, dt] for the cell whose state is set up in `sm`
// (c, cs(inert), kf, k0, lfc) and regs (c0, dc = 0).  Error control restates the
// rtol*|Y_k| + atol weights of ReactorNet::setTolerances (dfChemistryModel.C:414) in
// concentration units: sc_k = atol*rho/W_k + rtol*max(|c_k|, |c_k,new|); WRMS norm over the N
// species (CVODE's norm also counts the reactor's three constant unknowns).
template <int NP, int NSLOT>
// h_io (deltaTChem): a positive value on entry is the first trial step (warm start from the previous flow
// step: the reference carries a field of that name, dfChemistryModel.H, but never feeds it to CVODE); on
// exit it holds the step the controller would take next, ignoring the truncation at the end of the interval.
DFC_HD StepStats integrate_cell(const DevMech& m, double* sm, LaneRegs<NP, NSLOT>* regs, int dfc_lane,
                                double dt, double rtol, double atol_rho, int max_steps, double& h_io) {
    (void)dfc_lane;
    StepStats st{0, 0, kCellOk};
    const bool warm = h_io > 0.0;
    double h_free = dt;  // last step-size proposal that was not cut by the end of the interval
    double* c = sm + m.off_c;
    double* cs = sm + m.off_cs;
    double* vec = sm + m.off_vec;
    const int NA = m.NA;
    double t = 0.0, h = warm ? fmin(h_io, dt) : dt;
    bool after_reject = false;
    // absolute tolerance in concentration units, once per cell (not a global load + divide per attempt)
    DFC_LANES_BEGIN
    DFC_UNROLL
    for (int s = 0; s < NSLOT; ++s) {
        const int row = tl + kTeam * s;
        DFC_LR.atolc[s] = row < NA ? atol_rho / DFC_LDG(m.W + row) : 0.0;
    }
    DFC_LANES_END_NOSYNC
    while (true) {
        // base point: f0 = wdot(c), J = d wdot/d c (eval_wdot<true> also runs the Jacobian rows)
        eval_wdot<true>(m, sm, c, regs, dfc_lane);
        DFC_LANES_BEGIN
        DFC_UNROLL
        for (int s = 0; s < NSLOT; ++s) {
            const int row = tl + kTeam * s;
            DFC_LR.f0[s] = row < NA ? vec[row] : 0.0;
        }
        DFC_LANES_END
        if (st.accepted + st.rejected == 0 && !warm) {
            // first step size from the scaled rate norm (Hairer & Wanner II.4): h0 = K / ||f0 / sc||
            DFC_LANES_BEGIN
            double e2 = 0.0;
            DFC_UNROLL
            for (int s = 0; s < NSLOT; ++s) {
                const int row = tl + kTeam * s;
                if (row < NA) {
                    const double sc = DFC_LR.atolc[s] + rtol * fabs(c[row]);
                    const double e = DFC_LR.f0[s] / sc;
                    e2 = fma(e, e, e2);
                }
            }
            DFC_LR.red = e2;
            DFC_LANES_END_NOSYNC
            const double d1 = sqrt(team_sum(regs) / (double)m.N);
            // (bounded below: with atol -> 0 and a species at c = 0 the scaled norm is unbounded)
            if (d1 * dt > DFC_H0K) h = fmax(DFC_H0K / d1, 1e-3 * dt);
        }
        // vec shares storage with the derivative scratch: re-zero its padded tail
        DFC_LANES_BEGIN
        for (int j = NA + tl; j < NP; j += kTeam) vec[j] = 0.0;
        DFC_LANES_END
        bool accepted = false;
        double hs = 0.0;
        while (!accepted) {
            if (st.accepted + st.rejected >= max_steps) {
                st.status = kCellMaxSteps;
                return st;
            }
            hs = fmin(h, dt - t);
            if (!(hs > 1e-14 * dt)) {
                st.status = kCellUnderflow;
                return st;
            }
            const bool cut = hs < h;
            const double ih = 1.0 / hs;
            invert_matrix<NP, NSLOT>(m, sm, regs, dfc_lane, ih / kGamma);
            // Unstable-mode guard.  det(I/(h g) - J) = product of the pivots is negative iff an odd number of
            // real eigenvalues of J exceed 1/(h g): a chain-branching (explosive) mode with h*lambda > 4, which an
            // L-stable method damps instead of following (R(z) -> 0 for z -> +inf as well) while the embedded
            // error estimate stays small.  Such a step is never taken: h is cut until the mode is resolved.
            DFC_LANES_BEGIN
            double cnt = tl == 0 ? DFC_LR.tmp : 0.0;
            DFC_UNROLL
            for (int s = 0; s < LaneRegs<NP, NSLOT>::NROW; ++s) cnt += DFC_LR.invd[s] < 0.0 ? 1.0 : 0.0;
            DFC_LR.red = cnt;
            DFC_LANES_END_NOSYNC
            if (((int)team_sum(regs)) & 1) {
                ++st.rejected;
                after_reject = true;
                h = hs * kFacUnstable;
                continue;
            }
#ifdef __CUDA_ARCH__
#pragma unroll 1
#endif
            for (int i = 0; i < kStages; ++i) {
                if (i > 0) {
                    DFC_LANES_BEGIN
                    DFC_UNROLL
                    for (int s = 0; s < NSLOT; ++s) {
                        const int row = tl + kTeam * s;
                        if (row < NA) cs[row] = stage_point<NSLOT>(i, DFC_LR.k, s, c[row]);
                    }
                    DFC_LANES_END
                    eval_wdot<false>(m, sm, cs, regs, dfc_lane);
                }
                DFC_LANES_BEGIN
                DFC_UNROLL
                for (int s = 0; s < NSLOT; ++s) {
                    const int row = tl + kTeam * s;
                    if (row < NA) vec[row] = stage_rhs<NSLOT>(i, DFC_LR.k, s, (i == 0) ? DFC_LR.f0[s] : vec[row], ih);
                }
                DFC_LANES_END
                solve_stage<NP, NSLOT>(m, sm, regs, dfc_lane, vec, i);
            }
            // new solution = argument of the last stage + k_6; error estimate = k_6
            DFC_LANES_BEGIN
            double e2 = 0.0;
            DFC_UNROLL
            for (int s = 0; s < NSLOT; ++s) {
                const int row = tl + kTeam * s;
                if (row < NA) {
                    const double inc = stage_point<NSLOT>(5, DFC_LR.k, s, DFC_LR.k[kStages - 1][s]);
                    const double cn = c[row] + inc;
                    const double sc = DFC_LR.atolc[s] + rtol * fmax(fabs(c[row]), fabs(cn));
                    const double e = DFC_LR.k[kStages - 1][s] / sc;
                    e2 = fma(e, e, e2);
                }
            }
            DFC_LR.red = e2;
            DFC_LANES_END_NOSYNC
            double en = sqrt(team_sum(regs) / (double)m.N);
            if (!(en < 1e30)) en = 1e30;  // NaN / inf -> strong rejection
            double fac = kFacSafe / sqrt(sqrt(fmax(en, 1e-12)));
            if (en <= 1.0) {
                fac = fmin(fmax(fac, kFacMin), after_reject ? 1.0 : kFacMax);
                accepted = true;
                after_reject = false;
                ++st.accepted;
                t += hs;
                DFC_LANES_BEGIN
                DFC_UNROLL
                for (int s = 0; s < NSLOT; ++s) {
                    const int row = tl + kTeam * s;
                    if (row < NA) {
                        // the increment is recomputed (five FMAs) rather than kept live across the test;
                        // dc carries c(t) - c(0) for RR, c itself only feeds the rate evaluation
                        const double inc = stage_point<NSLOT>(5, DFC_LR.k, s, DFC_LR.k[kStages - 1][s]);
                        DFC_LR.dc[s] += inc;
                        c[row] += inc;
                    }
                }
                DFC_LANES_END
            } else {
                fac = fmin(fmax(fac, kFacMinReject), 1.0);
                ++st.rejected;
                after_reject = true;
            }
            h_free = (accepted && cut) ? fmax(h, hs * fac) : hs * fac;
            h = hs * fac;
        }
        if (t >= dt * (1.0 - 1e-14)) {
            h_io = h_free;
            return st;
        }
    }
}

}  // namespace dfchem
