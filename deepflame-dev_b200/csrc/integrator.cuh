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

// Integrate d c/dt = wdot(c) over [0, dt] for the cell whose state is set up in `sm`
// (c, cs(inert), kf, rkc, k0, lfc) and regs (c0, dc = 0).  Error control restates the
// rtol*|Y_k| + atol weights of ReactorNet::setTolerances (dfChemistryModel.C:414) in
// concentration units: sc_k = atol*rho/W_k + rtol*max(|c_k|, |c_k,new|); WRMS norm over the N
// species (CVODE's norm also counts the reactor's three constant unknowns).
template <int NP, int NSLOT>
DFC_HD StepStats integrate_cell(const DevMech& m, double* sm, LaneRegs<NP, NSLOT>* regs, int dfc_lane,
                                double dt, double rtol, double atol_rho, int max_steps) {
    (void)dfc_lane;
    StepStats st{0, 0, kCellOk};
    double* c = sm + m.off_c;
    double* cs = sm + m.off_cs;
    double* vec = sm + m.off_vec;
    double* red = sm + m.off_red;
    const int NA = m.NA;
    double t = 0.0, h = dt;
    bool after_reject = false;
    // zero the padded tail of vec once
    DFC_LANES_BEGIN
    for (int j = NA + tl; j < NP; j += kTeam) vec[j] = 0.0;
    DFC_LANES_END
    while (true) {
        // base point: f0 = wdot(c), J = d wdot/d c
        eval_wdot(m, sm, c, vec, dfc_lane, true);
        run_gather(m.jac, m.coef, sm + m.off_dq, sm + m.off_J, dfc_lane);
        DFC_LANES_BEGIN
        DFC_UNROLL
        for (int s = 0; s < NSLOT; ++s) {
            const int row = tl + kTeam * s;
            DFC_LR.f0[s] = row < NA ? vec[row] : 0.0;
        }
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
            const double ih = 1.0 / hs;
            gj_invert<NP, NSLOT>(m, sm, regs, dfc_lane, ih / kGamma);
            DFC_UNROLL
            for (int i = 0; i < kStages; ++i) {
                if (i > 0) {
                    DFC_LANES_BEGIN
                    DFC_UNROLL
                    for (int s = 0; s < NSLOT; ++s) {
                        const int row = tl + kTeam * s;
                        if (row < NA) {
                            double v = c[row];
                            DFC_UNROLL
                            for (int j = 0; j < i; ++j)
                                if (rodas_a(i, j) != 0.0) v = fma(rodas_a(i, j), DFC_LR.k[j][s], v);
                            cs[row] = v;
                        }
                    }
                    DFC_LANES_END
                    eval_wdot(m, sm, cs, vec, dfc_lane, false);
                }
                DFC_LANES_BEGIN
                DFC_UNROLL
                for (int s = 0; s < NSLOT; ++s) {
                    const int row = tl + kTeam * s;
                    if (row < NA) {
                        double v = (i == 0) ? DFC_LR.f0[s] : vec[row];
                        DFC_UNROLL
                        for (int j = 0; j < i; ++j)
                            if (rodas_c(i, j) != 0.0) v = fma(rodas_c(i, j) * ih, DFC_LR.k[j][s], v);
                        vec[row] = v;
                    }
                }
                DFC_LANES_END
                DFC_LANES_BEGIN
                double x[NSLOT];
                gj_apply<NP, NSLOT>(DFC_LR, vec, x);
                DFC_UNROLL
                for (int s = 0; s < NSLOT; ++s) DFC_LR.k[i][s] = x[s];
                DFC_LANES_END
            }
            // new solution = argument of the last stage + k_6; error estimate = k_6
            DFC_LANES_BEGIN
            double e2 = 0.0;
            DFC_UNROLL
            for (int s = 0; s < NSLOT; ++s) {
                const int row = tl + kTeam * s;
                if (row < NA) {
                    double inc = DFC_LR.k[kStages - 1][s];
                    DFC_UNROLL
                    for (int j = 0; j < kStages - 1; ++j) inc = fma(rodas_a(kStages - 1, j), DFC_LR.k[j][s], inc);
                    DFC_LR.inc[s] = inc;
                    const double cn = c[row] + inc;
                    const double sc = atol_rho / DFC_LDG(m.W + row) + rtol * fmax(fabs(c[row]), fabs(cn));
                    const double e = DFC_LR.k[kStages - 1][s] / sc;
                    e2 = fma(e, e, e2);
                }
            }
            red[tl] = e2;
            DFC_LANES_END
            double en = sqrt(team_sum(red) / (double)m.N);
            if (!(en < 1e30)) en = 1e30;  // NaN / inf -> strong rejection
            double fac = 0.9 / sqrt(sqrt(fmax(en, 1e-12)));
            fac = fmin(fmax(fac, 0.2), after_reject ? 1.0 : 6.0);
            if (en <= 1.0) {
                accepted = true;
                after_reject = false;
                ++st.accepted;
                t += hs;
                DFC_LANES_BEGIN
                DFC_UNROLL
                for (int s = 0; s < NSLOT; ++s) {
                    const int row = tl + kTeam * s;
                    if (row < NA) {
                        DFC_LR.dc[s] += DFC_LR.inc[s];
                        c[row] = DFC_LR.c0[s] + DFC_LR.dc[s];
                    }
                }
                DFC_LANES_END
            } else {
                ++st.rejected;
                after_reject = true;
            }
            h = hs * fac;
        }
        if (t >= dt * (1.0 - 1e-14)) return st;
    }
}

}  // namespace dfchem
