// One cell of dfChemistryModel::solveSingle (src/dfChemistryModel/dfChemistryModel.C:738-780)
// executed by one team: load + normalise Y (setState_TPY, :755), frozen-(T, rho) integration
// (:757-764), RR_i = (Ynew_i - Yraw_i)/dt*rho_OF and Qdot = -sum hc_i RR_i (:768-772).
// Shared by the CUDA kernels (kernels.cu) and the host emulation used in tests (tests/emu).
#pragma once
#include "integrator.cuh"

namespace dfchem {

struct SolveArgs {
    int nC;
    const double* T;
    const double* p;
    const double* rho;
    const double* dt;  // per cell or nullptr
    double dt_uniform;
    const double* Y;
    int ldY;
    double* RR;
    int ldRR;
    double* Qdot;
    double* cost;  // may be nullptr
    int* status;   // may be nullptr
    double rtol, atol;
    int max_steps;
    double* dtchem;  // optional in/out per-cell integrator step (deltaTChem): > 0 on entry = first trial step
};

// set up c (all N species), cs (inert tail), regs.c0/dc; returns rho_ct; ysum via reference
// once per team, before its first cell: Jacobian entries that no reaction touches are never
// written afterwards and stay zero for the whole launch
DFC_HD void clear_jacobian(const DevMech& m, double* sm, int dfc_lane) {
    (void)dfc_lane;
    DFC_LANES_BEGIN
    for (int i = tl; i < (m.NA + 1) * m.ldj; i += kTeam) sm[m.off_J + i] = 0.0;
    DFC_LANES_END
}

template <int NP, int NSLOT>
DFC_HD double load_cell_state(const DevMech& m, double* sm, LaneRegs<NP, NSLOT>* regs, int dfc_lane,
                              const double* Y, int ldY, int cell, double T, double p, double& ysum) {
    (void)dfc_lane;
    double* c = sm + m.off_c;
    double* cs = sm + m.off_cs;
    DFC_LANES_BEGIN
    double s = 0.0;
    for (int k = tl; k < m.N; k += kTeam) {
        const double y = Y[(size_t)DFC_LDG(m.perm + k) * ldY + cell];
        c[k] = y;  // raw mass fraction for now
        s += y;
    }
    DFC_LR.red = s;
    DFC_LANES_END
    ysum = team_sum(regs);
    const double rys = 1.0 / ysum;
    DFC_LANES_BEGIN
    double s = 0.0;
    for (int k = tl; k < m.N; k += kTeam) s += c[k] * rys / DFC_LDG(m.W + k);
    DFC_LR.red = s;
    DFC_LANES_END
    const double wbar = 1.0 / team_sum(regs);
    const double rho_ct = p * wbar / (kGasConstant * T);
    DFC_LANES_BEGIN
    for (int k = tl; k < m.N; k += kTeam) {
        const double ck = rho_ct * (c[k] * rys) / DFC_LDG(m.W + k);
        c[k] = ck;
        cs[k] = ck;
    }
    DFC_LANES_END
    DFC_LANES_BEGIN
    DFC_UNROLL
    for (int s = 0; s < NSLOT; ++s) {
        DFC_LR.dc[s] = 0.0;
    }
    DFC_LANES_END
    return rho_ct;
}

// constant slots the tables rely on: c[N] = cs[N] = 1 (absent factor), q[R] = dq[n_dq] = 0
// (row-program padding).  Called after setup_rate_constants, which uses dq as scratch.
DFC_HD void init_const_slots(const DevMech& m, double* sm, int dfc_lane) {
    (void)dfc_lane;
    DFC_LANES_BEGIN
    if (tl == 0) {
        sm[m.off_c + m.N] = 1.0;
        sm[m.off_cs + m.N] = 1.0;
        sm[m.off_c + m.N + 1] = 0.0;
        sm[m.off_cs + m.N + 1] = 0.0;
        sm[m.off_q + m.R] = 0.0;
        sm[m.off_dq + m.n_dq] = 0.0;
    }
    DFC_LANES_END
}

template <int NP, int NSLOT>
DFC_HD void solve_one_cell(const DevMech& m, const SolveArgs& a, double* sm, LaneRegs<NP, NSLOT>* regs,
                           int dfc_lane, int cell, StepStats& st) {
    const double T = a.T[cell], p = a.p[cell];
    const double dt = a.dt ? a.dt[cell] : a.dt_uniform;
    double ysum;
    const double rho_ct = load_cell_state<NP, NSLOT>(m, sm, regs, dfc_lane, a.Y, a.ldY, cell, T, p, ysum);
    setup_rate_constants(m, sm, T, dfc_lane);
    init_const_slots(m, sm, dfc_lane);
    st = StepStats{0, 0, kCellOk};
    double h_io = a.dtchem ? a.dtchem[cell] : 0.0;
    if (dt > 0.0) st = integrate_cell<NP, NSLOT>(m, sm, regs, dfc_lane, dt, a.rtol, a.atol * rho_ct, a.max_steps, h_io);
    // RR and Qdot
    const double rho_of = a.rho[cell];
    const double scale = dt > 0.0 ? rho_of / dt : 0.0;
    const double rys = 1.0 / ysum;
    DFC_LANES_BEGIN
    double qd = 0.0;
    DFC_UNROLL
    for (int s = 0; s < NSLOT; ++s) {
        const int k = tl + kTeam * s;
        if (k < m.N) {
            const int ko = DFC_LDG(m.perm + k);
            const double yraw = a.Y[(size_t)ko * a.ldY + cell];
            double dY = yraw * rys - yraw;  // normalisation quirk: RR is taken against the raw Y
            if (k < m.NA) dY += DFC_LR.dc[s] * DFC_LDG(m.W + k) / rho_ct;
            const double rr = dY * scale;
            a.RR[(size_t)ko * a.ldRR + cell] = rr;
            qd = fma(-DFC_LDG(m.hc + k), rr, qd);
        }
    }
    for (int k = tl + kTeam * NSLOT; k < m.N; k += kTeam) {  // inert tail beyond the row slots
        const int ko = DFC_LDG(m.perm + k);
        const double yraw = a.Y[(size_t)ko * a.ldY + cell];
        const double rr = (yraw * rys - yraw) * scale;
        a.RR[(size_t)ko * a.ldRR + cell] = rr;
        qd = fma(-DFC_LDG(m.hc + k), rr, qd);
    }
    DFC_LR.red = qd;
    DFC_LANES_END
    const double qdot = team_sum(regs);
    DFC_LANES_BEGIN
    if (tl == 0) {
        a.Qdot[cell] = qdot;
        if (a.cost) a.cost[cell] = (double)(st.accepted + st.rejected);
        if (a.status) a.status[cell] = st.status;
        if (a.dtchem && dt > 0.0) a.dtchem[cell] = h_io;
    }
    DFC_LANES_END
}

// calculateW (dfChemistryModel.C:897-923): wrate_i = wdot_i * W_i at (T, p, Y)
template <int NP, int NSLOT>
DFC_HD void wdot_one_cell(const DevMech& m, double* sm, LaneRegs<NP, NSLOT>* regs, int dfc_lane,
                          const double* T, const double* p, const double* Y, int ldY, double* wr,
                          int ldW, int cell) {
    double ysum;
    const double Tc = T[cell];
    load_cell_state<NP, NSLOT>(m, sm, regs, dfc_lane, Y, ldY, cell, Tc, p[cell], ysum);
    setup_rate_constants(m, sm, Tc, dfc_lane);
    init_const_slots(m, sm, dfc_lane);
    eval_wdot<false>(m, sm, sm + m.off_c, regs, dfc_lane);
    DFC_LANES_BEGIN
    for (int k = tl; k < m.N; k += kTeam)
        wr[(size_t)DFC_LDG(m.perm + k) * ldW + cell] = k < m.NA ? sm[m.off_vec + k] * DFC_LDG(m.W + k) : 0.0;
    DFC_LANES_END
}

// Net rates of progress q_r [kmol/m^3/s] at (T, p, Y) in the caller's reaction order: the
// getNetRatesOfProgress call of calculateRR (dfChemistryModel.C:925-987), which projects one
// reaction on one species as (nu''_i - nu'_i) q_r W_i.  rorder: internal -> caller's index.
// Optional outputs: qf / qr = forward / reverse rates of progress (PaSR.C:349-350) and tc = the PaSR
// "reactionRate" chemical time scale (PaSR.C:326-380): with wf_r = (sum of product nu) qf_r and
// wr_r = (sum of reactant nu) qr_r,  tc = sum(wf + wr) / sum(wf^2 + wr^2) * sum_i rho Y_i / W_i  (vGreat if 0),
// rho = the caller's (OpenFOAM) density, Y the raw mass fractions.  rxnu[2r], rxnu[2r+1]: those
// stoichiometric sums, internal reaction order.
template <int NP, int NSLOT>
DFC_HD void rop_one_cell(const DevMech& m, double* sm, LaneRegs<NP, NSLOT>* regs, int dfc_lane,
                         const double* T, const double* p, const double* Y, int ldY,
                         const int* rorder, double* q_out, double* qf_out, double* qr_out, int ldQ,
                         const double* rxnu, const double* rho, double* tc_out, int cell) {
    double ysum;
    const double Tc = T[cell];
    const double rho_ct = load_cell_state<NP, NSLOT>(m, sm, regs, dfc_lane, Y, ldY, cell, Tc, p[cell], ysum);
    setup_rate_constants(m, sm, Tc, dfc_lane);
    init_const_slots(m, sm, dfc_lane);
    double* cx = sm + m.off_c;
    DFC_LANES_BEGIN
    double s = 0.0;
    for (int k = tl; k < m.N; k += kTeam) s += cx[k];
    DFC_LR.red = s;
    DFC_LANES_END_NOSYNC
    const double ctot = team_sum(regs);
    DFC_LANES_BEGIN
    if (tl == 0) cx[m.N + 1] = ctot;
    DFC_LANES_END
    eval_reactions<false, true>(m, sm, cx, dfc_lane);
    DFC_LANES_BEGIN
    double sw = 0.0, sw2 = 0.0;
    for (int r = tl; r < m.R; r += kTeam) {
        const size_t o = (size_t)DFC_LDG(rorder + r) * ldQ + cell;
        const D2 fr = ld2(sm + m.off_dq + 2 * r);
        if (q_out) q_out[o] = sm[m.off_q + r];
        if (qf_out) qf_out[o] = fr.x;
        if (qr_out) qr_out[o] = fr.y;
        if (tc_out) {
            const double wf = DFC_LDG(rxnu + 2 * r) * fr.x, wr = DFC_LDG(rxnu + 2 * r + 1) * fr.y;
            sw += wf + wr;
            sw2 = fma(wf, wf, fma(wr, wr, sw2));
        }
    }
    DFC_LR.red = sw;
    DFC_LR.tmp = sw2;
    DFC_LANES_END
    if (tc_out) {
        const double sumW = team_sum(regs);
        DFC_LANES_BEGIN
        DFC_LR.red = DFC_LR.tmp;
        DFC_LANES_END_NOSYNC
        const double sumW2 = team_sum(regs);
        const double csum = rho[cell] * ysum * ctot / rho_ct;
        DFC_LANES_BEGIN
        if (tl == 0) tc_out[cell] = sumW2 == 0.0 ? 1.0e300 : sumW / sumW2 * csum;  // OpenFOAM's vGreat
        DFC_LANES_END
    }
}

}  // namespace dfchem
