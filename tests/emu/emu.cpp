// TEST INFRASTRUCTURE: sequential host emulation of the warp-per-cell kernels.
// Compiles the very same device source (csrc/chem_core.cuh, integrator.cuh, cell_driver.cuh)
// with a "lane block" = for-loop over the 32 lanes, so the kernel logic and the mechanism
// tables can be debugged and parity-checked against the oracle on a machine without a GPU.
// It is never linked into libdfchem_b200.so; only tests/ build and call it.
#include <cstdio>
#include <cstring>
#include <limits>
#include <stdexcept>
#include <string>
#include <vector>

#define DFC_EMU_BOUNDS 1
#include "../../deepflame-dev_b200/csrc/cell_driver.cuh"
#include "../../deepflame-dev_b200/csrc/lane_cell.cuh"
#include "../../deepflame-dev_b200/csrc/dispatch.hpp"
#include "../../deepflame-dev_b200/csrc/transport_fit.hpp"

using namespace dfchem;

// window of the cell block for the table-driven address checks (chem_core.cuh, DFC_CHECK_ADDR)
namespace dfchem {
thread_local EmuWindow dfc_emu_window = {nullptr, nullptr, nullptr};
}

namespace {
void oob(const void* p) {
    throw std::out_of_range("emulated cell block: table-driven address " +
                            std::to_string(static_cast<const char*>(p) - dfc_emu_window.lo) + " outside [0, " +
                            std::to_string(dfc_emu_window.hi - dfc_emu_window.lo) + ")");
}
struct Window {
    explicit Window(const std::vector<double>& sm) {
        dfc_emu_window = {reinterpret_cast<const char*>(sm.data()),
                          reinterpret_cast<const char*>(sm.data() + sm.size()), &oob};
    }
    ~Window() { dfc_emu_window = {nullptr, nullptr, nullptr}; }
};
struct Emu {
    Mechanism mech;
    HostTables ht;
    DevMech dm;
};
template <int NP, int NSLOT>
void solve_t(const DevMech& m, const SolveArgs& a, unsigned long long* counters) {
    std::vector<double> sm(m.cell_doubles, 0.0);
    Window win(sm);
    std::vector<LaneRegs<NP, NSLOT>> regs(kTeam);
    if (!load_wdot_regs(m, regs.data(), 0, 0u)) throw std::runtime_error("wdot schedule too large");
    for (int cell = 0; cell < a.nC; ++cell) {
        // the device never clears a cell block between cells: poison it, so that a read of a slot
        // this cell has not written shows up as NaN in the results
        std::fill(sm.begin(), sm.end(), std::numeric_limits<double>::quiet_NaN());
        clear_jacobian(m, sm.data(), 0);
        StepStats st;
        solve_one_cell<NP, NSLOT>(m, a, sm.data(), regs.data(), 0, cell, st);
        counters[0] += st.accepted;
        counters[1] += st.rejected;
        counters[2] += st.status != 0;
    }
}
template <int NP, int NSLOT>
void wdot_t(const DevMech& m, int nC, const double* T, const double* p, const double* Y, int ldY,
            double* wr, int ldW) {
    std::vector<double> sm(m.cell_doubles, std::numeric_limits<double>::quiet_NaN());
    Window win(sm);
    std::vector<LaneRegs<NP, NSLOT>> regs(kTeam);
    if (!load_wdot_regs(m, regs.data(), 0, 0u)) throw std::runtime_error("wdot schedule too large");
    for (int cell = 0; cell < nC; ++cell)
        wdot_one_cell<NP, NSLOT>(m, sm.data(), regs.data(), 0, T, p, Y, ldY, wr, ldW, cell);
}
// J[i*N + j] = d wdot_i / d c_j (caller species order, inert rows/cols zero) and the inverse check
template <int NP, int NSLOT>
void jac_t(const DevMech& m, double T, double p, const double* Y, double inv_hg, double* Jout,
           double* Ainv_out) {
    std::vector<double> sm(m.cell_doubles, std::numeric_limits<double>::quiet_NaN());
    Window win(sm);
    std::vector<LaneRegs<NP, NSLOT>> regs(kTeam);
    if (!load_wdot_regs(m, regs.data(), 0, 0u)) throw std::runtime_error("wdot schedule too large");
    double ysum;
    clear_jacobian(m, sm.data(), 0);
    load_cell_state<NP, NSLOT>(m, sm.data(), regs.data(), 0, Y, 1, 0, T, p, ysum);
    setup_rate_constants(m, sm.data(), T, 0);
    init_const_slots(m, sm.data(), 0);
    eval_wdot<true>(m, sm.data(), sm.data() + m.off_c, regs.data(), 0);
    const double* J = sm.data() + m.off_J;
    // dense J in caller order incl. the default-efficiency rank-one part and the special columns
    for (int i = 0; i < m.NA; ++i)
        for (int j = 0; j < m.N; ++j) {
            double v = J[m.NA * m.ldj + i];  // gsum: d/dc_j through [M] with default efficiency
            if (j < m.NA) v += J[j * m.ldj + i];
            Jout[m.perm[i] * m.N + m.perm[j]] = v;
        }
    if (Ainv_out) {
        invert_matrix<NP, NSLOT>(m, sm.data(), regs.data(), 0, inv_hg);
        // column j of the inverse = solve against the unit vector e_j (through the stage-0 slot)
        double* vec = sm.data() + m.off_vec;
        for (int j = 0; j < m.NA; ++j) {
            for (int i = 0; i < NP; ++i) vec[i] = i == j ? 1.0 : 0.0;
            solve_stage<NP, NSLOT>(m, sm.data(), regs.data(), 0, vec, 0);
            for (int i = 0; i < m.NA; ++i) Ainv_out[i * m.NA + j] = regs[i % kTeam].k[0][i / kTeam];
        }
    }
}
}  // namespace

extern "C" {
void* emu_create(const void* blob, size_t len) {
    try {
        Emu* e = new Emu;
        e->mech = Mechanism::deserialize(static_cast<const uint8_t*>(blob), len);
        e->ht = build_tables(e->mech);
        e->dm = e->ht.rebase(e->ht.blob.data());
        return e;
    } catch (const std::exception&) {
        return nullptr;
    }
}
void emu_destroy(void* h) { delete static_cast<Emu*>(h); }
// 1 when the address check fires for an offset one past a 4-double block (and not for the last one)
int emu_bounds_selftest() {
    std::vector<double> sm(4, 1.0);
    Window win(sm);
    try {
        (void)ldb(sm.data(), 24u);
    } catch (const std::exception&) {
        return 0;
    }
    try {
        (void)ldb(sm.data(), 32u);
    } catch (const std::out_of_range&) {
        return 1;
    }
    return 0;
}
void emu_info(void* h, int* out) {
    const DevMech& m = static_cast<Emu*>(h)->dm;
    int v[] = {m.N, m.NA, m.NP, m.R, m.n_msets, m.n_fall, m.n_pow, m.n_dq, m.n_part, m.cell_doubles,
               m.wdot.nA, 0, m.jac.nA + m.jac.nB, static_cast<Emu*>(h)->ht.jac_entries,
               static_cast<Emu*>(h)->ht.wdot_entries};
    std::memcpy(out, v, sizeof(v));
}
void emu_perm(void* h, int* out) {
    const Emu* e = static_cast<Emu*>(h);
    std::memcpy(out, e->ht.perm.data(), sizeof(int) * e->ht.perm.size());
}
int emu_solve(void* h, int nC, const double* T, const double* p, const double* rho, const double* dt,
              double dt_uniform, const double* Y, int ldY, double* RR, int ldRR, double* Qdot,
              double* cost, int* status, double rtol, double atol, int max_steps,
              unsigned long long* counters) {
    const DevMech& m = static_cast<Emu*>(h)->dm;
    SolveArgs a{nC, T, p, rho, dt, dt_uniform, Y, ldY, RR, ldRR, Qdot, cost, status, rtol, atol, max_steps};
    counters[0] = counters[1] = counters[2] = 0;
#define CALL(NP, NS) solve_t<NP, NS>(m, a, counters)
    try {
        DFC_DISPATCH_NP(m.NP, CALL)
    } catch (const std::exception& e) {
        std::fprintf(stderr, "emu: %s\n", e.what());
        return -1;
    }
#undef CALL
    return 0;
}
// lane-per-cell path (lane_cell.cuh) run one cell after another with lane stride 1; -2: mechanism not eligible
int emu_solve_lpc(void* h, int nC, const double* T, const double* p, const double* rho, const double* dt,
                  double dt_uniform, const double* Y, int ldY, double* RR, int ldRR, double* Qdot,
                  double* cost, int* status, double rtol, double atol, int max_steps,
                  unsigned long long* counters) {
    const Emu* e = static_cast<Emu*>(h);
    HostLpc lpc = build_lpc_tables(e->mech, e->ht.perm, e->dm.NA);
    if (!lpc.usable) return -2;
    const LpcMech m = lpc.rebase(lpc.blob.data());
    SolveArgs a{nC, T, p, rho, dt, dt_uniform, Y, ldY, RR, ldRR, Qdot, cost, status, rtol, atol, max_steps};
    counters[0] = counters[1] = counters[2] = 0;
    std::vector<double> L(m.lane_doubles);
#define CALL(NAP)                                                              \
    for (int cell = 0; cell < nC; ++cell) {                                    \
        std::fill(L.begin(), L.end(), std::numeric_limits<double>::quiet_NaN()); \
        LpcState S;                                                            \
        lpc_begin_cell<NAP>(m, L.data(), a, cell, S);                          \
        while (!lpc_attempt<NAP>(m, L.data(), S, rtol, atol, max_steps)) {     \
        }                                                                      \
        lpc_finish_cell<NAP>(m, L.data(), a, S);                               \
        counters[0] += S.acc;                                                  \
        counters[1] += S.rej;                                                  \
        counters[2] += S.status != 0;                                          \
    }
    DFC_DISPATCH_LPC(m.NAP, CALL)
#undef CALL
    return 0;
}
// layout of the lane-per-cell tables: {usable, NAP, lane_doubles, n_dq, f_rows, Jacobian ELL rows, table bytes, off_dq}
void emu_lpc_info(void* h, int* out) {
    const Emu* e = static_cast<Emu*>(h);
    HostLpc lpc = build_lpc_tables(e->mech, e->ht.perm, e->dm.NA);
    const LpcMech m = lpc.usable ? lpc.rebase(lpc.blob.data()) : LpcMech{};
    const int v[8] = {lpc.usable ? 1 : 0, m.NAP, m.lane_doubles, m.n_dq, m.f_rows, lpc.usable ? m.jrow[m.NAP] : 0, m.bytes, m.off_dq};
    std::memcpy(out, v, sizeof(v));
}
// the mechanism compiler's transport-fit generator (host code, no emulation involved); 0 ok, -2 no transport data
int emu_transport_fits(void* h, double* visc, double* cond, double* diff) {
    const Emu* e = static_cast<Emu*>(h);
    std::vector<double> v, c, d;
    if (!generate_transport_fits(e->mech, v, c, d)) return -2;
    std::copy(v.begin(), v.end(), visc);
    std::copy(c.begin(), c.end(), cond);
    std::copy(d.begin(), d.end(), diff);
    return 0;
}
int emu_wdot(void* h, int nC, const double* T, const double* p, const double* Y, int ldY, double* wr,
             int ldW) {
    const DevMech& m = static_cast<Emu*>(h)->dm;
#define CALL(NP, NS) wdot_t<NP, NS>(m, nC, T, p, Y, ldY, wr, ldW)
    try {
        DFC_DISPATCH_NP(m.NP, CALL)
    } catch (const std::exception& e) {
        std::fprintf(stderr, "emu: %s\n", e.what());
        return -1;
    }
#undef CALL
    return 0;
}
int emu_jacobian(void* h, double T, double p, const double* Y, double inv_hg, double* Jout,
                 double* Ainv_out) {
    const DevMech& m = static_cast<Emu*>(h)->dm;
#define CALL(NP, NS) jac_t<NP, NS>(m, T, p, Y, inv_hg, Jout, Ainv_out)
    try {
        DFC_DISPATCH_NP(m.NP, CALL)
    } catch (const std::exception& e) {
        std::fprintf(stderr, "emu: %s\n", e.what());
        return -1;
    }
#undef CALL
    return 0;
}
}
