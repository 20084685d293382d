// C ABI of libdfchem_b200.so (include/dfchem.h).  Host side: mechanism compile -> device tables,
// launch configuration, chunked H2D/compute/D2H pipeline for the host-pointer entry points.
// There is deliberately no CPU path here: every entry point runs the sm_100a kernels or fails.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/dfchem.h"
#include "dispatch.hpp"
#include "kernels.cuh"
#include "rebalance.cuh"
#include "transport.cuh"
#include "transport_fit.hpp"

using namespace dfchem;

namespace {
thread_local std::string g_create_error;

struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    cudaError_t ensure(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
        cudaError_t e = cudaMalloc(&p, bytes);
        if (e == cudaSuccess) cap = bytes;
        return e;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
};
constexpr int kSlots = 2;
constexpr int kRing = 256;  // launch slots: every kernel launch gets its own work-queue word and counter record
}  // namespace

struct dfchem_ctx {
    int device = 0;
    Mechanism mech;
    HostTables ht;
    void* d_blob = nullptr;
    DevMech dm{};
    double rtol = 1e-9, atol = 1e-15;  // dfChemistryModel.C:57-58
    int max_steps = 20000;             // CVODE default used by ReactorNet
    int sm_count = 0, warps = 0;  // warps (= cells in flight) per persistent CTA
    cudaStream_t streams[kSlots] = {nullptr, nullptr};
    DevBuf slot_in[kSlots], slot_out[kSlots], scratch;
    unsigned int* d_queue = nullptr;          // [kRing] work-queue words, one per launch (ring)
    unsigned long long* d_counters = nullptr; // [kRing][4] accepted, rejected, failed cells, (unused)
    unsigned long long seq = 0;               // launches issued so far (ring position)
    unsigned long long call_first = 0;        // first ring slot and number of solve launches of the last solve call
    unsigned long long call_n = 0;
    unsigned long long call_acc[3] = {0, 0, 0};  // counters already folded on the host (calls longer than the ring)
    unsigned long long launches = 0;
    double* dtchem = nullptr;                 // dfchem_bind_deltaTChem
    // in-library load balancer (rebalance.cuh)
    ncclComm_t comm = nullptr;
    bool comm_owned = false;
    int nranks = 1, rank = 0;
    cudaStream_t comm_stream = nullptr;
    cudaEvent_t ev[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    DevBuf bal_prefix, bal_scan_tmp, bal_small, bal_guest_in, bal_guest_out;
    int* h_counts = nullptr;                  // pinned [nranks^2]
    double bal_stats[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    bool bal_timed = false;
    int* d_rxsum = nullptr;                   // not used on the host; see d_rxnu
    double* d_rxnu = nullptr;                 // [2R] internal order: sum of product / reactant stoichiometric coefficients
    int* d_rorder = nullptr;                  // [R] internal -> caller's reaction index (rop kernel)
    DevBuf d_transport;                       // transport fit tables (dfchem_set_transport_fits)
    TransportTab tt{};
    int thermo_ctas = 0;                      // resident thermo_kernel CTAs per SM (occupancy query, cached)
    int thermo_pair_ctas = 0;                 // resident thermo_pair_kernel CTAs per SM
    int thermo_tma_ok = 0;                    // 0 unknown, 1 bulk-async thermo kernel usable, -1 tables + tiles exceed shared memory
    HostLpc lpc;                              // lane-per-cell tables (small mechanisms)
    void* d_lpc = nullptr;
    LpcMech lm{};
    int lpc_warps = 0;                        // > 0: dfchem_solve* run lpc_solve_kernel on batches of at least lpc_min_cells
    int lpc_min_cells = 0;
    bool call_lane = false;                   // the current solve call runs the lane kernel (begin_solve_call)
    bool have_call = false;
    std::string err, solver_name;
};

namespace {

int fail(dfchem_ctx* h, int code, const std::string& msg) {
    if (h) h->err = msg;
    g_create_error = msg;
    return code;
}
#define CUDA_TRY(h, call)                                                                      \
    do {                                                                                       \
        cudaError_t e_ = (call);                                                               \
        if (e_ != cudaSuccess)                                                                 \
            return fail(h, DFCHEM_ECUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); \
    } while (0)

size_t smem_bytes(const dfchem_ctx* h) {
    return (size_t)h->dm.hot_bytes + (size_t)h->warps * h->dm.cell_doubles * sizeof(double);
}

template <int NP, int NSLOT>
cudaError_t configure_t(dfchem_ctx* h) {
    if (h->dm.wdot.nA > wdot_rows_cap(NP)) return cudaErrorLaunchOutOfResources;
    int max_optin = 0;
    cudaError_t e = cudaDeviceGetAttribute(&max_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, h->device);
    if (e != cudaSuccess) return e;
    // persistent CTA, one per SM: as many warps (cells in flight) as shared memory allows
    const size_t cell = (size_t)h->dm.cell_doubles * sizeof(double);
    h->warps = (int)std::min<size_t>(max_warps(NP), ((size_t)max_optin - h->dm.hot_bytes) / cell);
    // measurement knob: fewer cells in flight per SM (how throughput scales with resident warps)
    if (const char* cap = std::getenv("DFCHEM_MAX_WARPS")) h->warps = std::max(1, std::min(h->warps, std::atoi(cap)));
    if (h->warps < 1) return cudaErrorLaunchOutOfResources;
    // the attribute is per kernel instantiation and shared by every handle using it: always the
    // device maximum, never this handle's size
    const size_t smem = smem_bytes(h);
    e = cudaFuncSetAttribute(solve_kernel<NP, NSLOT>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_optin);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(wdot_kernel<NP, NSLOT>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_optin);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(rop_kernel<NP, NSLOT>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_optin);
    if (e != cudaSuccess) return e;
    int nb = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, solve_kernel<NP, NSLOT>, h->warps * kTeam, smem);
    if (e != cudaSuccess) return e;
    return nb >= 1 ? cudaSuccess : cudaErrorLaunchOutOfResources;
}

size_t lpc_smem_bytes(const dfchem_ctx* h) {
    return (size_t)h->lm.bytes + (size_t)h->lpc_warps * kLpcLanes * h->lm.lane_doubles * sizeof(double);
}

template <int NP, int NSLOT>
void launch_solve_t(dfchem_ctx* h, const SolveArgs& a, unsigned int* queue, unsigned long long* cnt, cudaStream_t s) {
    const int grid = std::max(1, std::min(h->sm_count, (a.nC + h->warps - 1) / h->warps));
    solve_kernel<NP, NSLOT><<<grid, h->warps * kTeam, smem_bytes(h), s>>>(h->dm, a, queue, cnt);
}
template <int NP, int NSLOT>
void launch_wdot_t(dfchem_ctx* h, int nC, const double* T, const double* p, const double* Y, int ldY,
                   double* wr, int ldW, unsigned int* queue, cudaStream_t s) {
    const int grid = std::max(1, std::min(h->sm_count, (nC + h->warps - 1) / h->warps));
    wdot_kernel<NP, NSLOT><<<grid, h->warps * kTeam, smem_bytes(h), s>>>(h->dm, nC, T, p, Y, ldY, wr, ldW, queue);
}

template <int NP, int NSLOT>
void launch_rop_t(dfchem_ctx* h, int nC, const double* T, const double* p, const double* Y, int ldY,
                  double* q, double* qf, double* qr, int ldQ, const double* rho, double* tc, unsigned int* queue,
                  cudaStream_t s) {
    const int grid = std::max(1, std::min(h->sm_count, (nC + h->warps - 1) / h->warps));
    rop_kernel<NP, NSLOT><<<grid, h->warps * kTeam, smem_bytes(h), s>>>(h->dm, nC, T, p, Y, ldY, h->d_rorder, q, qf, qr, ldQ,
                                                                       h->d_rxnu, rho, tc, queue);
}

// Every launch takes the next slot of the ring: its own queue word and counter record, zeroed on the
// launching stream.  Launches of one handle on different streams therefore never share a counter.
int take_slot(dfchem_ctx* h, cudaStream_t s, unsigned int** queue, unsigned long long** cnt) {
    const size_t i = (size_t)(h->seq++ % kRing);
    *queue = h->d_queue + i;
    CUDA_TRY(h, cudaMemsetAsync(*queue, 0, sizeof(unsigned int), s));
    if (cnt) {
        *cnt = h->d_counters + 4 * i;
        CUDA_TRY(h, cudaMemsetAsync(*cnt, 0, 4 * sizeof(unsigned long long), s));
    }
    return DFCHEM_OK;
}

// nC: the cells of the whole call.  The kernel is chosen once per call, so that every chunk of a host-pointer call and the
// guest and resident launches of a balanced call run the same code (same bits for the same cell).
void begin_solve_call(dfchem_ctx* h, int nC) {
    h->call_lane = h->lpc_warps > 0 && nC >= h->lpc_min_cells;
    h->have_call = true;
    h->call_first = h->seq;
    h->call_n = 0;
    h->call_acc[0] = h->call_acc[1] = h->call_acc[2] = 0;
    h->launches = 0;
}

// sum the counter records of the launches of the last solve call (device must be idle for them)
int read_call_counters(dfchem_ctx* h, unsigned long long out[3]) {
    std::vector<unsigned long long> ring((size_t)kRing * 4);
    CUDA_TRY(h, cudaMemcpy(ring.data(), h->d_counters, ring.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    for (int k = 0; k < 3; ++k) out[k] = h->call_acc[k];
    for (unsigned long long j = 0; j < h->call_n; ++j) {
        const size_t i = (size_t)((h->call_first + j) % kRing);
        for (int k = 0; k < 3; ++k) out[k] += ring[4 * i + k];
    }
    return DFCHEM_OK;
}

int launch_solve(dfchem_ctx* h, const SolveArgs& a, cudaStream_t s) {
    unsigned int* queue = nullptr;
    unsigned long long* cnt = nullptr;
    // the ring position of the solve launches of one call must be contiguous: nothing else launches in between
    // (the handle is single-caller, include/dfchem.h)
    int rc = take_slot(h, s, &queue, &cnt);
    if (rc != DFCHEM_OK) return rc;
    if (h->call_lane) {
        const int per_cta = h->lpc_warps * kLpcLanes;
        const int grid = std::max(1, std::min(h->sm_count, (a.nC + per_cta - 1) / per_cta));
#define CALL(NAP) lpc_solve_kernel<NAP><<<grid, per_cta, lpc_smem_bytes(h), s>>>(h->lm, a, queue, cnt, h->mech.R <= 4 ? 16 : 4)
        DFC_DISPATCH_LPC(h->lm.NAP, CALL)
#undef CALL
    } else {
#define CALL(NP, NS) launch_solve_t<NP, NS>(h, a, queue, cnt, s)
        DFC_DISPATCH_NP(h->dm.NP, CALL)
#undef CALL
    }
    ++h->launches;
    ++h->call_n;
    CUDA_TRY(h, cudaGetLastError());
    return DFCHEM_OK;
}

int finish_create(dfchem_ctx* h, int device, dfchem_handle* out) {
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        delete h;
        return fail(nullptr, DFCHEM_ECUDA, "no CUDA device available (libdfchem_b200 has no CPU path)");
    }
    if (device < 0 || device >= ndev) {
        delete h;
        return fail(nullptr, DFCHEM_EINVAL, "device index out of range");
    }
    h->device = device;
    try {
        h->ht = build_tables(h->mech);
    } catch (const std::exception& e) {
        std::string msg = e.what();
        delete h;
        return fail(nullptr, DFCHEM_EUNSUPPORTED, msg);
    }
    auto bail = [&](const char* what, cudaError_t e) {
        std::string msg = std::string(what) + ": " + cudaGetErrorString(e);
        dfchem_destroy(h);
        return fail(nullptr, DFCHEM_ECUDA, msg);
    };
    cudaError_t e;
    if ((e = cudaSetDevice(device)) != cudaSuccess) return bail("cudaSetDevice", e);
    cudaDeviceProp prop;
    if ((e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess) return bail("cudaGetDeviceProperties", e);
    h->sm_count = prop.multiProcessorCount;
    if ((e = cudaMalloc(&h->d_blob, h->ht.blob.size())) != cudaSuccess) return bail("cudaMalloc(tables)", e);
    if ((e = cudaMemcpy(h->d_blob, h->ht.blob.data(), h->ht.blob.size(), cudaMemcpyHostToDevice)) != cudaSuccess)
        return bail("cudaMemcpy(tables)", e);
    h->dm = h->ht.rebase(h->d_blob);
    if ((e = cudaMalloc(&h->d_queue, kRing * sizeof(unsigned int))) != cudaSuccess) return bail("cudaMalloc", e);
    if ((e = cudaMalloc(&h->d_counters, (size_t)kRing * 4 * sizeof(unsigned long long))) != cudaSuccess) return bail("cudaMalloc", e);
    cudaMemset(h->d_counters, 0, (size_t)kRing * 4 * sizeof(unsigned long long));
    if (h->mech.R > 0) {
        if ((e = cudaMalloc(&h->d_rorder, h->mech.R * sizeof(int))) != cudaSuccess) return bail("cudaMalloc", e);
        if ((e = cudaMemcpy(h->d_rorder, h->ht.rorder.data(), h->mech.R * sizeof(int), cudaMemcpyHostToDevice)) != cudaSuccess)
            return bail("cudaMemcpy(rorder)", e);
        // PaSR reactionRate time scale (PaSR.C:355-378): per reaction sum of product / reactant stoichiometric coefficients
        std::vector<double> rxnu((size_t)2 * h->mech.R, 0.0);
        for (int rn = 0; rn < h->mech.R; ++rn) {
            const int r = h->ht.rorder[rn];
            for (int a = 0; a < kMaxSide; ++a) {
                if (h->mech.p_idx[r * kMaxSide + a] >= 0) rxnu[2 * rn] += h->mech.p_nu[r * kMaxSide + a];
                if (h->mech.r_idx[r * kMaxSide + a] >= 0) rxnu[2 * rn + 1] += h->mech.r_nu[r * kMaxSide + a];
            }
        }
        if ((e = cudaMalloc(&h->d_rxnu, rxnu.size() * sizeof(double))) != cudaSuccess) return bail("cudaMalloc", e);
        if ((e = cudaMemcpy(h->d_rxnu, rxnu.data(), rxnu.size() * sizeof(double), cudaMemcpyHostToDevice)) != cudaSuccess)
            return bail("cudaMemcpy(rxnu)", e);
    }
    for (int i = 0; i < kSlots; ++i)
        if ((e = cudaStreamCreateWithFlags(&h->streams[i], cudaStreamNonBlocking)) != cudaSuccess)
            return bail("cudaStreamCreate", e);
    e = cudaErrorInvalidValue;
#define CALL(NP, NS) e = configure_t<NP, NS>(h)
    DFC_DISPATCH_NP(h->dm.NP, CALL)
#undef CALL
    if (e != cudaSuccess) return bail("kernel configuration", e);
    // Lane-per-cell solve kernel (lane_cell.cuh, gather form with ELL lists).  Measured on B200 (profiles/r2/lane_final.json,
    // lane_refill_granularity.json; 1,048,576 cells, rtol 1e-6 / atol 1e-10) against the warp-per-cell kernel: c7_onestep 3.2e8 vs
    // 1.65e7 cells/s (20x), es80 1.70e7 vs 4.0e6 (4.2x), h2_li 6.4e6 vs 2.5e6 (2.6x).  Default: mechanisms with at most 8 active
    // species (padded sizes 4 / 6 / 8; the 12-wide instantiation spills and stays opt-in); DFCHEM_SOLVER=lane / warp forces either kernel.
    const char* solver_env = std::getenv("DFCHEM_SOLVER");
    const std::string solver = solver_env ? solver_env : "";
    constexpr int kLpcAutoActive = 8;
    if (solver == "lane" || (solver != "warp" && h->dm.NA <= kLpcAutoActive)) {
        h->lpc = build_lpc_tables(h->mech, h->ht.perm, h->dm.NA);
        if (h->lpc.usable) {
            int max_optin = 0;
            if ((e = cudaDeviceGetAttribute(&max_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, device)) != cudaSuccess)
                return bail("cudaDeviceGetAttribute", e);
            const size_t per_warp = (size_t)kLpcLanes * h->lpc.layout.lane_doubles * sizeof(double);
            const int max_w = h->lpc.layout.NAP <= 8 ? 8 : 4;   // the kernel's launch bounds (256 / 128 threads: 255 registers each)
            const int warps = (int)std::min<size_t>(max_w, ((size_t)max_optin - h->lpc.blob.size()) / per_warp);
            if (warps >= 1) {
                if ((e = cudaMalloc(&h->d_lpc, h->lpc.blob.size())) != cudaSuccess) return bail("cudaMalloc(lpc tables)", e);
                if ((e = cudaMemcpy(h->d_lpc, h->lpc.blob.data(), h->lpc.blob.size(), cudaMemcpyHostToDevice)) != cudaSuccess)
                    return bail("cudaMemcpy(lpc tables)", e);
                h->lm = h->lpc.rebase(h->d_lpc);
                e = cudaErrorInvalidValue;
#define CALL(NAP) e = cudaFuncSetAttribute(lpc_solve_kernel<NAP>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_optin)
                DFC_DISPATCH_LPC(h->lpc.layout.NAP, CALL)
#undef CALL
                if (e != cudaSuccess) return bail("cudaFuncSetAttribute(lpc_solve_kernel)", e);
                h->lpc_warps = warps;
                // A lane holds a cell for its whole integration: the lane kernel needs tens of thousands of cells to fill the
                // GPU (148 SMs x warps x 32 lanes) and a batch takes at least as long as its longest cell, which is 2-3x the
                // warp kernel's per-cell latency.  Measured (profiles/r2/small_batches_lane_vs_warp.json, ms per sub-step, lane /
                // warp): es80 1000 cells 4.1 / 1.9, 8192 cells 4.4 / 3.3, 16384 cells 4.4 / 5.3; h2_li 6.5 / 2.7, 6.6 / 4.8,
                // 6.7 / 8.2; c7_onestep (one reaction, 4.6 attempts) 0.12 / 0.16 at 1000 cells.  So small batches of a multi-
                // reaction mechanism (df0DFoam's 1000 cells, config C1) go one warp per cell.  DFCHEM_LANE_MIN_CELLS overrides.
                const char* mc = std::getenv("DFCHEM_LANE_MIN_CELLS");
                h->lpc_min_cells = mc ? std::max(0, std::atoi(mc)) : (solver == "lane" || h->mech.R <= 4 ? 0 : 12288);
            }
        }
    }
    *out = h;
    return DFCHEM_OK;
}

}  // namespace

extern "C" {

const char* dfchem_version(void) { return "dfchem-b200 0.1 (sm_100a)"; }
const char* dfchem_solver_name(dfchem_handle h) {
    if (!h) return "";
    const bool lane = h->have_call ? h->call_lane : h->lpc_warps > 0;
    h->solver_name = lane ? "lpc_solve_kernel<" + std::to_string(h->lm.NAP) + ">"
                          : "solve_kernel<" + std::to_string(h->dm.NP) + "," + std::to_string(h->dm.NP > 32 ? 2 : 1) + ">";
    return h->solver_name.c_str();
}

int dfchem_create(const char* mech_yaml_path, int device, dfchem_handle* out) {
    if (!mech_yaml_path || !out) return fail(nullptr, DFCHEM_EINVAL, "null argument");
    *out = nullptr;
    dfchem_ctx* h = new (std::nothrow) dfchem_ctx;
    if (!h) return fail(nullptr, DFCHEM_ENOMEM, "out of host memory");
    try {
        h->mech = compile_mechanism_file(mech_yaml_path);
    } catch (const std::exception& e) {
        std::string msg = e.what();
        delete h;
        return fail(nullptr, DFCHEM_EMECH, msg);
    }
    return finish_create(h, device, out);
}

int dfchem_create_from_blob(const void* blob, size_t len, int device, dfchem_handle* out) {
    if (!blob || !out) return fail(nullptr, DFCHEM_EINVAL, "null argument");
    *out = nullptr;
    dfchem_ctx* h = new (std::nothrow) dfchem_ctx;
    if (!h) return fail(nullptr, DFCHEM_ENOMEM, "out of host memory");
    try {
        h->mech = Mechanism::deserialize(static_cast<const uint8_t*>(blob), len);
    } catch (const std::exception& e) {
        std::string msg = e.what();
        delete h;
        return fail(nullptr, DFCHEM_EMECH, msg);
    }
    return finish_create(h, device, out);
}

void dfchem_destroy(dfchem_handle h) {
    if (!h) return;
    cudaSetDevice(h->device);
    for (int i = 0; i < kSlots; ++i) {
        if (h->streams[i]) cudaStreamDestroy(h->streams[i]);
        h->slot_in[i].release();
        h->slot_out[i].release();
    }
    h->scratch.release();
    if (h->d_blob) cudaFree(h->d_blob);
    if (h->d_queue) cudaFree(h->d_queue);
    if (h->d_counters) cudaFree(h->d_counters);
    if (h->d_rorder) cudaFree(h->d_rorder);
    if (h->d_lpc) cudaFree(h->d_lpc);
    if (h->comm && h->comm_owned && nccl_api().ok()) nccl_api().CommDestroy(h->comm);
    if (h->comm_stream) cudaStreamDestroy(h->comm_stream);
    for (cudaEvent_t e : h->ev)
        if (e) cudaEventDestroy(e);
    h->bal_prefix.release();
    h->bal_scan_tmp.release();
    h->bal_small.release();
    h->bal_guest_in.release();
    h->bal_guest_out.release();
    if (h->h_counts) cudaFreeHost(h->h_counts);
    if (h->d_rxnu) cudaFree(h->d_rxnu);
    h->d_transport.release();
    delete h;
}

const char* dfchem_last_error(dfchem_handle h) { return h ? h->err.c_str() : g_create_error.c_str(); }
int dfchem_species_count(dfchem_handle h) { return h ? h->mech.N : DFCHEM_EINVAL; }
int dfchem_reaction_count(dfchem_handle h) { return h ? h->mech.R : DFCHEM_EINVAL; }
const char* dfchem_species_name(dfchem_handle h, int i) {
    return (h && i >= 0 && i < h->mech.N) ? h->mech.species[i].c_str() : nullptr;
}
int dfchem_species_index(dfchem_handle h, const char* name) {
    if (!h || !name) return -1;
    for (int i = 0; i < h->mech.N; ++i)
        if (h->mech.species[i] == name) return i;
    return -1;
}
int dfchem_species_constants(dfchem_handle h, int which, double* out) {
    if (!h || !out || which < 0 || which > 1) return fail(h, DFCHEM_EINVAL, "bad argument");
    const std::vector<double>& v = which == 0 ? h->mech.W : h->mech.hc;
    std::copy(v.begin(), v.end(), out);
    return DFCHEM_OK;
}
int dfchem_set_tolerances(dfchem_handle h, double rtol, double atol) {
    if (!h || !(rtol > 0) || !(atol >= 0)) return fail(h, DFCHEM_EINVAL, "tolerances must be positive");
    h->rtol = rtol;
    h->atol = atol;
    return DFCHEM_OK;
}
int dfchem_set_max_steps(dfchem_handle h, int max_steps) {
    if (!h || max_steps < 1) return fail(h, DFCHEM_EINVAL, "max_steps must be >= 1");
    h->max_steps = max_steps;
    return DFCHEM_OK;
}

int dfchem_bind_deltaTChem(dfchem_handle h, double* deltaTChem) {
    if (!h) return DFCHEM_EINVAL;
    h->dtchem = deltaTChem;
    return DFCHEM_OK;
}

int dfchem_solve_dev(dfchem_handle h, int nC, const double* T, const double* p, const double* rho,
                     const double* dt, double dt_uniform, const double* Y, int ldY, double* RR,
                     int ldRR, double* Qdot, double* cost, int* status, void* stream) {
    if (!h) return DFCHEM_EINVAL;
    if (nC < 0 || (nC > 0 && (!T || !p || !rho || !Y || !RR || !Qdot)) || ldY < nC || ldRR < nC)
        return fail(h, DFCHEM_EINVAL, "dfchem_solve_dev: bad argument");
    CUDA_TRY(h, cudaSetDevice(h->device));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    begin_solve_call(h, nC);
    if (nC == 0) return DFCHEM_OK;
    SolveArgs a{nC, T, p, rho, dt, dt_uniform, Y, ldY, RR, ldRR, Qdot, cost, status, h->rtol, h->atol, h->max_steps, h->dtchem};
    return launch_solve(h, a, s);
}

namespace {
// Host-pointer chemistry sub-step: 2^18-cell chunks through two device staging slots on two streams
// (H2D, kernel, D2H overlap).  Y / RR are either species-major blocks (Y, ldY) or tables of N separately
// allocated arrays (Yptrs: the PtrList<volScalarField> of dfChemistryModel.H:105,121 as it is, no repack).
int solve_host(dfchem_ctx* h, int nC, const double* T, const double* p, const double* rho, const double* dt,
               double dt_uniform, const double* Y, int ldY, const double* const* Yptrs, double* RR, int ldRR,
               double* const* RRptrs, double* Qdot, double* cost, int* status) {
    CUDA_TRY(h, cudaSetDevice(h->device));
    begin_solve_call(h, nC);
    const int N = h->mech.N;
    const int chunk = 1 << 18;
    int rc = DFCHEM_OK;
    // per-slot device staging: in = T,p,rho,dt,dtchem,Y[N];  out = RR[N],Qdot,cost,status
    for (int i = 0, c0 = 0; c0 < nC && rc == DFCHEM_OK; ++i, c0 += chunk) {
        const int n = std::min(chunk, nC - c0);
        const int sl = i % kSlots;
        cudaStream_t s = h->streams[sl];
        const size_t ld = (size_t)chunk;
        auto body = [&]() -> int {
            if (h->call_n + 1 >= (unsigned long long)kRing) {  // ring about to wrap inside one call: fold on the host
                for (int q = 0; q < kSlots; ++q) CUDA_TRY(h, cudaStreamSynchronize(h->streams[q]));
                unsigned long long part[3];
                int r = read_call_counters(h, part);
                if (r != DFCHEM_OK) return r;
                for (int k = 0; k < 3; ++k) h->call_acc[k] = part[k];
                h->call_first = h->seq;
                h->call_n = 0;
            }
            CUDA_TRY(h, h->slot_in[sl].ensure((size_t)(N + 5) * ld * sizeof(double)));
            CUDA_TRY(h, h->slot_out[sl].ensure((size_t)(N + 3) * ld * sizeof(double)));
            double* din = static_cast<double*>(h->slot_in[sl].p);
            double* dout = static_cast<double*>(h->slot_out[sl].p);
            double *dT = din, *dp = din + ld, *drho = din + 2 * ld, *ddt = din + 3 * ld, *dtc = din + 4 * ld, *dY = din + 5 * ld;
            double *dRR = dout, *dQ = dout + (size_t)N * ld, *dcost = dQ + ld;
            int* dstat = reinterpret_cast<int*>(dcost + ld);
            const size_t nb = (size_t)n * sizeof(double);
            CUDA_TRY(h, cudaMemcpyAsync(dT, T + c0, nb, cudaMemcpyHostToDevice, s));
            CUDA_TRY(h, cudaMemcpyAsync(dp, p + c0, nb, cudaMemcpyHostToDevice, s));
            CUDA_TRY(h, cudaMemcpyAsync(drho, rho + c0, nb, cudaMemcpyHostToDevice, s));
            if (dt) CUDA_TRY(h, cudaMemcpyAsync(ddt, dt + c0, nb, cudaMemcpyHostToDevice, s));
            if (h->dtchem) CUDA_TRY(h, cudaMemcpyAsync(dtc, h->dtchem + c0, nb, cudaMemcpyHostToDevice, s));
            if (Yptrs) {
                for (int k = 0; k < N; ++k)
                    CUDA_TRY(h, cudaMemcpyAsync(dY + (size_t)k * ld, Yptrs[k] + c0, nb, cudaMemcpyHostToDevice, s));
            } else {
                CUDA_TRY(h, cudaMemcpy2DAsync(dY, ld * sizeof(double), Y + c0, (size_t)ldY * sizeof(double), nb, N,
                                              cudaMemcpyHostToDevice, s));
            }
            SolveArgs a{n, dT, dp, drho, dt ? ddt : nullptr, dt_uniform, dY, (int)ld, dRR, (int)ld, dQ,
                        cost ? dcost : nullptr, status ? dstat : nullptr, h->rtol, h->atol, h->max_steps,
                        h->dtchem ? dtc : nullptr};
            int r = launch_solve(h, a, s);
            if (r != DFCHEM_OK) return r;
            if (RRptrs) {
                for (int k = 0; k < N; ++k)
                    CUDA_TRY(h, cudaMemcpyAsync(RRptrs[k] + c0, dRR + (size_t)k * ld, nb, cudaMemcpyDeviceToHost, s));
            } else {
                CUDA_TRY(h, cudaMemcpy2DAsync(RR + c0, (size_t)ldRR * sizeof(double), dRR, ld * sizeof(double), nb, N,
                                              cudaMemcpyDeviceToHost, s));
            }
            CUDA_TRY(h, cudaMemcpyAsync(Qdot + c0, dQ, nb, cudaMemcpyDeviceToHost, s));
            if (cost) CUDA_TRY(h, cudaMemcpyAsync(cost + c0, dcost, nb, cudaMemcpyDeviceToHost, s));
            if (status) CUDA_TRY(h, cudaMemcpyAsync(status + c0, dstat, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost, s));
            if (h->dtchem) CUDA_TRY(h, cudaMemcpyAsync(h->dtchem + c0, dtc, nb, cudaMemcpyDeviceToHost, s));
            return DFCHEM_OK;
        };
        rc = body();
    }
    // both streams are drained before returning, on the error path as well: the caller's buffers and the
    // staging slots must not be in use by an outstanding copy once this function has returned
    std::string first_err = h->err;
    for (int sl = 0; sl < kSlots; ++sl) {
        cudaError_t e = cudaStreamSynchronize(h->streams[sl]);
        if (e != cudaSuccess && rc == DFCHEM_OK) rc = fail(h, DFCHEM_ECUDA, std::string("cudaStreamSynchronize: ") + cudaGetErrorString(e));
    }
    if (rc != DFCHEM_OK) {
        if (!first_err.empty() && rc != DFCHEM_ECUDA) h->err = first_err;
        return rc;
    }
    // the reference treats an integrator failure as fatal (CVODE errors propagate out of solveSingle,
    // dfChemistryModel.C:738-780): report it even when the caller passed no status array
    unsigned long long cnt[3];
    rc = read_call_counters(h, cnt);
    if (rc != DFCHEM_OK) return rc;
    if (cnt[2] != 0)
        return fail(h, DFCHEM_ECELLFAIL, "dfchem_solve: the integrator failed in " + std::to_string(cnt[2]) +
                                             " cell(s) (max steps or step-size underflow); RR/Qdot of those cells are partial");
    return DFCHEM_OK;
}
}  // namespace

int dfchem_solve(dfchem_handle h, int nC, const double* T, const double* p, const double* rho,
                 const double* dt, double dt_uniform, const double* Y, int ldY, double* RR, int ldRR,
                 double* Qdot, double* cost, int* status) {
    if (!h) return DFCHEM_EINVAL;
    if (nC < 0 || (nC > 0 && (!T || !p || !rho || !Y || !RR || !Qdot)) || ldY < nC || ldRR < nC)
        return fail(h, DFCHEM_EINVAL, "dfchem_solve: bad argument");
    return solve_host(h, nC, T, p, rho, dt, dt_uniform, Y, ldY, nullptr, RR, ldRR, nullptr, Qdot, cost, status);
}

int dfchem_solve_ptrs(dfchem_handle h, int nC, const double* T, const double* p, const double* rho,
                      const double* dt, double dt_uniform, const double* const* Yptrs, double* const* RRptrs,
                      double* Qdot, double* cost, int* status) {
    if (!h) return DFCHEM_EINVAL;
    bool ok = nC >= 0 && (nC == 0 || (T && p && rho && Yptrs && RRptrs && Qdot));
    if (ok && nC > 0)
        for (int k = 0; k < h->mech.N; ++k) ok = ok && Yptrs[k] && RRptrs[k];
    if (!ok) return fail(h, DFCHEM_EINVAL, "dfchem_solve_ptrs: bad argument");
    return solve_host(h, nC, T, p, rho, dt, dt_uniform, nullptr, 0, Yptrs, nullptr, 0, RRptrs, Qdot, cost, status);
}

int dfchem_last_counters(dfchem_handle h, unsigned long long out[4]) {
    if (!h || !out) return DFCHEM_EINVAL;
    CUDA_TRY(h, cudaSetDevice(h->device));
    int rc = read_call_counters(h, out);
    if (rc != DFCHEM_OK) return rc;
    out[3] = h->launches;
    return DFCHEM_OK;
}

int dfchem_wdot_dev(dfchem_handle h, int nC, const double* T, const double* p, const double* Y, int ldY,
                    double* wrate, int ldW, void* stream) {
    if (!h) return DFCHEM_EINVAL;
    if (nC < 0 || (nC > 0 && (!T || !p || !Y || !wrate)) || ldY < nC || ldW < nC)
        return fail(h, DFCHEM_EINVAL, "dfchem_wdot_dev: bad argument");
    if (nC == 0) return DFCHEM_OK;
    CUDA_TRY(h, cudaSetDevice(h->device));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    unsigned int* queue = nullptr;
    {
        int rc = take_slot(h, s, &queue, nullptr);
        if (rc != DFCHEM_OK) return rc;
    }
#define CALL(NP, NS) launch_wdot_t<NP, NS>(h, nC, T, p, Y, ldY, wrate, ldW, queue, s)
    DFC_DISPATCH_NP(h->dm.NP, CALL)
#undef CALL
    CUDA_TRY(h, cudaGetLastError());
    return DFCHEM_OK;
}

int dfchem_wdot(dfchem_handle h, int nC, const double* T, const double* p, const double* Y, int ldY,
                double* wrate, int ldW) {
    if (!h) return DFCHEM_EINVAL;
    if (nC < 0 || (nC > 0 && (!T || !p || !Y || !wrate)) || ldY < nC || ldW < nC)
        return fail(h, DFCHEM_EINVAL, "dfchem_wdot: bad argument");
    if (nC == 0) return DFCHEM_OK;
    CUDA_TRY(h, cudaSetDevice(h->device));
    const int N = h->mech.N;
    const size_t ld = (size_t)nC, nb = ld * sizeof(double);
    CUDA_TRY(h, h->scratch.ensure((2 * (size_t)N + 2) * nb));
    double* d = static_cast<double*>(h->scratch.p);
    double *dT = d, *dp = d + ld, *dY = d + 2 * ld, *dW = dY + (size_t)N * ld;
    cudaStream_t s = h->streams[0];
    CUDA_TRY(h, cudaMemcpyAsync(dT, T, nb, cudaMemcpyHostToDevice, s));
    CUDA_TRY(h, cudaMemcpyAsync(dp, p, nb, cudaMemcpyHostToDevice, s));
    CUDA_TRY(h, cudaMemcpy2DAsync(dY, nb, Y, (size_t)ldY * sizeof(double), nb, N, cudaMemcpyHostToDevice, s));
    int rc = dfchem_wdot_dev(h, nC, dT, dp, dY, nC, dW, nC, s);
    if (rc != DFCHEM_OK) return rc;
    CUDA_TRY(h, cudaMemcpy2DAsync(wrate, (size_t)ldW * sizeof(double), dW, nb, nb, N, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(h, cudaStreamSynchronize(s));
    return DFCHEM_OK;
}

// ---- dfchem_wdot without any device allocation (no counterpart in the reference; include/dfchem.h) ----------------
// The scratch of wdot_kernel (work-queue word) and the device staging of the host arrays live in one static device
// array.  In host code the NAME of that array is only a host-side stand-in: the device address comes from
// cudaGetSymbolAddress, once per device, with that device current.  First 256 bytes: queue word; rest: staging.
constexpr size_t kStaticHead = 256;
static __device__ __align__(256) unsigned char g_static_scratch[kStaticHead + DFCHEM_STATIC_SCRATCH_BYTES];

namespace {
constexpr int kStaticMaxDevices = 64;
std::mutex g_static_mutex;
unsigned char* g_static_ptr[kStaticMaxDevices] = {};

// device address of g_static_scratch on `device` (which must be current)
cudaError_t static_scratch_address(int device, unsigned char** out) {
    if (device < 0 || device >= kStaticMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_static_mutex);
    if (!g_static_ptr[device]) {
        void* ptr = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&ptr, g_static_scratch);
        if (e != cudaSuccess) return e;
        g_static_ptr[device] = static_cast<unsigned char*>(ptr);
    }
    *out = g_static_ptr[device];
    return cudaSuccess;
}
}  // namespace

int dfchem_wdot_static(dfchem_handle h, int nC, const double* T, const double* p, const double* Y, int ldY,
                       double* wrate, int ldW) {
    if (!h) return DFCHEM_EINVAL;
    if (nC < 0 || (nC > 0 && (!T || !p || !Y || !wrate)) || ldY < nC || ldW < nC)
        return fail(h, DFCHEM_EINVAL, "dfchem_wdot_static: bad argument");
    if (nC == 0) return DFCHEM_OK;
    const int N = h->mech.N;
    const size_t per_cell = (2 * (size_t)N + 2) * sizeof(double);
    if ((size_t)nC * per_cell > (size_t)DFCHEM_STATIC_SCRATCH_BYTES)
        return fail(h, DFCHEM_EINVAL,
                    "dfchem_wdot_static: " + std::to_string(nC) + " cells need " + std::to_string((size_t)nC * per_cell) +
                        " bytes of device staging, the static scratch holds " + std::to_string((size_t)DFCHEM_STATIC_SCRATCH_BYTES) +
                        " (at most " + std::to_string((size_t)DFCHEM_STATIC_SCRATCH_BYTES / per_cell) +
                        " cells per call for this mechanism); split the call or use dfchem_wdot");
    CUDA_TRY(h, cudaSetDevice(h->device));
    unsigned char* base = nullptr;
    CUDA_TRY(h, static_scratch_address(h->device, &base));
    unsigned int* queue = reinterpret_cast<unsigned int*>(base);
    double* d = reinterpret_cast<double*>(base + kStaticHead);
    const size_t ld = (size_t)nC, nb = ld * sizeof(double);
    double *dT = d, *dp = d + ld, *dY = d + 2 * ld, *dW = dY + (size_t)N * ld;
    cudaStream_t s = h->streams[0];
    // the array keeps the previous call's contents: the queue word must start at zero
    CUDA_TRY(h, cudaMemsetAsync(queue, 0, sizeof(unsigned int), s));
    CUDA_TRY(h, cudaMemcpyAsync(dT, T, nb, cudaMemcpyHostToDevice, s));
    CUDA_TRY(h, cudaMemcpyAsync(dp, p, nb, cudaMemcpyHostToDevice, s));
    CUDA_TRY(h, cudaMemcpy2DAsync(dY, nb, Y, (size_t)ldY * sizeof(double), nb, N, cudaMemcpyHostToDevice, s));
#define CALL(NP, NS) launch_wdot_t<NP, NS>(h, nC, dT, dp, dY, nC, dW, nC, queue, s)
    DFC_DISPATCH_NP(h->dm.NP, CALL)
#undef CALL
    CUDA_TRY(h, cudaGetLastError());
    CUDA_TRY(h, cudaMemcpy2DAsync(wrate, (size_t)ldW * sizeof(double), dW, nb, nb, N, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(h, cudaStreamSynchronize(s));
    return DFCHEM_OK;
}

int dfchem_correct_thermo_dev(dfchem_handle h, int n, const double* he, const double* p, const double* Y,
                              int ldY, double* T, double* psi, double* rho, double* cp, double* hai,
                              int ldH, void* stream) {
    if (!h) return DFCHEM_EINVAL;
    if (n < 0 || (n > 0 && (!he || !p || !Y || !T || !psi || !rho)) || ldY < n || (hai && ldH < n))
        return fail(h, DFCHEM_EINVAL, "dfchem_correct_thermo_dev: bad argument");
    if (n == 0) return DFCHEM_OK;
    CUDA_TRY(h, cudaSetDevice(h->device));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int N = h->mech.N;
    const size_t tsm = thermo_smem_bytes(N);
    if (tsm > 48 * 1024) return fail(h, DFCHEM_EUNSUPPORTED, "thermo kernel: species tables exceed 48 KB of shared memory");
    // Kernel choice (DFCHEM_THERMO = ldg | pair | tma; measured on B200, profiles/r2/thermo_variants.json, gri30 4M cells):
    //   ldg   one cell per thread, 8-byte loads, 32 warps/SM                          0.40 ms  (default)
    //   pair  two cells per thread, 16-byte loads, a third fewer instructions per cell 0.57 ms  (16 warps/SM at 128 registers)
    //   tma   species rows staged by the bulk-copy engine (cp.async.bulk + mbarrier)    0.60 ms  (8 warps/SM: one thread per
    //         cell and two 78 KB stages)
    // The kernel is latency-bound per warp; resident warps with loads in flight beat both fewer instructions and
    // copy-engine staging.  The two experiments stay selectable so that the comparison can be re-run.
    static const std::string variant = [] { const char* e = std::getenv("DFCHEM_THERMO"); return std::string(e ? e : "ldg"); }();
    int done = 0;
    auto aligned = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15u) == 0; };
    const bool rows_aligned = (ldY % 2) == 0 && aligned(Y) && aligned(he) && aligned(p) && aligned(T) && aligned(psi) && aligned(rho) &&
                              (!cp || aligned(cp)) && (!hai || (aligned(hai) && (ldH % 2) == 0));
    if (variant == "tma" && n >= kTmaTile && rows_aligned) {
        if (h->thermo_tma_ok == 0) {
            int max_optin = 0;
            CUDA_TRY(h, cudaDeviceGetAttribute(&max_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, h->device));
            h->thermo_tma_ok = thermo_tma_smem_bytes(N) <= (size_t)max_optin ? 1 : -1;
            if (h->thermo_tma_ok == 1)
                CUDA_TRY(h, cudaFuncSetAttribute(thermo_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, max_optin));
        }
        if (h->thermo_tma_ok == 1) {
            const int ntiles = n / kTmaTile;
            thermo_tma_kernel<<<std::min(ntiles, h->sm_count), kTmaTile, thermo_tma_smem_bytes(N), s>>>(
                h->dm, ntiles, he, p, Y, ldY, T, psi, rho, cp, hai, ldH);
            CUDA_TRY(h, cudaGetLastError());
            done = ntiles * kTmaTile;
        }
    } else if (variant == "pair" && n >= 2 && rows_aligned) {
        const int npairs = n / 2;
        // register budget / load batch of the pair kernel: (3 CTAs per SM, 6 loads) unless DFCHEM_THERMO_CFG=<minb><batch> says otherwise
        static const int cfg = [] { const char* e = std::getenv("DFCHEM_THERMO_CFG"); return e ? std::atoi(e) : 36; }();
        auto run = [&](auto kern) -> int {
            if (h->thermo_pair_ctas == 0) {
                int nb = 0;
                CUDA_TRY(h, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kern, 256, tsm));
                h->thermo_pair_ctas = std::max(1, nb);
            }
            static const int waves_env = [] { const char* e = std::getenv("DFCHEM_THERMO_WAVES"); return e ? std::max(1, std::atoi(e)) : 2; }();
            const int grid = std::min((npairs + 255) / 256, h->sm_count * h->thermo_pair_ctas * waves_env);
            kern<<<grid, 256, tsm, s>>>(h->dm, npairs, he, p, Y, ldY, T, psi, rho, cp, hai, ldH);
            CUDA_TRY(h, cudaGetLastError());
            return DFCHEM_OK;
        };
        int rc = DFCHEM_OK;
        switch (cfg) {
            case 26: rc = run(thermo_pair_kernel<2, 6>); break;
            case 28: rc = run(thermo_pair_kernel<2, 8>); break;
            case 34: rc = run(thermo_pair_kernel<3, 4>); break;
            case 44: rc = run(thermo_pair_kernel<4, 4>); break;
            case 43: rc = run(thermo_pair_kernel<4, 3>); break;
            default: rc = run(thermo_pair_kernel<3, 6>); break;
        }
        if (rc != DFCHEM_OK) return rc;
        done = 2 * npairs;
    }
    if (done < n) {
        const int nr = n - done;
        if (h->thermo_ctas == 0) {  // persistent CTAs: two waves of what is resident (the measured configuration)
            int nb = 0;
            CUDA_TRY(h, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, thermo_kernel, 256, tsm));
            h->thermo_ctas = std::max(1, nb);
        }
        static const int waves_env = [] { const char* e = std::getenv("DFCHEM_THERMO_WAVES"); return e ? std::max(1, std::atoi(e)) : 2; }();
        const int grid = std::min((nr + 255) / 256, h->sm_count * h->thermo_ctas * waves_env);
        thermo_kernel<<<grid, 256, tsm, s>>>(h->dm, nr, he + done, p + done, Y + done, ldY, T + done, psi + done, rho + done,
                                             cp ? cp + done : nullptr, hai ? hai + done : nullptr, ldH);
        CUDA_TRY(h, cudaGetLastError());
    }
    return DFCHEM_OK;
}

int dfchem_correct_thermo(dfchem_handle h, int n, const double* he, const double* p, const double* Y,
                          int ldY, double* T, double* psi, double* rho, double* cp, double* hai, int ldH) {
    if (!h) return DFCHEM_EINVAL;
    if (n < 0 || (n > 0 && (!he || !p || !Y || !T || !psi || !rho)) || ldY < n || (hai && ldH < n))
        return fail(h, DFCHEM_EINVAL, "dfchem_correct_thermo: bad argument");
    if (n == 0) return DFCHEM_OK;
    CUDA_TRY(h, cudaSetDevice(h->device));
    const int N = h->mech.N;
    const size_t ld = (size_t)n, nb = ld * sizeof(double);
    CUDA_TRY(h, h->scratch.ensure((2 * (size_t)N + 6) * nb));
    double* d = static_cast<double*>(h->scratch.p);
    double *dhe = d, *dp = d + ld, *dT = d + 2 * ld, *dpsi = d + 3 * ld, *drho = d + 4 * ld, *dcp = d + 5 * ld;
    double *dY = d + 6 * ld, *dH = dY + (size_t)N * ld;
    cudaStream_t s = h->streams[0];
    CUDA_TRY(h, cudaMemcpyAsync(dhe, he, nb, cudaMemcpyHostToDevice, s));
    CUDA_TRY(h, cudaMemcpyAsync(dp, p, nb, cudaMemcpyHostToDevice, s));
    CUDA_TRY(h, cudaMemcpyAsync(dT, T, nb, cudaMemcpyHostToDevice, s));
    CUDA_TRY(h, cudaMemcpy2DAsync(dY, nb, Y, (size_t)ldY * sizeof(double), nb, N, cudaMemcpyHostToDevice, s));
    int rc = dfchem_correct_thermo_dev(h, n, dhe, dp, dY, n, dT, dpsi, drho, cp ? dcp : nullptr,
                                       hai ? dH : nullptr, n, s);
    if (rc != DFCHEM_OK) return rc;
    CUDA_TRY(h, cudaMemcpyAsync(T, dT, nb, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(h, cudaMemcpyAsync(psi, dpsi, nb, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(h, cudaMemcpyAsync(rho, drho, nb, cudaMemcpyDeviceToHost, s));
    if (cp) CUDA_TRY(h, cudaMemcpyAsync(cp, dcp, nb, cudaMemcpyDeviceToHost, s));
    if (hai)
        CUDA_TRY(h, cudaMemcpy2DAsync(hai, (size_t)ldH * sizeof(double), dH, nb, nb, N, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(h, cudaStreamSynchronize(s));
    return DFCHEM_OK;
}

int dfchem_h_from_T_dev(dfchem_handle h, int n, const double* T, const double* Y, int ldY, double* he,
                        void* stream) {
    if (!h) return DFCHEM_EINVAL;
    if (n < 0 || (n > 0 && (!T || !Y || !he)) || ldY < n) return fail(h, DFCHEM_EINVAL, "dfchem_h_from_T_dev: bad argument");
    if (n == 0) return DFCHEM_OK;
    CUDA_TRY(h, cudaSetDevice(h->device));
    enthalpy_kernel<<<std::min((n + 255) / 256, h->sm_count * 8), 256, thermo_smem_bytes(h->mech.N), static_cast<cudaStream_t>(stream)>>>(
        h->dm, n, T, Y, ldY, he);
    CUDA_TRY(h, cudaGetLastError());
    return DFCHEM_OK;
}

int dfchem_h_from_T(dfchem_handle h, int n, const double* T, const double* Y, int ldY, double* he) {
    if (!h) return DFCHEM_EINVAL;
    if (n < 0 || (n > 0 && (!T || !Y || !he)) || ldY < n) return fail(h, DFCHEM_EINVAL, "dfchem_h_from_T: bad argument");
    if (n == 0) return DFCHEM_OK;
    CUDA_TRY(h, cudaSetDevice(h->device));
    const int N = h->mech.N;
    const size_t ld = (size_t)n, nb = ld * sizeof(double);
    CUDA_TRY(h, h->scratch.ensure(((size_t)N + 2) * nb));
    double* d = static_cast<double*>(h->scratch.p);
    cudaStream_t s = h->streams[0];
    CUDA_TRY(h, cudaMemcpyAsync(d, T, nb, cudaMemcpyHostToDevice, s));
    CUDA_TRY(h, cudaMemcpy2DAsync(d + 2 * ld, nb, Y, (size_t)ldY * sizeof(double), nb, N, cudaMemcpyHostToDevice, s));
    int rc = dfchem_h_from_T_dev(h, n, d, d + 2 * ld, n, d + ld, s);
    if (rc != DFCHEM_OK) return rc;
    CUDA_TRY(h, cudaMemcpyAsync(he, d + ld, nb, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(h, cudaStreamSynchronize(s));
    return DFCHEM_OK;
}

// ---- rates of progress (calculateRR, dfChemistryModel.C:925-987; PaSR.C:326-385) ------------------
namespace {
int rop_dev_impl(dfchem_ctx* h, const char* who, int nC, const double* T, const double* p, const double* Y, int ldY,
                 double* q, double* qf, double* qr, int ldQ, const double* rho, double* tc, cudaStream_t s) {
    if (nC < 0 || (nC > 0 && (!T || !p || !Y)) || ldY < nC || ((q || qf || qr) && ldQ < nC) || (tc && !rho))
        return fail(h, DFCHEM_EINVAL, std::string(who) + ": bad argument");
    if (nC == 0 || h->mech.R == 0) return DFCHEM_OK;
    CUDA_TRY(h, cudaSetDevice(h->device));
    unsigned int* queue = nullptr;
    int rc = take_slot(h, s, &queue, nullptr);
    if (rc != DFCHEM_OK) return rc;
#define CALL(NP, NS) launch_rop_t<NP, NS>(h, nC, T, p, Y, ldY, q, qf, qr, ldQ, rho, tc, queue, s)
    DFC_DISPATCH_NP(h->dm.NP, CALL)
#undef CALL
    CUDA_TRY(h, cudaGetLastError());
    return DFCHEM_OK;
}

// host-pointer wrapper: stages T, p, rho, Y in; up to three [R][nC] blocks and tc out
int rop_host_impl(dfchem_ctx* h, const char* who, int nC, const double* T, const double* p, const double* rho,
                  const double* Y, int ldY, double* q, double* qf, double* qr, int ldQ, double* tc) {
    if (nC < 0 || (nC > 0 && (!T || !p || !Y)) || ldY < nC || ((q || qf || qr) && ldQ < nC) || (tc && !rho))
        return fail(h, DFCHEM_EINVAL, std::string(who) + ": bad argument");
    if (nC == 0 || h->mech.R == 0) return DFCHEM_OK;
    CUDA_TRY(h, cudaSetDevice(h->device));
    const int N = h->mech.N, R = h->mech.R;
    const size_t ld = (size_t)nC, nb = ld * sizeof(double);
    CUDA_TRY(h, h->scratch.ensure(((size_t)N + 3 * (size_t)R + 4) * nb));
    double* d = static_cast<double*>(h->scratch.p);
    double *dT = d, *dp = d + ld, *drho = d + 2 * ld, *dtc = d + 3 * ld, *dY = d + 4 * ld, *dQ = dY + (size_t)N * ld,
           *dQf = dQ + (size_t)R * ld, *dQr = dQf + (size_t)R * ld;
    cudaStream_t s = h->streams[0];
    CUDA_TRY(h, cudaMemcpyAsync(dT, T, nb, cudaMemcpyHostToDevice, s));
    CUDA_TRY(h, cudaMemcpyAsync(dp, p, nb, cudaMemcpyHostToDevice, s));
    if (rho) CUDA_TRY(h, cudaMemcpyAsync(drho, rho, nb, cudaMemcpyHostToDevice, s));
    CUDA_TRY(h, cudaMemcpy2DAsync(dY, nb, Y, (size_t)ldY * sizeof(double), nb, N, cudaMemcpyHostToDevice, s));
    int rc = rop_dev_impl(h, who, nC, dT, dp, dY, nC, q ? dQ : nullptr, qf ? dQf : nullptr, qr ? dQr : nullptr, nC,
                          rho ? drho : nullptr, tc ? dtc : nullptr, s);
    if (rc == DFCHEM_OK) {
        if (q) CUDA_TRY(h, cudaMemcpy2DAsync(q, (size_t)ldQ * sizeof(double), dQ, nb, nb, R, cudaMemcpyDeviceToHost, s));
        if (qf) CUDA_TRY(h, cudaMemcpy2DAsync(qf, (size_t)ldQ * sizeof(double), dQf, nb, nb, R, cudaMemcpyDeviceToHost, s));
        if (qr) CUDA_TRY(h, cudaMemcpy2DAsync(qr, (size_t)ldQ * sizeof(double), dQr, nb, nb, R, cudaMemcpyDeviceToHost, s));
        if (tc) CUDA_TRY(h, cudaMemcpyAsync(tc, dtc, nb, cudaMemcpyDeviceToHost, s));
    }
    CUDA_TRY(h, cudaStreamSynchronize(s));
    return rc;
}
}  // namespace

int dfchem_rop_dev(dfchem_handle h, int nC, const double* T, const double* p, const double* Y, int ldY,
                   double* q, int ldQ, void* stream) {
    if (!h) return DFCHEM_EINVAL;
    if (nC > 0 && !q) return fail(h, DFCHEM_EINVAL, "dfchem_rop_dev: bad argument");
    return rop_dev_impl(h, "dfchem_rop_dev", nC, T, p, Y, ldY, q, nullptr, nullptr, ldQ, nullptr, nullptr,
                        static_cast<cudaStream_t>(stream));
}
int dfchem_rop(dfchem_handle h, int nC, const double* T, const double* p, const double* Y, int ldY,
               double* q, int ldQ) {
    if (!h) return DFCHEM_EINVAL;
    if (nC > 0 && !q) return fail(h, DFCHEM_EINVAL, "dfchem_rop: bad argument");
    return rop_host_impl(h, "dfchem_rop", nC, T, p, nullptr, Y, ldY, q, nullptr, nullptr, ldQ, nullptr);
}
int dfchem_rop_fwd_rev_dev(dfchem_handle h, int nC, const double* T, const double* p, const double* Y, int ldY,
                           double* qf, double* qr, int ldQ, void* stream) {
    if (!h) return DFCHEM_EINVAL;
    if (nC > 0 && (!qf || !qr)) return fail(h, DFCHEM_EINVAL, "dfchem_rop_fwd_rev_dev: bad argument");
    return rop_dev_impl(h, "dfchem_rop_fwd_rev_dev", nC, T, p, Y, ldY, nullptr, qf, qr, ldQ, nullptr, nullptr,
                        static_cast<cudaStream_t>(stream));
}
int dfchem_rop_fwd_rev(dfchem_handle h, int nC, const double* T, const double* p, const double* Y, int ldY,
                       double* qf, double* qr, int ldQ) {
    if (!h) return DFCHEM_EINVAL;
    if (nC > 0 && (!qf || !qr)) return fail(h, DFCHEM_EINVAL, "dfchem_rop_fwd_rev: bad argument");
    return rop_host_impl(h, "dfchem_rop_fwd_rev", nC, T, p, nullptr, Y, ldY, nullptr, qf, qr, ldQ, nullptr);
}
int dfchem_tc_dev(dfchem_handle h, int nC, const double* T, const double* p, const double* rho, const double* Y,
                  int ldY, double* tc, void* stream) {
    if (!h) return DFCHEM_EINVAL;
    if (nC > 0 && (!tc || !rho)) return fail(h, DFCHEM_EINVAL, "dfchem_tc_dev: bad argument");
    return rop_dev_impl(h, "dfchem_tc_dev", nC, T, p, Y, ldY, nullptr, nullptr, nullptr, nC, rho, tc,
                        static_cast<cudaStream_t>(stream));
}
int dfchem_tc(dfchem_handle h, int nC, const double* T, const double* p, const double* rho, const double* Y, int ldY,
              double* tc) {
    if (!h) return DFCHEM_EINVAL;
    if (nC > 0 && (!tc || !rho)) return fail(h, DFCHEM_EINVAL, "dfchem_tc: bad argument");
    return rop_host_impl(h, "dfchem_tc", nC, T, p, rho, Y, ldY, nullptr, nullptr, nullptr, nC, tc);
}

int dfchem_reaction_stoich(dfchem_handle h, int reaction, double* nu_net) {
    if (!h || !nu_net || reaction < 0 || reaction >= h->mech.R) return fail(h, DFCHEM_EINVAL, "dfchem_reaction_stoich: bad argument");
    const Mechanism& m = h->mech;
    std::fill(nu_net, nu_net + m.N, 0.0);
    for (int a = 0; a < kMaxSide; ++a) {
        const int ri = m.r_idx[reaction * kMaxSide + a], pi = m.p_idx[reaction * kMaxSide + a];
        if (ri >= 0) nu_net[ri] -= m.r_nu[reaction * kMaxSide + a];
        if (pi >= 0) nu_net[pi] += m.p_nu[reaction * kMaxSide + a];
    }
    return DFCHEM_OK;
}

// ---- transport (correctThermo mu / alpha / rhoD, dfChemistryModel.C:513-541) -----------------------
int dfchem_set_transport_fits(dfchem_handle h, const double* visc, const double* cond, const double* diff) {
    if (!h || !visc || !cond || !diff) return fail(h, DFCHEM_EINVAL, "dfchem_set_transport_fits: bad argument");
    CUDA_TRY(h, cudaSetDevice(h->device));
    const int N = h->mech.N;
    const std::vector<double>& W = h->mech.W;
    const size_t nW = (size_t)N, nV = (size_t)N * kDiffStride, nK = (size_t)N * N * 2, nD = (size_t)N * N * kDiffStride;
    const size_t padW = nW + (nW & 1);  // keeps the fit rows 16-byte aligned
    std::vector<double> t(padW + 2 * nV + nK + nD, 0.0);
    double *tW = t.data(), *tv = tW + padW, *tc = tv + nV, *tk = tc + nV, *td = tk + nK;
    for (int i = 0; i < N; ++i) {
        tW[i] = W[i];
        for (int k = 0; k < 5; ++k) {
            tv[i * kDiffStride + k] = visc[i * 5 + k];
            tc[i * kDiffStride + k] = cond[i * 5 + k];
        }
        for (int j = 0; j < N; ++j) {
            // dfThermo.cu:40-43 (viscosity_conatant1/2), 1/sqrt(8) folded into the first
            tk[((size_t)i * N + j) * 2 + 0] = std::pow(1.0 + W[i] / W[j], -0.5) / 2.8284271247461903;
            tk[((size_t)i * N + j) * 2 + 1] = std::pow(W[j] / W[i], 0.25);
            for (int k = 0; k < 5; ++k) td[((size_t)i * N + j) * kDiffStride + k] = diff[((size_t)i * N + j) * 5 + k];
        }
    }
    CUDA_TRY(h, h->d_transport.ensure(t.size() * sizeof(double)));
    CUDA_TRY(h, cudaMemcpy(h->d_transport.p, t.data(), t.size() * sizeof(double), cudaMemcpyHostToDevice));
    const double* base = static_cast<const double*>(h->d_transport.p);
    h->tt.N = N;
    h->tt.W = base;
    h->tt.visc = base + (tv - tW);
    h->tt.cond = base + (tc - tW);
    h->tt.wilke = base + (tk - tW);
    h->tt.diff = base + (td - tW);
    // per-function attribute shared by every handle: always the device maximum, never this handle's size
    int max_optin = 0;
    CUDA_TRY(h, cudaDeviceGetAttribute(&max_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, h->device));
    if (transport_smem_bytes(N) > (size_t)max_optin)
        return fail(h, DFCHEM_EUNSUPPORTED, "transport kernel: species columns exceed the shared memory of one CTA");
    CUDA_TRY(h, cudaFuncSetAttribute(transport_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, max_optin));
    return DFCHEM_OK;
}

// Fits generated by the mechanism compiler from the species' Lennard-Jones data (transport_fit.cpp); optionally also
// returned to the caller (visc[N*5], cond[N*5], diff[N*N*5], any may be NULL).
int dfchem_generate_transport_fits(dfchem_handle h, double* visc_out, double* cond_out, double* diff_out) {
    if (!h) return DFCHEM_EINVAL;
    std::vector<double> visc, cond, diff;
    if (!generate_transport_fits(h->mech, visc, cond, diff))
        return fail(h, DFCHEM_EMECH, "dfchem_generate_transport_fits: the mechanism has no complete `transport:` data");
    if (visc_out) std::copy(visc.begin(), visc.end(), visc_out);
    if (cond_out) std::copy(cond.begin(), cond.end(), cond_out);
    if (diff_out) std::copy(diff.begin(), diff.end(), diff_out);
    return dfchem_set_transport_fits(h, visc.data(), cond.data(), diff.data());
}

// thermo_<mech>.txt of the reference (src_gpu/dfThermo.cu:361-435): int N; double W[N];
// N x 15 NASA [Tmid, high7, low7]; N x 5 viscosity; N x 5 conductivity; N x N x 5 binary diffusion
int dfchem_load_transport_file(dfchem_handle h, const char* path) {
    if (!h || !path) return fail(h, DFCHEM_EINVAL, "dfchem_load_transport_file: bad argument");
    FILE* fp = std::fopen(path, "rb");
    if (!fp) return fail(h, DFCHEM_EMECH, std::string("cannot open transport coefficient file ") + path);
    const int N = h->mech.N;
    int n_file = 0;
    std::vector<double> W(N), nasa((size_t)N * 15), visc((size_t)N * 5), cond((size_t)N * 5), diff((size_t)N * N * 5);
    bool ok = std::fread(&n_file, sizeof(int), 1, fp) == 1 && n_file == N;
    ok = ok && std::fread(W.data(), sizeof(double), W.size(), fp) == W.size();
    ok = ok && std::fread(nasa.data(), sizeof(double), nasa.size(), fp) == nasa.size();
    ok = ok && std::fread(visc.data(), sizeof(double), visc.size(), fp) == visc.size();
    ok = ok && std::fread(cond.data(), sizeof(double), cond.size(), fp) == cond.size();
    ok = ok && std::fread(diff.data(), sizeof(double), diff.size(), fp) == diff.size();
    std::fclose(fp);
    if (!ok) return fail(h, DFCHEM_EMECH, "transport coefficient file: species count differs from the mechanism or file truncated");
    for (int i = 0; i < N; ++i)
        if (std::fabs(W[i] - h->mech.W[i]) > 1e-9 * h->mech.W[i])
            return fail(h, DFCHEM_EMECH, "transport coefficient file: molecular weights differ from the mechanism (species order?)");
    return dfchem_set_transport_fits(h, visc.data(), cond.data(), diff.data());
}

int dfchem_transport_dev(dfchem_handle h, int n, const double* T, const double* p, const double* rho,
                         const double* cp, const double* Y, int ldY, double* mu, double* alpha,
                         double* rhoD, int ldD, int unity_lewis, void* stream) {
    if (!h) return DFCHEM_EINVAL;
    const bool need_cp = alpha || (rhoD && unity_lewis), need_rho = rhoD && !unity_lewis;
    if (n < 0 || (n > 0 && (!T || !Y || (need_cp && !cp) || (need_rho && (!rho || !p)))) || ldY < n || (rhoD && ldD < n))
        return fail(h, DFCHEM_EINVAL, "dfchem_transport_dev: bad argument");
    if (h->tt.N != h->mech.N) return fail(h, DFCHEM_EINVAL, "dfchem_transport: no transport fits loaded (dfchem_set_transport_fits)");
    if (n == 0) return DFCHEM_OK;
    CUDA_TRY(h, cudaSetDevice(h->device));
    const size_t smem = transport_smem_bytes(h->mech.N);
    // CTAs per SM: as many as shared memory and registers allow (8).  A sweep (profiles/r1, gri30, 4M cells)
    // of 3/4/5/6/8 CTAs per SM gave 28.9/27.2/23.2/22.2/19.4 ms: the kernel is latency-bound, so resident
    // warps count for more than keeping the N^2 x 64-byte fit tables L1-resident.
    static const int per_sm_env = [] { const char* e = std::getenv("DFCHEM_TR_CTAS"); return e ? std::atoi(e) : 0; }();
    int per_sm = (int)std::max<size_t>(1, std::min<size_t>(8, (200 * 1024) / smem));
    if (per_sm_env > 0) per_sm = per_sm_env;
    const int grid = std::min((n + kTransportCells - 1) / kTransportCells, h->sm_count * per_sm);
    transport_kernel<<<grid, kTransportBlock, smem, static_cast<cudaStream_t>(stream)>>>(h->tt, n, T, p, rho, cp, Y, ldY, mu, alpha,
                                                                                         rhoD, ldD, unity_lewis);
    CUDA_TRY(h, cudaGetLastError());
    return DFCHEM_OK;
}

int dfchem_transport(dfchem_handle h, int n, const double* T, const double* p, const double* rho, const double* cp,
                     const double* Y, int ldY, double* mu, double* alpha, double* rhoD, int ldD, int unity_lewis) {
    if (!h) return DFCHEM_EINVAL;
    const bool need_cp = alpha || (rhoD && unity_lewis), need_rho = rhoD && !unity_lewis;
    if (n < 0 || (n > 0 && (!T || !Y || (need_cp && !cp) || (need_rho && (!rho || !p)))) || ldY < n || (rhoD && ldD < n))
        return fail(h, DFCHEM_EINVAL, "dfchem_transport: bad argument");
    if (n == 0) return DFCHEM_OK;
    CUDA_TRY(h, cudaSetDevice(h->device));
    const int N = h->mech.N;
    const size_t ld = (size_t)n, nb = ld * sizeof(double);
    CUDA_TRY(h, h->scratch.ensure((2 * (size_t)N + 6) * nb));
    double* d = static_cast<double*>(h->scratch.p);
    double *dT = d, *dp = d + ld, *drho = d + 2 * ld, *dcp = d + 3 * ld, *dmu = d + 4 * ld, *dal = d + 5 * ld;
    double *dY = d + 6 * ld, *dD = dY + (size_t)N * ld;
    cudaStream_t s = h->streams[0];
    CUDA_TRY(h, cudaMemcpyAsync(dT, T, nb, cudaMemcpyHostToDevice, s));
    if (p) CUDA_TRY(h, cudaMemcpyAsync(dp, p, nb, cudaMemcpyHostToDevice, s));
    if (rho) CUDA_TRY(h, cudaMemcpyAsync(drho, rho, nb, cudaMemcpyHostToDevice, s));
    if (cp) CUDA_TRY(h, cudaMemcpyAsync(dcp, cp, nb, cudaMemcpyHostToDevice, s));
    CUDA_TRY(h, cudaMemcpy2DAsync(dY, nb, Y, (size_t)ldY * sizeof(double), nb, N, cudaMemcpyHostToDevice, s));
    int rc = dfchem_transport_dev(h, n, dT, p ? dp : nullptr, rho ? drho : nullptr, cp ? dcp : nullptr, dY, n,
                                  mu ? dmu : nullptr, alpha ? dal : nullptr, rhoD ? dD : nullptr, n, unity_lewis, s);
    if (rc != DFCHEM_OK) return rc;
    if (mu) CUDA_TRY(h, cudaMemcpyAsync(mu, dmu, nb, cudaMemcpyDeviceToHost, s));
    if (alpha) CUDA_TRY(h, cudaMemcpyAsync(alpha, dal, nb, cudaMemcpyDeviceToHost, s));
    if (rhoD)
        CUDA_TRY(h, cudaMemcpy2DAsync(rhoD, (size_t)ldD * sizeof(double), dD, nb, nb, N, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(h, cudaStreamSynchronize(s));
    return DFCHEM_OK;
}

// ---- boundary faces of correctThermo (dfChemistryModel.C:546-727; device precedent src_gpu/dfThermo.cu:603-657) -------
int dfchem_correct_thermo_patches_dev(dfchem_handle h, int n_patches, const int* patch_size, const int* patch_type,
                                      int n_boundary, double* b_he, const double* b_p, const double* b_Y, double* b_T,
                                      double* b_psi, double* b_rho, double* b_cp, double* b_mu, double* b_alpha,
                                      double* b_rhoD, int unity_lewis, int n_cells, const int* face_cell,
                                      const double* T, const double* he, const double* psi, const double* rho,
                                      const double* mu, const double* alpha, const double* rhoD, void* stream) {
    if (!h) return DFCHEM_EINVAL;
    if (n_patches < 0 || (n_patches > 0 && (!patch_size || !patch_type)) || n_boundary < 0 || !b_he || !b_p || !b_Y || !b_T ||
        !b_psi || !b_rho || !b_cp)
        return fail(h, DFCHEM_EINVAL, "dfchem_correct_thermo_patches_dev: bad argument");
    const bool transport = b_mu || b_alpha || b_rhoD;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    int offset = 0;
    for (int i = 0; i < n_patches; ++i) {
        const int sz = patch_size[i], ty = patch_type[i];
        if (sz < 0 || ty < 0 || ty > 2) return fail(h, DFCHEM_EINVAL, "dfchem_correct_thermo_patches_dev: bad patch table");
        const int span = ty == DFCHEM_PATCH_PROCESSOR ? 2 * sz : sz;
        if (offset + span > n_boundary) return fail(h, DFCHEM_EINVAL, "dfchem_correct_thermo_patches_dev: patches exceed n_boundary");
        if (sz > 0) {
            int rc;
            if (ty == DFCHEM_PATCH_FIXED_T) {   // T is given: he <- h(T, Y) first (dfChemistryModel.C:572-589); the Newton solve then returns T
                rc = dfchem_h_from_T_dev(h, sz, b_T + offset, b_Y + offset, n_boundary, b_he + offset, stream);
                if (rc != DFCHEM_OK) return rc;
            }
            rc = dfchem_correct_thermo_dev(h, sz, b_he + offset, b_p + offset, b_Y + offset, n_boundary, b_T + offset, b_psi + offset,
                                           b_rho + offset, b_cp + offset, nullptr, n_boundary, stream);
            if (rc != DFCHEM_OK) return rc;
            if (transport) {
                rc = dfchem_transport_dev(h, sz, b_T + offset, b_p + offset, b_rho + offset, b_cp + offset, b_Y + offset, n_boundary,
                                          b_mu ? b_mu + offset : nullptr, b_alpha ? b_alpha + offset : nullptr,
                                          b_rhoD ? b_rhoD + offset : nullptr, n_boundary, unity_lewis, stream);
                if (rc != DFCHEM_OK) return rc;
            }
            if (ty == DFCHEM_PATCH_PROCESSOR) {   // internal side of a processor patch = the owner cells' values
                if (!face_cell || !T || !he || !psi || !rho || n_cells <= 0)
                    return fail(h, DFCHEM_EINVAL, "dfchem_correct_thermo_patches_dev: processor patch needs the internal fields and face_cell");
                PatchRows rows{};
                int nr = 0;
                auto add = [&](const double* in, double* bnd) {
                    if (in && bnd) {
                        rows.in[nr] = in;
                        rows.bnd[nr] = bnd;
                        ++nr;
                    }
                };
                add(T, b_T); add(he, b_he); add(psi, b_psi); add(rho, b_rho); add(mu, b_mu); add(alpha, b_alpha);
                const bool sp = rhoD && b_rhoD;
                patch_internal_side_kernel<<<std::max(1, std::min((sz + 255) / 256, h->sm_count * 4)), 256, 0, s>>>(
                    rows, nr, sz, offset, face_cell, rhoD, b_rhoD, sp ? h->mech.N : 0, (size_t)n_cells, (size_t)n_boundary);
                CUDA_TRY(h, cudaGetLastError());
            }
        }
        offset += span;
    }
    return DFCHEM_OK;
}

int dfchem_bench_fp64_peak(int device, double* tflops) {
    if (!tflops) return DFCHEM_EINVAL;
    if (cudaSetDevice(device) != cudaSuccess) return DFCHEM_ECUDA;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return DFCHEM_ECUDA;
    double* d = nullptr;
    if (cudaMalloc(&d, sizeof(double)) != cudaSuccess) return DFCHEM_ECUDA;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int grid = prop.multiProcessorCount * 8, iters = 20000;
    double best = 0.0;
    for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(e0);
        fp64_peak_kernel<<<grid, 256>>>(d, iters, 1.0 + rep);
        cudaEventRecord(e1);
        if (cudaEventSynchronize(e1) != cudaSuccess) break;
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        const double flops = 2.0 * 16.0 * (double)iters * 256.0 * grid;
        if (rep > 0 && ms > 0) best = std::max(best, flops / (ms * 1e-3) / 1e12);
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(d);
    *tflops = best;
    return cudaGetLastError() == cudaSuccess && best > 0 ? DFCHEM_OK : DFCHEM_ECUDA;
}

// allAverage / headTail rules of LoadBalancer::getOperations / getOperationsRedezVous
// (loadBalancing/LoadBalancer.C:124-180, :183-232): the same code the device planner runs (rebalance.cuh).
int dfchem_balance_plan2(const double* loads, int n, int algorithm, double* send) {
    if (!loads || !send || n < 1 || n > kMaxRanks || algorithm < 0 || algorithm > 1) return DFCHEM_EINVAL;
    for (int me = 0; me < n; ++me) balance_plan_row(loads, n, me, algorithm, send + (size_t)me * n);
    return DFCHEM_OK;
}
int dfchem_balance_plan(const double* loads, int n, double* send) { return dfchem_balance_plan2(loads, n, 0, send); }
// Host form of the device planner (same code): cells rank `me` sends to every rank, given all loads and its own
// per-cell cost.  Used by the CPU tests of the N > 1 protocol; dfchem_solve_balanced_dev runs it on the device.
int dfchem_balance_counts(const double* loads, int n, int me, int algorithm, const double* cost, int nC, int* counts_row) {
    if (!loads || !cost || !counts_row || n < 1 || n > kMaxRanks || me < 0 || me >= n || nC < 1 || algorithm < 0 || algorithm > 1)
        return DFCHEM_EINVAL;
    std::vector<double> prefix((size_t)nC);
    double acc = 0.0;
    for (int i = 0; i < nC; ++i) prefix[i] = (acc += cost[i]);
    double send[kMaxRanks];
    balance_plan_row(loads, n, me, algorithm, send);
    balance_split_row(send, n, prefix.data(), nC, counts_row);
    return DFCHEM_OK;
}

#define NCCL_TRY(h, call)                                                                            \
    do {                                                                                             \
        ncclResult_t r_ = (call);                                                                    \
        if (r_ != ncclSuccess)                                                                       \
            return fail(h, DFCHEM_ENCCL, std::string(#call) + ": " + nccl_api().GetErrorString(r_)); \
    } while (0)

int dfchem_comm_unique_id(void* id128) {
    if (!id128) return DFCHEM_EINVAL;
    if (!nccl_api().ok()) return fail(nullptr, DFCHEM_ENCCL, "libnccl.so.2 not found");
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    ncclUniqueId id;
    if (nccl_api().GetUniqueId(&id) != ncclSuccess) return fail(nullptr, DFCHEM_ENCCL, "ncclGetUniqueId failed");
    std::memcpy(id128, &id, sizeof(id));
    return DFCHEM_OK;
}

static int comm_common(dfchem_ctx* h) {
    if (!h->comm_stream) CUDA_TRY(h, cudaStreamCreateWithFlags(&h->comm_stream, cudaStreamNonBlocking));
    for (cudaEvent_t& e : h->ev)
        if (!e) CUDA_TRY(h, cudaEventCreate(&e));
    if (!h->h_counts) CUDA_TRY(h, cudaMallocHost(&h->h_counts, sizeof(int) * (size_t)kMaxRanks * kMaxRanks));
    return DFCHEM_OK;
}

int dfchem_comm_init(dfchem_handle h, int nranks, int rank, const void* id128) {
    if (!h || !id128 || nranks < 1 || nranks > kMaxRanks || rank < 0 || rank >= nranks)
        return fail(h, DFCHEM_EINVAL, "dfchem_comm_init: bad argument");
    if (!nccl_api().ok()) return fail(h, DFCHEM_ENCCL, "libnccl.so.2 not found");
    CUDA_TRY(h, cudaSetDevice(h->device));
    ncclUniqueId id;
    std::memcpy(&id, id128, sizeof(id));
    NCCL_TRY(h, nccl_api().CommInitRank(&h->comm, nranks, id, rank));
    h->comm_owned = true;
    h->nranks = nranks;
    h->rank = rank;
    return comm_common(h);
}

int dfchem_comm_attach(dfchem_handle h, void* nccl_comm, int nranks, int rank) {
    if (!h || !nccl_comm || nranks < 1 || nranks > kMaxRanks || rank < 0 || rank >= nranks)
        return fail(h, DFCHEM_EINVAL, "dfchem_comm_attach: bad argument");
    if (!nccl_api().ok()) return fail(h, DFCHEM_ENCCL, "libnccl.so.2 not found");
    CUDA_TRY(h, cudaSetDevice(h->device));
    h->comm = static_cast<ncclComm_t>(nccl_comm);
    h->comm_owned = false;
    h->nranks = nranks;
    h->rank = rank;
    return comm_common(h);
}

// One balanced chemistry sub-step (solve_CVODE with loadbalancing active, dfChemistryModel.C:1031-1082).
int dfchem_solve_balanced_dev(dfchem_handle h, int nC, const double* T, const double* p, const double* rho,
                              const double* dt, double dt_uniform, const double* Y, int ldY, double* RR, int ldRR,
                              double* Qdot, const double* cost_prev, double* cost, int* status, int algorithm,
                              void* stream) {
    if (!h) return DFCHEM_EINVAL;
    if (nC < 1 || !T || !p || !rho || !Y || !RR || !Qdot || !cost || ldY < nC || ldRR < nC || algorithm < 0 || algorithm > 1)
        return fail(h, DFCHEM_EINVAL, "dfchem_solve_balanced_dev: bad argument (nC >= 1 and a cost array are required)");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    for (double& v : h->bal_stats) v = 0.0;
    h->bal_timed = false;
    if (!h->comm || h->nranks == 1 || !cost_prev)   // nothing to balance against / no cost yet: the plain sub-step
        return dfchem_solve_dev(h, nC, T, p, rho, dt, dt_uniform, Y, ldY, RR, ldRR, Qdot, cost, status, stream);
    if (h->dtchem) return fail(h, DFCHEM_EUNSUPPORTED, "dfchem_solve_balanced_dev: unbind deltaTChem first (migrated cells carry no step history)");
    CUDA_TRY(h, cudaSetDevice(h->device));
    NcclApi& nc = nccl_api();
    const int n = h->nranks, me = h->rank, N = h->mech.N;
    begin_solve_call(h, nC);

    // ---- plan on the device: prefix sum of the cost, all-gather of the loads, pairing + split, all-gather of counts
    CUDA_TRY(h, cudaEventRecord(h->ev[0], s));
    CUDA_TRY(h, h->bal_prefix.ensure((size_t)nC * sizeof(double)));
    CUDA_TRY(h, h->bal_small.ensure(sizeof(double) * kMaxRanks + sizeof(int) * (size_t)(kMaxRanks + kMaxRanks * kMaxRanks)));
    double* d_prefix = static_cast<double*>(h->bal_prefix.p);
    double* d_loads = static_cast<double*>(h->bal_small.p);
    int* d_row = reinterpret_cast<int*>(d_loads + kMaxRanks);
    int* d_counts = d_row + kMaxRanks;
    size_t tmp_bytes = 0;
    CUDA_TRY(h, cub::DeviceScan::InclusiveSum(nullptr, tmp_bytes, cost_prev, d_prefix, nC, s));
    CUDA_TRY(h, h->bal_scan_tmp.ensure(tmp_bytes));
    CUDA_TRY(h, cub::DeviceScan::InclusiveSum(h->bal_scan_tmp.p, tmp_bytes, cost_prev, d_prefix, nC, s));
    NCCL_TRY(h, nc.AllGather(d_prefix + (nC - 1), d_loads, 1, ncclDouble, h->comm, s));
    plan_split_kernel<<<1, 32, 0, s>>>(d_loads, n, me, algorithm, d_prefix, nC, d_row);
    CUDA_TRY(h, cudaGetLastError());
    NCCL_TRY(h, nc.AllGather(d_row, d_counts, n, ncclInt, h->comm, s));
    CUDA_TRY(h, cudaMemcpyAsync(h->h_counts, d_counts, sizeof(int) * (size_t)n * n, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(h, cudaStreamSynchronize(s));   // the one host synchronisation: cell counts size the exchanges
    int send_cnt[kMaxRanks], recv_cnt[kMaxRanks], send_off[kMaxRanks], recv_off[kMaxRanks];
    int n_send = 0, n_recv = 0;
    for (int j = 0; j < n; ++j) {
        send_cnt[j] = h->h_counts[(size_t)me * n + j];
        recv_cnt[j] = h->h_counts[(size_t)j * n + me];
        send_off[j] = n_send;
        recv_off[j] = n_recv;
        n_send += send_cnt[j];
        n_recv += recv_cnt[j];
    }
    if (n_send > 0 && n_recv > 0) return fail(h, DFCHEM_EINVAL, "balance plan: a rank may be sender or receiver, not both");
    CUDA_TRY(h, cudaEventRecord(h->ev[1], s));

    // ---- exchange: N + 4 row slices per peer, straight out of the caller's arrays into the guest block
    const size_t gl = (size_t)std::max(n_recv, 1);
    CUDA_TRY(h, h->bal_guest_in.ensure((size_t)(N + 4) * gl * sizeof(double)));
    CUDA_TRY(h, h->bal_guest_out.ensure(((size_t)(N + 2) * gl) * sizeof(double) + gl * sizeof(int)));
    double* gin = static_cast<double*>(h->bal_guest_in.p);     // rows T, p, rho, dt, Y[N]; leading dimension gl
    double* gout = static_cast<double*>(h->bal_guest_out.p);   // rows RR[N], Qdot, cost; then int status[gl]
    int* gstat = reinterpret_cast<int*>(gout + (size_t)(N + 2) * gl);
    cudaStream_t cs = h->comm_stream;
    CUDA_TRY(h, cudaStreamWaitEvent(cs, h->ev[1], 0));
    CUDA_TRY(h, cudaEventRecord(h->ev[2], cs));
    NCCL_TRY(h, nc.GroupStart());
    for (int j = 0; j < n; ++j) {
        if (send_cnt[j] > 0) {
            const size_t o = (size_t)send_off[j], c = (size_t)send_cnt[j];
            NCCL_TRY(h, nc.Send(T + o, c, ncclDouble, j, h->comm, cs));
            NCCL_TRY(h, nc.Send(p + o, c, ncclDouble, j, h->comm, cs));
            NCCL_TRY(h, nc.Send(rho + o, c, ncclDouble, j, h->comm, cs));
            if (dt) NCCL_TRY(h, nc.Send(dt + o, c, ncclDouble, j, h->comm, cs));
            for (int k = 0; k < N; ++k) NCCL_TRY(h, nc.Send(Y + (size_t)k * ldY + o, c, ncclDouble, j, h->comm, cs));
        }
        if (recv_cnt[j] > 0) {
            const size_t o = (size_t)recv_off[j], c = (size_t)recv_cnt[j];
            NCCL_TRY(h, nc.Recv(gin + o, c, ncclDouble, j, h->comm, cs));
            NCCL_TRY(h, nc.Recv(gin + gl + o, c, ncclDouble, j, h->comm, cs));
            NCCL_TRY(h, nc.Recv(gin + 2 * gl + o, c, ncclDouble, j, h->comm, cs));
            if (dt) NCCL_TRY(h, nc.Recv(gin + 3 * gl + o, c, ncclDouble, j, h->comm, cs));
            for (int k = 0; k < N; ++k) NCCL_TRY(h, nc.Recv(gin + (size_t)(4 + k) * gl + o, c, ncclDouble, j, h->comm, cs));
        }
    }
    NCCL_TRY(h, nc.GroupEnd());
    CUDA_TRY(h, cudaEventRecord(h->ev[3], cs));

    // ---- guests first (their results travel home while the residents are integrated), then the residents
    int rc = DFCHEM_OK;
    CUDA_TRY(h, cudaStreamWaitEvent(s, h->ev[3], 0));
    if (n_recv > 0) {
        SolveArgs ga{n_recv, gin, gin + gl, gin + 2 * gl, dt ? gin + 3 * gl : nullptr, dt_uniform, gin + 4 * gl, (int)gl,
                     gout, (int)gl, gout + (size_t)N * gl, gout + (size_t)(N + 1) * gl, gstat, h->rtol, h->atol, h->max_steps, nullptr};
        rc = launch_solve(h, ga, s);
        if (rc != DFCHEM_OK) return rc;
    }
    CUDA_TRY(h, cudaEventRecord(h->ev[4], s));
    const int n_own = nC - n_send;
    if (n_own > 0) {
        const size_t o = (size_t)n_send;
        SolveArgs oa{n_own, T + o, p + o, rho + o, dt ? dt + o : nullptr, dt_uniform, Y + o, ldY, RR + o, ldRR, Qdot + o,
                     cost + o, status ? status + o : nullptr, h->rtol, h->atol, h->max_steps, nullptr};
        rc = launch_solve(h, oa, s);
        if (rc != DFCHEM_OK) return rc;
    }
    CUDA_TRY(h, cudaEventRecord(h->ev[5], s));

    // ---- results return to the owner, straight into its RR / Qdot / cost / status arrays
    CUDA_TRY(h, cudaStreamWaitEvent(cs, h->ev[4], 0));
    NCCL_TRY(h, nc.GroupStart());
    for (int j = 0; j < n; ++j) {
        if (recv_cnt[j] > 0) {   // guests of rank j go home
            const size_t o = (size_t)recv_off[j], c = (size_t)recv_cnt[j];
            for (int k = 0; k < N; ++k) NCCL_TRY(h, nc.Send(gout + (size_t)k * gl + o, c, ncclDouble, j, h->comm, cs));
            NCCL_TRY(h, nc.Send(gout + (size_t)N * gl + o, c, ncclDouble, j, h->comm, cs));
            NCCL_TRY(h, nc.Send(gout + (size_t)(N + 1) * gl + o, c, ncclDouble, j, h->comm, cs));
            NCCL_TRY(h, nc.Send(gstat + o, c, ncclInt, j, h->comm, cs));
        }
        if (send_cnt[j] > 0) {
            const size_t o = (size_t)send_off[j], c = (size_t)send_cnt[j];
            for (int k = 0; k < N; ++k) NCCL_TRY(h, nc.Recv(RR + (size_t)k * ldRR + o, c, ncclDouble, j, h->comm, cs));
            NCCL_TRY(h, nc.Recv(Qdot + o, c, ncclDouble, j, h->comm, cs));
            NCCL_TRY(h, nc.Recv(cost + o, c, ncclDouble, j, h->comm, cs));
            // without a caller status array the migrated cells' status lands in the prefix-sum buffer (idle by now)
            NCCL_TRY(h, nc.Recv(status ? status + o : reinterpret_cast<int*>(d_prefix) + o, c, ncclInt, j, h->comm, cs));
        }
    }
    NCCL_TRY(h, nc.GroupEnd());
    CUDA_TRY(h, cudaEventRecord(h->ev[6], cs));
    CUDA_TRY(h, cudaStreamWaitEvent(s, h->ev[6], 0));
    CUDA_TRY(h, cudaEventRecord(h->ev[7], s));
    h->bal_stats[0] = n_send;
    h->bal_stats[1] = n_recv;
    h->bal_stats[2] = (double)n_send * ((N + 3 + (dt ? 1 : 0)) * 8.0) + (double)n_recv * ((N + 2) * 8.0 + 4.0);   // bytes this rank sent
    h->bal_timed = true;
    return DFCHEM_OK;
}

// [0] cells sent, [1] guest cells received, [2] bytes sent by this rank (both directions), then milliseconds of the
// last balanced call: [3] plan (prefix sum, two all-gathers, pairing, count copy), [4] outbound exchange,
// [5] guest solve, [6] resident solve, [7] whole call.  The caller must have synchronised the stream.
int dfchem_balance_stats(dfchem_handle h, double out[8]) {
    if (!h || !out) return DFCHEM_EINVAL;
    if (h->bal_timed) {
        float ms = 0.f;
        auto el = [&](int a, int b) { return cudaEventElapsedTime(&ms, h->ev[a], h->ev[b]) == cudaSuccess ? (double)ms : -1.0; };
        h->bal_stats[3] = el(0, 1);
        h->bal_stats[4] = el(2, 3);
        h->bal_stats[5] = el(1, 4);
        h->bal_stats[6] = el(4, 5);
        h->bal_stats[7] = el(0, 7);
        h->bal_timed = false;
    }
    std::copy(h->bal_stats, h->bal_stats + 8, out);
    return DFCHEM_OK;
}

}  // extern "C"
