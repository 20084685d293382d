// Cost-weighted migration of cells between the GPUs of one box, inside the library: the reference's chemistry
// load balancer (src/dfChemistryModel/loadBalancing/) with an NCCL data plane over NVLink.
//
//   LoadBalancerBase::updateState   LoadBalancer.C:29-56      all-gather of per-rank loads (sum of the previous
//                                                              step's per-cell cost)          -> ncclAllGather, 1 double
//   LoadBalancer::getOperations     LoadBalancer.C:124-180    allAverage pairing             -> balance_plan_row()
//   getOperationsRedezVous          LoadBalancer.C:183-232    headTail pairing               -> balance_plan_row()
//   timesToProblemCounts            LoadBalancer.C:97-122     load units -> leading cells    -> plan_split_kernel
//   LoadBalancerBase::balance       LoadBalancerBase.H:143-189 send the leading cells         -> grouped ncclSend/ncclRecv
//   getRemaining                    LoadBalancerBase.H:205-225 the cells that stay            -> pointer offset, no copy
//   LoadBalancerBase::unbalance     LoadBalancerBase.H:227-255 solutions return to the owner  -> grouped ncclSend/ncclRecv
//
// B200-first differences: no AoS ChemistryProblem records and no packing at all -- a migrated range of cells is
// N + 4 contiguous row slices of the caller's species-major arrays, each sent by its own ncclSend inside one group and
// received straight into the guest block (and, on the way back, straight into the owner's RR / Qdot / cost / status
// arrays); the planner runs on the device (prefix sum of the cost, pairing, binary search), so one 4 n^2-byte
// device-to-host copy of the cell counts is the only host synchronisation; guests are solved first, and their results
// travel home on a second stream while the resident cells are being integrated.
#pragma once
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>

#include <cub/device/device_scan.cuh>

namespace dfchem {

constexpr int kMaxRanks = 64;

// ---- the pairing rules, shared by the host entry point (dfchem_balance_plan) and the device planner ------------
// loads[n] -> send[to] for rank `me` (algorithm 0: allAverage, 1: headTail).  Every rank evaluates the same global
// pairing and keeps its own row; operations below 1 % of the mean load are dropped, as in the reference.
__host__ __device__ inline void balance_plan_row(const double* loads, int n, int me, int algorithm, double* send) {
    int order[kMaxRanks];
    double v[kMaxRanks];
    double mean = 0.0;
    for (int i = 0; i < n; ++i) {
        mean += loads[i];
        send[i] = 0.0;
    }
    mean /= n;
    for (int i = 0; i < n; ++i) {  // stable insertion sort by load, ascending
        int j = i;
        while (j > 0 && loads[order[j - 1]] > loads[i]) {
            order[j] = order[j - 1];
            --j;
        }
        order[j] = i;
    }
    for (int i = 0; i < n; ++i) v[i] = loads[order[i]];
    if (algorithm == 1) {  // headTail: i-th heaviest gives half the difference to the i-th lightest
        for (int i = 0; i < n / 2; ++i) {
            const int hi = n - 1 - i, lo = i;
            const double x = 0.5 * (v[hi] - v[lo]);
            if (x > 0.01 * mean && order[hi] == me) send[order[lo]] += x;
        }
        return;
    }
    int lo = 0, hi = n - 1;  // allAverage: lightest receiver, heaviest sender
    while (lo < hi) {
        const double give = v[hi] - mean, take = mean - v[lo];
        if (!(give > 0.0) || !(take > 0.0)) break;
        const double x = give < take ? give : take;
        if (x > 0.01 * mean && order[hi] == me) send[order[lo]] += x;
        v[hi] -= x;
        v[lo] += x;
        if (v[hi] - mean <= 1e-12 * mean) --hi;
        if (mean - v[lo] <= 1e-12 * mean) ++lo;
    }
}

// timesToProblemCounts: prefix = inclusive prefix sum of the per-cell cost (index order, as the reference walks its
// problem list); destination j receives the next cells until their summed cost reaches its share.  At least one
// cell always stays.  counts_row[n]: cells this rank sends to each rank.
__host__ __device__ inline void balance_split_row(const double* send, int n, const double* prefix, int nC, int* counts_row) {
    int start = 0;
    const int limit = nC - 1;
    for (int j = 0; j < n; ++j) {
        int cnt = 0;
        if (send[j] > 0.0 && start < limit) {
            const double target = (start > 0 ? prefix[start - 1] : 0.0) + send[j];
            int a = start, b = limit;  // first index whose inclusive prefix exceeds the target
            while (a < b) {
                const int mid = (a + b) >> 1;
                if (prefix[mid] <= target) a = mid + 1; else b = mid;
            }
            cnt = a - start;
            start = a;
        }
        counts_row[j] = cnt;
    }
}

__global__ void plan_split_kernel(const double* loads, int n, int me, int algorithm, const double* prefix, int nC,
                                  int* counts_row) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    double send[kMaxRanks];
    balance_plan_row(loads, n, me, algorithm, send);
    balance_split_row(send, n, prefix, nC, counts_row);
}

// ---- NCCL, resolved at run time: the process may already hold a libnccl (torch's); that one is used ---------------
struct NcclApi {
    void* lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    bool ok() const { return lib != nullptr; }
};
inline NcclApi& nccl_api() {
    static NcclApi api = [] {
        NcclApi a;
        void* lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);
        if (!lib) lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!lib) return a;
#define DFC_NCCL_SYM(field, name) a.field = reinterpret_cast<decltype(a.field)>(dlsym(lib, name))
        DFC_NCCL_SYM(GetUniqueId, "ncclGetUniqueId");
        DFC_NCCL_SYM(CommInitRank, "ncclCommInitRank");
        DFC_NCCL_SYM(CommDestroy, "ncclCommDestroy");
        DFC_NCCL_SYM(AllGather, "ncclAllGather");
        DFC_NCCL_SYM(Send, "ncclSend");
        DFC_NCCL_SYM(Recv, "ncclRecv");
        DFC_NCCL_SYM(GroupStart, "ncclGroupStart");
        DFC_NCCL_SYM(GroupEnd, "ncclGroupEnd");
        DFC_NCCL_SYM(GetErrorString, "ncclGetErrorString");
#undef DFC_NCCL_SYM
        if (a.GetUniqueId && a.CommInitRank && a.CommDestroy && a.AllGather && a.Send && a.Recv && a.GroupStart &&
            a.GroupEnd && a.GetErrorString)
            a.lib = lib;
        return a;
    }();
    return api;
}

}  // namespace dfchem
