// Flat tables of the LANE-PER-CELL chemistry kernel (lane_cell.cuh): the small-mechanism path
// (H2 / one-step mechanisms: a handful of active species), where a whole warp per cell would leave
// most lanes idle in the matrix phases.  One lane integrates one cell from start to end; the 32
// lanes of a warp walk the SAME table entry at the same time, so every table read is a broadcast
// and the decode is paid once per 32 cells.  Second output of the host mechanism compiler next to
// dev_tables.hpp; the rate laws are the same (Cantera 2.6 GasKinetics).
#pragma once
#include <cstdint>
#include <vector>

#include "mechanism.hpp"

namespace dfchem {

constexpr int kLpcMaxActive = 12;    // mechanisms with more active species use the warp-per-cell kernels
constexpr int kLpcMaxNet = 2 * kMaxSide;

// One reaction, read uniformly by the warp (all fields in internal species numbering, active first).
struct LpcReaction {
    int32_t kind;               // ReactionKind
    int32_t rev;                // reversible
    int32_t nr, np;             // concentration factors on the forward / reverse side
    int32_t rsp[kMaxSide];      // forward factors: species
    int32_t psp[kMaxSide];      // reverse factors: species (reversible reactions only)
    int32_t simple;             // every order is 1, 2 or 3 (no pow() needed)
    int32_t mset;               // third-body efficiency set (-1: none)
    int32_t fall;               // falloff record (-1: none)
    int32_t nnet;               // species with non-zero net stoichiometric coefficient
    int32_t nsp[kLpcMaxNet];    // ... their indices (all < NA)
    int32_t caller_index;       // index in the Mechanism record (rates-of-progress output order)
    int32_t pad_;
    double rord[kMaxSide];      // forward orders (the `orders` entry, else the stoichiometric coefficient)
    double pord[kMaxSide];      // reverse orders = product stoichiometric coefficients
    double nnu[kLpcMaxNet];     // net stoichiometric coefficients nu'' - nu'
    double lnA, b, Ta;          // k = exp(lnA + b ln T - Ta/T)   (high-pressure limit for falloff)
};

// Device/host view.  Per-lane arrays live in shared memory as X[i * LS + lane] (LS = 32 on the
// device, 1 in the host emulation); offsets below are in doubles per lane.
struct LpcMech {
    int N, NA, R, n_msets, n_fall;
    int off_c, off_cs, off_kf, off_rkc, off_k0, off_lfc, off_f, off_J, off_A, off_k, off_f0, off_dc, off_x,
        lane_doubles;
    // tables (one contiguous blob; staged into shared memory by every CTA)
    const uint8_t* base;
    int bytes;
    const LpcReaction* reac;   // [R]
    const double* eff;         // [n_msets * N] dense efficiencies, internal order
    const double* fall;        // [n_fall * 8] ln A0, b0, Ta0, Troe A, T3, T1, T2, is_troe
    const double* W;           // [N]
    const double* Tmid;        // [N]
    const double* nasa;        // [N * 14]
    const double* hc;          // [N]
    const int* perm;           // [N] internal -> caller
};

struct HostLpc {
    bool usable = false;        // NA <= kLpcMaxActive and every reaction representable
    LpcMech layout{};           // pointer fields hold byte offsets into blob until rebase()
    std::vector<uint8_t> blob;
    LpcMech rebase(const void* base) const;
};

// perm: internal -> caller species index with the NA active species first (HostTables::perm)
HostLpc build_lpc_tables(const Mechanism& m, const std::vector<int>& perm, int NA);

}  // namespace dfchem
