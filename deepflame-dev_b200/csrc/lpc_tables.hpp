// Flat tables of the LANE-PER-CELL chemistry kernel (lane_cell.cuh): the small-mechanism path
// (H2 / one-step mechanisms: a handful of active species), where a whole warp per cell would leave
// most lanes idle in the matrix phases.  One lane integrates one cell from start to end; the 32
// lanes of a warp walk the SAME table entry at the same time, so every table read is a broadcast
// and the decode is paid once per 32 cells.  Second output of the host mechanism compiler next to
// dev_tables.hpp; the rate laws are the same (Cantera 2.6 GasKinetics).
//
// GATHER FORM (round 2, second version): the reaction loop only STORES (q_r and the partial
// derivatives d q_r / d c_j go to per-lane shared-memory columns); wdot_i and J_ij are then summed
// from per-target entry lists in loops whose TARGET index is a compile-time constant (the kernel is
// templated on the padded species count), so the accumulators live in registers and no shared-memory
// read-modify-write chain exists anywhere.  The lists are stored in ELL form (third version): one row
// serves all NAP targets at once, which gives the lane NAP independent load -> multiply-add chains
// per iteration instead of one list walked after the other.
#pragma once
#include <cstdint>
#include <vector>

#include "mechanism.hpp"

namespace dfchem {

constexpr int kLpcMaxActive = 12;    // mechanisms with more active species use the warp-per-cell kernels
constexpr int kLpcMaxSpecies = 60;   // species indices are 6-bit fields of the reaction record (index N = "one")

// Reaction record, 16 bytes, read uniformly by the warp.  A side holds up to 3 concentration factors; an
// integer order nu is expanded into nu repeated factors; an absent factor is the index N, whose concentration
// slot holds 1.0.  Reactions that do not fit (fractional orders, more than 3 factors) use the pow path:
// factors = distinct species, powtab[pow_id*9 + {0..2 forward order, 3..5 unused, 6..8 reverse order}].
//   x = kind | rev<<2 | pow<<3 | rf0<<8 | rf1<<14 | rf2<<20
//   y = pf0 | pf1<<6 | pf2<<12 | mset<<20            (mset 0xFF: none)
//   z = fall_id | pow_id<<8 | caller_index<<16
//   w = dq_off   first slot of this reaction in the per-lane derivative column; slot order: present reactant
//                factors, present product factors (reversible only), d q/d[M] (three-body / falloff only)
struct LpcRx {
    uint32_t x, y, z, w;
};

// Device/host view.  Per-lane arrays live in shared memory as X[i * LS + lane] (LS = 32 on the
// device, 1 in the host emulation); offsets below are in doubles per lane.  The arrays whose size depends on
// the padded active-species count only come first, at compile-time offsets (lane_cell.cuh: LpcFixed<NAP>).
struct LpcMech {
    int N, NA, NAP, R, n_msets, n_fall, n_dq;
    int off_c, off_cs, off_kf, off_rkc, off_k0, off_lfc, off_M, off_q, off_dq, lane_doubles;
    // tables (one contiguous blob; staged into shared memory by every CTA)
    const uint8_t* base;
    int bytes;
    const LpcRx* rx;           // [R]
    const double* powtab;      // [n_pow * 9]
    const double* eff;         // [n_msets * N] dense efficiencies, internal order
    const double* fall;        // [n_fall * 8] ln A0, b0, Ta0, Troe A, T3, T1, T2, is_troe
    const double* arrh;        // [R * 3] ln A, b, Ta
    // gather lists in ELL form: row k of a list set holds ONE entry for each of the NAP targets (padded with
    // coefficient 0 and the index of a slot that always holds 0.0), so a gather is `rows` iterations of NAP
    // independent multiply-adds.  Indices are 16-bit, two per word: target t -> word t/2, half t%2.
    int f_rows;                // wdot: targets = species, entries (reaction, nu); zero slot = q[q_zero]
    int q_zero;
    const uint32_t* fidx;      // [f_rows][NAP/2]
    const double* fcoef;       // [f_rows][NAP]
    const int* jrow;           // [NAP + 1] Jacobian: matrix row i uses ELL rows jrow[i] .. jrow[i+1], targets = columns,
    const uint32_t* jidx;      //   entries (dq slot, coefficient); zero slot = dq[n_dq]
    const double* jcoef;
    const int* gptr;           // [R + 1] equilibrium-constant lists: entries gidx (species), gcoef (net nu)
    const int* gidx;
    const double* gcoef;
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

constexpr int lpc_pad(int NA) { return NA <= 4 ? 4 : NA <= 6 ? 6 : NA <= 8 ? 8 : 12; }
// up to 6 x 6 the LU factors of a step attempt stay in the lane's registers for the six stage solves (72 registers);
// larger matrices keep them in the fixed region
constexpr bool lpc_lu_in_regs(int NAP) { return NAP <= 6; }
// per-lane doubles of the fixed region: J, [LU,] six stage vectors, f0, dc
constexpr int lpc_fixed_doubles(int NAP) { return (lpc_lu_in_regs(NAP) ? 1 : 2) * NAP * NAP + 8 * NAP; }
// room for the derivative column where it aliases dead arrays of the fixed region, starting at offset NAP*NAP: the
// stage vectors and f0 (LU in registers) or the LU factors and the stage vectors
constexpr int lpc_alias_room(int NAP) { return lpc_lu_in_regs(NAP) ? 7 * NAP : NAP * NAP + 6 * NAP; }

// perm: internal -> caller species index with the NA active species first (HostTables::perm)
HostLpc build_lpc_tables(const Mechanism& m, const std::vector<int>& perm, int NA);

}  // namespace dfchem
