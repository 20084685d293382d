// Flat device tables built from a Mechanism by the host mechanism compiler (build_tables.cpp)
// and consumed by the CUDA kernels (chem_core.cuh).  This is the "flat device tables of
// Arrhenius, third-body, falloff (Troe/Lindemann) and NASA-7 data" of the north star; it plays
// the role of the Cantera Kinetics/ThermoPhase objects the reference reaches through
// CanteraMixture (src/dfCanteraMixture/CanteraMixture.H:105-115).
//
// Species are renumbered internally: the NA "active" species (those that appear in at least
// one reaction) come first, the inert ones (N2, AR, HE ... : dC/dt == 0 exactly) last.  Only the
// active block enters the per-cell linear algebra.  perm[internal] = caller (Cantera) index.
#pragma once
#include <cstdint>
#include <vector>

#include "mechanism.hpp"

namespace dfchem {

constexpr int kTeam = 32;  // lanes cooperating on one cell (one warp)

struct alignas(8) U2 {
    uint32_t x, y;
};
struct alignas(16) U4 {
    uint32_t x, y, z, w;
};

// Lane-blocked "segmented gather" program.  Lane l executes, for it = 0..n_iter-1, the entry
// ent[(it/4)*4*kTeam + l*4 + it%4] (four consecutive entries of a lane are one 16-byte load):
//   acc += coef[e.coef] * src[e.src];  if (e.store) { dst[e.dst] = acc; acc = 0; }
// Every destination is written by exactly one lane, so no atomics are needed.  n_iter % 4 == 0.
//   bits 0-11 src | bits 12-18 coef id | bit 19 store | bits 20-31 dst
struct GatherProg {
    int n_iter;
    const uint32_t* ent;
};

// Reaction descriptor (16 B).  A side holds up to 3 concentration factors; an integer
// stoichiometric coefficient/order nu is expanded into nu repeated factors ("2 OH" -> OH, OH),
// so products and the product rule for derivatives need no exponent.  Reactions that do not fit
// (fractional `orders`, more than 3 molecules on a side) use the pow path: factors are the
// distinct species and powtab[pow_id*9 + {0..2 order, 3..5 reactant nu, 6..8 product nu}].
//   x = rf0 | rf1<<8 | rf2<<16 | pf0<<24          (0xFF = empty)
//   y = pf1 | pf2<<8 | flags<<16 | mset<<24        flags: bits0-1 kind, bit2 reversible, bit3 pow
//   z = fall_id | pow_id<<16
//   w = dq_off   first slot of this reaction in the per-cell derivative scratch; slot order:
//                present reactant factors, present product factors (reversible only),
//                d q/d[M] (three-body / falloff only)
constexpr uint32_t kFlagRev = 4u, kFlagPow = 8u;

struct DevMech {
    int N, NA, NP, R, n_msets, n_fall, n_pow, n_dq, n_part, ldj;  // NP: padded NA (kernel template)
    // per-cell shared-memory layout, offsets in doubles from the cell's base (all even)
    int off_c, off_cs, off_kf, off_rkc, off_k0, off_lfc, off_M, off_q, off_dq, off_J, off_vec,
        off_row, off_part, off_red, cell_doubles;
    // hot tables (rdesc, coef, efficiency sets, gather programs) are one contiguous prefix of the
    // table blob; the kernels stage it in shared memory once per CTA (rebase_hot)
    const uint8_t* hot_base;
    int hot_bytes;
    // species tables, internal order
    const int* perm;       // [N]
    const double* W;       // [N]
    const double* Tmid;    // [N]
    const double* nasa;    // [N*14]  low-range a0..a6 (T <= Tmid), high-range a0..a6
    const double* hc;      // [N]
    // reactions
    const U4* rdesc;       // [R]
    const double* arrh;    // [R*3]  ln A, b, Ea/R   (k_inf for falloff)
    const double* fall;    // [n_fall*8] ln A0, b0, Ea0/R, Troe A, T3, T1, T2, is_troe
    const double* powtab;  // [n_pow*9]
    // third-body efficiency sets: M_s = dflt_s*Ctot + sum_j (eff_sj - dflt_s) c_j
    const double* ms_dflt; // [n_msets]
    const int* ms_ptr;     // [n_msets+1]
    const int* ms_idx;     // species (internal)
    const double* ms_val;  // eff - dflt
    // gather programs
    const double* coef;    // distinct coefficients, coef[0] = 0
    GatherProg wdotA;      // q[R] -> part[n_part]
    GatherProg wdotB;      // part -> w[NA]
    GatherProg jac;        // dq[n_dq] -> J[(NA+1)*ldj]; column NA holds gsum_i = sum_r nu_ir dflt_r dq_r/d[M]
};

// Host-side owner of the table storage.  `blob` is one contiguous byte array; DevMech pointers
// are produced by rebase() against either the host copy (emulator/tests) or a device copy.
struct HostTables {
    DevMech layout{};            // pointer fields hold byte offsets into blob until rebase()
    std::vector<uint8_t> blob;
    std::vector<int> perm;       // internal -> caller
    int jac_entries = 0, wdot_entries = 0;
    DevMech rebase(const void* base) const;
};

HostTables build_tables(const Mechanism& m);

}  // namespace dfchem
