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

// Lane-blocked "row program": a sparse matrix-vector product scheduled at mechanism-compile time.
// Lane l executes rows i = 0..n_rows-1; row i of lane l is the 16-byte record ent[i*kTeam + l]
// (four entries, one LDS.128) plus the word dst[i*kTeam + l].  The row sums its four entries and
// stores (or, with the accumulate flag, adds) the sum to *(double*)(dst_base + (dst & 0xFFFFFF)).
//   type-A rows (i < nA): value = +/- src            entry = src_byte_offset | sign << 31
//   type-B rows (i >= nA): value = coef[id] * src    entry = src_byte_offset | id << 16
// Integer stoichiometric coefficients are expanded into repeated +/-1 entries, so the common case
// needs no coefficient load and no multiply.  Padding entries point at a slot that holds 0.0
// (type A) or use coef[0] = 0 (type B); idle rows write 0 to a per-lane dump slot.  Rows that share a
// destination are scheduled on the same lane in program order, so `accumulate` needs no atomics.
constexpr uint32_t kRowAcc = 0x80000000u;
struct RowProg {
    int nA, nB;
    const U4* ent;
    const uint32_t* dst;
};

// Reaction descriptor (16 B).  A side holds up to 3 concentration factors; an integer
// stoichiometric coefficient/order nu is expanded into nu repeated factors ("2 OH" -> OH, OH),
// so products and the product rule for derivatives need no exponent.  An absent factor is the
// index N: slot N of the concentration array always holds 1.0.  Reactions that do not fit
// (fractional `orders`, more than 3 molecules on a side) use the pow path: factors are the
// distinct species and powtab[pow_id*9 + {0..2 order, 3..5 reactant nu, 6..8 product nu}].
// Factor fields are stored pre-scaled (index*8 = byte offset of the double):
//   x = kind | rev<<2 | rf0<<3 | rf1<<11 | rf2<<19 | pow<<27
//   y =                 pf0<<3 | pf1<<11 | pf2<<19
//   z = mset | fall_id<<8 | pow_id<<20
//   w = dq_off   first slot of this reaction in the per-cell derivative scratch; slot order:
//                present reactant factors, present product factors (reversible only),
//                d q/d[M] (three-body / falloff only)
constexpr uint32_t kFlagRev = 4u, kFlagPow = 1u << 27;

// wdot = S q as a register-resident schedule: every lane keeps its rows' shared-memory offsets in
// registers for the whole launch (they are constants of the mechanism), so a row costs four
// loads, four adds and a store with no decoding.  Row = (q[s0] + q[s1]) - (q[s2] + q[s3]); integer
// stoichiometric coefficients are expanded into repeated entries and unused slots point at a
// slot that holds 0.0.  Stage A rows accumulate into at most four partial sums per species
// (consecutive rows of one partial sit on one lane); every species has exactly four partials,
// part[4i .. 4i+3], so stage B is the fixed pattern w_i = their sum and needs no table.
// Table record (uint16 x 6) per (row, lane): s0..s3 and dst = double index relative to the cell
// base, flags: bit0 = start a new partial (reset the accumulator).  Rows that do not finish a
// partial store to the lane's dump slot.
struct WdotProg {
    int nA, nB;            // stage-A rows per lane (nB unused, 0)
    const uint16_t* tab;   // [nA * kTeam * 6]
};
constexpr int kWdotRowsMax = 16;

struct DevMech {
    int N, NA, NP, R, n_msets, n_fall, n_pow, n_dq, n_part, ldj;  // NP: padded NA (kernel template)
    // per-cell shared-memory layout, offsets in doubles from the cell's base (all even)
    // c, cs: [N + 2] concentrations, slot N = 1.0 (absent factor), slot N+1 = sum of all N
    // kf: [2R] interleaved (kf_r, 1/Kc_r);  q: [R + 1], slot R = 0.0;  dq: [n_dq + 1], last = 0.0
    int off_c, off_cs, off_kf, off_k0, off_lfc, off_M, off_q, off_dq, off_J, off_vec,
        off_row, off_part, off_red, cell_doubles;  // off_red: kTeam dump slots
    // hot tables (rdesc, coef, row programs) are one contiguous prefix of the
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
    // row programs
    const double* coef;    // distinct coefficients, coef[0] = 0
    // third-body sets, evaluated by the reaction's own lane: M_s = sum_t val_t * c[off_t] over the
    // entries ms_rng[s] = first | count << 16; entry = {val (2 words), byte offset of c_j, 0};
    // the default efficiency enters as the entry (dflt_s, slot N+1 = total concentration)
    const U4* ms_ent;
    const uint32_t* ms_rng;
    WdotProg wdot;         // q[R+1] -> part[n_part] -> w[NA]  (cold table: read once per launch)
    RowProg jac;           // dq[n_dq+1] -> J[(NA+1)*ldj]; column NA holds gsum_i = sum_r nu_ir dflt_r dq_r/d[M]
};

// Host-side owner of the table storage.  `blob` is one contiguous byte array; DevMech pointers
// are produced by rebase() against either the host copy (emulator/tests) or a device copy.
struct HostTables {
    DevMech layout{};            // pointer fields hold byte offsets into blob until rebase()
    std::vector<uint8_t> blob;
    std::vector<int> perm;       // internal -> caller
    std::vector<int> rorder;     // internal reaction index -> index in the Mechanism record (caller's)
    int jac_entries = 0, wdot_entries = 0;
    DevMech rebase(const void* base) const;
};

HostTables build_tables(const Mechanism& m);

}  // namespace dfchem
