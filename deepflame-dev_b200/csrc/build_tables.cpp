// Second stage of the host mechanism compiler: Mechanism -> flat device tables (dev_tables.hpp).
// Builds the reaction descriptors, third-body efficiency sets, and the conflict-free, load-
// balanced gather schedules the warp-per-cell kernels use for wdot = S q and for the analytic
// Jacobian d wdot / d c (the arithmetic Cantera's GasKinetics does per cell for the reference,
// reached from src/dfChemistryModel/dfChemistryModel.C:748-764 and :916).
#include <algorithm>
#include <cmath>
#include <cstring>
#include <map>
#include <stdexcept>

#include "dev_tables.hpp"

namespace dfchem {
namespace {

struct GEntry {
    int src;
    double coef;
};
struct GTarget {
    int dst;
    std::vector<GEntry> e;
};

struct CoefTable {
    std::vector<double> v{0.0};
    int id(double c) {
        for (size_t i = 1; i < v.size(); ++i)
            if (v[i] == c) return (int)i;
        v.push_back(c);
        if (v.size() > 128) throw std::runtime_error("tables: more than 128 distinct schedule coefficients");
        return (int)v.size() - 1;
    }
};

struct HostRowProg {
    int nA = 0, nB = 0;
    std::vector<U4> ent;
    std::vector<uint32_t> dst;
};

// Rows of one target: type-A rows (+/-1 entries, integer coefficients expanded) then type-B rows.
struct TargetRows {
    int dst;
    std::vector<U4> a, b;
};

TargetRows make_rows(const GTarget& t, CoefTable& ct, int zero_src) {
    TargetRows tr;
    tr.dst = t.dst;
    std::vector<uint32_t> ea, eb;
    for (const GEntry& e : t.e) {
        if (e.src < 0 || e.src > 0x1FFF) throw std::runtime_error("tables: row-program source out of range");
        const uint32_t off = (uint32_t)e.src * 8u;
        const double ac = std::fabs(e.coef);
        if (ac == std::round(ac) && ac >= 1.0 && ac <= 4.0 && zero_src >= 0) {
            for (int k = 0; k < (int)ac; ++k) ea.push_back(off | (e.coef < 0 ? 0x80000000u : 0u));
        } else if (e.coef != 0.0) {
            eb.push_back(off | ((uint32_t)ct.id(e.coef) << 16));
        }
    }
    auto rows = [](std::vector<uint32_t>& v, uint32_t pad, std::vector<U4>& out) {
        while (v.size() % 4) v.push_back(pad);
        for (size_t i = 0; i < v.size(); i += 4) out.push_back(U4{v[i], v[i + 1], v[i + 2], v[i + 3]});
    };
    rows(ea, zero_src >= 0 ? (uint32_t)zero_src * 8u : 0u, tr.a);
    rows(eb, 0u, tr.b);
    return tr;
}

// Whole targets onto kTeam lanes, longest first onto the least loaded lane (a type-B row costs
// about 1.75 type-A rows).  All rows of a target stay on one lane: first row stores, rest add.
// Idle rows (lanes with fewer rows than the longest) write their zero sum to dump slot
// dst_limit + lane, so the interpreter needs no "skip" test.
HostRowProg pack_rows(std::vector<TargetRows> targets, int dst_limit) {
    auto cost = [](const TargetRows& t) { return 4 * (int)t.a.size() + 7 * (int)t.b.size(); };
    std::stable_sort(targets.begin(), targets.end(),
                     [&](const TargetRows& x, const TargetRows& y) { return cost(x) > cost(y); });
    struct Lane {
        std::vector<U4> a, b;
        std::vector<uint32_t> da, db;
        int load = 0;
    };
    std::vector<Lane> lane(kTeam);
    for (const TargetRows& t : targets) {
        if (t.a.empty() && t.b.empty()) continue;
        if (t.dst < 0 || t.dst >= dst_limit) throw std::runtime_error("tables: row-program destination out of range");
        int best = 0;
        for (int l = 1; l < kTeam; ++l)
            if (lane[l].load < lane[best].load) best = l;
        Lane& L = lane[best];
        L.load += cost(t);
        bool first = true;
        for (const U4& r : t.a) {
            L.a.push_back(r);
            L.da.push_back((uint32_t)t.dst * 8u | (first ? 0u : kRowAcc));
            first = false;
        }
        for (const U4& r : t.b) {
            L.b.push_back(r);
            L.db.push_back((uint32_t)t.dst * 8u | (first ? 0u : kRowAcc));
            first = false;
        }
    }
    HostRowProg p;
    for (auto& L : lane) {
        p.nA = std::max(p.nA, (int)L.a.size());
        p.nB = std::max(p.nB, (int)L.b.size());
    }
    const int n = p.nA + p.nB;
    p.ent.assign((size_t)std::max(n, 1) * kTeam, U4{0, 0, 0, 0});
    p.dst.assign((size_t)std::max(n, 1) * kTeam, 0u);
    for (int l = 0; l < kTeam; ++l) {
        for (int i = 0; i < std::max(n, 1); ++i) p.dst[(size_t)i * kTeam + l] = (uint32_t)(dst_limit + l) * 8u;
        for (size_t i = 0; i < lane[l].a.size(); ++i) {
            p.ent[i * kTeam + l] = lane[l].a[i];
            p.dst[i * kTeam + l] = lane[l].da[i];
        }
        for (size_t i = 0; i < lane[l].b.size(); ++i) {
            p.ent[(p.nA + i) * kTeam + l] = lane[l].b[i];
            p.dst[(p.nA + i) * kTeam + l] = lane[l].db[i];
        }
    }
    return p;
}

bool is_int(double x) { return std::fabs(x - std::round(x)) < 1e-12; }

struct Blob {
    std::vector<uint8_t> b;
    template <class T>
    size_t add(const std::vector<T>& v) {
        size_t off = (b.size() + 15) & ~size_t(15);
        b.resize(off + std::max<size_t>(v.size(), 1) * sizeof(T), 0);
        if (!v.empty()) std::memcpy(b.data() + off, v.data(), v.size() * sizeof(T));
        return off;
    }
};

template <class T>
const T* as_ptr(size_t off) {
    return reinterpret_cast<const T*>(off);
}
template <class T>
const T* shift(const T* off, const void* base) {
    return reinterpret_cast<const T*>(reinterpret_cast<const uint8_t*>(base) + reinterpret_cast<size_t>(off));
}

int choose_np(int NA) {
    static const int sizes[] = {8, 16, 20, 24, 32, 33, 34, 35, 36, 40, 48, 64};
    for (int s : sizes)
        if (NA <= s) return s;
    throw std::runtime_error("tables: more than 64 active species is not supported by the compiled kernel set");
}

}  // namespace

DevMech HostTables::rebase(const void* base) const {
    DevMech d = layout;
    d.hot_base = shift(d.hot_base, base);
    d.perm = shift(d.perm, base);
    d.W = shift(d.W, base);
    d.Tmid = shift(d.Tmid, base);
    d.nasa = shift(d.nasa, base);
    d.hc = shift(d.hc, base);
    d.rdesc = shift(d.rdesc, base);
    d.arrh = shift(d.arrh, base);
    d.fall = shift(d.fall, base);
    d.powtab = shift(d.powtab, base);
    d.coef = shift(d.coef, base);
    d.ms_ent = shift(d.ms_ent, base);
    d.ms_rng = shift(d.ms_rng, base);
    d.wdot.tab = shift(d.wdot.tab, base);
    d.jac.ent = shift(d.jac.ent, base);
    d.jac.dst = shift(d.jac.dst, base);
    return d;
}

HostTables build_tables(const Mechanism& m) {
    const int N = m.N, R = m.R;
    if (N > 254) throw std::runtime_error("tables: more than 254 species");
    HostTables ht;

    // ---- species renumbering: active first ------------------------------------------------
    // active = some reaction changes it (non-zero net stoichiometric coefficient); species that
    // only act as colliders (H + O2 + N2 <=> HO2 + N2) stay constant and are treated as inert
    std::vector<char> active(N, 0);
    for (int r = 0; r < R; ++r) {
        std::map<int, double> net;
        for (int j = 0; j < kMaxSide; ++j) {
            if (m.r_idx[r * kMaxSide + j] >= 0) net[m.r_idx[r * kMaxSide + j]] -= m.r_nu[r * kMaxSide + j];
            if (m.p_idx[r * kMaxSide + j] >= 0) net[m.p_idx[r * kMaxSide + j]] += m.p_nu[r * kMaxSide + j];
        }
        for (auto& kv : net)
            if (kv.second != 0.0) active[kv.first] = 1;
    }
    std::vector<int> perm, inv(N, -1);
    for (int k = 0; k < N; ++k)
        if (active[k]) perm.push_back(k);
    const int NA = (int)perm.size();
    for (int k = 0; k < N; ++k)
        if (!active[k]) perm.push_back(k);
    for (int i = 0; i < N; ++i) inv[perm[i]] = i;
    const int NP = choose_np(std::max(NA, 1));
    const int ldj = std::max(NA, 1);

    std::vector<double> W(N), Tmid(N), nasa((size_t)N * 14), hc(N);
    for (int i = 0; i < N; ++i) {
        int k = perm[i];
        W[i] = m.W[k];
        Tmid[i] = m.Tmid[k];
        hc[i] = m.hc[k];
        for (int a = 0; a < 7; ++a) {
            nasa[(size_t)i * 14 + a] = m.nasa_lo[(size_t)k * 7 + a];
            nasa[(size_t)i * 14 + 7 + a] = m.nasa_hi[(size_t)k * 7 + a];
        }
    }

    // ---- third-body efficiency sets --------------------------------------------------------
    std::vector<double> ms_dflt, ms_val;
    std::vector<int> ms_ptr{0}, ms_idx;
    std::vector<std::vector<double>> set_rows;  // internal order
    std::vector<int> rx_mset(R, 0);
    for (int r = 0; r < R; ++r) {
        if (m.kind[r] == kElementary) continue;
        std::vector<double> row(N);
        for (int i = 0; i < N; ++i) row[i] = m.eff[(size_t)r * N + perm[i]];
        int s = -1;
        for (size_t q = 0; q < set_rows.size(); ++q)
            if (set_rows[q] == row) s = (int)q;
        if (s < 0) {
            s = (int)set_rows.size();
            set_rows.push_back(row);
            std::map<double, int> cnt;
            for (double e : row) ++cnt[e];
            double dflt = 1.0;
            int best = 0;
            for (auto& kv : cnt)
                if (kv.second > best || (kv.second == best && kv.first == 1.0)) best = kv.second, dflt = kv.first;
            ms_dflt.push_back(dflt);
            for (int i = 0; i < N; ++i)
                if (row[i] != dflt) ms_idx.push_back(i), ms_val.push_back(row[i] - dflt);
            ms_ptr.push_back((int)ms_idx.size());
        }
        rx_mset[r] = s;
    }
    const int n_msets = (int)set_rows.size();
    if (n_msets > 255) throw std::runtime_error("tables: more than 255 distinct efficiency sets");

    // ---- reaction descriptors, derivative slots, gather targets ---------------------------
    std::vector<U4> rdesc(R);
    std::vector<double> arrh((size_t)R * 3), fall, powtab;
    std::vector<GTarget> wt(NA);                       // wdot targets (one per active species)
    std::map<std::pair<int, int>, GTarget> jt;         // Jacobian targets keyed (col j, row i)
    int n_fall = 0, n_pow = 0, n_dq = 0;
    auto jac_add = [&](int i, int j, int src, double coef) {
        GTarget& t = jt[{j, i}];
        t.dst = j * ldj + i;
        t.e.push_back({src, coef});
    };
    // internal reaction order: falloff first, then three-body, then elementary, so that the
    // expensive (log10/exp10) falloff path is confined to the first warp iterations
    std::vector<int> rorder(R);
    for (int r = 0; r < R; ++r) rorder[r] = r;
    std::stable_sort(rorder.begin(), rorder.end(), [&](int a, int b) { return m.kind[a] > m.kind[b]; });
    for (int rn = 0; rn < R; ++rn) {
        const int r = rorder[rn];  // index into the Mechanism record; rn = internal index
        const int* ri = &m.r_idx[r * kMaxSide];
        const int* pi = &m.p_idx[r * kMaxSide];
        const double* rnu = &m.r_nu[r * kMaxSide];
        const double* rord = &m.r_ord[r * kMaxSide];
        const double* pnu = &m.p_nu[r * kMaxSide];
        const bool rev = m.reversible[r] != 0;
        const int kind = m.kind[r];
        // fast path?
        bool fast = true;
        int nrf = 0, npf = 0;
        for (int j = 0; j < kMaxSide; ++j) {
            if (ri[j] >= 0) {
                if (!is_int(rord[j]) || rord[j] < 0) fast = false;
                if (rev && rord[j] != rnu[j]) throw std::runtime_error("tables: reversible reaction with explicit orders");
                nrf += (int)std::lround(rord[j]);
            }
            if (pi[j] >= 0 && rev) {
                if (!is_int(pnu[j])) fast = false;
                npf += (int)std::lround(pnu[j]);
            }
        }
        if (nrf > 3 || npf > 3) fast = false;
        const uint8_t kNone = (uint8_t)N;  // slot N of the concentration arrays holds 1.0
        uint8_t rf[3] = {kNone, kNone, kNone}, pf[3] = {kNone, kNone, kNone};
        uint32_t flags = (uint32_t)kind | (rev ? kFlagRev : 0u);
        int pow_id = 0;
        if (fast) {
            int a = 0, b = 0;
            for (int j = 0; j < kMaxSide; ++j) {
                if (ri[j] >= 0)
                    for (int c = 0; c < (int)std::lround(rord[j]); ++c) rf[a++] = (uint8_t)inv[ri[j]];
                if (pi[j] >= 0 && rev)
                    for (int c = 0; c < (int)std::lround(pnu[j]); ++c) pf[b++] = (uint8_t)inv[pi[j]];
            }
        } else {
            flags |= kFlagPow;
            pow_id = n_pow++;
            powtab.resize((size_t)n_pow * 9, 0.0);
            for (int j = 0; j < kMaxSide; ++j) {
                if (ri[j] >= 0) {
                    rf[j] = (uint8_t)inv[ri[j]];
                    powtab[(size_t)pow_id * 9 + j] = rord[j];
                    powtab[(size_t)pow_id * 9 + 3 + j] = rnu[j];
                }
                if (pi[j] >= 0 && rev) {
                    pf[j] = (uint8_t)inv[pi[j]];
                    powtab[(size_t)pow_id * 9 + 6 + j] = pnu[j];
                }
            }
        }
        int fall_id = 0;
        if (kind == kFalloffLindemann || kind == kFalloffTroe) {
            fall_id = n_fall++;
            if (!(m.A0[r] > 0)) throw std::runtime_error("tables: falloff low-pressure A must be positive");
            const double* t = &m.troe[(size_t)r * 4];
            double rec[8] = {std::log(m.A0[r]), m.b0[r], m.Ta0[r], t[0], t[1], t[2], t[3],
                             kind == kFalloffTroe ? 1.0 : 0.0};
            fall.insert(fall.end(), rec, rec + 8);
        }
        if (m.A[r] < 0) throw std::runtime_error("tables: negative pre-exponential factor");
        arrh[(size_t)rn * 3 + 0] = m.A[r] > 0 ? std::log(m.A[r]) : -INFINITY;
        arrh[(size_t)rn * 3 + 1] = m.b[r];
        arrh[(size_t)rn * 3 + 2] = m.Ta[r];

        // net stoichiometry (internal active indices)
        std::map<int, double> net;
        for (int j = 0; j < kMaxSide; ++j) {
            if (ri[j] >= 0) net[inv[ri[j]]] -= rnu[j];
            if (pi[j] >= 0) net[inv[pi[j]]] += pnu[j];
        }
        for (auto& kv : net)
            if (kv.second != 0.0) wt[kv.first].e.push_back({rn, kv.second});

        // derivative slots + Jacobian entries
        const int dq_off = n_dq;
        std::vector<std::pair<int, int>> deps;  // (species j, slot)
        for (int a = 0; a < 3; ++a)
            if (rf[a] != kNone) deps.push_back({rf[a], n_dq++});
        for (int a = 0; a < 3; ++a)
            if (pf[a] != kNone) deps.push_back({pf[a], n_dq++});
        int mslot = -1;
        if (kind != kElementary) mslot = n_dq++;
        for (auto& kv : net) {
            if (kv.second == 0.0) continue;
            const int i = kv.first;
            for (auto& d : deps)
                if (d.first < NA) jac_add(i, d.first, d.second, kv.second);  // inert columns never move
            if (mslot >= 0) {
                const int s = rx_mset[r];
                if (ms_dflt[s] != 0.0) jac_add(i, NA, mslot, kv.second * ms_dflt[s]);  // gsum column
                for (int q = ms_ptr[s]; q < ms_ptr[s + 1]; ++q)
                    if (ms_idx[q] < NA) jac_add(i, ms_idx[q], mslot, kv.second * ms_val[q]);
            }
        }
        U4 d;
        if (fall_id > 0xFFF || pow_id > 0xFFF) throw std::runtime_error("tables: too many falloff / pow reactions");
        d.x = flags | ((uint32_t)rf[0] << 3) | ((uint32_t)rf[1] << 11) | ((uint32_t)rf[2] << 19);
        d.y = ((uint32_t)pf[0] << 3) | ((uint32_t)pf[1] << 11) | ((uint32_t)pf[2] << 19);
        d.z = (uint32_t)rx_mset[r] | ((uint32_t)fall_id << 8) | ((uint32_t)pow_id << 20);
        d.w = (uint32_t)dq_off;
        rdesc[rn] = d;
    }

    // ---- row programs ----------------------------------------------------------------------
    CoefTable ct;
    // third-body sets: M_s = dflt_s * ctot + sum_j (eff_sj - dflt_s) c_j ; ctot lives in slot N+1
    std::vector<U4> ms_ent;
    std::vector<uint32_t> ms_rng;
    auto ms_push = [&](double val, int idx) {
        uint32_t w[2];
        std::memcpy(w, &val, 8);
        ms_ent.push_back(U4{w[0], w[1], (uint32_t)idx * 8u, 0u});
    };
    for (int sidx = 0; sidx < n_msets; ++sidx) {
        const size_t first = ms_ent.size();
        if (ms_dflt[sidx] != 0.0) ms_push(ms_dflt[sidx], N + 1);
        for (int q = ms_ptr[sidx]; q < ms_ptr[sidx + 1]; ++q) ms_push(ms_val[q], ms_idx[q]);
        if (first > 0xFFFF) throw std::runtime_error("tables: efficiency table too large");
        ms_rng.push_back((uint32_t)first | ((uint32_t)(ms_ent.size() - first) << 16));
    }
    // (wdot schedule: built after the layout below, it needs the cell offsets)
    std::vector<TargetRows> jrows;
    for (auto& kv : jt) {
        ht.jac_entries += (int)kv.second.e.size();
        jrows.push_back(make_rows(kv.second, ct, n_dq));
    }
    HostRowProg progJ = pack_rows(jrows, (NA + 1) * ldj);

    // ---- per-cell shared-memory layout (doubles; every offset even for 16-byte access) -----
    DevMech& L = ht.layout;
    L.N = N; L.NA = NA; L.NP = NP; L.R = R; L.n_msets = n_msets; L.n_fall = n_fall; L.n_pow = n_pow;
    const int n_part = 4 * std::max(NA, 1);  // at most four partial sums per species
    L.n_dq = n_dq; L.n_part = n_part; L.ldj = ldj;
    int off = 0;
    auto take = [&](int n) { int o = off; off += (n + 1) & ~1; return o; };
    L.off_c = take(N + 2);
    L.off_cs = take(N + 2);
    L.off_kf = take(2 * R);
    L.off_k0 = take(n_fall);
    L.off_lfc = take(n_fall);
    L.off_M = 0;  // unused: every reaction lane evaluates its own third-body concentration
    L.off_q = take(R + 1);
    // dq is dead once the Jacobian row program has run (inside eval_wdot<true>, before the wdot
    // rows); the partial sums, the stage vector and the Gauss-Jordan pivot-row buffer reuse it
    const int dq_need = std::max(std::max(n_dq + 1, N), 2 * R + 2);  // also holds the (forward, reverse) rate pairs of rop_one_cell
    const int alias_need = ((std::max(n_part, 1) + 1) & ~1) + ((NP + 1) & ~1) + 2 * (NP + 6);  // row: 2 pivot records
    L.off_dq = take(std::max(dq_need, alias_need));
    {
        int o = L.off_dq;
        L.off_part = o; o += (std::max(n_part, 1) + 1) & ~1;
        L.off_vec = o; o += (NP + 1) & ~1;
        L.off_row = o;
    }
    L.off_J = take((NA + 1) * ldj + kTeam);  // + dump slots of the Jacobian row program
    L.off_red = take(kTeam);  // dump slots of the wdot schedule
    L.cell_doubles = off;

    // ---- wdot schedule (see WdotProg) ------------------------------------------------------
    std::vector<uint16_t> wtab;
    int wnA = 0, wnB = 0;
    {
        struct WRow { uint16_t s[4]; uint16_t dst; uint16_t flags; };
        const uint16_t zero = (uint16_t)(L.off_q + R);
        // rows of each species: [+,+,-,-]
        std::vector<std::vector<WRow>> srows(NA);
        for (int i = 0; i < NA; ++i) {
            std::vector<uint16_t> pos, neg;
            for (const GEntry& e : wt[i].e) {
                ht.wdot_entries += 1;
                const double ac = std::fabs(e.coef);
                if (ac != std::round(ac) || ac > 32.0)
                    throw std::runtime_error("tables: non-integer net stoichiometric coefficients are not supported");
                for (int k = 0; k < (int)ac; ++k) (e.coef > 0 ? pos : neg).push_back((uint16_t)(L.off_q + e.src));
            }
            const size_t nr = std::max((pos.size() + 1) / 2, (neg.size() + 1) / 2);
            for (size_t r2 = 0; r2 < nr; ++r2) {
                WRow w;
                for (int k = 0; k < 2; ++k) {
                    w.s[k] = 2 * r2 + k < pos.size() ? pos[2 * r2 + k] : zero;
                    w.s[2 + k] = 2 * r2 + k < neg.size() ? neg[2 * r2 + k] : zero;
                }
                w.dst = 0;
                w.flags = 0;
                srows[i].push_back(w);
            }
        }
        // split every species into <= 4 groups of consecutive rows; a group = one partial sum on one lane
        struct Group { int species, part; std::vector<WRow> rows; };
        std::vector<Group> groups;
        std::vector<std::vector<int>> parts_of(NA);
        for (int i = 0; i < NA; ++i) {
            // exactly four partial sums per species (all-zero rows pad the short ones): stage B is then
            // the fixed pattern w_i = part[4i] + part[4i+1] + part[4i+2] + part[4i+3]
            while (srows[i].size() < 4) {
                WRow w;
                for (int k = 0; k < 4; ++k) w.s[k] = zero;
                w.dst = 0;
                w.flags = 0;
                srows[i].push_back(w);
            }
            const int nr = (int)srows[i].size();
            const int ng = 4;
            for (int gI = 0; gI < ng; ++gI) {
                Group G;
                G.species = i;
                G.part = 4 * i + gI;
                const int lo = nr * gI / ng, hi = nr * (gI + 1) / ng;
                G.rows.assign(srows[i].begin() + lo, srows[i].begin() + hi);
                parts_of[i].push_back(G.part);
                groups.push_back(G);
            }
        }
        std::stable_sort(groups.begin(), groups.end(),
                         [](const Group& x, const Group& y) { return x.rows.size() > y.rows.size(); });
        std::vector<std::vector<WRow>> laneA(kTeam), laneB(kTeam);
        for (const Group& G : groups) {
            int best = 0;
            for (int l = 1; l < kTeam; ++l)
                if (laneA[l].size() < laneA[best].size()) best = l;
            for (size_t r2 = 0; r2 < G.rows.size(); ++r2) {
                WRow w = G.rows[r2];
                w.flags = r2 == 0 ? 1 : 0;
                w.dst = r2 + 1 == G.rows.size() ? (uint16_t)(L.off_part + G.part) : (uint16_t)(L.off_red + best);
                laneA[best].push_back(w);
            }
        }
        for (int l = 0; l < kTeam; ++l) {
            wnA = std::max(wnA, (int)laneA[l].size());
            wnB = std::max(wnB, (int)laneB[l].size());
        }
        if (wnA > kWdotRowsMax)
            throw std::runtime_error("tables: wdot schedule exceeds the register budget of the kernels");
        if (off > 0xFFFF) throw std::runtime_error("tables: per-cell block too large for 16-bit offsets");
        wtab.assign((size_t)(wnA + wnB) * kTeam * 6, 0);
        auto put = [&](int row, int l, const WRow& w) {
            uint16_t* t = &wtab[((size_t)row * kTeam + l) * 6];
            for (int k = 0; k < 4; ++k) t[k] = w.s[k];
            t[4] = w.dst;
            t[5] = w.flags;
        };
        WRow idle;
        for (int k = 0; k < 4; ++k) idle.s[k] = zero;
        idle.flags = 1;
        for (int l = 0; l < kTeam; ++l) {
            idle.dst = (uint16_t)(L.off_red + l);
            for (int r2 = 0; r2 < wnA; ++r2) put(r2, l, r2 < (int)laneA[l].size() ? laneA[l][r2] : idle);
            for (int r2 = 0; r2 < wnB; ++r2) put(wnA + r2, l, r2 < (int)laneB[l].size() ? laneB[l][r2] : idle);
        }
    }

    // ---- blob ------------------------------------------------------------------------------
    Blob B;
    ht.perm = perm;
    ht.rorder = rorder;
    // hot prefix (staged in shared memory by the kernels)
    L.rdesc = as_ptr<U4>(B.add(rdesc));
    L.coef = as_ptr<double>(B.add(ct.v));
    auto add_prog = [&](const HostRowProg& hp, RowProg& out) {
        out.nA = hp.nA;
        out.nB = hp.nB;
        out.ent = as_ptr<U4>(B.add(hp.ent));
        out.dst = as_ptr<uint32_t>(B.add(hp.dst));
    };
    L.ms_ent = as_ptr<U4>(B.add(ms_ent));
    L.ms_rng = as_ptr<uint32_t>(B.add(ms_rng));
    add_prog(progJ, L.jac);
    L.hot_base = as_ptr<uint8_t>(0);
    L.hot_bytes = (int)((B.b.size() + 15) & ~size_t(15));
    // cold tables (read once per cell from global memory)
    L.perm = as_ptr<int>(B.add(perm));
    L.W = as_ptr<double>(B.add(W));
    L.Tmid = as_ptr<double>(B.add(Tmid));
    L.nasa = as_ptr<double>(B.add(nasa));
    L.hc = as_ptr<double>(B.add(hc));
    L.arrh = as_ptr<double>(B.add(arrh));
    L.fall = as_ptr<double>(B.add(fall));
    L.powtab = as_ptr<double>(B.add(powtab));
    L.wdot.nA = wnA;
    L.wdot.nB = wnB;
    L.wdot.tab = as_ptr<uint16_t>(B.add(wtab));
    ht.blob = std::move(B.b);
    return ht;
}

}  // namespace dfchem
