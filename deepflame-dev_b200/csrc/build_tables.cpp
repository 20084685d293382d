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

// Longest-processing-time bin packing of whole targets onto kTeam lanes; returns the
// lane-blocked program (see GatherProg).
std::vector<uint32_t> pack_program(std::vector<GTarget> targets, CoefTable& ct, int& n_iter) {
    std::stable_sort(targets.begin(), targets.end(),
                     [](const GTarget& a, const GTarget& b) { return a.e.size() > b.e.size(); });
    std::vector<std::vector<uint32_t>> lane(kTeam);
    for (const GTarget& t : targets) {
        if (t.e.empty()) continue;
        int best = 0;
        for (int l = 1; l < kTeam; ++l)
            if (lane[l].size() < lane[best].size()) best = l;
        if (t.dst < 0 || t.dst > 0xFFF) throw std::runtime_error("tables: gather destination out of range");
        for (size_t i = 0; i < t.e.size(); ++i) {
            if (t.e[i].src < 0 || t.e[i].src > 0xFFF) throw std::runtime_error("tables: gather source out of range");
            lane[best].push_back((uint32_t)t.e[i].src | ((uint32_t)ct.id(t.e[i].coef) << 12) |
                                 ((i + 1 == t.e.size()) ? (1u << 19) : 0u) | ((uint32_t)t.dst << 20));
        }
    }
    n_iter = 0;
    for (auto& l : lane) n_iter = std::max<int>(n_iter, (int)l.size());
    n_iter = (n_iter + 3) & ~3;
    std::vector<uint32_t> out((size_t)n_iter * kTeam, 0u);
    for (int l = 0; l < kTeam; ++l)
        for (size_t i = 0; i < lane[l].size(); ++i) out[(i / 4) * 4 * kTeam + l * 4 + i % 4] = lane[l][i];
    return out;
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
    static const int sizes[] = {8, 16, 20, 24, 32, 34, 40, 48, 64};
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
    d.ms_dflt = shift(d.ms_dflt, base);
    d.ms_ptr = shift(d.ms_ptr, base);
    d.ms_idx = shift(d.ms_idx, base);
    d.ms_val = shift(d.ms_val, base);
    d.coef = shift(d.coef, base);
    d.wdotA.ent = shift(d.wdotA.ent, base);
    d.wdotB.ent = shift(d.wdotB.ent, base);
    d.jac.ent = shift(d.jac.ent, base);
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
        uint8_t rf[3] = {255, 255, 255}, pf[3] = {255, 255, 255};
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
            if (rf[a] != 255) deps.push_back({rf[a], n_dq++});
        for (int a = 0; a < 3; ++a)
            if (pf[a] != 255) deps.push_back({pf[a], n_dq++});
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
        d.x = rf[0] | (rf[1] << 8) | (rf[2] << 16) | ((uint32_t)pf[0] << 24);
        d.y = pf[1] | (pf[2] << 8) | (flags << 16) | ((uint32_t)rx_mset[r] << 24);
        d.z = (uint32_t)fall_id | ((uint32_t)pow_id << 16);
        d.w = (uint32_t)dq_off;
        rdesc[rn] = d;
    }

    // ---- gather programs -------------------------------------------------------------------
    CoefTable ct;
    // wdot stage A: split each species' entries into chunks so lanes are balanced
    int total = 0;
    for (auto& t : wt) total += (int)t.e.size();
    ht.wdot_entries = total;
    const int chunk = std::max(4, (total + 2 * kTeam - 1) / (2 * kTeam));
    std::vector<GTarget> partA, tgtB(NA);
    int n_part = 0;
    for (int i = 0; i < NA; ++i) {
        tgtB[i].dst = i;
        for (size_t a = 0; a < wt[i].e.size(); a += chunk) {
            GTarget p;
            p.dst = n_part;
            p.e.assign(wt[i].e.begin() + a, wt[i].e.begin() + std::min(wt[i].e.size(), a + chunk));
            partA.push_back(p);
            tgtB[i].e.push_back({n_part, 1.0});
            ++n_part;
        }
        if (tgtB[i].e.empty()) {  // active species with zero net production everywhere
            tgtB[i].e.push_back({0, 0.0});
        }
    }
    int itA = 0, itB = 0, itJ = 0;
    std::vector<uint32_t> progA = pack_program(partA, ct, itA);
    // stage B: coefficient 0.0 entries must still store -> give them a real id
    std::vector<uint32_t> progB = pack_program(tgtB, ct, itB);
    std::vector<GTarget> jtargets;
    for (auto& kv : jt) {
        jtargets.push_back(kv.second);
        ht.jac_entries += (int)kv.second.e.size();
    }
    std::vector<uint32_t> progJ = pack_program(jtargets, ct, itJ);

    // ---- per-cell shared-memory layout (doubles; every offset even for 16-byte access) -----
    DevMech& L = ht.layout;
    L.N = N; L.NA = NA; L.NP = NP; L.R = R; L.n_msets = n_msets; L.n_fall = n_fall; L.n_pow = n_pow;
    L.n_dq = n_dq; L.n_part = n_part; L.ldj = ldj;
    int off = 0;
    auto take = [&](int n) { int o = off; off += (n + 1) & ~1; return o; };
    L.off_c = take(N);
    L.off_cs = take(N);
    L.off_kf = take(R);
    L.off_rkc = take(R);
    L.off_k0 = take(n_fall);
    L.off_lfc = take(n_fall);
    L.off_M = take(n_msets);
    L.off_q = take(R);
    L.off_dq = take(std::max(n_dq, N));  // also used as g[N] scratch during set-up
    L.off_J = take((NA + 1) * ldj);
    L.off_vec = take(NP);
    L.off_row = take(2 * NP);
    L.off_part = take(std::max(n_part, 1));
    L.off_red = take(kTeam);
    L.cell_doubles = off;

    // ---- blob ------------------------------------------------------------------------------
    Blob B;
    ht.perm = perm;
    // hot prefix (staged in shared memory by the kernels)
    L.rdesc = as_ptr<U4>(B.add(rdesc));
    L.coef = as_ptr<double>(B.add(ct.v));
    L.ms_dflt = as_ptr<double>(B.add(ms_dflt));
    L.ms_val = as_ptr<double>(B.add(ms_val));
    L.ms_ptr = as_ptr<int>(B.add(ms_ptr));
    L.ms_idx = as_ptr<int>(B.add(ms_idx));
    L.wdotA.n_iter = itA;
    L.wdotA.ent = as_ptr<uint32_t>(B.add(progA));
    L.wdotB.n_iter = itB;
    L.wdotB.ent = as_ptr<uint32_t>(B.add(progB));
    L.jac.n_iter = itJ;
    L.jac.ent = as_ptr<uint32_t>(B.add(progJ));
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
    ht.blob = std::move(B.b);
    return ht;
}

}  // namespace dfchem
