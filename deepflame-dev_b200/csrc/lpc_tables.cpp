// Mechanism -> tables of the lane-per-cell kernel (lpc_tables.hpp).
#include "lpc_tables.hpp"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <map>
#include <stdexcept>

namespace dfchem {
namespace {
template <class T>
size_t blob_add(std::vector<uint8_t>& b, const std::vector<T>& v) {
    size_t off = (b.size() + 15) & ~size_t(15);
    b.resize(off + std::max<size_t>(v.size(), 1) * sizeof(T), 0);
    if (!v.empty()) std::memcpy(b.data() + off, v.data(), v.size() * sizeof(T));
    return off;
}
template <class T>
const T* shift(const T* off, const void* base) {
    return reinterpret_cast<const T*>(reinterpret_cast<const uint8_t*>(base) + reinterpret_cast<size_t>(off));
}
bool is_int(double x) { return std::fabs(x - std::round(x)) < 1e-12; }
}  // namespace

LpcMech HostLpc::rebase(const void* base) const {
    LpcMech d = layout;
    d.base = static_cast<const uint8_t*>(base);
    d.rx = shift(d.rx, base);
    d.powtab = shift(d.powtab, base);
    d.eff = shift(d.eff, base);
    d.fall = shift(d.fall, base);
    d.arrh = shift(d.arrh, base);
    d.fidx = shift(d.fidx, base);
    d.fcoef = shift(d.fcoef, base);
    d.jrow = shift(d.jrow, base);
    d.jidx = shift(d.jidx, base);
    d.jcoef = shift(d.jcoef, base);
    d.gptr = shift(d.gptr, base);
    d.gidx = shift(d.gidx, base);
    d.gcoef = shift(d.gcoef, base);
    d.W = shift(d.W, base);
    d.Tmid = shift(d.Tmid, base);
    d.nasa = shift(d.nasa, base);
    d.hc = shift(d.hc, base);
    d.perm = shift(d.perm, base);
    return d;
}

HostLpc build_lpc_tables(const Mechanism& m, const std::vector<int>& perm, int NA) {
    HostLpc h;
    const int N = m.N, R = m.R;
    if (NA < 1 || NA > kLpcMaxActive || N > kLpcMaxSpecies || R > 0xFFFF || (int)perm.size() != N) return h;
    const int NAP = lpc_pad(NA);
    std::vector<int> inv(N, -1);
    for (int i = 0; i < N; ++i) inv[perm[i]] = i;

    std::vector<LpcRx> rx(R);
    std::vector<std::vector<double>> sets;
    std::vector<double> fall, powtab, arrh((size_t)R * 3);
    std::vector<std::vector<std::pair<int, double>>> flist(NAP), jlist((size_t)NAP * NAP);
    std::vector<int> gptr{0}, gidx;
    std::vector<double> gcoef;
    int n_fall = 0, n_pow = 0, n_dq = 0;
    for (int r = 0; r < R; ++r) {
        const int* ri = &m.r_idx[r * kMaxSide];
        const int* pi = &m.p_idx[r * kMaxSide];
        const double* rnu = &m.r_nu[r * kMaxSide];
        const double* rord = &m.r_ord[r * kMaxSide];
        const double* pnu = &m.p_nu[r * kMaxSide];
        const bool rev = m.reversible[r] != 0;
        const int kind = m.kind[r];
        bool fast = true;
        int nrf = 0, npf = 0;
        for (int j = 0; j < kMaxSide; ++j) {
            if (ri[j] >= 0) {
                if (!is_int(rord[j]) || rord[j] < 0) fast = false;
                if (rev && rord[j] != rnu[j]) return h;  // reversible with explicit orders: not representable
                nrf += (int)std::lround(rord[j]);
            }
            if (pi[j] >= 0 && rev) {
                if (!is_int(pnu[j])) fast = false;
                npf += (int)std::lround(pnu[j]);
            }
        }
        if (nrf > 3 || npf > 3) fast = false;
        const uint32_t kNone = (uint32_t)N;
        uint32_t rf[3] = {kNone, kNone, kNone}, pf[3] = {kNone, kNone, kNone};
        int pow_id = 0;
        if (fast) {
            int a = 0, b = 0;
            for (int j = 0; j < kMaxSide; ++j) {
                if (ri[j] >= 0)
                    for (int c = 0; c < (int)std::lround(rord[j]); ++c) rf[a++] = (uint32_t)inv[ri[j]];
                if (pi[j] >= 0 && rev)
                    for (int c = 0; c < (int)std::lround(pnu[j]); ++c) pf[b++] = (uint32_t)inv[pi[j]];
            }
        } else {
            pow_id = n_pow++;
            if (pow_id > 0xFF) return h;
            powtab.resize((size_t)n_pow * 9, 0.0);
            for (int j = 0; j < kMaxSide; ++j) {
                if (ri[j] >= 0 && rord[j] != 0.0) {  // a zero order contributes the factor 1 (Cantera's C_AnyN)
                    rf[j] = (uint32_t)inv[ri[j]];
                    powtab[(size_t)pow_id * 9 + j] = rord[j];
                }
                if (pi[j] >= 0 && rev) {
                    pf[j] = (uint32_t)inv[pi[j]];
                    powtab[(size_t)pow_id * 9 + 6 + j] = pnu[j];
                }
            }
        }
        int fall_id = 0, mset = 0xFF;
        if (kind != kElementary) {
            std::vector<double> row(N);
            for (int i = 0; i < N; ++i) row[i] = m.eff[(size_t)r * N + perm[i]];
            int s = -1;
            for (size_t q = 0; q < sets.size(); ++q)
                if (sets[q] == row) s = (int)q;
            if (s < 0) {
                s = (int)sets.size();
                sets.push_back(row);
            }
            if (s >= 0xFF) return h;
            mset = s;
        }
        if (kind == kFalloffLindemann || kind == kFalloffTroe) {
            if (!(m.A0[r] > 0) || n_fall > 0xFF) return h;
            const double* t = &m.troe[(size_t)r * 4];
            const double rec[8] = {std::log(m.A0[r]), m.b0[r], m.Ta0[r], t[0], t[1], t[2], t[3], kind == kFalloffTroe ? 1.0 : 0.0};
            fall.insert(fall.end(), rec, rec + 8);
            fall_id = n_fall++;
        }
        if (m.A[r] < 0) return h;
        arrh[(size_t)r * 3 + 0] = m.A[r] > 0 ? std::log(m.A[r]) : -INFINITY;
        arrh[(size_t)r * 3 + 1] = m.b[r];
        arrh[(size_t)r * 3 + 2] = m.Ta[r];

        std::map<int, double> net;
        for (int j = 0; j < kMaxSide; ++j) {
            if (ri[j] >= 0) net[inv[ri[j]]] -= rnu[j];
            if (pi[j] >= 0) net[inv[pi[j]]] += pnu[j];
        }
        for (auto& kv : net) {
            if (kv.second == 0.0) continue;
            if (kv.first >= NA) return h;
            flist[kv.first].push_back({r, kv.second});
            gidx.push_back(kv.first);
            gcoef.push_back(kv.second);
        }
        gptr.push_back((int)gidx.size());

        // derivative slots and Jacobian entries
        const int dq_off = n_dq;
        std::vector<std::pair<int, int>> deps;  // (species j, slot)
        for (int a = 0; a < 3; ++a)
            if (rf[a] != kNone) deps.push_back({(int)rf[a], n_dq++});
        for (int a = 0; a < 3; ++a)
            if (pf[a] != kNone) deps.push_back({(int)pf[a], n_dq++});
        int mslot = -1;
        if (kind != kElementary) mslot = n_dq++;
        for (auto& kv : net) {
            if (kv.second == 0.0) continue;
            const int i = kv.first;
            for (auto& d : deps)
                if (d.first < NA) jlist[(size_t)i * NAP + d.first].push_back({d.second, kv.second});
            if (mslot >= 0)
                for (int j = 0; j < NA; ++j)
                    if (sets[mset][j] != 0.0) jlist[(size_t)i * NAP + j].push_back({mslot, kv.second * sets[mset][j]});
        }
        LpcRx& x = rx[r];
        x.x = (uint32_t)kind | (rev ? 4u : 0u) | (fast ? 0u : 8u) | (rf[0] << 8) | (rf[1] << 14) | (rf[2] << 20);
        x.y = pf[0] | (pf[1] << 6) | (pf[2] << 12) | ((uint32_t)mset << 20);
        x.z = (uint32_t)fall_id | ((uint32_t)pow_id << 8) | ((uint32_t)r << 16);
        x.w = (uint32_t)dq_off;
    }
    if (n_dq > 0xFFFF) return h;
    std::vector<double> eff;
    for (auto& s : sets) eff.insert(eff.end(), s.begin(), s.end());
    // ELL form of the gather lists (lpc_tables.hpp): `rows` x NAP entries, padded with (zero slot, 0.0)
    const int q_zero = std::max(R, N);
    auto ell = [&](const std::vector<std::pair<int, double>>* lists, int zero, std::vector<uint32_t>& idx,
                   std::vector<double>& coef) {
        size_t rows = 0;
        for (int t = 0; t < NAP; ++t) rows = std::max(rows, lists[t].size());
        for (size_t k = 0; k < rows; ++k) {
            for (int t = 0; t < NAP; t += 2) {
                uint32_t w = 0;
                for (int u = 0; u < 2; ++u) {
                    const bool has = k < lists[t + u].size();
                    w |= (uint32_t)(has ? lists[t + u][k].first : zero) << (16 * u);
                    coef.push_back(has ? lists[t + u][k].second : 0.0);
                }
                idx.push_back(w);
            }
        }
        return (int)rows;
    };
    std::vector<uint32_t> fidx, jidx;
    std::vector<double> fcoef, jcoef;
    std::vector<int> jrow{0};
    if (q_zero > 0xFFFF) return h;
    const int f_rows = ell(flist.data(), q_zero, fidx, fcoef);
    for (int i = 0; i < NAP; ++i) jrow.push_back(jrow.back() + ell(jlist.data() + (size_t)i * NAP, n_dq, jidx, jcoef));
    std::vector<double> W(N), Tmid(N), nasa((size_t)N * 14), hc(N);
    for (int i = 0; i < N; ++i) {
        const int k = perm[i];
        W[i] = m.W[k];
        Tmid[i] = m.Tmid[k];
        hc[i] = m.hc[k];
        for (int a = 0; a < 7; ++a) {
            nasa[(size_t)i * 14 + a] = m.nasa_lo[(size_t)k * 7 + a];
            nasa[(size_t)i * 14 + 7 + a] = m.nasa_hi[(size_t)k * 7 + a];
        }
    }

    LpcMech& L = h.layout;
    L.N = N; L.NA = NA; L.NAP = NAP; L.R = R; L.n_msets = (int)sets.size(); L.n_fall = n_fall; L.n_dq = n_dq;
    // variable region after the fixed one; the derivative column is only alive during the base evaluation, when
    // the LU factors and the stage vectors are dead: it aliases them when it fits
    int off = lpc_fixed_doubles(NAP);
    auto take = [&](int n) { int o = off; off += n; return o; };
    L.off_c = take(N + 1);      // slot N holds 1.0 (absent factor)
    L.off_cs = take(N + 1);
    L.off_kf = take(R);
    L.off_rkc = take(R);
    L.off_k0 = take(n_fall);
    L.off_lfc = take(n_fall);
    L.off_M = take((int)sets.size());   // third-body concentration of every efficiency set, per evaluation
    L.off_q = take(std::max(R, N) + 1);   // doubles as the g/RT scratch of the set-up; last slot: 0.0 (list padding)
    L.f_rows = f_rows;
    L.q_zero = q_zero;
    L.off_dq = n_dq + 1 <= lpc_alias_room(NAP) ? NAP * NAP : take(n_dq + 1);   // slot n_dq: 0.0 (list padding)
    L.lane_doubles = off;
    L.rx = reinterpret_cast<const LpcRx*>(blob_add(h.blob, rx));
    L.powtab = reinterpret_cast<const double*>(blob_add(h.blob, powtab));
    L.eff = reinterpret_cast<const double*>(blob_add(h.blob, eff));
    L.fall = reinterpret_cast<const double*>(blob_add(h.blob, fall));
    L.arrh = reinterpret_cast<const double*>(blob_add(h.blob, arrh));
    L.fidx = reinterpret_cast<const uint32_t*>(blob_add(h.blob, fidx));
    L.fcoef = reinterpret_cast<const double*>(blob_add(h.blob, fcoef));
    L.jrow = reinterpret_cast<const int*>(blob_add(h.blob, jrow));
    L.jidx = reinterpret_cast<const uint32_t*>(blob_add(h.blob, jidx));
    L.jcoef = reinterpret_cast<const double*>(blob_add(h.blob, jcoef));
    L.gptr = reinterpret_cast<const int*>(blob_add(h.blob, gptr));
    L.gidx = reinterpret_cast<const int*>(blob_add(h.blob, gidx));
    L.gcoef = reinterpret_cast<const double*>(blob_add(h.blob, gcoef));
    L.W = reinterpret_cast<const double*>(blob_add(h.blob, W));
    L.Tmid = reinterpret_cast<const double*>(blob_add(h.blob, Tmid));
    L.nasa = reinterpret_cast<const double*>(blob_add(h.blob, nasa));
    L.hc = reinterpret_cast<const double*>(blob_add(h.blob, hc));
    L.perm = reinterpret_cast<const int*>(blob_add(h.blob, perm));
    h.blob.resize((h.blob.size() + 15) & ~size_t(15), 0);
    L.bytes = (int)h.blob.size();
    L.base = nullptr;
    h.usable = true;
    return h;
}

}  // namespace dfchem
