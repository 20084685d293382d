// Mechanism -> tables of the lane-per-cell kernel (lpc_tables.hpp).
#include "lpc_tables.hpp"

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
}  // namespace

LpcMech HostLpc::rebase(const void* base) const {
    LpcMech d = layout;
    d.base = static_cast<const uint8_t*>(base);
    d.reac = shift(d.reac, base);
    d.eff = shift(d.eff, base);
    d.fall = shift(d.fall, base);
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
    if (NA < 1 || NA > kLpcMaxActive || (int)perm.size() != N) return h;
    std::vector<int> inv(N, -1);
    for (int i = 0; i < N; ++i) inv[perm[i]] = i;

    std::vector<LpcReaction> reac(R);
    std::vector<std::vector<double>> sets;
    std::vector<double> fall;
    int n_fall = 0;
    for (int r = 0; r < R; ++r) {
        LpcReaction& x = reac[r];
        std::memset(&x, 0, sizeof(x));
        x.kind = m.kind[r];
        x.rev = m.reversible[r] ? 1 : 0;
        x.mset = x.fall = -1;
        x.caller_index = r;
        x.simple = 1;
        std::map<int, double> net;
        for (int j = 0; j < kMaxSide; ++j) {
            const int ri = m.r_idx[r * kMaxSide + j], pi = m.p_idx[r * kMaxSide + j];
            if (ri >= 0) {
                const double o = m.r_ord[r * kMaxSide + j];
                net[inv[ri]] -= m.r_nu[r * kMaxSide + j];
                if (x.rev && o != m.r_nu[r * kMaxSide + j]) return h;  // reversible with explicit orders: not representable
                if (o != 0.0) {  // a zero order contributes the factor 1 (Cantera's C_AnyN)
                    x.rsp[x.nr] = inv[ri];
                    x.rord[x.nr] = o;
                    ++x.nr;
                    if (!(o == 1.0 || o == 2.0 || o == 3.0)) x.simple = 0;
                }
            }
            if (pi >= 0) {
                const double nu = m.p_nu[r * kMaxSide + j];
                net[inv[pi]] += nu;
                if (x.rev) {
                    x.psp[x.np] = inv[pi];
                    x.pord[x.np] = nu;
                    ++x.np;
                    if (!(nu == 1.0 || nu == 2.0 || nu == 3.0)) x.simple = 0;
                }
            }
        }
        for (auto& kv : net) {
            if (kv.second == 0.0) continue;
            if (kv.first >= NA || x.nnet >= kLpcMaxNet) return h;
            x.nsp[x.nnet] = kv.first;
            x.nnu[x.nnet] = kv.second;
            ++x.nnet;
        }
        if (m.A[r] < 0) return h;
        x.lnA = m.A[r] > 0 ? std::log(m.A[r]) : -INFINITY;
        x.b = m.b[r];
        x.Ta = m.Ta[r];
        if (x.kind != kElementary) {
            std::vector<double> row(N);
            for (int i = 0; i < N; ++i) row[i] = m.eff[(size_t)r * N + perm[i]];
            int s = -1;
            for (size_t q = 0; q < sets.size(); ++q)
                if (sets[q] == row) s = (int)q;
            if (s < 0) {
                s = (int)sets.size();
                sets.push_back(row);
            }
            x.mset = s;
        }
        if (x.kind == kFalloffLindemann || x.kind == kFalloffTroe) {
            if (!(m.A0[r] > 0)) return h;
            const double* t = &m.troe[(size_t)r * 4];
            const double rec[8] = {std::log(m.A0[r]), m.b0[r], m.Ta0[r], t[0], t[1], t[2], t[3],
                                   x.kind == kFalloffTroe ? 1.0 : 0.0};
            fall.insert(fall.end(), rec, rec + 8);
            x.fall = n_fall++;
        }
    }
    std::vector<double> eff;
    for (auto& s : sets) eff.insert(eff.end(), s.begin(), s.end());
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
    L.N = N; L.NA = NA; L.R = R; L.n_msets = (int)sets.size(); L.n_fall = n_fall;
    int off = 0;
    auto take = [&](int n) { int o = off; off += n; return o; };
    L.off_c = take(N);
    L.off_cs = take(N);
    L.off_kf = take(R);
    L.off_rkc = take(R);
    L.off_k0 = take(n_fall);
    L.off_lfc = take(n_fall);
    L.off_f = take(NA);
    L.off_J = take(NA * NA);
    L.off_A = take(NA * NA);
    L.off_k = take(6 * NA);
    L.off_f0 = take(NA);
    L.off_dc = take(NA);
    L.off_x = take(std::max(NA, N));  // stage right-hand side; doubles as the g/RT scratch of the set-up
    L.lane_doubles = off;
    L.reac = reinterpret_cast<const LpcReaction*>(blob_add(h.blob, reac));
    L.eff = reinterpret_cast<const double*>(blob_add(h.blob, eff));
    L.fall = reinterpret_cast<const double*>(blob_add(h.blob, fall));
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
