// Mixture transport properties of correctThermo (dfChemistryModel.C:513-541): mu (Wilke rule on the
// fitted species viscosities), alpha = lambda_mix / cp, rhoD_i = rho * D_i^mix (or alpha for
// UnityLewis).  The arithmetic is the one Cantera's MixTransport evaluates from its polynomial
// fits in ln T, which the reference restates on the device in src_gpu/dfThermo.cu:54-71 (mole
// fractions), :73-86 (1, lnT .. lnT^4), :111-151 (viscosity), :172-206 (conductivity / cp),
// :208-255 (mixture-averaged diffusion); the fits themselves come in the reference's
// thermo_<mech>.txt layout (dfThermo.cu:361-435).  One launch replaces the reference's five
// kernels and their global scratch arrays (mole fractions, T polynomial, species viscosities and
// conductivities): per cell those live in registers and two shared-memory columns.
//
// Bound: FP64 CUDA cores (N^2 pair terms per cell: 11 DFMA-class operations and one reciprocal per
// diffusion pair, 4 per Wilke pair); HBM traffic is (N+4) reads + (N+2) writes of 8 bytes.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

namespace dfchem {

constexpr int kTransportBlock = 128;                       // 4 warps
constexpr int kTransportCells = 32;                        // cells per CTA pass: lane = cell
constexpr int kTransportSplit = kTransportBlock / kTransportCells;  // warps sharing one group of cells
constexpr int kDiffStride = 6;  // 5 coefficients + 1 pad: three aligned 16-byte loads per pair

// device table, caller's (Cantera) species order; all pointers into one allocation
struct TransportTab {
    int N = 0;
    const double* W = nullptr;      // [N]
    const double* visc = nullptr;   // [N][6]   sqrt(mu_i / sqrt(T)) = sum_j visc[i][j] lnT^j
    const double* cond = nullptr;   // [N][6]   lambda_i / sqrt(T)   = sum_j cond[i][j] lnT^j
    const double* wilke = nullptr;  // [N][N][2] {(1 + W_i/W_j)^-1/2 / sqrt(8), (W_j/W_i)^1/4}
    const double* diff = nullptr;   // [N][N][6] D_ij p / T^1.5      = sum_k diff[i][j][k] lnT^k
};

__device__ __forceinline__ double tr_poly(const double* __restrict__ c, double l1, double l2, double l3, double l4) {
    const double2 c01 = __ldg(reinterpret_cast<const double2*>(c));
    const double2 c23 = __ldg(reinterpret_cast<const double2*>(c) + 1);
    const double c4 = __ldg(c + 4);
    // the reference's summation order: ((((c0*1 + c1 l) + c2 l^2) + c3 l^3) + c4 l^4)
    return fma(c4, l4, fma(c23.y, l3, fma(c23.x, l2, fma(c01.y, l1, c01.x))));
}

// Mapping: a CTA pass handles 32 consecutive cells, lane = cell, and its 4 warps split the species
// index i (warp w owns i = w, w+4, ...): every table address is warp-uniform (one L1 wavefront per
// load), every global access is a coalesced 256-byte row, and a cell's N^2 pair terms run on four
// threads.  Shared memory per CTA: X[N][32] mole fractions, rsv[N][32] reciprocal sqrt-viscosities
// and 6 x 4 x 32 partial sums -- 2N+24 columns of 256 bytes, so ten or more CTAs fit an SM for
// gri30 and registers (<= 64 per thread) bound the occupancy, not shared memory.
__global__ void __launch_bounds__(kTransportBlock, 8)
transport_kernel(const TransportTab t, int n, const double* __restrict__ T, const double* __restrict__ p,
                 const double* __restrict__ rho, const double* __restrict__ cp, const double* __restrict__ Y,
                 int ldY, double* __restrict__ mu, double* __restrict__ alpha, double* __restrict__ rhoD,
                 int ldD, int unity_lewis) {
    extern __shared__ double2 dfc_tr_smem[];
    constexpr int C = kTransportCells, S = kTransportSplit;
    const int N = t.N;
    const int lane = threadIdx.x % C, w = threadIdx.x / C;
    double* X = reinterpret_cast<double*>(dfc_tr_smem) + lane;  // X[i * C]
    double* rsv = X + (size_t)N * C;                            // rsv[i * C]
    double* part = rsv + (size_t)N * C;                         // part[(q * S + w) * C], q < 6
    const int groups = (n + C - 1) / C;
    for (int g = blockIdx.x; g < groups; g += gridDim.x) {       // block-uniform trip count (barriers inside)
        const int c = g * C + lane;
        const bool live = c < n;
        const int cc = live ? c : n - 1;                         // tail lanes recompute the last cell, store nothing
        const double Tc = T[cc];
        const double l1 = log(Tc), l2 = l1 * l1, l3 = l1 * l2, l4 = l2 * l2;
        const double sqT = sqrt(Tc);
        // mole fractions and mean molecular weight (dfThermo.cu:54-71)
        double s = 0.0;
        for (int i = w; i < N; i += S) {
            const double y = Y[(size_t)i * ldY + cc];
            X[i * C] = y;
            s += y / __ldg(t.W + i);
        }
        part[w * C] = s;
        __syncthreads();
        s = 0.0;
#pragma unroll
        for (int k = 0; k < S; ++k) s += part[k * C];            // same order in every warp: identical bits
        double wbar = 0.0, sum_l = 0.0, sum_il = 0.0;
        for (int i = w; i < N; i += S) {
            const double wi = __ldg(t.W + i);
            const double x = X[i * C] / (wi * s);
            X[i * C] = x;
            wbar = fma(x, wi, wbar);
            rsv[i * C] = __drcp_rn(tr_poly(t.visc + i * kDiffStride, l1, l2, l3, l4));
            const double lam = tr_poly(t.cond + i * kDiffStride, l1, l2, l3, l4) * sqT;
            sum_l = fma(x, lam, sum_l);
            sum_il += x / lam;
        }
        part[(1 * S + w) * C] = wbar;
        part[(2 * S + w) * C] = sum_l;
        part[(3 * S + w) * C] = sum_il;
        __syncthreads();
        wbar = sum_l = sum_il = 0.0;
#pragma unroll
        for (int k = 0; k < S; ++k) {
            wbar += part[(1 * S + k) * C];
            sum_l += part[(2 * S + k) * C];
            sum_il += part[(3 * S + k) * C];
        }
        // alpha = lambda_mix / cp_mass (dfThermo.cu:172-206)
        double al = 0.0;
        if (alpha || unity_lewis) {
            al = 0.5 * (sum_l + 1.0 / sum_il) / cp[cc];
            if (alpha && live && w == 0) alpha[c] = al;
        }
        // Wilke mixture viscosity (dfThermo.cu:111-151): this warp's rows i
        if (mu) {
            double mix = 0.0;
            for (int i = w; i < N; i += S) {
                const double sv = tr_poly(t.visc + i * kDiffStride, l1, l2, l3, l4);
                const double2* wk = reinterpret_cast<const double2*>(t.wilke) + (size_t)i * N;
                double den = 0.0;
#pragma unroll 4
                for (int j = 0; j < N; ++j) {
                    const double2 k = __ldg(wk + j);
                    const double tmp = fma(sv * rsv[j * C], k.y, 1.0);
                    den = fma(X[j * C] * k.x, tmp * tmp, den);
                }
                mix += X[i * C] * (sv * sv) / den;
            }
            part[(4 * S + w) * C] = mix;
        }
        if (rhoD) {
            if (unity_lewis) {  // dfChemistryModel.C:521-527
                if (live)
                    for (int i = w; i < N; i += S) rhoD[(size_t)i * ldD + c] = al;
            } else {
                // mixture-averaged diffusion, mass-flux form (dfThermo.cu:208-255; getMixDiffCoeffsMass)
                const double powT = Tc * sqT, rho_div_p = rho[cc] / p[cc];
                for (int i = w; i < N; i += S) {
                    const double xi = X[i * C];
                    const double* dr = t.diff + (size_t)i * N * kDiffStride;
                    double s1 = 0.0, s2 = 0.0;
#pragma unroll 4
                    for (int j = 0; j < N; ++j) {
                        if (j == i) continue;
                        const double r = __drcp_rn(tr_poly(dr + j * kDiffStride, l1, l2, l3, l4) * powT);
                        const double xr = X[j * C] * r;
                        s1 += xr;
                        s2 = fma(xr, __ldg(t.W + j), s2);
                    }
                    s2 *= xi / (wbar - xi * __ldg(t.W + i));
                    if (live) rhoD[(size_t)i * ldD + c] = (xi + 1e-10 > 1.0) ? 0.0 : rho_div_p / (s1 + s2);
                }
            }
        }
        __syncthreads();  // mu partials complete; X / rsv / part free for the next group
        if (mu && w == 0 && live) {
            double mix = 0.0;
#pragma unroll
            for (int k = 0; k < S; ++k) mix += part[(4 * S + k) * C];
            mu[c] = mix * sqT;
        }
        __syncthreads();
    }
}

inline size_t transport_smem_bytes(int N) { return (size_t)(2 * N + 6 * kTransportSplit) * kTransportCells * sizeof(double); }

}  // namespace dfchem
