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

constexpr int kTransportBlock = 128;
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

// thread per cell; shared memory: X[N][BLK] mole fractions, rsv[N][BLK] reciprocal sqrt-viscosities
__global__ void __launch_bounds__(kTransportBlock)
transport_kernel(const TransportTab t, int n, const double* __restrict__ T, const double* __restrict__ p,
                 const double* __restrict__ rho, const double* __restrict__ cp, const double* __restrict__ Y,
                 int ldY, double* __restrict__ mu, double* __restrict__ alpha, double* __restrict__ rhoD,
                 int ldD, int unity_lewis) {
    extern __shared__ double2 dfc_tr_smem[];
    constexpr int B = kTransportBlock;
    const int N = t.N;
    double* X = reinterpret_cast<double*>(dfc_tr_smem) + threadIdx.x;
    double* rsv = X + (size_t)N * B;
    // every thread touches only its own shared-memory column: no block barrier anywhere
    for (int c = blockIdx.x * B + threadIdx.x; c < n; c += gridDim.x * B) {
        const double Tc = T[c];
        const double l1 = log(Tc), l2 = l1 * l1, l3 = l1 * l2, l4 = l2 * l2;
        const double sqT = sqrt(Tc);
        // mole fractions and mean molecular weight (dfThermo.cu:54-71)
        double s = 0.0;
        for (int i = 0; i < N; ++i) {
            const double y = Y[(size_t)i * ldY + c];
            X[i * B] = y;
            s += y / __ldg(t.W + i);
        }
        double wbar = 0.0, sum_l = 0.0, sum_il = 0.0;
        for (int i = 0; i < N; ++i) {
            const double w = __ldg(t.W + i);
            const double x = X[i * B] / (w * s);
            X[i * B] = x;
            wbar = fma(x, w, wbar);
            rsv[i * B] = __drcp_rn(tr_poly(t.visc + i * kDiffStride, l1, l2, l3, l4));
            const double lam = tr_poly(t.cond + i * kDiffStride, l1, l2, l3, l4) * sqT;
            sum_l = fma(x, lam, sum_l);
            sum_il += x / lam;
        }
        // Wilke mixture viscosity (dfThermo.cu:111-151)
        if (mu) {
            double mix = 0.0;
            for (int i = 0; i < N; ++i) {
                const double sv = tr_poly(t.visc + i * kDiffStride, l1, l2, l3, l4);
                const double2* wk = reinterpret_cast<const double2*>(t.wilke) + (size_t)i * N;
                double den = 0.0;
#pragma unroll 4
                for (int j = 0; j < N; ++j) {
                    const double2 k = __ldg(wk + j);
                    const double tmp = fma(sv * rsv[j * B], k.y, 1.0);
                    den = fma(X[j * B] * k.x, tmp * tmp, den);
                }
                mix += X[i * B] * (sv * sv) / den;
            }
            mu[c] = mix * sqT;
        }
        // alpha = lambda_mix / cp_mass (dfThermo.cu:172-206)
        double al = 0.0;
        if (alpha || unity_lewis) {
            al = 0.5 * (sum_l + 1.0 / sum_il) / cp[c];
            if (alpha) alpha[c] = al;
        }
        if (!rhoD) continue;
        if (unity_lewis) {  // dfChemistryModel.C:521-527
            for (int i = 0; i < N; ++i) rhoD[(size_t)i * ldD + c] = al;
            continue;
        }
        // mixture-averaged diffusion, mass-flux form (dfThermo.cu:208-255; getMixDiffCoeffsMass)
        const double powT = Tc * sqT, rho_div_p = rho[c] / p[c];
        for (int i = 0; i < N; ++i) {
            const double xi = X[i * B];
            if (xi + 1e-10 > 1.0) {
                rhoD[(size_t)i * ldD + c] = 0.0;
                continue;
            }
            const double* dr = t.diff + (size_t)i * N * kDiffStride;
            double s1 = 0.0, s2 = 0.0;
#pragma unroll 4
            for (int j = 0; j < N; ++j) {
                if (j == i) continue;
                const double r = __drcp_rn(tr_poly(dr + j * kDiffStride, l1, l2, l3, l4) * powT);
                const double xr = X[j * B] * r;
                s1 += xr;
                s2 = fma(xr, __ldg(t.W + j), s2);
            }
            s2 *= xi / (wbar - xi * __ldg(t.W + i));
            rhoD[(size_t)i * ldD + c] = rho_div_p / (s1 + s2);
        }
    }
}

inline size_t transport_smem_bytes(int N) { return (size_t)2 * N * kTransportBlock * sizeof(double); }

}  // namespace dfchem
