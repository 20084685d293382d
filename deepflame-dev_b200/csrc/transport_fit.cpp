// Transport-fit generator of the host mechanism compiler: species viscosities, conductivities and
// binary diffusion coefficients from the Lennard-Jones / Stockmayer data in the mechanism file, as the 5-term
// polynomials in ln T that `dfThermo` reads from thermo_<mech>.txt (src_gpu/dfThermo.cu:361-435) and that the CPU
// path gets from Cantera at run time (viscosity(), thermalConductivity(), getMixDiffCoeffsMass,
// dfChemistryModel.C:513-541).  Restates Cantera 2.6 GasTransport::fitProperties / getTransport formulas:
//   eta_k   = 5/16 sqrt(pi m_k kB T) / (pi sigma_k^2 Omega22*(T*, delta*))               fitted as sqrt(eta)/T^(1/4)
//   D_ij p  = 3/16 sqrt(2 pi (kB T)^3 / m_ij) / (pi sigma_ij^2 Omega11*(T*, delta*))     fitted as D_ij p / T^(3/2)
//   lambda_k = eta_k/W_k R (f_trans 3/2 + f_rot c_rot + f_vib c_vib)   (modified Eucken, Parker rotational relaxation) fitted as lambda/sqrt(T)
// with the combination rules (polar - non-polar induction correction xi) of GasTransport::setupCollisionParameters,
// 50 temperatures between the phase's min and max temperature, relative least squares, degree 4.
// ONE deliberate difference: Cantera interpolates the Monchick-Mason (1961) tables of the reduced collision integrals,
// which are not in the reference tree; here Omega22* and Omega11* come from the Neufeld-Janzen-Aziz (1972) closed forms
// of the same Lennard-Jones integrals (0.1 % of the tables at delta* = 0) plus Brokaw's polar term 0.2 delta*^2/T*
// (0.19 for Omega11*).  Measured against the one Cantera-made file in the tree (ES80, tests/test_transport_fit.py):
// species viscosities within 0.5 %, conductivities 1 %, binary diffusion coefficients 1 % over 300-3000 K.
#include <cmath>
#include <stdexcept>
#include <vector>

#include "mechanism.hpp"
#include "transport_fit.hpp"

namespace dfchem {
namespace {
constexpr double kBoltz = 1.380649e-23, kAvogadro = 6.02214076e26, kEps0 = 8.8541878128e-12, kPi = 3.14159265358979323846;
constexpr double kDebye = 3.33564095198152e-30;  // C m

double omega22(double ts, double ds) {
    return 1.16145 * std::pow(ts, -0.14874) + 0.52487 * std::exp(-0.77320 * ts) + 2.16178 * std::exp(-2.43787 * ts) + 0.2 * ds * ds / ts;
}
double omega11(double ts, double ds) {
    return 1.06036 * std::pow(ts, -0.15610) + 0.19300 * std::exp(-0.47635 * ts) + 1.03587 * std::exp(-1.52996 * ts) +
           1.76474 * std::exp(-3.89411 * ts) + 0.19 * ds * ds / ts;
}
double frot(double ts) {
    return 1.0 + 0.5 * std::pow(kPi, 1.5) / std::sqrt(ts) + (0.25 * kPi * kPi + 2.0) / ts + std::pow(kPi, 1.5) / std::pow(ts, 1.5);
}

// weighted least squares of degree 4 in x: minimises sum w (y - p(x))^2; returns monomial coefficients c0..c4
void polyfit4(const std::vector<double>& x, const std::vector<double>& y, const std::vector<double>& w, double* c) {
    const int n = (int)x.size();
    long double xm = 0, xh = 0;
    for (double v : x) xm += v;
    xm /= n;
    for (double v : x) xh = std::max(xh, (long double)std::fabs(v - (double)xm));
    long double A[5][6] = {};
    for (int k = 0; k < n; ++k) {
        long double z = (x[k] - xm) / xh, p[5] = {1, z, z * z, z * z * z, z * z * z * z};
        for (int i = 0; i < 5; ++i) {
            for (int j = 0; j < 5; ++j) A[i][j] += w[k] * p[i] * p[j];
            A[i][5] += w[k] * p[i] * y[k];
        }
    }
    for (int i = 0; i < 5; ++i) {  // Gauss-Jordan with partial pivoting
        int piv = i;
        for (int r = i + 1; r < 5; ++r)
            if (fabsl(A[r][i]) > fabsl(A[piv][i])) piv = r;
        for (int j = 0; j < 6; ++j) std::swap(A[i][j], A[piv][j]);
        for (int r = 0; r < 5; ++r) {
            if (r == i) continue;
            const long double f = A[r][i] / A[i][i];
            for (int j = i; j < 6; ++j) A[r][j] -= f * A[i][j];
        }
    }
    long double a[5];
    for (int i = 0; i < 5; ++i) a[i] = A[i][5] / A[i][i];
    // p(z), z = (x - xm)/xh  ->  monomials in x
    long double out[5] = {0, 0, 0, 0, 0};
    static const int binom[5][5] = {{1, 0, 0, 0, 0}, {1, 1, 0, 0, 0}, {1, 2, 1, 0, 0}, {1, 3, 3, 1, 0}, {1, 4, 6, 4, 1}};
    for (int k = 0; k < 5; ++k) {
        const long double s = a[k] / powl(xh, k);
        for (int j = 0; j <= k; ++j) out[j] += s * binom[k][j] * powl(-xm, k - j);
    }
    for (int i = 0; i < 5; ++i) c[i] = (double)out[i];
}
}  // namespace

bool generate_transport_fits(const Mechanism& m, std::vector<double>& visc, std::vector<double>& cond, std::vector<double>& diff) {
    const int N = m.N;
    if ((int)m.trans.size() != N * 6) return false;
    std::vector<double> mass(N), eps(N), sig(N), dip(N), alpha(N), zrot(N), crot(N);
    double Tmin = 0.0, Tmax = 1e30;
    for (int k = 0; k < N; ++k) {
        const double* t = &m.trans[(size_t)k * 6];
        crot[k] = t[0] == 0.0 ? 0.0 : (t[0] == 1.0 ? 1.0 : 1.5);
        eps[k] = t[1] * kBoltz;
        sig[k] = t[2] * 1e-10;
        dip[k] = t[3] * kDebye;
        alpha[k] = t[4] * 1e-30;
        zrot[k] = t[5];
        mass[k] = m.W[k] / kAvogadro;
        if (!(eps[k] > 0) || !(sig[k] > 0)) return false;
        Tmin = std::max(Tmin, m.Tlo[k]);
        Tmax = std::min(Tmax, m.Thi[k]);
    }
    if (!(Tmax > Tmin)) return false;
    // pair parameters (GasTransport::setupCollisionParameters)
    std::vector<double> mred((size_t)N * N), dij((size_t)N * N), eij((size_t)N * N), del((size_t)N * N);
    for (int i = 0; i < N; ++i)
        for (int j = 0; j < N; ++j) {
            double d = 0.5 * (sig[i] + sig[j]), e = std::sqrt(eps[i] * eps[j]);
            const bool pi_ = dip[i] > 0, pj = dip[j] > 0;
            if (pi_ != pj) {  // polar - non-polar: induction correction
                const int kp = pi_ ? i : j, kn = pi_ ? j : i;
                const double d3np = std::pow(sig[kn], 3), d3pl = std::pow(sig[kp], 3);
                const double alpha_star = alpha[kn] / d3np;
                const double mu_p_star = dip[kp] / std::sqrt(4 * kPi * kEps0 * d3pl * eps[kp]);
                const double xi = 1.0 + 0.25 * alpha_star * mu_p_star * mu_p_star * std::sqrt(eps[kp] / eps[kn]);
                d *= std::pow(xi, -1.0 / 6.0);
                e *= xi * xi;
            }
            mred[(size_t)i * N + j] = mass[i] * mass[j] / (mass[i] + mass[j]);
            dij[(size_t)i * N + j] = d;
            eij[(size_t)i * N + j] = e;
            del[(size_t)i * N + j] = 0.5 * dip[i] * dip[j] / (4 * kPi * kEps0 * e * d * d * d);
        }
    const int np = 50;
    std::vector<double> tlog(np), Ts(np), y(np), w(np), y2(np), w2(np);
    for (int n = 0; n < np; ++n) {
        Ts[n] = Tmin + (Tmax - Tmin) * n / (np - 1);
        tlog[n] = std::log(Ts[n]);
    }
    visc.assign((size_t)N * 5, 0.0);
    cond.assign((size_t)N * 5, 0.0);
    diff.assign((size_t)N * N * 5, 0.0);
    for (int k = 0; k < N; ++k) {
        const size_t kk = (size_t)k * N + k;
        const double ts298 = kBoltz * 298.0 / eps[k], fz298 = frot(ts298);
        for (int n = 0; n < np; ++n) {
            const double t = Ts[n], sqT = std::sqrt(t);
            const double* a = t <= m.Tmid[k] ? &m.nasa_lo[(size_t)k * 7] : &m.nasa_hi[(size_t)k * 7];
            const double cp_R = a[0] + t * (a[1] + t * (a[2] + t * (a[3] + t * a[4])));
            const double tstar = kBoltz * t / eps[k];
            const double om22 = omega22(tstar, del[kk]), om11 = omega11(tstar, del[kk]);
            const double dself = 3.0 / 16.0 * std::sqrt(2.0 * kPi / mred[kk]) * std::pow(kBoltz * t, 1.5) / (kPi * sig[k] * sig[k] * om11);
            const double eta = 5.0 / 16.0 * std::sqrt(kPi * mass[k] * kBoltz * t) / (om22 * kPi * sig[k] * sig[k]);
            const double f_int = m.W[k] / (kGasConstant * t) * dself / eta;
            const double A = 2.5 - f_int, B = zrot[k] * fz298 / frot(tstar) + 2.0 / kPi * (5.0 / 3.0 * crot[k] + f_int);
            const double c1 = 2.0 / kPi * A / B;
            const double cv_int = cp_R - 2.5 - crot[k];
            const double f_rot = f_int * (1.0 + c1), f_trans = 2.5 * (1.0 - c1 * crot[k] / 1.5);
            const double lam = eta / m.W[k] * kGasConstant * (f_trans * 1.5 + f_rot * crot[k] + f_int * cv_int);
            y[n] = std::sqrt(eta / sqT);          // sqrt(eta)/T^(1/4)
            w[n] = 1.0 / (y[n] * y[n]);
            y2[n] = lam / sqT;
            w2[n] = 1.0 / (y2[n] * y2[n]);
        }
        polyfit4(tlog, y, w, &visc[(size_t)k * 5]);
        polyfit4(tlog, y2, w2, &cond[(size_t)k * 5]);
    }
    for (int i = 0; i < N; ++i)
        for (int j = i; j < N; ++j) {
            const size_t ij = (size_t)i * N + j;
            for (int n = 0; n < np; ++n) {
                const double t = Ts[n], tstar = kBoltz * t / eij[ij];
                const double d = 3.0 / 16.0 * std::sqrt(2.0 * kPi / mred[ij]) * std::pow(kBoltz * t, 1.5) /
                                 (kPi * dij[ij] * dij[ij] * omega11(tstar, del[ij]));
                y[n] = d / std::pow(t, 1.5);
                w[n] = 1.0 / (y[n] * y[n]);
            }
            polyfit4(tlog, y, w, &diff[ij * 5]);
            for (int c = 0; c < 5; ++c) diff[((size_t)j * N + i) * 5 + c] = diff[ij * 5 + c];
        }
    return true;
}

}  // namespace dfchem
