/* TEST / BASELINE INFRASTRUCTURE (oracle) -- never linked into libdfchem_b200.so.
 *
 * Plain-C restatement of the reference's CPU path for one chemistry sub-step, used (a) as the
 * timed CPU baseline of bench.py (`cpu_baseline`, `--impl reference`) and (b) as a second,
 * independent checker beside the NumPy oracle.  Per cell it does the work
 * dfChemistryModel::solveSingle (src/dfChemistryModel/dfChemistryModel.C:738-780) makes Cantera
 * and CVODE do: fresh integrator, normalise Y (setState_TPY :755), integrate
 * dY_k/dt = wdot_k W_k / rho at frozen (T, rho) with a variable-order (1-5) BDF/NDF method,
 * Newton iteration, dense LU with partial pivoting and a finite-difference Jacobian (the CVODE
 * configuration ReactorNet uses), RMS error norm with rtol*|Y|+atol weights,
 * then RR_i = (Ynew_i - Yraw_i)/dt*rho_OF and Qdot = -sum hc_i RR_i (:768-772).
 * The step/order controller follows the published quasi-constant-step NDF algorithm of
 * Shampine & Reichelt ("The MATLAB ODE suite", 1997).  Cantera and SUNDIALS are absent third-party
 * dependencies (libcantera 2.6), so this is a port, not the reference binary: "parity unpinned".
 * Rate laws: SURVEY.md Appendix A.2-A.7.  Input: the DFCHEMB1 mechanism blob.
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define RGAS 8314.46261815324
#define ONE_ATM 101325.0
#define MAXS 3
#define SMALL 1e-300
#define BIG 1e300

typedef struct {
    int N, R, ne;
    const double *W, *Tmid, *lo, *hi, *hc;
    const int32_t *kind, *rev, *ridx, *pidx;
    const double *A, *b, *Ta, *A0, *b0, *Ta0, *troe, *rnu, *rord, *pnu, *eff;
    int *has_eff; /* [R] 1 if any efficiency != 1 */
    uint8_t* copy;
} Mech;

Mech* oc_load(const void* blob, size_t len) {
    if (len < 24 || memcmp(blob, "DFCHEMB1", 8)) return NULL;
    Mech* m = (Mech*)calloc(1, sizeof(Mech));
    m->copy = (uint8_t*)malloc(len);
    memcpy(m->copy, blob, len);
    const int32_t* h = (const int32_t*)(m->copy + 8);
    m->N = h[0]; m->R = h[1]; m->ne = h[2];
    if (h[3] != MAXS) { free(m->copy); free(m); return NULL; }
    const int N = m->N, R = m->R;
    const uint8_t* p = m->copy + 24 + 32 * (size_t)N + 8 * (size_t)m->ne;
#define TAKE_D(name, n) m->name = (const double*)p; p += 8 * (size_t)(n)
#define TAKE_I(name, n) m->name = (const int32_t*)p; p += 4 * (size_t)(n)
    p += 8 * (size_t)N * m->ne; /* comp */
    TAKE_D(W, N); p += 8 * (size_t)N; /* Tlo */
    TAKE_D(Tmid, N); p += 8 * (size_t)N; /* Thi */
    TAKE_D(lo, 7 * N); TAKE_D(hi, 7 * N); TAKE_D(hc, N);
    TAKE_I(kind, R); TAKE_I(rev, R); TAKE_I(ridx, MAXS * R); TAKE_I(pidx, MAXS * R);
    TAKE_D(A, R); TAKE_D(b, R); TAKE_D(Ta, R); TAKE_D(A0, R); TAKE_D(b0, R); TAKE_D(Ta0, R);
    TAKE_D(troe, 4 * R); TAKE_D(rnu, MAXS * R); TAKE_D(rord, MAXS * R); TAKE_D(pnu, MAXS * R);
    TAKE_D(eff, (size_t)R * N);
    /* optional trailing section "TRN1": N x 6 transport data of the fit generator (not used by the chemistry port) */
    if (len - (size_t)(p - m->copy) >= 8 && !memcmp(p, "TRN1", 4)) p += 8 + 8 * 6 * (size_t)N;
    if ((size_t)(p - m->copy) != len) { free(m->copy); free(m); return NULL; }
    m->has_eff = (int*)calloc(R, sizeof(int));
    for (int r = 0; r < R; ++r)
        for (int k = 0; k < N; ++k)
            if (m->eff[(size_t)r * N + k] != 1.0) m->has_eff[r] = 1;
    return m;
}
void oc_free(Mech* m) { if (m) { free(m->has_eff); free(m->copy); free(m); } }
int oc_species(const Mech* m) { return m->N; }

/* per-cell constants of the frozen-T sub-step */
typedef struct { double *kf, *k0, *lfc, *rkc; } Rates;

static double powi(double c, double o) { /* Cantera's rule, SURVEY A.6 */
    if (o == 1.0) return c;
    if (o == 2.0) return c * c;
    if (o == 3.0) return c * c * c;
    if (o == 0.0) return 1.0;
    return c > 0.0 ? pow(c, o) : 0.0;
}

void oc_rate_constants(const Mech* m, double T, Rates* rc, double* g) {
    const double lT = log(T), rT = 1.0 / T, lc0 = log(ONE_ATM / (RGAS * T));
    for (int k = 0; k < m->N; ++k) {
        const double* a = (T <= m->Tmid[k] ? m->lo : m->hi) + 7 * k;
        const double h = a[0] + T * (a[1] / 2 + T * (a[2] / 3 + T * (a[3] / 4 + T * a[4] / 5))) + a[5] / T;
        const double s = a[0] * lT + T * (a[1] + T * (a[2] / 2 + T * (a[3] / 3 + T * a[4] / 4))) + a[6];
        g[k] = h - s;
    }
    for (int r = 0; r < m->R; ++r) {
        rc->kf[r] = m->A[r] * exp(m->b[r] * lT - m->Ta[r] * rT);
        rc->k0[r] = 0.0; rc->lfc[r] = 0.0; rc->rkc[r] = 0.0;
        if (m->kind[r] >= 2) {
            rc->k0[r] = m->A0[r] * exp(m->b0[r] * lT - m->Ta0[r] * rT);
            if (m->kind[r] == 3) {
                const double* t = m->troe + 4 * r;
                double fc = (1 - t[0]) * exp(-T / t[1]) + t[0] * exp(-T / t[2]);
                if (isfinite(t[3])) fc += exp(-t[3] / T);
                rc->lfc[r] = log10(fmax(fc, SMALL));
            }
        }
        if (m->rev[r]) {
            double dG = 0, dn = 0;
            for (int j = 0; j < MAXS; ++j) {
                if (m->ridx[MAXS * r + j] >= 0) { dG -= m->rnu[MAXS * r + j] * g[m->ridx[MAXS * r + j]]; dn -= m->rnu[MAXS * r + j]; }
                if (m->pidx[MAXS * r + j] >= 0) { dG += m->pnu[MAXS * r + j] * g[m->pidx[MAXS * r + j]]; dn += m->pnu[MAXS * r + j]; }
            }
            rc->rkc[r] = fmin(exp(dG - dn * lc0), BIG);
        }
    }
}

/* net molar production rates wdot[N] at concentrations C[N] */
void oc_wdot_eval(const Mech* m, const Rates* rc, const double* C, double* w) {
    const int N = m->N;
    double ctot = 0;
    for (int k = 0; k < N; ++k) { ctot += C[k]; w[k] = 0.0; }
    for (int r = 0; r < m->R; ++r) {
        double kf = rc->kf[r];
        if (m->kind[r] != 0) {
            double M = ctot;
            if (m->has_eff[r]) {
                M = 0;
                const double* e = m->eff + (size_t)r * N;
                for (int k = 0; k < N; ++k) M += e[k] * C[k];
            }
            if (m->kind[r] == 1) kf *= M;
            else {
                const double pr = M * rc->k0[r] / (kf + SMALL);
                double F = 1.0;
                if (m->kind[r] == 3) {
                    const double lf = rc->lfc[r], lpr = log10(fmax(pr, SMALL));
                    const double cc = -0.4 - 0.67 * lf, nn = 0.75 - 1.27 * lf;
                    const double f1 = (lpr + cc) / (nn - 0.14 * (lpr + cc));
                    F = pow(10.0, lf / (1.0 + f1 * f1));
                }
                kf = kf * (pr / (1.0 + pr)) * F;
            }
        }
        double qf = kf, qr = kf * rc->rkc[r];
        for (int j = 0; j < MAXS; ++j) {
            const int a = m->ridx[MAXS * r + j], b = m->pidx[MAXS * r + j];
            if (a >= 0) qf *= powi(C[a], m->rord[MAXS * r + j]);
            if (b >= 0 && qr != 0.0) qr *= powi(C[b], m->pnu[MAXS * r + j]);
        }
        const double q = qf - qr;
        for (int j = 0; j < MAXS; ++j) {
            const int a = m->ridx[MAXS * r + j], b = m->pidx[MAXS * r + j];
            if (a >= 0) w[a] -= m->rnu[MAXS * r + j] * q;
            if (b >= 0) w[b] += m->pnu[MAXS * r + j] * q;
        }
    }
}
