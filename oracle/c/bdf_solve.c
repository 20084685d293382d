/* TEST / BASELINE INFRASTRUCTURE (oracle), part 2: the CVODE-equivalent integrator and the batch
 * driver.  Compiled as one translation unit together with the kinetics (oracle_c.c).  See the
 * header of oracle_c.c for what this restates and why it is a port ("parity unpinned"). */
#include "oracle_c.c"

#define MAX_ORDER 5
#define NEWTON_MAXITER 4

typedef struct {
    const Mech* m;
    Rates rc;
    double rho, *scaleW; /* scaleW[k] = W_k / rho:  dY/dt = wdot * W / rho, C = Y / scaleW */
    double *C, *w;
    long nfev;
} Cell;

static void rhs(Cell* c, const double* y, double* f) {
    const int N = c->m->N;
    for (int k = 0; k < N; ++k) c->C[k] = y[k] / c->scaleW[k];
    oc_wdot_eval(c->m, &c->rc, c->C, c->w);
    for (int k = 0; k < N; ++k) f[k] = c->w[k] * c->scaleW[k];
    ++c->nfev;
}

static double rms(const double* v, const double* scale, int n) {
    double s = 0;
    for (int i = 0; i < n; ++i) { const double e = v[i] / scale[i]; s += e * e; }
    return sqrt(s / n);
}

/* dense LU with partial pivoting, row-major a[n*n] */
static int lu_factor(double* a, int* piv, int n) {
    for (int k = 0; k < n; ++k) {
        int p = k; double mx = fabs(a[k * n + k]);
        for (int i = k + 1; i < n; ++i) if (fabs(a[i * n + k]) > mx) { mx = fabs(a[i * n + k]); p = i; }
        piv[k] = p;
        if (mx == 0.0) return -1;
        if (p != k) for (int j = 0; j < n; ++j) { double t = a[k * n + j]; a[k * n + j] = a[p * n + j]; a[p * n + j] = t; }
        const double r = 1.0 / a[k * n + k];
        for (int i = k + 1; i < n; ++i) {
            const double l = a[i * n + k] * r;
            a[i * n + k] = l;
            if (l != 0.0) for (int j = k + 1; j < n; ++j) a[i * n + j] -= l * a[k * n + j];
        }
    }
    return 0;
}
static void lu_solve(const double* a, const int* piv, int n, double* b) {
    for (int k = 0; k < n; ++k) { if (piv[k] != k) { double t = b[k]; b[k] = b[piv[k]]; b[piv[k]] = t; }
        for (int i = k + 1; i < n; ++i) b[i] -= a[i * n + k] * b[k]; }
    for (int k = n - 1; k >= 0; --k) { for (int j = k + 1; j < n; ++j) b[k] -= a[k * n + j] * b[j]; b[k] /= a[k * n + k]; }
}

/* difference-quotient Jacobian (CVODE's internal dense DQ Jacobian): n extra RHS evaluations */
static void fd_jacobian(Cell* c, const double* y, const double* f0, const double* scale, double* J, double* yt, double* ft) {
    const int n = c->m->N;
    memcpy(yt, y, sizeof(double) * n);
    for (int j = 0; j < n; ++j) {
        const double d = fmax(1.4901161193847656e-8 * fabs(y[j]), 1e-3 * scale[j]);
        yt[j] = y[j] + d;
        rhs(c, yt, ft);
        for (int i = 0; i < n; ++i) J[i * n + j] = (ft[i] - f0[i]) / d;
        yt[j] = y[j];
    }
}

/* rescale the backward-difference array D for a step-size change by `factor` */
static void change_D(double* D, int order, double factor, int n) {
    double R[6][6], U[6][6], RU[6][6], tmp[6];
    for (int which = 0; which < 2; ++which) {
        const double fac = which == 0 ? factor : 1.0;
        double(*M)[6] = which == 0 ? R : U;
        for (int j = 0; j <= order; ++j) M[0][j] = 1.0;
        for (int i = 1; i <= order; ++i) {
            M[i][0] = 0.0;
            for (int j = 1; j <= order; ++j) M[i][j] = M[i - 1][j] * ((i - 1 - fac * j) / i);
        }
        for (int i = 1; i <= order; ++i) M[i][0] = 0.0;
        M[0][0] = 1.0;
    }
    for (int i = 0; i <= order; ++i)
        for (int j = 0; j <= order; ++j) { double s = 0; for (int k = 0; k <= order; ++k) s += R[i][k] * U[k][j]; RU[i][j] = s; }
    for (int v = 0; v < n; ++v) {
        for (int i = 0; i <= order; ++i) { double s = 0; for (int k = 0; k <= order; ++k) s += RU[k][i] * D[k * n + v]; tmp[i] = s; }
        for (int i = 0; i <= order; ++i) D[i * n + v] = tmp[i];
    }
}

/* integrate y' = f(y) from 0 to tend; returns 0 ok, 1 max steps, 2 step underflow */
static int bdf_integrate(Cell* c, double* y, double tend, double rtol, double atol, int max_steps, long* nsteps, long* nlu) {
    const int n = c->m->N;
    static const double kappa[6] = {0, -0.1850, -1.0 / 9, -0.0823, -0.0415, 0};
    double gamma[7], alpha[7], errc[7];
    gamma[0] = 0;
    for (int k = 1; k <= 6; ++k) gamma[k] = gamma[k - 1] + 1.0 / k;
    for (int k = 0; k <= 5; ++k) { alpha[k] = (1 - kappa[k]) * gamma[k]; errc[k] = kappa[k] * gamma[k] + 1.0 / (k + 1); }
    errc[6] = 1.0 / 7;
    double* buf = (double*)malloc(sizeof(double) * (size_t)(8 * n + 12 * n + 2 * n * n));
    double *D = buf, *f = D + 8 * n, *scale = f + n, *ypred = scale + n, *psi = ypred + n, *d = psi + n,
           *ynew = d + n, *dy = ynew + n, *yt = dy + n, *ft = yt + n, *err = ft + n, *f1 = err + n, *y1 = f1 + n,
           *J = y1 + n, *LU = J + n * n;
    int* piv = (int*)malloc(sizeof(int) * n);
    const double newton_tol = fmax(10 * 2.220446049250313e-16 / rtol, fmin(0.03, sqrt(rtol)));
    double t = 0.0;
    rhs(c, y, f);
    for (int i = 0; i < n; ++i) scale[i] = atol + rtol * fabs(y[i]);
    /* initial step (Hairer-Wanner heuristic) */
    double h_abs;
    {
        const double d0 = rms(y, scale, n), d1 = rms(f, scale, n);
        const double h0 = (d0 < 1e-5 || d1 < 1e-5) ? 1e-6 * tend : 0.01 * d0 / d1;
        for (int i = 0; i < n; ++i) y1[i] = y[i] + h0 * f[i];
        rhs(c, y1, f1);
        for (int i = 0; i < n; ++i) err[i] = f1[i] - f[i];
        const double d2 = rms(err, scale, n) / h0;
        const double h1 = (d1 <= 1e-15 && d2 <= 1e-15) ? fmax(1e-6 * tend, h0 * 1e-3) : sqrt(0.01 / fmax(d1, d2));
        h_abs = fmin(fmin(100 * h0, h1), tend);
    }
    memset(D, 0, sizeof(double) * 8 * n);
    for (int i = 0; i < n; ++i) { D[i] = y[i]; D[n + i] = f[i] * h_abs; }
    int order = 1, n_equal = 0, have_lu = 0, current_jac = 1, status = 0;
    /* CVODE's Newton-matrix policy (SUNDIALS cvode.c: DGMAX 0.3, MSBP 20, THRESH 1.5): the factored matrix
     * I - c J is kept across steps and step-size changes until c = h/alpha has drifted by more than 30 % from the
     * value it was factored with, 20 steps have passed, or the iteration fails; the correction is scaled by
     * 2/(1 + c/c_lu) when c != c_lu; the step size is only raised when the proposed ratio is at least 1.5. */
    double cc_lu = 0.0;
    long steps_since_lu = 0;
    fd_jacobian(c, y, f, scale, J, yt, ft);
    while (t < tend * (1 - 1e-14)) {
        if (*nsteps >= max_steps) { status = 1; break; }
        int accepted = 0, n_iter = 0;
        double safety = 0.9, error_norm = 0, t_new = t, h = h_abs;
        while (!accepted) {
            if (h_abs < 1e-14 * tend) { status = 2; goto done; }
            h = h_abs; t_new = t + h;
            if (t_new > tend) { t_new = tend; change_D(D, order, (t_new - t) / h_abs, n); n_equal = 0; }
            h = t_new - t; h_abs = h;
            for (int i = 0; i < n; ++i) {
                double s = 0, ps = 0;
                for (int k = 0; k <= order; ++k) s += D[k * n + i];
                for (int k = 1; k <= order; ++k) ps += D[k * n + i] * gamma[k];
                ypred[i] = s; psi[i] = ps / alpha[order]; scale[i] = atol + rtol * fabs(s);
            }
            const double cc = h / alpha[order];
            if (have_lu && (fabs(cc / cc_lu - 1.0) > 0.3 || steps_since_lu >= 20)) have_lu = 0;
            int converged = 0;
            while (!converged) {
                if (!have_lu) {
                    for (int i = 0; i < n * n; ++i) LU[i] = -cc * J[i];
                    for (int i = 0; i < n; ++i) LU[i * n + i] += 1.0;
                    if (lu_factor(LU, piv, n) != 0) { break; }
                    have_lu = 1; ++*nlu; cc_lu = cc; steps_since_lu = 0;
                }
                const double crate_scale = 2.0 / (1.0 + cc / cc_lu);
                /* simplified Newton */
                memset(d, 0, sizeof(double) * n);
                memcpy(ynew, ypred, sizeof(double) * n);
                double dy_old = -1;
                for (n_iter = 0; n_iter < NEWTON_MAXITER;) {
                    rhs(c, ynew, ft);
                    int finite = 1;
                    for (int i = 0; i < n; ++i) { if (!isfinite(ft[i])) finite = 0; dy[i] = cc * ft[i] - psi[i] - d[i]; }
                    ++n_iter;
                    if (!finite) break;
                    lu_solve(LU, piv, n, dy);
                    if (crate_scale != 1.0) for (int i = 0; i < n; ++i) dy[i] *= crate_scale;
                    const double dyn = rms(dy, scale, n);
                    double rate = -1;
                    if (dy_old >= 0) {
                        rate = dyn / dy_old;
                        if (rate >= 1 || pow(rate, NEWTON_MAXITER - (n_iter - 1)) / (1 - rate) * dyn > newton_tol) break;
                    }
                    for (int i = 0; i < n; ++i) { ynew[i] += dy[i]; d[i] += dy[i]; }
                    if (dyn == 0 || (rate >= 0 && rate / (1 - rate) * dyn < newton_tol)) { converged = 1; break; }
                    dy_old = dyn;
                }
                if (!converged) {
                    if (have_lu && cc_lu != cc) { have_lu = 0; continue; }   /* stale matrix: refactor at the current c first */
                    if (current_jac) break;
                    rhs(c, ypred, ft);
                    fd_jacobian(c, ypred, ft, scale, J, yt, f1);
                    have_lu = 0; current_jac = 1;
                }
            }
            if (!converged) { h_abs *= 0.5; change_D(D, order, 0.5, n); n_equal = 0; have_lu = 0; continue; }
            safety = 0.9 * (2 * NEWTON_MAXITER + 1) / (2.0 * NEWTON_MAXITER + n_iter);
            for (int i = 0; i < n; ++i) { scale[i] = atol + rtol * fabs(ynew[i]); err[i] = errc[order] * d[i]; }
            error_norm = rms(err, scale, n);
            if (error_norm > 1) {
                const double factor = fmax(0.2, safety * pow(error_norm, -1.0 / (order + 1)));
                h_abs *= factor; change_D(D, order, factor, n); n_equal = 0;
            } else accepted = 1;
        }
        ++*nsteps; ++n_equal; ++steps_since_lu;
        t = t_new; current_jac = 0;
        memcpy(y, ynew, sizeof(double) * n);
        for (int i = 0; i < n; ++i) { D[(order + 2) * n + i] = d[i] - D[(order + 1) * n + i]; D[(order + 1) * n + i] = d[i]; }
        for (int k = order; k >= 0; --k) for (int i = 0; i < n; ++i) D[k * n + i] += D[(k + 1) * n + i];
        if (n_equal < order + 1) continue;
        double em = INFINITY, ep = INFINITY;
        if (order > 1) { for (int i = 0; i < n; ++i) err[i] = errc[order - 1] * D[order * n + i]; em = rms(err, scale, n); }
        if (order < MAX_ORDER) { for (int i = 0; i < n; ++i) err[i] = errc[order + 1] * D[(order + 2) * n + i]; ep = rms(err, scale, n); }
        const double fm = pow(fmax(em, 1e-300), -1.0 / order), f0 = pow(fmax(error_norm, 1e-300), -1.0 / (order + 1)),
                     fp = pow(fmax(ep, 1e-300), -1.0 / (order + 2));
        double best = f0; int delta = 0;
        if (fm > best) { best = fm; delta = -1; }
        if (fp > best) { best = fp; delta = 1; }
        order += delta;
        double factor = fmin(10.0, safety * best);
        if (factor < 1.5 && delta == 0) continue;   /* THRESH: keep h (and the matrix) unless the gain is worth it */
        if (factor < 1.5) factor = 1.0;
        h_abs *= factor; if (factor != 1.0) change_D(D, order, factor, n); n_equal = 0;
    }
done:
    free(buf); free(piv);
    return status;
}
