/* TEST / BASELINE INFRASTRUCTURE (oracle), part 3: batch driver (solveList / solveSingle loop of
 * dfChemistryModel.C:826-841 and :738-780) on host threads, plus calculateW and thermo helpers.
 * One translation unit: driver.c -> bdf_solve.c -> oracle_c.c.  Exports (C ABI, ctypes):
 *   oc_load / oc_free / oc_species
 *   oc_solve_batch   one chemistry sub-step for nC cells on `nthreads` threads (one cell per task,
 *                    which is what one-MPI-rank-per-core gives the reference)
 *   oc_wdot_batch    wrate_i = wdot_i W_i
 */
#include "bdf_solve.c"

typedef struct {
    const Mech* m;
    int nC, ldY, ldRR, max_steps;
    const double *T, *p, *rho, *dt, *Y;
    double dt_uniform, rtol, atol;
    double *RR, *Qdot, *cost;
    int* status;
    volatile int* next;
    long nsteps, nfev, nlu;
} Job;

static void solve_cell(const Job* j, int cell, Cell* c, double* y, double* yraw, double* g, long* nsteps, long* nlu) {
    const Mech* m = j->m;
    const int N = m->N;
    double ysum = 0, rw = 0;
    for (int k = 0; k < N; ++k) { yraw[k] = j->Y[(size_t)k * j->ldY + cell]; ysum += yraw[k]; }
    for (int k = 0; k < N; ++k) { y[k] = yraw[k] / ysum; rw += y[k] / m->W[k]; }
    const double T = j->T[cell], dt = j->dt ? j->dt[cell] : j->dt_uniform;
    c->rho = j->p[cell] / (rw * RGAS * T); /* setState_TPY density */
    for (int k = 0; k < N; ++k) c->scaleW[k] = m->W[k] / c->rho;
    oc_rate_constants(m, T, &c->rc, g);
    long ns = 0, nl = 0;
    int st = 0;
    if (dt > 0) st = bdf_integrate(c, y, dt, j->rtol, j->atol, j->max_steps, &ns, &nl);
    double q = 0;
    for (int k = 0; k < N; ++k) {
        const double rr = dt > 0 ? (y[k] - yraw[k]) / dt * j->rho[cell] : 0.0;
        j->RR[(size_t)k * j->ldRR + cell] = rr;
        q -= m->hc[k] * rr;
    }
    j->Qdot[cell] = q;
    if (j->cost) j->cost[cell] = (double)ns;
    if (j->status) j->status[cell] = st;
    *nsteps += ns; *nlu += nl;
}

static void* worker(void* arg) {
    Job* j = (Job*)arg;
    const Mech* m = j->m;
    const int N = m->N, R = m->R;
    Cell c;
    c.m = m; c.nfev = 0;
    double* buf = (double*)malloc(sizeof(double) * (size_t)(4 * R + 6 * N));
    c.rc.kf = buf; c.rc.k0 = buf + R; c.rc.lfc = buf + 2 * R; c.rc.rkc = buf + 3 * R;
    c.scaleW = buf + 4 * R; c.C = c.scaleW + N; c.w = c.C + N;
    double *y = c.w + N, *yraw = y + N, *g = yraw + N;
    long ns = 0, nl = 0;
    for (;;) {
        const int base = __sync_fetch_and_add(j->next, 16);
        if (base >= j->nC) break;
        const int end = base + 16 < j->nC ? base + 16 : j->nC;
        for (int cell = base; cell < end; ++cell) solve_cell(j, cell, &c, y, yraw, g, &ns, &nl);
    }
    __sync_fetch_and_add(&j->nsteps, ns);
    __sync_fetch_and_add(&j->nfev, c.nfev);
    __sync_fetch_and_add(&j->nlu, nl);
    free(buf);
    return NULL;
}

/* stats[3]: total BDF steps, RHS evaluations, LU factorisations */
int oc_solve_batch(const Mech* m, int nC, const double* T, const double* p, const double* rho, const double* dt,
                   double dt_uniform, const double* Y, int ldY, double rtol, double atol, int max_steps, double* RR,
                   int ldRR, double* Qdot, double* cost, int* status, int nthreads, long* stats) {
    if (!m || nC < 0 || nthreads < 1) return -1;
    volatile int next = 0;
    Job j = {m, nC, ldY, ldRR, max_steps, T, p, rho, dt, Y, dt_uniform, rtol, atol, RR, Qdot, cost, status, &next, 0, 0, 0};
    if (nthreads > 256) nthreads = 256;
    pthread_t th[256];
    for (int i = 1; i < nthreads; ++i) pthread_create(&th[i], NULL, worker, &j);
    worker(&j);
    for (int i = 1; i < nthreads; ++i) pthread_join(th[i], NULL);
    if (stats) { stats[0] = j.nsteps; stats[1] = j.nfev; stats[2] = j.nlu; }
    return 0;
}

int oc_wdot_batch(const Mech* m, int nC, const double* T, const double* p, const double* Y, int ldY, double* wr, int ldW) {
    const int N = m->N, R = m->R;
    double* buf = (double*)malloc(sizeof(double) * (size_t)(4 * R + 4 * N));
    Rates rc = {buf, buf + R, buf + 2 * R, buf + 3 * R};
    double *C = buf + 4 * R, *w = C + N, *g = w + N, *yn = g + N;
    for (int c = 0; c < nC; ++c) {
        double ysum = 0, rw = 0;
        for (int k = 0; k < N; ++k) ysum += Y[(size_t)k * ldY + c];
        for (int k = 0; k < N; ++k) { yn[k] = Y[(size_t)k * ldY + c] / ysum; rw += yn[k] / m->W[k]; }
        const double rho = p[c] / (rw * RGAS * T[c]);
        for (int k = 0; k < N; ++k) C[k] = rho * yn[k] / m->W[k];
        oc_rate_constants(m, T[c], &rc, g);
        oc_wdot_eval(m, &rc, C, w);
        for (int k = 0; k < N; ++k) wr[(size_t)k * ldW + c] = w[k] * m->W[k];
    }
    free(buf);
    return 0;
}
