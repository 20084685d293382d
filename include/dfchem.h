/* dfchem.h -- C ABI of libdfchem_b200.so: B200-native per-cell stiff chemistry integration
 * and thermo update, the drop-in for the cell-local hot path of DeepFlame's dfChemistryModel.
 *
 * Every entry point names the reference interface it replaces (paths relative to the
 * reference repo show-me-code/deepflame-dev):
 *
 *   dfchem_create            Cantera::newSolution(mechanismFile,"") in CanteraMixture
 *                            src/dfCanteraMixture/CanteraMixture.C:56-68, and the constants the
 *                            dfChemistryModel ctor caches (species order :234-237, hc_ :335-338,
 *                            relTol/absTol defaults :57-58).  Flat-ABI precedent:
 *                            opencc_ode_init(), applications/solvers/dfLowMachFoam/createFields_GPU.H:61
 *   dfchem_solve[_dev]       dfChemistryModel::solve_CVODE body between getProblems and
 *                            updateReactionRates, src/dfChemistryModel/dfChemistryModel.C:1010-1083,
 *                            i.e. solveSingle :738-780 for every cell.  Flat-ABI precedent:
 *                            opencc_ode_all(), applications/solvers/dfLowMachFoam/YEqn.H:45-74;
 *                            device-pointer precedent: dfChemistrySolver::Inference,
 *                            src_gpu/dfChemistrySolver.cu:129-206 (SoA species-major, caller stream)
 *   dfchem_wdot[_dev]        dfChemistryModel::calculateW, dfChemistryModel.C:897-923
 *   dfchem_correct_thermo[_dev]  dfChemistryModel::correctThermo cell loop, dfChemistryModel.C:418-543
 *                            (T from (h,p,Y), psi, rho, cp, hai); device precedent
 *                            dfThermo::correctThermo, src_gpu/dfThermo.cu:572-671
 *   dfchem_h_from_T[_dev]    dfChemistryModel::updateEnergy, dfChemistryModel.H:326-352, and the
 *                            fixesValue(T) patches of correctThermo :572-589; device precedent
 *                            dfThermo::updateEnergy src_gpu/dfThermo.cu:557-570
 *   dfchem_correct_thermo_patches_dev   boundary loop of correctThermo, dfChemistryModel.C:546-727; device precedent
 *                            src_gpu/dfThermo.cu:603-657 (per-patch launches, processor patches with 2x entries)
 *   dfchem_transport[_dev]   transport part of correctThermo, dfChemistryModel.C:513-541 (mu, alpha,
 *                            rhoD; UnityLewis :521-527); device precedent calculateViscosityGPU /
 *                            calculateThermoConductivityGPU / calculateRhoDGPU,
 *                            src_gpu/dfThermo.cu:502-536 (kernels :111-255)
 *   dfchem_generate_transport_fits   Cantera GasTransport::fitProperties behind newTransportMgr (CanteraMixture.C:68)
 *   dfchem_load_transport_file / dfchem_set_transport_fits   dfThermo::setConstantValue +
 *                            initCoeffsfromBinaryFile, src_gpu/dfThermo.cu:361-435 (thermo_<mech>.txt)
 *   dfchem_rop[_dev], dfchem_reaction_stoich   dfChemistryModel::calculateRR,
 *                            dfChemistryModel.C:925-987 (getNetRatesOfProgress + projection)
 *   dfchem_solve_ptrs        the same, taking PtrList<volScalarField> Y_ / RR_ as N separate arrays
 *                            (dfChemistryModel.H:105,121) instead of getProblems' AoS repack (:784-823)
 *   dfchem_bind_deltaTChem   volScalarField deltaTChem_ (dfChemistryModel.H; written :871, returned :894)
 *   dfchem_rop_fwd_rev[_dev], dfchem_tc[_dev]   PaSR chemistry time scale "reactionRate",
 *                            src/dfCombustionModels/PaSR/PaSR.C:326-385
 *   dfchem_solve_balanced_dev, dfchem_comm_*   LoadBalancerBase::balance / getRemaining / unbalance,
 *                            loadBalancing/LoadBalancerBase.H:143-255, updateState LoadBalancer.C:29-56
 *   dfchem_balance_plan[2]   LoadBalancer::getOperations (allAverage) / getOperationsRedezVous (headTail),
 *                            src/dfChemistryModel/loadBalancing/LoadBalancer.C:124-180
 *
 * Layout: species-major SoA, element (species i, cell c) at  X[i*ld + c]  (the layout of
 * dfMatrixDataBase::d_y, src_gpu/dfMatrixDataBase.cu:358, and of N separate OpenFOAM fields).
 * All arithmetic is FP64.  Units are SI-kmol (K, Pa, kg/m3, s, J/kg, kg/m3/s, J/m3/s).
 * Every function returns 0 on success or a negative DFCHEM_E* code; dfchem_last_error() gives
 * the message.  There is no CPU fallback: without a usable CUDA device dfchem_create fails.
 */
#ifndef DFCHEM_H
#define DFCHEM_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct dfchem_ctx* dfchem_handle;

enum {
    DFCHEM_OK = 0,
    DFCHEM_EINVAL = -1,    /* bad argument */
    DFCHEM_EMECH = -2,     /* mechanism file missing / outside the supported dialect */
    DFCHEM_ECUDA = -3,     /* CUDA runtime error (no device, launch failure, ...) */
    DFCHEM_ENOMEM = -4,    /* host or device allocation failed */
    DFCHEM_EUNSUPPORTED = -5, /* mechanism too large for the compiled kernel set */
    DFCHEM_ENCCL = -7,       /* NCCL missing or a collective failed (load balancer) */
    DFCHEM_ECELLFAIL = -6    /* dfchem_solve / dfchem_solve_ptrs: the integrator failed in at least one cell (max steps
                              * or step underflow); every output array is written, the failed cells hold the partial
                              * result and a non-zero status.  The reference aborts here (uncaught CVODE error in
                              * solveSingle, dfChemistryModel.C:738-780): the OpenFOAM wrapper turns it into FatalError */
};

/* per-cell status written by dfchem_solve* (ChemistrySolution has no such field; the reference
 * lets CVODE failures propagate as uncaught exceptions, dfChemistryModel.C:738-780) */
enum { DFCHEM_CELL_OK = 0, DFCHEM_CELL_MAXSTEPS = 1, DFCHEM_CELL_STEP_UNDERFLOW = 2 };

/* Create a context on CUDA device `device` from a Cantera-YAML mechanism file. */
int dfchem_create(const char* mech_yaml_path, int device, dfchem_handle* out);
/* Same, from the flat DFCHEMB1 blob written by the mechanism compiler (dfchem_mechc). */
int dfchem_create_from_blob(const void* blob, size_t len, int device, dfchem_handle* out);
void dfchem_destroy(dfchem_handle h);
const char* dfchem_last_error(dfchem_handle h); /* h may be NULL: error of the failed create */

int dfchem_species_count(dfchem_handle h);
int dfchem_reaction_count(dfchem_handle h);
const char* dfchem_species_name(dfchem_handle h, int i); /* Cantera order = YAML phases[0].species */
int dfchem_species_index(dfchem_handle h, const char* name); /* -1 when absent */
/* constants (length N, host): which = 0: W [kg/kmol], 1: hc = Hf298SS/W [J/kg] */
int dfchem_species_constants(dfchem_handle h, int which, double* out);

/* Integrator options; defaults mirror odeCoeffs of CanteraTorchProperties
 * (relTol 1e-9, absTol 1e-15, dfChemistryModel.C:57-58; ReactorNet max steps 20000). */
int dfchem_set_tolerances(dfchem_handle h, double rtol, double atol);
int dfchem_set_max_steps(dfchem_handle h, int max_steps);

/* Chemistry sub-step for nC cells (solveSingle semantics, dfChemistryModel.C:738-780):
 *   Yn = Y/sum(Y);  rho_ct = p*Wbar/(R*T);  integrate dY/dt = wdot*W/rho_ct at frozen (T, rho_ct)
 *   over dt;  RR_i = (Ynew_i - Y_i)/dt * rho;  Qdot = -sum_i hc_i*RR_i.
 * dt: per-cell array (length nC) or NULL to use dt_uniform.  rho is the caller's (OpenFOAM's)
 * density used only for the RR scaling.  cost (optional, may be NULL) receives the number of
 * integrator step attempts per cell: it plays the role of ChemistrySolution::cpuTime
 * (ChemistrySolution.H:44-68) as the load-balancer's input.  status optional (may be NULL).
 * Host-pointer form: blocking; copies through pinned staging buffers owned by the handle. */
int dfchem_solve(dfchem_handle h, int nC, const double* T, const double* p, const double* rho,
                 const double* dt, double dt_uniform, const double* Y, int ldY,
                 double* RR, int ldRR, double* Qdot, double* cost, int* status);
/* Device-pointer form: stream-ordered on `stream` (a cudaStream_t passed as void*), no host
 * synchronisation inside.  All pointers are device pointers on the handle's device. */
int dfchem_solve_dev(dfchem_handle h, int nC, const double* T, const double* p, const double* rho,
                     const double* dt, double dt_uniform, const double* Y, int ldY,
                     double* RR, int ldRR, double* Qdot, double* cost, int* status, void* stream);

/* Pointer-table form of dfchem_solve: Yptrs[i] / RRptrs[i] (i < N, host) are the N separately allocated
 * arrays of PtrList<volScalarField> Y_ / RR_ (dfChemistryModel.H:105,121) as they are -- no repack into one
 * block on the host; each species row is staged by its own asynchronous copy. */
int dfchem_solve_ptrs(dfchem_handle h, int nC, const double* T, const double* p, const double* rho,
                      const double* dt, double dt_uniform, const double* const* Yptrs,
                      double* const* RRptrs, double* Qdot, double* cost, int* status);

/* deltaTChem (dfChemistryModel.H `deltaTChem_`; the reference carries the field but never feeds it to
 * CVODE: every cell cold-starts, dfChemistryModel.C:871,894).  Binding an array of nC doubles makes the
 * following dfchem_solve* calls (a) start every cell whose entry is > 0 with that first trial step instead
 * of the cold-start heuristic and (b) write back the step the controller would take next.  The pointer is
 * interpreted like the other arrays of the call that uses it: host memory for dfchem_solve /
 * dfchem_solve_ptrs, device memory for dfchem_solve_dev.  NULL unbinds.  The returned chemistry time step
 * of solve() stays `great` (updateReactionRates, dfChemistryModel.C:894). */
int dfchem_bind_deltaTChem(dfchem_handle h, double* deltaTChem);

/* Instantaneous net production rates wrate_i = wdot_i*W_i [kg/m3/s] at (T, p, Y)
 * (calculateW, dfChemistryModel.C:897-923). */
int dfchem_wdot(dfchem_handle h, int nC, const double* T, const double* p, const double* Y,
                int ldY, double* wrate, int ldW);
int dfchem_wdot_dev(dfchem_handle h, int nC, const double* T, const double* p, const double* Y,
                    int ldY, double* wrate, int ldW, void* stream);

/* dfchem_wdot for callers that may not allocate device memory (CUDA graph capture around the call site, tight
 * serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same arguments (host pointers) and same results as dfchem_wdot, but the kernel's scratch -- the
 * work-queue word and the device staging of T, p, Y and wrate, (2N + 2) doubles per cell -- lives in a file-scope
 * static device array of DFCHEM_STATIC_SCRATCH_BYTES that every GPU holding the library keeps until exit; the call
 * makes no cudaMalloc / cudaFree and uses no allocating library routine.  A call whose staging does not fit
 * returns DFCHEM_EINVAL (dfchem_last_error names the largest nC for this mechanism).  All handles of a process on
 * one device share the array: calls that overlap on the GPU (two handles, two host threads) race; the call
 * itself is synchronous, so one caller at a time is safe.  The queue word is zeroed by every call; the staging is
 * fully overwritten before it is read. */
#define DFCHEM_STATIC_SCRATCH_BYTES (8u << 20)
int dfchem_wdot_static(dfchem_handle h, int nC, const double* T, const double* p, const double* Y,
                       int ldY, double* wrate, int ldW);

/* correctThermo cell loop (dfChemistryModel.C:418-543) without the transport properties:
 * T: in = initial guess, out = solution of h(T, Y/sum Y) = he (Newton, cp slope, |dT|<=100 K
 * per iteration, stop |dT| < 1e-9*T);  psi = Wbar/(R T);  rho = p*psi;  cp [J/kg/K] optional;
 * hai[i*ld + c] = h_i(T)/W_i [J/kg] optional.  n may be cells or boundary faces. */
int dfchem_correct_thermo(dfchem_handle h, int n, const double* he, const double* p,
                          const double* Y, int ldY, double* T, double* psi, double* rho,
                          double* cp, double* hai, int ldH);
int dfchem_correct_thermo_dev(dfchem_handle h, int n, const double* he, const double* p,
                              const double* Y, int ldY, double* T, double* psi, double* rho,
                              double* cp, double* hai, int ldH, void* stream);

/* he = h_mass(T, Y/sum Y) [J/kg]  (updateEnergy, dfChemistryModel.H:326-352). */
int dfchem_h_from_T(dfchem_handle h, int n, const double* T, const double* Y, int ldY, double* he);
int dfchem_h_from_T_dev(dfchem_handle h, int n, const double* T, const double* Y, int ldY,
                        double* he, void* stream);

/* Net rates of progress q[r*ldQ + c] [kmol/m^3/s] of every reaction r (order of the mechanism
 * file) at (T, p, Y): the CanteraKinetics_->getNetRatesOfProgress call of calculateRR
 * (dfChemistryModel.C:925-987).  calculateRR(reactionI, speciei) is nu_net[speciei] * q[reactionI]
 * * W[speciei] with nu_net from dfchem_reaction_stoich (products minus reactants,
 * dfChemistryModel.C:967-982). */
int dfchem_rop(dfchem_handle h, int nC, const double* T, const double* p, const double* Y,
               int ldY, double* q, int ldQ);
int dfchem_rop_dev(dfchem_handle h, int nC, const double* T, const double* p, const double* Y,
                   int ldY, double* q, int ldQ, void* stream);
int dfchem_reaction_stoich(dfchem_handle h, int reaction, double* nu_net /* [N] */);
/* Forward and reverse rates of progress qf, qr [r*ldQ + c] (getFwdRatesOfProgress / getRevRatesOfProgress at
 * setState_TPX(T, p, rho Y_i/W_i), src/dfCombustionModels/PaSR/PaSR.C:341-350), and the PaSR "reactionRate"
 * chemical time scale tc [s] built from them (PaSR.C:352-378): tc = sum_r (wf_r + wr_r) / sum_r (wf_r^2 + wr_r^2)
 * * sum_i rho Y_i / W_i, with wf_r = (sum of product nu) qf_r, wr_r = (sum of reactant nu) qr_r; vGreat = 1e300
 * where nothing reacts.  rho = the caller's density.  (The "formationRate" scale, laminar.C:73-102, is
 * max_i Y_i/|RR_i| of the solve() result and needs no entry point.) */
int dfchem_rop_fwd_rev(dfchem_handle h, int nC, const double* T, const double* p, const double* Y,
                       int ldY, double* qf, double* qr, int ldQ);
int dfchem_rop_fwd_rev_dev(dfchem_handle h, int nC, const double* T, const double* p, const double* Y,
                           int ldY, double* qf, double* qr, int ldQ, void* stream);
int dfchem_tc(dfchem_handle h, int nC, const double* T, const double* p, const double* rho,
              const double* Y, int ldY, double* tc);
int dfchem_tc_dev(dfchem_handle h, int nC, const double* T, const double* p, const double* rho,
                  const double* Y, int ldY, double* tc, void* stream);

/* Transport part of correctThermo (dfChemistryModel.C:513-541): mu = Wilke mixture viscosity
 * [Pa s], alpha = lambda_mix / cp [kg/m/s], rhoD[i*ldD + c] = rho * D_i^mix [kg/m/s] (mass-flux
 * mixture-averaged form, getMixDiffCoeffsMass) or = alpha when unity_lewis != 0
 * (dfChemistryModel.C:521-527).  The species fits are Cantera's polynomials in ln T, supplied in
 * the layout of the reference's thermo_<mech>.txt (src_gpu/dfThermo.cu:361-435): visc[N][5],
 * cond[N][5], diff[N][N][5]; evaluation follows src_gpu/dfThermo.cu:111-255.  Any of mu, alpha,
 * rhoD may be NULL; cp [J/kg/K] (from dfchem_correct_thermo) is read for alpha, rho and p for
 * rhoD. */
int dfchem_set_transport_fits(dfchem_handle h, const double* visc, const double* cond,
                              const double* diff);
int dfchem_load_transport_file(dfchem_handle h, const char* thermo_txt_path);
/* Fit generator of the mechanism compiler (what Cantera's GasTransport::fitProperties does for the reference at
 * start-up): builds the same three fit tables from the `transport:` blocks of the mechanism file (geometry, well-depth,
 * diameter, dipole, polarizability, rotational-relaxation) and installs them; the tables are also copied out when the
 * pointers are non-NULL.  Collision integrals: Neufeld closed forms + Brokaw polar term instead of Cantera's
 * Monchick-Mason table interpolation (tables absent from the tree): within 1 % of the Cantera-made ES80 file. */
int dfchem_generate_transport_fits(dfchem_handle h, double* visc /* [N*5] */, double* cond /* [N*5] */,
                                   double* diff /* [N*N*5] */);
int dfchem_transport(dfchem_handle h, int n, const double* T, const double* p, const double* rho,
                     const double* cp, const double* Y, int ldY, double* mu, double* alpha,
                     double* rhoD, int ldD, int unity_lewis);
int dfchem_transport_dev(dfchem_handle h, int n, const double* T, const double* p,
                         const double* rho, const double* cp, const double* Y, int ldY, double* mu,
                         double* alpha, double* rhoD, int ldD, int unity_lewis, void* stream);

/* Boundary faces of correctThermo (dfChemistryModel.C:546-727), device arrays laid out like the reference's
 * dfMatrixDataBase boundary fields (src_gpu/dfThermo.cu:603-657): all patches back to back in arrays of n_boundary
 * entries (species-major with leading dimension n_boundary), patch i occupying patch_size[i] entries -- twice that for
 * a processor patch: [neighbour side | internal side].  Per patch: DFCHEM_PATCH_FIXED_T: T is given, he <- h(T, Y)
 * (:572-589); otherwise T <- solve h(T, Y) = he; then psi, rho, cp and, when the output pointers are given, mu, alpha,
 * rhoD; a processor patch then copies T, he, psi, rho, mu, alpha, rhoD of the owner cells (face_cell) into its internal
 * side (correct_internal_boundary_field_scalar, src_gpu/dfMatrixOpBase.cu:387-399).  patch_size / patch_type are host
 * arrays, everything else device memory; stream-ordered, no host synchronisation. */
enum { DFCHEM_PATCH_GENERIC = 0, DFCHEM_PATCH_FIXED_T = 1, DFCHEM_PATCH_PROCESSOR = 2 };
int dfchem_correct_thermo_patches_dev(dfchem_handle h, int n_patches, const int* patch_size, const int* patch_type,
                                      int n_boundary, double* b_he, const double* b_p, const double* b_Y, double* b_T,
                                      double* b_psi, double* b_rho, double* b_cp, double* b_mu, double* b_alpha,
                                      double* b_rhoD, int unity_lewis, int n_cells, const int* face_cell,
                                      const double* T, const double* he, const double* psi, const double* rho,
                                      const double* mu, const double* alpha, const double* rhoD, void* stream);

/* Counters of the last dfchem_solve* call, summed over its cells: [0] accepted steps, [1] rejected steps,
 * [2] cells with status != 0, [3] kernel launches issued by the last call.  After dfchem_solve_dev the
 * caller must have synchronised the stream first; the _dev form cannot return DFCHEM_ECELLFAIL (it does not
 * wait for the kernel), so its caller checks [2] here or the status array.
 * Threading: a handle is single-caller (like dfChemistryModel, whose scratch members are not re-entrant);
 * its _dev entry points may be issued on several streams -- every launch owns a private work-queue word and
 * counter record -- but the counters reported here are those of the most recent solve call only. */
int dfchem_last_counters(dfchem_handle h, unsigned long long out[4]);

/* Load-balancer plan, restating LoadBalancer::getOperations (allAverage rule,
 * loadBalancing/LoadBalancer.C:124-180) for a cost-weighted all-to-all-v between GPUs:
 * loads[n_ranks] -> send[from*n_ranks + to] = load units to move (0 on the diagonal);
 * operations smaller than 1 % of the mean load are dropped as in the reference. */
int dfchem_balance_plan(const double* loads, int n_ranks, double* send);
/* The same with the pairing rule chosen: algorithm 0 = allAverage, 1 = headTail (getOperationsRedezVous,
 * LoadBalancer.C:183-232: the i-th heaviest rank gives half its excess over the i-th lightest). */
int dfchem_balance_plan2(const double* loads, int n_ranks, int algorithm, double* send);
/* timesToProblemCounts (LoadBalancer.C:97-122) for rank `me`: counts_row[to] = how many leading cells of its batch
 * (index order) go to each rank, given every rank's load and its own per-cell cost; at least one cell stays.  Host
 * form of the planner dfchem_solve_balanced_dev runs on the device. */
int dfchem_balance_counts(const double* loads, int n_ranks, int me, int algorithm, const double* cost, int nC,
                          int* counts_row);

/* In-library load balancer (loadBalancing/LoadBalancerBase.H:143-255 balance / getRemaining / unbalance) over NCCL:
 * one rank per GPU, each with its own handle.  The communicator is either created here from an id that rank 0
 * obtained with dfchem_comm_unique_id and handed to the others by whatever the caller has (MPI_Bcast in the
 * reference's dfNcclBase, src_gpu/dfNcclBase.cu:40-56), or an existing ncclComm_t is attached. */
int dfchem_comm_unique_id(void* id128 /* 128 bytes out */);
int dfchem_comm_init(dfchem_handle h, int n_ranks, int rank, const void* id128);
int dfchem_comm_attach(dfchem_handle h, void* nccl_comm, int n_ranks, int rank);
/* Balanced chemistry sub-step, device pointers, stream-ordered except for ONE host wait on 4 n_ranks^2 bytes (the
 * cell counts that size the exchanges).  cost_prev [nC]: the per-cell cost of the previous flow step (the `cost`
 * output of the previous call; NULL = plain dfchem_solve_dev); cost [nC] is required.  Loads are all-gathered, the
 * chosen rule pairs senders with receivers, a sender ships the leading cells of its batch (index order, as
 * timesToProblemCounts walks the problem list, LoadBalancer.C:97-122) until their summed cost reaches the share,
 * guests are integrated where they land and their RR / Qdot / cost / status return into the owner's arrays: results
 * are bit-identical to the unbalanced call.  Collective: every rank of the communicator must call it. */
int dfchem_solve_balanced_dev(dfchem_handle h, int nC, const double* T, const double* p, const double* rho,
                              const double* dt, double dt_uniform, const double* Y, int ldY, double* RR,
                              int ldRR, double* Qdot, const double* cost_prev, double* cost, int* status,
                              int algorithm, void* stream);
/* Of the last balanced call (stream synchronised by the caller): [0] cells sent, [1] guest cells received,
 * [2] bytes sent by this rank, then ms: [3] plan, [4] outbound exchange, [5] guest solve, [6] resident solve,
 * [7] whole call. */
int dfchem_balance_stats(dfchem_handle h, double out[8]);

/* Measurement helper: FP64 DFMA throughput of `device` in TFLOP/s (16 independent chains per
 * thread, best of 4 timed launches).  It is the denominator bench.py reports "% of FP64 peak"
 * against, because MEASURED_PEAKS.json carries no FP64 figure. */
int dfchem_bench_fp64_peak(int device, double* tflops);

const char* dfchem_version(void);
/* Name of the kernel that the last dfchem_solve* call of this handle launched -- before the first call: the one it launches on
 * large batches -- ("solve_kernel<34,2>": warp per cell;
 * "lpc_solve_kernel<6>": lane per cell, mechanisms with at most 8 active species or DFCHEM_SOLVER=lane; batches below 12288 cells
 * of such a mechanism still run one warp per cell, which finishes a small batch sooner).  For measurement records; no reference
 * counterpart. */
const char* dfchem_solver_name(dfchem_handle h);

#ifdef __cplusplus
}
#endif
#endif /* DFCHEM_H */
