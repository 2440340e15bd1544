/* iemic_b200.h -- C ABI of the B200-native I-EMIC hot path (libiemic_b200.so).
 *
 * Plain C: pointers, sizes and scalars only; no torch, Epetra or CUDA types in any signature (streams and
 * NCCL ids cross as void* / byte arrays).  Every function returns 0 on success and a negative code on
 * error; it never aborts the process, so Continuation::reset() (src/continuation/Continuation.H:198-214 of
 * the reference) keeps working.  The message of the last error on the calling thread is returned by
 * iemic_b200_last_error().
 *
 * Which reference interface each entry point replaces (paths relative to the reference checkout):
 *
 *   thcm_b200_create / _destroy        THCM::THCM + Fortran init_     src/ocean/THCM.C:176-800, src/ocean/usrc.F90:6-139
 *   thcm_b200_set_par / _get_par       setparcs_ / getparcs_          src/ocean/usrc.F90:163-196  (THCM::setParameter, THCM.C:1883)
 *   thcm_b200_set_landmask             set_landmask_                  src/ocean/usrc.F90:336-395
 *   thcm_b200_rhs  (device)            THCM::evaluate(sol,rhs,false)  src/ocean/THCM.C:949-1033 -> Fortran rhs_  usrc.F90:506-586
 *   thcm_b200_jacobian (device)        THCM::evaluate(sol,null,true)  src/ocean/THCM.C:1035-1192 -> Fortran matrix_ usrc.F90:432-504
 *   thcm_b200_graph / _csr             THCM::CreateMaximalGraph       src/ocean/THCM.C:2241-2539; THCM::getJacobian THCM.H:146
 *   thcm_b200_diagB                    THCM::DiagB (localDiagB_)      src/ocean/THCM.C:1150-1153, assemble.F90:18-54
 *   thcm_b200_rhs_host / _jacobian_host  the same two calls with HOST buffers (what Ocean::computeRHS /
 *                                      Ocean::computeJacobian do with Epetra vectors, src/ocean/Ocean.C:1267-1300)
 *   thcm_b200_spmv (device) / _spmv_host  Epetra_CrsMatrix::Apply on the Jacobian: Ocean::applyMatrix src/ocean/Ocean.C:1352-1357
 *   thcm_b200_set_values               Epetra_CrsMatrix::ReplaceMyValues (load a matrix with the same graph)
 *   gmres_b200_solve (device) / _host  GMRESSolver<Model,VectorPtr>::solve  src/gmressolver/GMRESSolver.H:81-255
 *                                      (Ocean::solve, src/ocean/Ocean.C:1060-1138, is the live caller of the same job)
 *   vec_b200_*                         the Epetra_Vector subset Continuation/GMRESSolver use (Update, Dot, Norm2, Scale,
 *                                      PutScalar): src/continuation/ContinuationDecl.H:15-46, src/gmressolver/GMRESSolverDecl.H:7-16
 *   thcm_b200_comm_*                   TRIOS::Domain (Decomp2D, ghosted maps, Import/Export)  src/trios/TRIOS_Domain.C:81-195, 517-664
 *
 * Layout conventions (identical to the reference, src/trios/THCMdefs.H:22-23, src/ocean/matetc.F90:123-144):
 *   unknown index (0-based) = 6*((k*m + j)*n + i) + var,  var = 0..5 for u,v,w,p,T,S, i fastest, k=0 the bottom layer;
 *   land mask: int[(n+2)(m+2)(l+2)], i fastest, border ring included, 0 ocean / 1 land / 3 periodic (par.F90:78-81);
 *   CSR: 0-based int32 rowptr/col, fp64 values, columns sorted inside each row, static CreateMaximalGraph pattern.
 * All arithmetic is fp64.
 */
#ifndef IEMIC_B200_H
#define IEMIC_B200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct thcm_b200 thcm_b200_t;

/* Same fields, in the same order, as the argument list of Fortran init (usrc.F90:6-139) and the "THCM"
 * parameter list consumed at THCM.C:189-268.  Angles in radians. */
typedef struct {
  int n, m, l;
  double xmin, xmax, ymin, ymax;
  int periodic;
  double hdim, qz;
  double alphaT, alphaS;
  int ih, vmix, tap, rho_mixing, coriolis_on, forcing_type;
  int tres, sres, iza, ite, its;
  int coupled_T, coupled_S;
  int int_sign;             /* "Salinity Integral Sign" */
  int intcond_i, intcond_j; /* "Integral row coordinate i/j", 0-based, -1 = last; intcond_i = -2: no integral row even
                             * with SRES=0 (the Fortran layer's view, used by the seam-S1 adaptor thcm_fortran_abi.h) */
  int fix_pressure_points;
} thcm_b200_config;

enum { IEMIC_B200_OK = 0, IEMIC_B200_EINVAL = -1, IEMIC_B200_ECUDA = -2, IEMIC_B200_ENOTSUP = -3, IEMIC_B200_ENCCL = -4, IEMIC_B200_ECALLBACK = -5 };
enum { GMRES_B200_MGS = 0, GMRES_B200_CGS2 = 1, GMRES_B200_CGS2F = 2 };
enum { GMRES_B200_PREC_NONE = 0, GMRES_B200_PREC_BLOCKJACOBI = 1, GMRES_B200_PREC_CALLBACK = 2 };

const char* iemic_b200_last_error(void);
const char* iemic_b200_version(void);
int iemic_b200_device_count(void);

/* ---- model handle (one per GPU / per process, like the THCM singleton, THCM.H:76-83) ---- */
int thcm_b200_create(const thcm_b200_config* cfg, const int* landm, thcm_b200_t** out);
/* Distributed variant: rank `rank` of `nranks` owns one block of the reference's 2-D lon x lat decomposition
 * (TRIOS::Domain::Decomp2D).  nccl_id = the 128-byte ncclUniqueId made by thcm_b200_comm_unique_id on rank 0
 * and distributed by the host program (MPI_Bcast / torch.distributed); NULL with nranks==1. */
int thcm_b200_create_dist(const thcm_b200_config* cfg, const int* landm, int rank, int nranks,
                          const void* nccl_id, thcm_b200_t** out);
int thcm_b200_comm_unique_id(void* nccl_id_out /* 128 bytes */);
int thcm_b200_destroy(thcm_b200_t* h);
int thcm_b200_set_stream(thcm_b200_t* h, void* cuda_stream);
int thcm_b200_sync(thcm_b200_t* h);

int thcm_b200_set_par(thcm_b200_t* h, int idx1, double value);
int thcm_b200_get_par(thcm_b200_t* h, int idx1, double* value);
int thcm_b200_set_landmask(thcm_b200_t* h, const int* landm);
int thcm_b200_set_intcond_correction(thcm_b200_t* h, double corr); /* THCM.C:1013 intCorrection_ */

/* sizes: local (this rank's) rows, stored vector length incl. halo cells, non-zeros; global rows */
int thcm_b200_sizes(thcm_b200_t* h, int64_t* nrows_local, int64_t* nvec_local, int64_t* nnz_local, int64_t* nrows_global);
/* decomposition of this rank: owned block in global 0-based cell coordinates and the process grid */
int thcm_b200_block(thcm_b200_t* h, int* i0, int* j0, int* nloc, int* mloc, int* npN, int* npM);
/* global unknown index of every entry of a stored vector (owned rows first, then halo cells; -1 = unused halo slot) */
int thcm_b200_global_ids(thcm_b200_t* h, int64_t* gid /* nvec_local */);

/* ---- assembly ---- */
int thcm_b200_rhs(thcm_b200_t* h, const double* d_x, double* d_F);   /* device pointers; d_x has nvec_local entries */
int thcm_b200_jacobian(thcm_b200_t* h, const double* d_x);            /* fills the handle's CSR values and diag(B) */
int thcm_b200_halo_exchange(thcm_b200_t* h, double* d_x);
/* how the exchange steps travel: 0 one GPU (nothing to exchange), 1 NCCL send/recv + allreduce, 2 peer memory over
 * NVLink (CUDA IPC; halo push/pull kernels, reductions finished inside the reducing kernels).  Mode 2 is chosen when
 * every rank can map every other rank's buffers; IEMIC_P2P=0 in the environment forces mode 1. */
/* THCM::evaluateB (src/ocean/THCM.C:1195-1235): diag(B) alone into d_diagB (nrows_local doubles; NULL: the handle's own array
 * that thcm_b200_csr returns) -- fillcolB (assemble.F90:18-54) with the fixed-pressure and integral rows zeroed. */
int thcm_b200_evaluate_b(thcm_b200_t* h, double* d_diagB);

/* ---- Coupling to an atmosphere / sea-ice model ("Coupled Temperature" = 1, "Coupled Salinity" = 1 is refused) ----
 * The ocean side of Ocean::synchronize (src/ocean/Ocean.C:1443-1493): the neighbouring model's interface fields go in with
 * thcm_b200_insert_field (THCM::setAtmosphereT/Q/A, setSeaIceM, src/ocean/THCM.C:1375-1462 -> m_inserts, src/ocean/inserts.F90:12-166:
 * whole n x m surface, i fastest, every rank passes the whole field), and take effect at the following set_*_parameters /
 * set_par call, which reruns forcing and lin exactly like the Fortran (src/ocean/usrc.F90:237-333, 163-181):
 *   surface T forcing per cell   (src/ocean/forcing.F90:73-96)        -> forcing_cells_kernel
 *   surface T row coefficients   (src/ocean/usrc.F90:725-741)         -> the assembly kernels (mask-dependent centre entry, (T,S) entry)
 * The structs are the bind(C) types of usrc.F90:250-269 / :303-311 = Atmosphere::CommPars / SeaIce::CommPars. */
typedef struct { double tdim, qdim, nuq, eta, dqso, dqsi, dqdt, Eo0, Ei0, Cs, t0o, t0i, a0, da, tauf, tauc, comb, albf; } thcm_b200_atmos_pars;
typedef struct { double zeta, a0, Lf, s0, rhoo, qvar, q0; } thcm_b200_seaice_pars;
enum { THCM_B200_FIELD_ATMOS_T = 0, THCM_B200_FIELD_ATMOS_Q = 1, THCM_B200_FIELD_ATMOS_A = 2, THCM_B200_FIELD_ATMOS_P = 3,
       THCM_B200_FIELD_SEAICE_Q = 4, THCM_B200_FIELD_SEAICE_M = 5, THCM_B200_FIELD_SEAICE_G = 6 };
int thcm_b200_insert_field(thcm_b200_t* h, int which, const double* field /* n*m */);
int thcm_b200_set_atmos_parameters(thcm_b200_t* h, const thcm_b200_atmos_pars* pars);   /* set_atmos_parameters_, usrc.F90:237-293 */
int thcm_b200_set_seaice_parameters(thcm_b200_t* h, const thcm_b200_seaice_pars* pars); /* set_seaice_parameters_, usrc.F90:296-333 */
int thcm_b200_getdeps(thcm_b200_t* h, double* out7);  /* getdeps_, usrc.F90:201-219: Ooa, Os, nus, eta, lvsc, qdim, pQSnd */
int thcm_b200_get_suno(thcm_b200_t* h, double* suno /* n*m */);                         /* THCM::getSunO */
/* m_probe::get_derivatives, coupled_T part (src/ocean/probe.F90:381-444, THCM::getDerivatives THCM.C:1596-1626): dF_T/dM_seaice
 * on the owned surface points, (j-j0)*nloc + (i-i0) (one GPU: the whole surface); d_x is a device state vector. */
int thcm_b200_get_dftdm(thcm_b200_t* h, const double* d_x, double* dftdm /* nloc*mloc, host */);
/* Ocean::getBlock(Atmosphere) / getBlock(SeaIce) (src/ocean/Ocean.C:1538-1746), the T rows, in compact form: rows[r] is the global
 * ocean row of the r-th non-empty row, its entries are jco/co[beg[r] .. beg[r+1]) in the reference's push order, columns are
 * Atmosphere::interface_row / SeaIce::interface_row (dof 3 / 4).  Capacity: n*m rows, 3*n*m entries.  Utils::CRSMat over all
 * ocean rows follows by scattering beg. */
int thcm_b200_atmos_block(thcm_b200_t* h, int* rows, int* beg, int* jco, double* co, int* nrows_out);
int thcm_b200_seaice_block(thcm_b200_t* h, const double* d_x, int* rows, int* beg, int* jco, double* co, int* nrows_out);
/* CouplingBlock<Ocean, Model> (src/coupledmodel/CouplingBlock.H): the block on the device; entries of a row are kept sorted by column
 * as Epetra's FillComplete leaves them.  apply: d_out = C d_in over the rank's ocean rows (accumulate = 0, CouplingBlock::applyMatrix
 * :206) or d_out += C d_in (accumulate = 1: the out.Update(1.0, z, 1.0) of CoupledModel::applyMatrix, CoupledModel.C:436-470, without
 * the temporary).  d_in: the neighbouring model's state (ncols doubles, replicated on every rank), on the handle's stream. */
typedef struct coupling_b200 coupling_b200_t;
int coupling_b200_create(thcm_b200_t* h, int nrows, const int* rows, const int* beg, const int* jco, const double* co, int64_t ncols,
                         coupling_b200_t** out);
int coupling_b200_apply(coupling_b200_t* b, const double* d_in, double* d_out, int accumulate);
void coupling_b200_destroy(coupling_b200_t* b);

int thcm_b200_exchange_mode(thcm_b200_t* h);             /* fills the halo part of a stored vector (no-op on 1 GPU) */
/* residual (d_F may be NULL) and/or Jacobian on one state; on several GPUs the halo exchange of d_x runs concurrently with
 * the tiles that read no halo cell (NCCL on a second stream), the rim tiles follow.  Replaces the sequence
 * Solve2Assembly + rhs_/matrix_ of THCM::evaluate (THCM.C:972-1058). */
int thcm_b200_assemble(thcm_b200_t* h, double* d_x, double* d_F, int jacobian);
int thcm_b200_csr(thcm_b200_t* h, const int** d_rowptr, const int** d_col, double** d_val, const double** d_diagB);
int thcm_b200_graph_host(thcm_b200_t* h, int* rowptr, int* col);      /* host copies of the static graph */
int thcm_b200_values_host(thcm_b200_t* h, double* val, double* diagB);
int thcm_b200_set_values_host(thcm_b200_t* h, const double* val);
int thcm_b200_forcing_host(thcm_b200_t* h, double* frc);
int thcm_b200_intcond_host(thcm_b200_t* h, double* coeff, int* rowintcon);
/* host-buffer entry points: what the reference's Ocean::computeRHS / computeJacobian do on Epetra vectors */
int thcm_b200_rhs_host(thcm_b200_t* h, const double* x, double* F);
int thcm_b200_jacobian_host(thcm_b200_t* h, const double* x, double* val /* may be NULL */, double* diagB /* may be NULL */);
/* One Newton-step assembly on ONE state, as Continuation::newtonCorrector does through computeRHS + computeJacobian on
 * the model's internal state_ (src/continuation/Continuation.H:654,714): x is uploaded once, F (may be NULL) is
 * downloaded while the Jacobian kernel runs, the Jacobian stays on the device for thcm_b200_spmv / gmres_b200_solve. */
int thcm_b200_assemble_host(thcm_b200_t* h, const double* x, double* F, int jacobian);

/* ---- sparse matrix-vector product y = J x on the handle's CSR ---- */
int thcm_b200_spmv(thcm_b200_t* h, double* d_x /* halo part is overwritten when nranks>1 */, double* d_y);
int thcm_b200_spmv_host(thcm_b200_t* h, const double* x, double* y);

/* ---- vector kernels (Epetra_Vector subset): all device pointers, length = nrows_local, results global ---- */
int vec_b200_dot(thcm_b200_t* h, const double* d_a, const double* d_b, double* result);
/* No counterpart in the reference: added for callers that must not allocate (CUDA graph
 * capture, tight serving loops).  Same product as vec_b200_dot on ONE GPU, without a handle: d_out[0] = sum_i d_a[i]*d_b[i] is
 * left on the device, the call is asynchronous on `stream` (a cudaStream_t, may be 0), and nothing is allocated anywhere -- the
 * reduction scratch is a static 16.6 KB device array inside the library that every GPU loading it holds until exit.
 * 0 <= n <= 2^30, otherwise IEMIC_B200_EINVAL.  All calls share that array: calls that overlap on one GPU (different streams,
 * threads or processes' contexts apart) race; order them on one stream. */
int vec_b200_dot_noalloc(const double* d_a, const double* d_b, int64_t n, double* d_out, void* stream);
int vec_b200_norm2(thcm_b200_t* h, const double* d_a, double* result);
int vec_b200_norminf(thcm_b200_t* h, const double* d_a, double* result); /* Epetra_Vector::NormInf, global over the ranks */
/* owned part of a vector between host memory and the device, ordered on the handle's stream (synchronous for the caller) */
int thcm_b200_upload(thcm_b200_t* h, double* d_dst, const double* src, int64_t n);
int thcm_b200_download(thcm_b200_t* h, double* dst, const double* d_src, int64_t n);
int vec_b200_update(thcm_b200_t* h, double a, const double* d_A, double b, double* d_this); /* this = a*A + b*this */
int vec_b200_scale(thcm_b200_t* h, double a, double* d_this);
int vec_b200_putscalar(thcm_b200_t* h, double a, double* d_this);

int vec_b200_multiply(thcm_b200_t* h, double a, const double* d_A, const double* d_B, double b, double* d_this); /* this = a*A.*B + b*this (Epetra Multiply, Ocean.C:1213,1241) */
int vec_b200_reciprocal(thcm_b200_t* h, const double* d_A, double* d_this);                                      /* this = 1./A (Ocean.C:1195) */

/* ---- diagonal 6x6 blocks of the handle's current Jacobian: scaling and the fixed point-block-Jacobi preconditioner ----
 * thcm_b200_average_block      m_scaling::average_block (src/ocean/scaling.F90:29-62) followed by the SumAll of
 *                              THCM::RecomputeScaling (src/ocean/THCM.C:1711-1718): db[ii + 6*jj] (host, column-major like the
 *                              Fortran array) = mean over all OCEAN cells of J(6c+ii, 6c+jj).  Rows the C++ glue overwrites
 *                              (integral condition, pressure fix) enter with the values the Fortran built, as in the reference.
 *                              On several GPUs the mean is the global one (sums and counts are all-reduced): independent of the
 *                              decomposition and equal to the reference on one rank; the reference's own multi-rank result is
 *                              the mean of per-rank means over ghosted sub-domains (THCM.C:1713-1718) and depends on the rank count.
 * thcm_b200_recompute_scaling  THCM::RecomputeScaling (THCM.C:1694-1750): m_scaling::scal + compute (scaling.F90:69-99, 198-284),
 *                              reciprocals and T/S averaging; fills the device vectors d_rowscal / d_colscal (nrows_local each,
 *                              either may be NULL; land rows get 1) = THCM::getRowScaling / getColScaling, and optionally the
 *                              six per-unknown factors on the host (rs6 / cs6).  IEMIC_B200_EINVAL if the block is singular.
 * thcm_b200_left_scale         Epetra_CrsMatrix::LeftScale on the Jacobian values (Ocean::scaleProblem / unscaleProblem,
 *                              src/ocean/Ocean.C:1207, 1238): val(row,:) *= d_scal[row].
 * thcm_b200_block_jacobi       d_blockinv[36*c ..] = inverse (row-major) of the 6x6 diagonal block of owned cell c of the matrix
 *                              as stored (the operator gmres_b200_solve applies); singular blocks become the identity and are
 *                              counted in *nsingular (may be NULL: no synchronisation then). */
int thcm_b200_average_block(thcm_b200_t* h, double* db36);
/* m_scaling::scal (scaling.F90:198-284) alone, host arithmetic on 36 numbers: db36 column-major -> THCM-convention factors
 * dr6 / dc6 (Trilinos uses their reciprocals).  IEMIC_B200_EINVAL: block singular to working precision, outputs untouched. */
int thcm_b200_scal(const double* db36, double* dr6, double* dc6);
int thcm_b200_recompute_scaling(thcm_b200_t* h, double* d_rowscal, double* d_colscal, double* rs6, double* cs6);
int thcm_b200_left_scale(thcm_b200_t* h, const double* d_scal);
int thcm_b200_block_jacobi(thcm_b200_t* h, double* d_blockinv, int* nsingular);
/* z = blockdiag(d_blockinv) v on the handle's stream: the apply of that preconditioner alone (Model::applyPrecon, src/utils/Model.H:52-118) */
int vec_b200_block_jacobi_apply(thcm_b200_t* h, const double* d_blockinv, const double* d_v, double* d_z);

/* ---- GMRES(m) on the handle's Jacobian (GMRESSolver.H:81-255) ----
 * ortho: GMRES_B200_MGS reproduces the reference's modified Gram-Schmidt loop (:177-181); GMRES_B200_CGS2 is the
 * fused two-pass classical Gram-Schmidt (three reductions per step: h1, h2, norm); GMRES_B200_CGS2F is the same
 * projection with two reductions per step (the second also returns <w,w>, the new norm follows from Pythagoras and
 * the normalisation is fused into the second update: 6 kernels and 2 allreduces per Arnoldi step).  prec: right (flexible) preconditioner; BLOCKJACOBI takes nrows_local/6
 * inverted 6x6 blocks, row-major, in d_blockinv; CALLBACK applies the function registered with gmres_b200_set_precond.
 * hist[it] receives resid_ after inner iteration it, and the entry behind the last one the explicit residual after the last
 * cycle: the array must hold maxit + 2 doubles.  check_every: 1 tests convergence every iteration like the reference; 0 never leaves the
 * device until the cycle ends (fixed iteration count).  Return value: 0 converged, 1 not converged (like solve()),
 * 2 accuracy limit: in the CGS2 modes the explicit residual at the end of a cycle (GMRESSolver.H:228-242) is compared
 * with the recurrence residual the reference accepts on; when it is more than max(10 tol, 100 x recurrence, min(100 eps |R||y|/|b|, 1e-8))
 * the better of the full update and the one that keeps only the leading Krylov columns whose estimated residual is still
 * attainable (iemic_b200/csrc/gmres_guard.h) is taken and, like GMRESSolver.H:213-246, the solve restarts from the explicit
 * residual; a second trip in a row, or a cycle without progress -- a singular Jacobian with a right-hand side at the rounding
 * floor of F -- ends the solve with that iterate in d_x.
 * The MGS mode keeps the reference's behaviour to the letter.  IEMIC_GMRES_GUARD=0 in the environment switches it off. */
int gmres_b200_solve(thcm_b200_t* h, const double* d_b, double* d_x, int m, int maxit, double tol,
                     int ortho, int prec, const double* d_blockinv, int check_every,
                     double* hist, int* iters);
/* Right (flexible) preconditioner supplied by the caller: z = M^-1 v, both device vectors of n = nrows_local doubles, enqueued on
 * `stream` (the solver's cudaStream_t) -- the hook Ocean::solve uses through setRightPrec (src/ocean/Ocean.C:972-975) and
 * GMRESSolver through model_.applyPrecon (GMRESSolver.H:153-172); a TRIOS / MRILU apply living on the host or on another device
 * synchronises inside.  Return 0, anything else aborts the solve with IEMIC_B200_ECALLBACK.  Used when prec = GMRES_B200_PREC_CALLBACK. */
typedef int (*gmres_b200_precond_fn)(void* user, const double* d_v, double* d_z, int64_t n, void* stream);
int gmres_b200_set_precond(thcm_b200_t* h, gmres_b200_precond_fn fn, void* user);
int gmres_b200_solve_host(thcm_b200_t* h, const double* b, double* x, int m, int maxit, double tol,
                          int ortho, int prec, const double* blockinv, int check_every,
                          double* hist, int* iters);

/* ---- instrumentation: number of kernels this handle launched since creation; device time helpers ---- */
int thcm_b200_launch_count(thcm_b200_t* h, int64_t* launches);
/* average device time (ms) of the last timed kernel groups, measured with CUDA events on the handle's stream:
 * which = 0 rhs, 1 jacobian, 2 spmv, 3 orthogonalisation (per Arnoldi step), accumulated since _timing_reset */
int thcm_b200_timing_enable(thcm_b200_t* h, int on);
int thcm_b200_timing_reset(thcm_b200_t* h);
int thcm_b200_timing_get(thcm_b200_t* h, int which, double* total_ms, int64_t* count);

/* ---- plain device-memory helpers for hosts without their own allocator ---- */
int iemic_b200_malloc(void** d_ptr, int64_t bytes);
int iemic_b200_free(void* d_ptr);
int iemic_b200_memcpy_h2d(void* d_dst, const void* src, int64_t bytes);
int iemic_b200_memcpy_d2h(void* dst, const void* d_src, int64_t bytes);

#ifdef __cplusplus
}
#endif
#endif
