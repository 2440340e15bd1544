/* thcm_oracle.h -- C ABI of the CPU oracle (TEST INFRASTRUCTURE, never shipped).
 *
 * The oracle is a structure-faithful C++ restatement of the reference's Fortran THCM
 * kernels (src/ocean/{usrc,spf,boundary,assemble,matetc,forcing,grid,mix_imp}.F90) plus
 * the C++ glue of src/ocean/THCM.C (static graph, sign flip, integral condition) and of
 * src/gmressolver/GMRESSolver.H.  It materialises the dense stencil tensor An(27,6,6,n,m,l)
 * exactly like the reference does.
 *
 * PARITY STATUS: the reference itself cannot be built in this image (no Fortran compiler,
 * no MPI, no Trilinos) and its tests hold no stored RHS/Jacobian vectors.  The oracle is pinned
 * (a) by the KNOWN ANSWER of the reference's own integration test src/tests/intt_2dmoc.C:96-97,
 * 127-128 -- psi_max = 14.7 +- 0.1 Sv, psi_min = 0 +- 1e-4 and the mirror symmetry of the two
 * branches after the chain of continuations on the 3x6x6 2-D MOC case (tests/test_known_answer_2dmoc.py
 * gets 14.704 Sv), which exercises lin, nlin_rhs, nlin_jac, boundaries, forcing, the integral
 * condition and Newton-consistency of the Jacobian together -- and tests/continuation_oracle.cpp reaches the same numbers
 * (14.7043 Sv, 0, symmetry 5.7e-5) with the reference's own UNMODIFIED Continuation<Model> (src/continuation/Continuation.H,
 * compiled in place) driving this oracle through all seven legs of that test (tests/test_continuation_oracle.py); (a') by the GOLDEN STATE the reference itself wrote,
 * test/ocean/ocean_reference.h5 (end of the 16^3 gateway continuation of src/tests/reft_ocean.C: land mask, periodic,
 * forcing type 2, convective adjustment with P_VC = 9.95): the oracle's residual at that state with the parameters stored
 * beside it is ||F||_2 = 7.7e-6 against ||F(0)||_2 = 19.8 -- 1e-11 in the momentum rows, 1e-15 in continuity -- i.e. a
 * root of the reference's F is a root of the oracle's F, and conversely Newton on the oracle's F and J -- from that state, or
 * from the zero state through a continuation (tests/golden/reft_from_scratch.py) -- ends on the golden vector: every u, v, w,
 * T, S value within 4e-6, the per-field norms reft_ocean.C:62-79 compares within 6e-5 of its 1e-3 (tests/test_reference_state.py, fixture
 * tests/golden/ocean_reference.npz made by tests/golden/make_fixtures.py); and (b) by the invariants I1-I5 of
 * SURVEY.md section 8c (tests/test_oracle_invariants.py, tests/test_dist_cpu.py).  The GMRES restatement is pinned bit for
 * bit against the reference's GMRESSolver.H itself, compiled and run here (tests/test_gmres_reference_pin.py).  Not pinned by
 * any reference number: the Jacobian values themselves (consistency with F only: I4 and the directional
 * difference at the golden state).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may call anything declared here.
 */
#ifndef THCM_ORACLE_H
#define THCM_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct thcm_oracle thcm_oracle;

/* Mirrors the argument list of Fortran init (usrc.F90:6-139) + m_global::initialize flags. */
typedef struct {
  int n, m, l;
  double xmin, xmax, ymin, ymax; /* radians (THCM.C:201-204 converts degrees) */
  int periodic;
  double hdim, qz;
  double alphaT, alphaS;
  int ih, vmix, tap, rho_mixing, coriolis_on, forcing_type;
  int tres, sres, iza, ite, its;
  int coupled_T, coupled_S;
  int int_sign;          /* "Salinity Integral Sign", THCM.C:254 */
  int intcond_i, intcond_j; /* "Integral row coordinate i/j" (0-based, -1 = N-1 / M-1) */
  int fix_pressure_points;
} thcm_oracle_config;

thcm_oracle* thcm_oracle_create(const thcm_oracle_config* cfg, const int* landm /* (n+2)(m+2)(l+2), i fastest */);
void thcm_oracle_destroy(thcm_oracle*);

void thcm_oracle_set_par(thcm_oracle*, int idx1, double value); /* setparcs, usrc.F90:163 */
double thcm_oracle_get_par(thcm_oracle*, int idx1);
int thcm_oracle_ndim(thcm_oracle*);

/* Fortran-level entry points (usrc.F90:432, :506): value-dependent CSR, 1-based. */
void thcm_oracle_fortran_rhs(thcm_oracle*, const double* un, double* B);
void thcm_oracle_fortran_matrix(thcm_oracle*, const double* un);
int64_t thcm_oracle_fortran_nnz(thcm_oracle*);
void thcm_oracle_fortran_csr(thcm_oracle*, int* begA, int* jcoA, double* coA, double* coB);
void thcm_oracle_frc(thcm_oracle*, double* frc); /* current Frc (after boundaries masking) */

/* THCM.C level (THCM::evaluate :949-1192): static CreateMaximalGraph pattern, 0-based. */
int64_t thcm_oracle_graph_nnz(thcm_oracle*);
void thcm_oracle_graph(thcm_oracle*, int* rowptr /* ndim+1 */, int* col /* nnz */);
void thcm_oracle_eval_rhs(thcm_oracle*, const double* x, double* F);
void thcm_oracle_eval_jac(thcm_oracle*, const double* x, double* val /* nnz */, double* diagB /* ndim */);
int thcm_oracle_rowintcon(thcm_oracle*);
void thcm_oracle_intcond_coeff(thcm_oracle*, double* c /* ndim */);
void thcm_oracle_set_intcond_correction(thcm_oracle*, double corr);
int64_t thcm_oracle_dropped_entries(thcm_oracle*); /* Fortran entries outside the static graph in the last eval_jac */

/* Coupling to an atmosphere / sea-ice model: usrc.F90:237-333 (pars in the order of the bind(C) structs),
 * inserts.F90:12-166 (which: 0 tatm, 1 qatm, 2 albe, 3 patm, 4 qsa, 5 msi, 6 gsi; takes effect at the next set_* call as in the
 * Fortran), getdeps usrc.F90:201-219 (Ooa, Os, nus, eta, lvsc, qdim, pQSnd), suno(j), probe.F90:381-444 (dftdm).
 * Pinned by the known answer of the reference's own coupling test (src/tests/test_coupled.C:402-447: the T-row sum of C12) and by
 * differences of this oracle's F with respect to the inserted fields (tests/test_coupled_cpu.py); no stored vectors exist for it. */
void thcm_oracle_set_atmos_parameters(thcm_oracle*, const double* pars /* 18 */);
void thcm_oracle_set_seaice_parameters(thcm_oracle*, const double* pars /* 7 */);
void thcm_oracle_insert(thcm_oracle*, int which, const double* field /* n*m */);
void thcm_oracle_getdeps(thcm_oracle*, double* out /* 7 */);
void thcm_oracle_suno(thcm_oracle*, double* suno /* m */);
void thcm_oracle_dftdm(thcm_oracle*, const double* un, double* dftdm /* n*m */);

/* m_scaling::average_block / scal (scaling.F90:29-62, 198-284) on the Fortran CSR of the last fortran_matrix / eval_jac
 * call, and THCM::RecomputeScaling (THCM.C:1694-1750) on one process.  db, mat: column-major 6x6. */
void thcm_oracle_average_block(thcm_oracle*, double* db /* 36 */);
int thcm_oracle_scal(const double* mat /* 36 */, double* dr /* 6 */, double* dc /* 6 */); /* 0: singular block */
int thcm_oracle_recompute_scaling(thcm_oracle*, double* rowScaling /* ndim */, double* colScaling /* ndim */);

/* matAvec-order CSR SpMV (matetc.F90:147-166), 0-based CSR. */
void thcm_oracle_spmv(int nrows, const int* rowptr, const int* col, const double* val, const double* x, double* y);

/* GMRESSolver.H:81-255 restated for plain vectors; prec: 0 none, 1 = 6x6 point-block Jacobi
 * (blocks given inverted, row-major, nrows/6 of them) applied as right (flexible) preconditioner.
 * hist[it] = resid_ after inner iteration it (length >= maxit+2).  Returns GMRESSolver::solve() code. */
int thcm_oracle_gmres(int nrows, const int* rowptr, const int* col, const double* val,
                      const double* b, double* x, int m, int maxit, double tol,
                      int prec, const double* blockinv, double* hist, int* iters);

#ifdef __cplusplus
}
#endif
#endif
