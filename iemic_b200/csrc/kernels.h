// kernels.h -- launch interface between the C ABI (capi.cu) and the sm_100a kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "thcm_rows.cuh"
#include "thcm_cell.cuh"
#include "thcm_coupled.cuh"
#include "peer.cuh"

namespace thcm {

struct AsmArgs {
  Scalars S;
  Tables T;
  const double* x;        // stored state vector (owned cells, then halo ring)
  const uint64_t* sig;    // per stored cell
  const int* rowptr;      // static graph
  const double* frc;      // forcing, already masked (rhs only)
  double* val;            // CSR values (jacobian only)
  double* diagB;          // mass matrix diagonal (jacobian only)
  double* F;              // residual (rhs only)
  SpecialRows sp;
  int ntx, nty, ntz;
  int part;               // 0 every tile, 1 only tiles that read no halo cell, 2 only tiles on the block's rim
  int kz0, nkz;           // restrict the launch to nkz tile layers starting at layer kz0 (nkz = 0: all layers)
};
int asm_tile_levels();    // model levels per tile layer
void launch_diagb(const Scalars& S, const uint64_t* sig, const SpecialRows& sp, double* diagB, cudaStream_t st); // THCM::evaluateB
// surface rows of Frc from the per-latitude profiles (device copy of Host::fprof) after a parameter change
void launch_forcing_surface(const Scalars& S, const uint64_t* sig, const double* fprof, double* frc, cudaStream_t st);
void launch_jac(const AsmArgs& a, cudaStream_t st);
void launch_rhs(const AsmArgs& a, cudaStream_t st);
void launch_asm(const AsmArgs& a, cudaStream_t st); // Jacobian and residual of the same state in one pass over the staged tiles
// ---- row N2: coupling to an atmosphere / sea-ice model (coupled.cu) ----
void launch_forcing_cells(const Scalars& S, const uint64_t* sig, const double* ftcell /* nloc*mloc */, double* frc, cudaStream_t st);
struct DftdmArgs {
  Scalars S; const uint64_t* sig; const double* x;
  const double *tatm, *albe, *qatm;  // whole surface, (j-1)*n + (i-1), or null (zero field)
  const double* suno;                // (m + 2), Fortran j
  DftdmScal c;
  double* out;                       // nloc*mloc
};
void launch_dftdm(const DftdmArgs& a, cudaStream_t st);
void launch_coupling_apply(int nrows, const int* rows, const int* beg, const int* jco, const double* co, const double* x, double* y, int accumulate,
                           cudaStream_t st);
int jac_smem_bytes();
int rhs_smem_bytes();
int asm_smem_bytes();

// integral condition (THCM.C:1006-1018): F[row] = sign * (sum_c coef[c]*x[6c+5] - corr); partial sums per block,
// finished (after the optional allreduce) by intcond_finish.
void launch_intcond_dot(const double* coefS, const double* x, long long ncell, double* partial /* 1 double */, cudaStream_t st);
void launch_intcond_finish(const double* dotval, double sign, double corr, double* F, int row, int rowPfix1, int rowPfix2, cudaStream_t st);
void launch_intcond_add(const double* dotval, double sign, double* y, int row, cudaStream_t st); // y[row] += sign * dotval (other ranks' share of the dense row)

// ---- CSR SpMV (spmv.cu) ----
// Most cells of the static THCM graph are "regular": all 27 neighbours stored, so the 104 columns of the cell's six rows
// are the cell's own first column plus one fixed table of offsets.  Blocks made of such cells compute their column
// indices instead of loading them (the plan is verified against the real column array on the host); everything else --
// domain edges, the periodic seam, halo neighbours, the dense integral row -- takes the generic CSR path.
constexpr int SPMV_CELL_NNZ = 104;
struct SpmvTab { int off[SPMV_CELL_NNZ]; }; // column of entry q of a regular cell, relative to 6 * cell
struct SpmvPlan {
  int nblocks;
  const int* blk_row;            // device, nblocks+1 row boundaries; a block with a single "long" row is reduced block-wide
  int nnz_cap;                   // max non-zeros staged per block
  const int* blk_p0;             // device, first non-zero of every block (regular blocks skip the rowptr loads); may be null
  const unsigned char* blk_reg;  // device, 1 = block of whole regular cells; may be null (all generic)
  SpmvTab tab;                   // column offsets of a regular cell, relative to 6 * cell
};
void launch_spmv(const SpmvPlan& p, const int* rowptr, const int* col, const double* val, const double* x, double* y, cudaStream_t st);
// structured product over the static THCM graph (spmv_tile.cu): columns of complete cells are compile-time tile offsets
void launch_spmv_tile(const Scalars& S, const int* rowptr, const int* col, const double* val, const double* x, double* y, int long_row, cudaStream_t st);
int spmv_tile_long_row();
int spmv_nnz_cap();
int spmv_rows_cap();
int spmv_regular_cells(); // regular cells per block

// ---- diagonal 6x6 blocks of the Jacobian: THCM scaling, point-block Jacobi, row scaling (scaling.cu) ----
struct DiagArgs {
  const int* rowptr;
  const unsigned int* dinfo;  // per row: bits 0..25 offset of the first centre-cell entry inside the row, bits 26..31 unknowns present
  const double* val;
  const uint64_t* sig;
  long long ncell;            // owned cells
  SpecialRows sp;             // rows overwritten by the C++ glue and the side buffer holding their Fortran-level centre entries
  int stored_rows;            // 1: take those rows as stored (the matrix GMRES sees); 0: as the Fortran built them (m_scaling)
};
struct ScalTab { double r[6], c[6]; }; // per-unknown factors on ocean cells (land rows: 1)
int average_block_partials(); // doubles of workspace launch_average_block needs
// out37[ii*6+jj] = sum over owned OCEAN cells of J(6c+ii,6c+jj), out37[36] = their number (scaling.F90:29-62)
void launch_average_block(const DiagArgs& a, double* partial, unsigned int* ticket, double* out37, cudaStream_t st);
void launch_scaling_fill(const uint64_t* sig, long long nrows, const ScalTab& t, double* rowscal, double* colscal, cudaStream_t st);
// binv[c] = inverse of the diagonal block of cell c, row-major (identity where the block is singular, counted in *nsingular)
void launch_block_inverse(const DiagArgs& a, double* binv, int* nsingular, cudaStream_t st);
void launch_left_scale(const int* rowptr, double* val, const double* s, long long nrows, cudaStream_t st); // val(row,:) *= s[row]
void launch_multiply(double a, const double* A, const double* B, double b, double* y, long long n, cudaStream_t st); // y = a*A.*B + b*y
void launch_reciprocal(const double* A, double* y, long long n, cudaStream_t st);

// ---- vector / GMRES kernels (gmres.cu) ----
struct RedBuf { double* partial; unsigned int* ticket; int nblk_max; }; // workspace for block-partial reductions
int red_blocks();
// Arnoldi bookkeeping on the device (single thread): column j of H from h1 (+h2), norm from nrm2 (squared),
// Givens rotations (GMRESSolver.H:190-200), residual estimate hist[j] = |s[j+1]| / bnorm.
struct GivensArgs {
  double* H; int ldh;          // (m+1) x m, column major
  double* cs; double* sn; double* s;
  const double* h1; const double* h2; // h2 may be null
  const double* nrm2;          // squared norm of w (or norm if !sq)
  int sq;
  int pyth;                    // norm from h2: sqrt(h2[j+1] - sum_k h2[k]^2), h2[j+1] = <w,w> before the second projection
  const double* bnorm;         // device scalar: ||b||
  double* hist;
  double* hnext;               // receives H[j+1][j] (the norm) for the normalisation kernel
  int j;
};
// out[k] = <V_k, w>, k = 0..nv-1 (V_k = V + k*ldv), deterministic two-stage reduction finished by the last block.
// With pr.nranks > 0 the finishing block also sums over the ranks through peer memory (peer.cuh); this holds for every
// reducing kernel below.
void launch_multi_dot(const double* V, long long ldv, int nv, const double* w, long long n, double* out, RedBuf rb, cudaStream_t st,
                      const PeerRed& pr = PeerRed{}, const GivensArgs* fused_givens = nullptr /* run this step's bookkeeping in the finishing block */);
// w -= sum_k h[k] V_k ; optionally hacc[k] += h[k] ; optionally nrm2_out = ||w_new||^2
void launch_multi_axpy(const double* V, long long ldv, int nv, const double* h, double* w, long long n,
                       double* nrm2_out, RedBuf rb, cudaStream_t st, const double* inv_scale = nullptr /* w /= *inv_scale afterwards */,
                       const PeerRed& pr = PeerRed{});
// Fused CGS2 passes 2+3: w -= sum_k h[k] V_k, then out[k] = <V_k, w> (k < nv), out[nv] = <w, w> from the same read of the basis
// (nv <= multi_axpy_dot_max_nv()); finishing block as in launch_multi_dot.
int multi_axpy_dot_max_nv();
void launch_multi_axpy_dot(const double* V, long long ldv, int nv, const double* h, double* w, long long n, double* out, RedBuf rb, cudaStream_t st,
                           const PeerRed& pr = PeerRed{}, const GivensArgs* fused_givens = nullptr);
// MGS fused step: if (hprev) w -= hprev[0]*vprev ; out = <w, vnext> (vnext may be null: then out = <w,w>)
void launch_axpy_dot(const double* hprev, const double* vprev, const double* vnext, double* w, long long n, double* out, RedBuf rb, cudaStream_t st,
                     const PeerRed& pr = PeerRed{});
void launch_dot(const double* a, const double* b, long long n, double* out, RedBuf rb, cudaStream_t st, const PeerRed& pr = PeerRed{});
// same product with the reduction scratch in a static device array (no allocation anywhere); returns 0 or a cudaError_t
int launch_dot_noalloc(const double* a, const double* b, long long n, double* out, cudaStream_t st);
long long dot_noalloc_max_n();
void launch_absmax(const double* a, long long n, double* out /* 1 double */, cudaStream_t st);
// y = alpha * x with alpha = (inv ? 1/s[0] : s[0]) read on the device ; s may be sqrt'ed first (sq: s[0] holds a squared norm)
void launch_scale_copy(const double* x, double* y, long long n, const double* s, int inv, int sq, cudaStream_t st);
void launch_update(double a, const double* A, double b, double* y, long long n, cudaStream_t st); // y = a*A + b*y
void launch_fill(double a, double* y, long long n, cudaStream_t st);
// x += sum_k yk[k] * Z_k
void launch_multi_axpy_plus(const double* Z, long long ldz, int nv, const double* yk, double* x, long long n, cudaStream_t st);
// z = blockdiag(Binv) * v, 6x6 blocks
void launch_block_jacobi(const double* binv, const double* v, double* z, long long nblocks, cudaStream_t st);
// r = b - r  (r holds A x on entry)
void launch_residual(const double* b, double* r, long long n, cudaStream_t st);

void launch_givens(const GivensArgs& g, cudaStream_t st);

// One GMRES(m) cycle in one cooperative launch (gmres_cycle.cu): single GPU, fused CGS2, no or block-Jacobi preconditioner.
// The basis starts with v_0 in V; the kernel builds up to maxcols further columns and stops early once hist[i] < tol (check).
struct CycleArgs {
  const int* rowptr; const int* col; const double* val;
  const double* binv;          // 6x6 inverse diagonal blocks or null
  double* V; double* Z;        // basis (m+1 vectors of ld) ; preconditioned basis (with binv)
  long long ld, n;
  int long_row;                // row summed by a whole block (dense salinity-integral row) or -1
  double* part; int ldp;       // gmres_cycle_partials(m) doubles ; ldp is set by the launcher
  int m, maxcols;
  double tol; int check;
  int* out_cols; int* stop;    // device ints: columns built ; *stop must be 0 at launch
  double* h1; double* h2;      // the buffers gv.h1 / gv.h2 point at
  GivensArgs gv;
};
// The same cycle for a tiny problem inside one thread-block cluster, basis in distributed shared memory (gmres_cluster.cu);
// plain CGS2 (explicit norm) or the fused variant.  R, RP, KG, npad are filled in by the launcher.
struct ClusterArgs {
  const int* rowptr; const int* col; const double* val;
  const double* binv;
  double* V; double* Z;
  long long ld, n;
  int m, maxcols, fused;
  double tol; int check;
  int* out_cols; int* stop;
  double* H; int ldh;
  double* cs; double* sn; double* s; double* hist;
  const double* bnorm;
  int R, RP, KG, npad;
};
int gmres_cluster_fits(long long n, int m);
int launch_gmres_cluster(ClusterArgs a, cudaStream_t st); // 0, a cudaError_t, or -1: does not fit
long long gmres_cycle_max_rows();
int gmres_cycle_partials(int m);
int launch_gmres_cycle(CycleArgs a, cudaStream_t st); // 0 or a cudaError_t
// y = H(0:k,0:k)^-1 s(0:k)   (GMRESSolver::backSolve, :258-289)
void launch_backsolve(const double* H, int ldh, const double* s, int k, double* y, cudaStream_t st);
// beta = sqrt(nrm2) ; s[0] = beta ; s[1..m] = 0 ; bnorm keeps ||b|| (set separately)
void launch_cycle_init(const double* nrm2, double* s, int m, double* beta_out, cudaStream_t st);
void launch_sqrt_store(const double* nrm2, double* out, cudaStream_t st);

} // namespace thcm
