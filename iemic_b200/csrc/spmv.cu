// spmv.cu -- CSR sparse matrix-vector product y = A x for sm_100a.
//
// Replaces Epetra_CrsMatrix::Apply as called by Ocean::applyMatrix (src/ocean/Ocean.C:1352-1357 of the reference) and
// the serial Fortran matAvec (src/ocean/matetc.F90:147-166).
//
// The THCM rows are short (7..24 non-zeros, 17.3 on average), so the kernel streams the matrix instead of walking
// rows: a thread block owns a run of consecutive rows holding at most NNZ_CAP non-zeros (plan built once on the host),
// loads that run's values and column indices with 16-byte, fully coalesced, evict-first loads, multiplies by the
// gathered x (read-only path; x has strong stencil locality and lives in L1/L2), parks the products in shared
// memory and lets one thread per row add its products in storage order.  Products and sums are rounded
// separately and rows are summed left to right, exactly like matAvec, so the result is bit-identical to the
// reference's serial loop and independent of the launch configuration.
// A row longer than NNZ_CAP (the dense salinity-integral row, src/ocean/THCM.C:2494-2513) gets a block to itself
// and a block-wide tree reduction.
// The kernel used when IEMIC_SPMV_TILE=0 (spmv_kernel_r; the default is spmv_tile.cu) additionally knows the graph's "regular" cells (kernels.h): blocks of whole regular
// cells compute their columns and row boundaries instead of loading them and keep their products transposed in shared
// memory.  Measured at 1 degree: 0.586 ms generic -> 0.509 ms (4.9 TB/s of the 12 B/nnz model; real DRAM traffic is lower).
// HBM traffic: 12 B per non-zero (8 B where the columns are computed) + 4 B rowptr + 8 B y per row + x once.
#include "kernels.h"
#include <cstdlib>

namespace thcm {
namespace {
constexpr int SP_THREADS = 256;
constexpr int ROWS_CAP = SP_THREADS;
constexpr unsigned FULL = 0xffffffffu;

// NNZ_CAP: products staged per block (8 B of shared memory each); MINB: resident blocks per SM the register budget allows
template <int NNZ_CAP, int MINB>
__global__ void __launch_bounds__(SP_THREADS, MINB) spmv_kernel(const int* __restrict__ blk_row, const int* __restrict__ rowptr,
                                                          const int* __restrict__ col, const double* __restrict__ val,
                                                          const double* __restrict__ x, double* __restrict__ y) {
  __shared__ __align__(16) double prod[NNZ_CAP + 8];
  const int r0 = blk_row[blockIdx.x], r1 = blk_row[blockIdx.x + 1];
  const int nr = r1 - r0;
  const int p0 = __ldg(rowptr + r0), p1 = __ldg(rowptr + r1);
  if (nr == 1 && p1 - p0 > NNZ_CAP) { // long row
    double s = 0.0;
    for (int p = p0 + threadIdx.x; p < p1; p += SP_THREADS) s += __ldcs(val + p) * __ldg(x + __ldcs(col + p));
    for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(FULL, s, o);
    if ((threadIdx.x & 31) == 0) prod[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
      double t = 0.0;
      for (int w = 0; w < SP_THREADS / 32; w++) t += prod[w];
      y[r0] = t;
    }
    return;
  }
  const int pa = p0 & ~3; // 16-byte aligned start for the int4 / double2 loads (arrays are padded by 4 entries)
  int4 c[NNZ_CAP / (4 * SP_THREADS)];
  double2 va[NNZ_CAP / (4 * SP_THREADS)], vb[NNZ_CAP / (4 * SP_THREADS)];
#pragma unroll
  for (int it = 0; it < NNZ_CAP / (4 * SP_THREADS); it++) {
    const int p = pa + 4 * (threadIdx.x + SP_THREADS * it);
    if (p < p1) {
      c[it] = __ldcs(reinterpret_cast<const int4*>(col + p));
      va[it] = __ldcs(reinterpret_cast<const double2*>(val + p));
      vb[it] = __ldcs(reinterpret_cast<const double2*>(val + p + 2));
    }
  }
#pragma unroll
  for (int it = 0; it < NNZ_CAP / (4 * SP_THREADS); it++) {
    const int p = pa + 4 * (threadIdx.x + SP_THREADS * it);
    if (p < p1) {
      double2 o0, o1;
      // entries before p0 / after p1 belong to other blocks' rows: never summed, but keep the gathers in range
      o0.x = (p + 0 >= p0 && p + 0 < p1) ? va[it].x * __ldg(x + c[it].x) : 0.0;
      o0.y = (p + 1 >= p0 && p + 1 < p1) ? va[it].y * __ldg(x + c[it].y) : 0.0;
      o1.x = (p + 2 >= p0 && p + 2 < p1) ? vb[it].x * __ldg(x + c[it].z) : 0.0;
      o1.y = (p + 3 >= p0 && p + 3 < p1) ? vb[it].y * __ldg(x + c[it].w) : 0.0;
      *reinterpret_cast<double2*>(prod + (p - pa)) = o0;
      *reinterpret_cast<double2*>(prod + (p - pa) + 2) = o1;
    }
  }
  __syncthreads();
  if (threadIdx.x < nr) {
    const int q0 = __ldg(rowptr + r0 + threadIdx.x) - pa, q1 = __ldg(rowptr + r0 + threadIdx.x + 1) - pa;
    double s = 0.0;
    for (int q = q0; q < q1; q++) s = __dadd_rn(s, prod[q]);
    y[r0 + threadIdx.x] = s;
  }
}

// Scalar-lane flavour: lane = consecutive non-zero, so one gather instruction touches the x entries of ~2 adjacent rows
// (a dozen 32-byte sectors) instead of the ~7 rows a 4-per-thread vector layout spreads it over: the L1 gather
// wavefronts, not HBM, bound this kernel.
template <int NNZ_CAP, int MINB, int NT>
__global__ void __launch_bounds__(NT, MINB) spmv_kernel_s(const int* __restrict__ blk_row, const int* __restrict__ rowptr,
                                                            const int* __restrict__ col, const double* __restrict__ val,
                                                            const double* __restrict__ x, double* __restrict__ y) {
  __shared__ double prod[NNZ_CAP];
  const int r0 = blk_row[blockIdx.x], r1 = blk_row[blockIdx.x + 1];
  const int nr = r1 - r0;
  const int p0 = __ldg(rowptr + r0), p1 = __ldg(rowptr + r1);
  if (nr == 1 && p1 - p0 > NNZ_CAP) { // long row
    double s = 0.0;
    for (int p = p0 + threadIdx.x; p < p1; p += NT) s += __ldcs(val + p) * __ldg(x + __ldcs(col + p));
    for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(FULL, s, o);
    if ((threadIdx.x & 31) == 0) prod[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
      double t = 0.0;
      for (int w = 0; w < NT / 32; w++) t += prod[w];
      y[r0] = t;
    }
    return;
  }
  constexpr int NIT = NNZ_CAP / NT;
  int c[NIT];
  double v[NIT];
#pragma unroll
  for (int it = 0; it < NIT; it++) {
    const int p = p0 + threadIdx.x + NT * it;
    if (p < p1) { c[it] = __ldcs(col + p); v[it] = __ldcs(val + p); }
  }
#pragma unroll
  for (int it = 0; it < NIT; it++) {
    const int p = p0 + threadIdx.x + NT * it;
    if (p < p1) prod[p - p0] = v[it] * __ldg(x + c[it]);
  }
  __syncthreads();
  if (threadIdx.x < nr) {
    const int q0 = __ldg(rowptr + r0 + threadIdx.x) - p0, q1 = __ldg(rowptr + r0 + threadIdx.x + 1) - p0;
    double s = 0.0;
    for (int q = q0; q < q1; q++) s = __dadd_rn(s, prod[q]);
    y[r0 + threadIdx.x] = s;
  }
}

// Regular-cell flavour of the scalar-lane kernel: a block of whole regular cells knows its columns (6*cell + tab[q]) and
// its row boundaries (fixed 24/22/7/11/20/20 split), so the value loads and the x gathers are issued together instead
// of behind the rowptr and column loads: two dependent memory stages instead of four, and 4 B per non-zero less traffic.
// Generic blocks run the code of spmv_kernel_s.  Same products, same summation order: bit-identical results.
__device__ __forceinline__ int reg_row_off(int a) { // first entry of row a (U,V,W,P,T,S) inside a cell's 104 values
  return a == 0 ? 0 : (a == 1 ? 24 : (a == 2 ? 46 : (a == 3 ? 53 : (a == 4 ? 64 : 84))));
}
template <int NNZ_CAP, int MINB, int NT>
__global__ void __launch_bounds__(NT, MINB) spmv_kernel_r(const int* __restrict__ blk_row, const int* __restrict__ blk_p0,
                                                            const unsigned char* __restrict__ blk_reg, const __grid_constant__ SpmvTab tab,
                                                            const int* __restrict__ rowptr, const int* __restrict__ col,
                                                            const double* __restrict__ val, const double* __restrict__ x,
                                                            double* __restrict__ y) {
  // Regular blocks park the products transposed -- entry j of row r at prod[j * RS + r], RS odd -- so that the row sums
  // (thread = row, all threads at the same j) and the product stores are both free of bank conflicts; the plain layout
  // cost the generic kernel 2.4 shared-memory wavefronts per useful one and the L1 data pipe is what saturates here.
  constexpr int G = NNZ_CAP / SPMV_CELL_NNZ, RS = 6 * G + 1, PSIZE = 24 * RS > NNZ_CAP ? 24 * RS : NNZ_CAP;
  __shared__ double prod[PSIZE];
  __shared__ int tabs[SPMV_CELL_NNZ];   // column offset of entry q of a regular cell
  __shared__ short slot[SPMV_CELL_NNZ]; // j * RS + a: where entry q goes in `prod` (before adding 6 * cell)
  constexpr int NIT = NNZ_CAP / NT;
  const int r0 = blk_row[blockIdx.x], r1 = blk_row[blockIdx.x + 1];
  const int nr = r1 - r0;
  const int pstart = blk_p0[blockIdx.x]; // requested together with the row bounds and the flag: one latency, not two
  if (blk_reg[blockIdx.x]) {
    const int p0 = pstart;
    for (int t = threadIdx.x; t < SPMV_CELL_NNZ; t += NT) {
      tabs[t] = tab.off[t];
      const int a = t >= 84 ? 5 : (t >= 64 ? 4 : (t >= 53 ? 3 : (t >= 46 ? 2 : (t >= 24 ? 1 : 0))));
      slot[t] = (short)((t - reg_row_off(a)) * RS + a);
    }
    const int nn = (nr / 6) * SPMV_CELL_NNZ, c0 = 6 * (r0 / 6);
    // branch-free on purpose (indices clamped instead of predicated): all value loads and all gathers of a thread are
    // in flight together; with per-iteration branches the compiler consumes each gather before issuing the next
    double v[NIT], xv[NIT];
#pragma unroll
    for (int it = 0; it < NIT; it++) v[it] = __ldcs(val + p0 + min(threadIdx.x + NT * it, nn - 1));
    __syncthreads();
    int where[NIT];
#pragma unroll
    for (int it = 0; it < NIT; it++) {
      const int d = min(threadIdx.x + NT * it, nn - 1);
      const int cell = d / SPMV_CELL_NNZ, q = d - cell * SPMV_CELL_NNZ;
      xv[it] = __ldg(x + c0 + 6 * cell + tabs[q]);
      where[it] = slot[q] + 6 * cell; // (computing this by arithmetic instead was measured: register spills, slower)
    }
#pragma unroll
    for (int it = 0; it < NIT; it++)
      if (threadIdx.x + NT * it < nn) prod[where[it]] = v[it] * xv[it];
    __syncthreads();
    if (threadIdx.x < nr) {
      const int a = threadIdx.x % 6;
      const int len = (a == 5 ? SPMV_CELL_NNZ : reg_row_off(a + 1)) - reg_row_off(a);
      double s = 0.0;
      for (int j = 0; j < len; j++) s = __dadd_rn(s, prod[j * RS + threadIdx.x]);
      y[r0 + threadIdx.x] = s;
    }
    return;
  }
  const int p0 = pstart, p1 = __ldg(rowptr + r1);
  if (nr == 1 && p1 - p0 > NNZ_CAP) { // long row
    double s = 0.0;
    for (int p = p0 + threadIdx.x; p < p1; p += NT) s += __ldcs(val + p) * __ldg(x + __ldcs(col + p));
    for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(FULL, s, o);
    if ((threadIdx.x & 31) == 0) prod[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
      double t = 0.0;
      for (int w = 0; w < NT / 32; w++) t += prod[w];
      y[r0] = t;
    }
    return;
  }
  int c[NIT];
  double v[NIT];
#pragma unroll
  for (int it = 0; it < NIT; it++) {
    const int p = p0 + threadIdx.x + NT * it;
    if (p < p1) { c[it] = __ldcs(col + p); v[it] = __ldcs(val + p); }
  }
#pragma unroll
  for (int it = 0; it < NIT; it++) {
    const int p = p0 + threadIdx.x + NT * it;
    if (p < p1) prod[p - p0] = v[it] * __ldg(x + c[it]);
  }
  __syncthreads();
  if (threadIdx.x < nr) {
    const int q0 = __ldg(rowptr + r0 + threadIdx.x) - p0, q1 = __ldg(rowptr + r0 + threadIdx.x + 1) - p0;
    double s = 0.0;
    for (int q = q0; q < q1; q++) s = __dadd_rn(s, prod[q]);
    y[r0 + threadIdx.x] = s;
  }
}

// (Staging the nine neighbour lines of x in shared memory instead of gathering from global memory was measured too:
// 0.56-0.80 ms against 0.51 ms -- the extra barrier and shared-memory traffic cost more than the gathers.)

// variant table; IEMIC_SPMV_VARIANT selects one at start-up (tuning aid), default = the best measured on B200
struct Variant { int cap, minb, nt; };
const Variant kVariants[] = {{4096, 3, 256}, {2048, 4, 256}, {2048, 6, 256}, {1024, 8, 256}, {512, 16, 128}, {1024, 12, 128}, {1024, 8, 256}, {256, 32, 64}, {512, 24, 64}, {256, 16, 128}};
int g_variant = -1;
int variant() {
  if (g_variant < 0) {
    const char* e = getenv("IEMIC_SPMV_VARIANT");
    int v = e ? atoi(e) : 4;
    g_variant = (v >= 0 && v < 10) ? v : 4;
  }
  return g_variant;
}
} // namespace

// regular-cell kernel configurations (IEMIC_SPMV_RVAR, tuning aid): {products staged, blocks/SM, threads}; they replace
// the default generic variant (4) whenever the plan carries regular blocks
static const int kRegular[][3] = {{512, 16, 128}, {1024, 8, 256}, {1024, 12, 128}, {1024, 16, 128}, {2048, 4, 256}, {2048, 8, 256}, {1024, 16, 64}, {1536, 8, 128}};
static int g_rvar = -1;
static int rvariant() {
  if (g_rvar < 0) {
    const char* e = getenv("IEMIC_SPMV_RVAR");
    int v = e ? atoi(e) : 2;
    g_rvar = (v >= 0 && v < 8) ? v : 2;
  }
  return g_rvar;
}
static bool regular_on() { return variant() == 4; }
int spmv_nnz_cap() { return (regular_on() ? kRegular[rvariant()][0] : kVariants[variant()].cap) - 4; } // an aligned start may shift a run by up to 3 entries
int spmv_rows_cap() { return regular_on() ? kRegular[rvariant()][2] : kVariants[variant()].nt; }       // one thread per row in the reduction phase
int spmv_regular_cells() { return regular_on() ? kRegular[rvariant()][0] / SPMV_CELL_NNZ : 0; }

void launch_spmv(const SpmvPlan& p, const int* rowptr, const int* col, const double* val, const double* x, double* y, cudaStream_t st) {
  if (p.nblocks <= 0) return;
  if (p.blk_reg && p.blk_p0 && regular_on()) {
#define RLAUNCH(CAP, MINB, NT) spmv_kernel_r<CAP, MINB, NT><<<p.nblocks, NT, 0, st>>>(p.blk_row, p.blk_p0, p.blk_reg, p.tab, rowptr, col, val, x, y)
    switch (rvariant()) {
      case 0: RLAUNCH(512, 16, 128); break;
      case 1: RLAUNCH(1024, 8, 256); break;
      case 2: RLAUNCH(1024, 12, 128); break;
      case 3: RLAUNCH(1024, 16, 128); break;
      case 4: RLAUNCH(2048, 4, 256); break;
      case 5: RLAUNCH(2048, 8, 256); break;
      case 6: RLAUNCH(1024, 16, 64); break;
      default: RLAUNCH(1536, 8, 128); break;
    }
#undef RLAUNCH
    return;
  }
  switch (variant()) {
    case 0: spmv_kernel<4096, 3><<<p.nblocks, SP_THREADS, 0, st>>>(p.blk_row, rowptr, col, val, x, y); break;
    case 1: spmv_kernel<2048, 4><<<p.nblocks, SP_THREADS, 0, st>>>(p.blk_row, rowptr, col, val, x, y); break;
    case 2: spmv_kernel<2048, 6><<<p.nblocks, SP_THREADS, 0, st>>>(p.blk_row, rowptr, col, val, x, y); break;
    case 3: spmv_kernel<1024, 8><<<p.nblocks, SP_THREADS, 0, st>>>(p.blk_row, rowptr, col, val, x, y); break;
    case 4: spmv_kernel_s<512, 16, 128><<<p.nblocks, 128, 0, st>>>(p.blk_row, rowptr, col, val, x, y); break;
    case 5: spmv_kernel_s<1024, 12, 128><<<p.nblocks, 128, 0, st>>>(p.blk_row, rowptr, col, val, x, y); break;
    case 6: spmv_kernel_s<1024, 8, 256><<<p.nblocks, 256, 0, st>>>(p.blk_row, rowptr, col, val, x, y); break;
    case 7: spmv_kernel_s<256, 32, 64><<<p.nblocks, 64, 0, st>>>(p.blk_row, rowptr, col, val, x, y); break;
    case 8: spmv_kernel_s<512, 24, 64><<<p.nblocks, 64, 0, st>>>(p.blk_row, rowptr, col, val, x, y); break;
    default: spmv_kernel_s<256, 16, 128><<<p.nblocks, 128, 0, st>>>(p.blk_row, rowptr, col, val, x, y); break;
  }
}

} // namespace thcm
