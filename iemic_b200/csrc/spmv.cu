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
// HBM traffic: 12 B per non-zero + 4 B rowptr + 8 B y per row + x once.
#include "kernels.h"

namespace thcm {
namespace {
constexpr int SP_THREADS = 256;
constexpr int NNZ_CAP = 4096;  // products staged per block (32 KB of shared memory)
constexpr int ROWS_CAP = SP_THREADS;
constexpr unsigned FULL = 0xffffffffu;

__global__ void __launch_bounds__(SP_THREADS) spmv_kernel(const int* __restrict__ blk_row, const int* __restrict__ rowptr,
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
} // namespace

int spmv_nnz_cap() { return NNZ_CAP - 4; } // the aligned start may shift the run by up to 3 entries
int spmv_rows_cap() { return ROWS_CAP; }

void launch_spmv(const SpmvPlan& p, const int* rowptr, const int* col, const double* val, const double* x, double* y, cudaStream_t st) {
  if (p.nblocks > 0) spmv_kernel<<<p.nblocks, SP_THREADS, 0, st>>>(p.blk_row, rowptr, col, val, x, y);
}

} // namespace thcm
