// gmres_cycle.cu -- one whole GMRES(m) cycle in ONE cooperative kernel, for problems small enough to live in L2.
//
// GMRESSolver::solve (src/gmressolver/GMRESSolver.H:142-213) runs, per inner iteration, a preconditioner apply, a matrix apply,
// the orthogonalisation, the Givens update and a convergence test.  The large-grid path of this library spends 4..6 kernel
// launches per iteration on that (and one device -> host round trip per convergence test), which is what bounds the small
// configurations.  The tiniest ones (2-D MOC, 648 unknowns) run inside one cluster (gmres_cluster.cu); this kernel takes the
// sizes between that and IEMIC_GMRES_CYCLE_ROWS (20 000 rows: measured 28.3 k against 18.3 k iterations/s at 17 094 rows; at
// the 262 656 rows of the 4-degree grid it loses, 9.5 k against 11.2 k, and is not selected).  The blocks of one cooperative
// launch walk through the iterations themselves, separated by grid-wide barriers:
//
//   [z_i = B^-1 v_i]  |  w = A z_i  |  h1 = V^T w (block partials)  |  w -= V h1 ; h2 = V^T w, <w,w> (partials)  |
//   w -= V h2 ; v_{i+1} = w / sqrt(<w,w> - |h2|^2) ; block 0: Hessenberg column, Givens, residual estimate, stop flag
//
// = 3 barriers per iteration (4 with the block-Jacobi preconditioner), no launch, no host.  The arithmetic is the fused CGS2 of
// gmres.cu (same projections, Pythagoras norm, givens_step); the matrix apply sums a row in eight interleaved partial sums (a
// different association than matAvec; the explicit residual that accepts a cycle is computed by the driver's bit-exact product).
// Reductions are deterministic: per-block partial sums in global memory, added by every block in the same fixed tree.
// Rows written by other blocks (the SpMV input) are read with ld.global.cg: L1 is not coherent between SMs within a launch.
// Single GPU only (no halo exchange between the iterations).
#include <cooperative_groups.h>
#include <cstdlib>
#include "kernels.h"
#include "gmres_givens.cuh"

namespace cg = cooperative_groups;

namespace thcm {
namespace {
constexpr int CT = 256;
constexpr unsigned FULL = 0xffffffffu;

__device__ __forceinline__ double warp_total(double v) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(FULL, v, o);
  return v;
}
// part[k] = sum over this block's rows [r0, r1) of V_k[r] * w[r], k = 0..nk-1 (vector k == kw is w itself): one warp per k, the
// lanes stride over the rows, no block barrier
// (partials are stored vector-major, part[k * ldp + block], so that the sum over the blocks reads consecutive words)
__device__ __forceinline__ void block_dots(const double* V, long long ld, const double* w, int nk, int kw, long long r0, long long r1, double* part, int ldp) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int k = warp; k < nk; k += CT / 32) {
    const double* vk = (k == kw) ? w : V + (size_t)k * ld;
    double acc = 0.0;
    for (long long r = r0 + lane; r < r1; r += 32) acc = fma(vk[r], w[r], acc);
    acc = warp_total(acc);
    if (lane == 0) part[(size_t)k * ldp + blockIdx.x] = acc;
  }
}
// hs[k] = sum over the blocks of part[k][.], k < nk: a warp per k, lanes stride over the blocks, fixed tree -- the same number
// in every block
__device__ __forceinline__ void sum_partials(const double* part, int ldp, int nblk, int nk, double* hs) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int k = warp; k < nk; k += CT / 32) {
    double s = 0.0;
    for (int b = lane; b < nblk; b += 32) s += __ldcg(part + (size_t)k * ldp + b);
    s = warp_total(s);
    if (lane == 0) hs[k] = s;
  }
}

__global__ void __launch_bounds__(CT) gmres_cycle_kernel(const CycleArgs a) {
  cg::grid_group grid = cg::this_grid();
  extern __shared__ double hs[];          // a.m + 2 coefficients
  __shared__ double wred[CT / 32];
  const int nblk = gridDim.x;
  // every block owns a contiguous chunk of rows for the whole cycle
  // (a multiple of 16 rows: no 128-byte line of a basis vector is shared between two blocks)
  const long long chunk = ((a.n + nblk - 1) / nblk + 15) / 16 * 16;
  const long long r0 = (long long)blockIdx.x * chunk, r1 = r0 + chunk < a.n ? r0 + chunk : a.n;
  double* part1 = a.part;
  double* part2 = a.part + (size_t)(a.m + 2) * a.ldp;
  for (int i = 0; i < a.maxcols; i++) {
    const double* vi = a.V + (size_t)i * a.ld;
    double* w = a.V + (size_t)(i + 1) * a.ld;
    const double* xin = vi;
    if (a.binv) { // z_i = blockdiag(Binv) v_i, 6x6 blocks row-major (block_jacobi_kernel of gmres.cu)
      double* zi = a.Z + (size_t)i * a.ld;
      for (long long r = r0 + threadIdx.x; r < r1; r += CT) {
        const long long c = r / 6;
        const int q = (int)(r - 6 * c);
        const double* B = a.binv + 36 * c + 6 * q;
        const double* v = vi + 6 * c;
        double s = 0.0;
#pragma unroll
        for (int k = 0; k < 6; k++) s = fma(B[k], __ldcg(v + k), s);
        zi[r] = s;
      }
      xin = zi;
      grid.sync();
    }
    // w = A xin: 8 lanes per row, lane q takes the entries q, q+8, ... of the row (coalesced), the eight sums are added in lane order
    {
      const int slot = threadIdx.x >> 3, seg = threadIdx.x & 7;
      for (long long rb = r0; rb < r1; rb += CT / 8) {
        const long long r = rb + slot;
        double s = 0.0;
        if (r < r1 && r != a.long_row) {
          const int p0 = a.rowptr[r], p1 = a.rowptr[r + 1];
          for (int p = p0 + seg; p < p1; p += 8) s = __dadd_rn(s, __dmul_rn(__ldg(a.val + p), __ldcg(xin + __ldg(a.col + p))));
        }
        double tot = s;
        for (int d = 1; d < 8; d++) { const double o = __shfl_sync(FULL, s, d, 8); tot = __dadd_rn(tot, o); }
        if (seg == 0 && r < r1 && r != a.long_row) w[r] = tot;
      }
    }
    if (a.long_row >= r0 && a.long_row < r1) { // dense row: the whole block, fixed order (spmv_longrow_kernel of spmv_tile.cu)
      const int p0 = a.rowptr[a.long_row], p1 = a.rowptr[a.long_row + 1];
      double s = 0.0;
      for (int p = p0 + threadIdx.x; p < p1; p += CT) s = fma(a.val[p], __ldcg(xin + a.col[p]), s);
      s = warp_total(s);
      if ((threadIdx.x & 31) == 0) wred[threadIdx.x >> 5] = s;
      __syncthreads();
      if (threadIdx.x == 0) {
        double t = 0.0;
        for (int q = 0; q < CT / 32; q++) t += wred[q];
        w[a.long_row] = t;
      }
    }
    __syncthreads(); // the block's rows of w are complete (each block only reads its own rows of w below)
    // first projection: block partials of <V_k, w>
    block_dots(a.V, a.ld, w, i + 1, -1, r0, r1, part1, a.ldp);
    grid.sync();
    sum_partials(part1, a.ldp, nblk, i + 1, hs);
    __syncthreads();
    if (blockIdx.x == 0) for (int k = threadIdx.x; k <= i; k += CT) a.h1[k] = hs[k];
    // w -= V h1 on this block's rows, then the second projection and <w,w> from the updated rows
    for (long long r = r0 + threadIdx.x; r < r1; r += CT) {
      double x = w[r];
      for (int k = 0; k <= i; k++) x = fma(-hs[k], a.V[(size_t)k * a.ld + r], x);
      w[r] = x;
    }
    __syncthreads();
    block_dots(a.V, a.ld, w, i + 2, i + 1, r0, r1, part2, a.ldp);
    grid.sync();
    sum_partials(part2, a.ldp, nblk, i + 2, hs);
    __syncthreads();
    // ||w - V h2||^2 = <w,w> - |h2|^2 for orthonormal V (givens_step computes the same number for the Hessenberg matrix)
    double hh = 0.0;
    for (int k = 0; k <= i; k++) hh = fma(hs[k], hs[k], hh);
    const double ww = hs[i + 1], d2 = ww - hh;
    const double nrm = (d2 > 1.0e-28 * ww) ? sqrt(d2) : 0.0;
    const double scale = (nrm > 0.0) ? nrm : 1.0;
    for (long long r = r0 + threadIdx.x; r < r1; r += CT) {
      double x = w[r];
      for (int k = 0; k <= i; k++) x = fma(-hs[k], a.V[(size_t)k * a.ld + r], x);
      w[r] = x / scale;
    }
    if (blockIdx.x == 0) {
      for (int k = threadIdx.x; k <= i + 1; k += CT) a.h2[k] = hs[k];
      __syncthreads();
      if (threadIdx.x == 0) {
        GivensArgs g = a.gv;
        g.j = i;
        givens_step(g);
        *a.out_cols = i + 1;
        if (a.check && a.gv.hist[i] < a.tol) *a.stop = 1;
        __threadfence();
      }
    }
    grid.sync();
    if (*(volatile int*)a.stop) break; // written before the barrier, read by everybody after it: all blocks agree
  }
}
} // namespace

// Largest problem the cycle kernel takes (rows); above it the launch-per-kernel path with its bandwidth-tuned kernels wins
// (measured: 262 656 rows 9.2 k against 12.1 k iterations/s, 1 536 rows 21.0 k against 19.2 k)
long long gmres_cycle_max_rows() {
  static long long lim = -1;
  if (lim < 0) { const char* e = getenv("IEMIC_GMRES_CYCLE_ROWS"); lim = e ? atoll(e) : 20000; }
  return lim;
}
int gmres_cycle_partials(int m) { // doubles of workspace
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  return 2 * ((2 * sms + 31) / 32 * 32) * (m + 2);
}
// returns 0 or a cudaError_t; a.part must hold gmres_cycle_partials(a.m) doubles
int launch_gmres_cycle(CycleArgs a, cudaStream_t st) {
  int dev = 0, sms = 148, coop = 0, per_sm = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
  if (!coop) return (int)cudaErrorNotSupported;
  const size_t smem = (size_t)(a.m + 2) * sizeof(double);
  cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, gmres_cycle_kernel, CT, smem);
  if (e != cudaSuccess) return (int)e;
  if (per_sm < 1) return (int)cudaErrorLaunchOutOfResources;
  long long want = (a.n + CT - 1) / CT;            // one row per thread where the grid allows it
  const long long cap = (long long)sms * (per_sm < 2 ? per_sm : 2);
  if (want > cap) want = cap;
  if (want < 1) want = 1;
  a.ldp = (int)((want + 31) / 32 * 32); // partials: (m + 2) rows of ldp blocks, twice
  void* args[] = {(void*)&a};
  return (int)cudaLaunchCooperativeKernel((const void*)gmres_cycle_kernel, dim3((unsigned)want), dim3(CT), args, smem, st);
}

} // namespace thcm
