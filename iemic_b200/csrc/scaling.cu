// scaling.cu -- work on the 6x6 diagonal blocks of the device-resident Jacobian:
//
//   average_block_kernel   m_scaling::average_block (src/ocean/scaling.F90:29-62): sum of the diagonal blocks of all OCEAN
//                          cells and their count; THCM::RecomputeScaling (src/ocean/THCM.C:1694-1750) turns the average
//                          into six row and six column factors (host, thcm::scal in thcm_host.cpp)
//   scaling_fill_kernel    m_scaling::compute (scaling.F90:69-99) + the reciprocal / T-S averaging of THCM.C:1726-1745
//   block_inverse_kernel   inverted 6x6 point blocks = the fixed point-block-Jacobi preconditioner of gmres_b200_solve
//   left_scale_kernel      Epetra_CrsMatrix::LeftScale as Ocean::scaleProblem uses it (src/ocean/Ocean.C:1207, 1238)
//   multiply / reciprocal  Epetra_MultiVector::Multiply / Reciprocal on the vectors (Ocean.C:1195, 1213, 1241)
//
// All of it streams: one pass over 6 (of 104) values per cell plus 4 bytes of position data per row.
#include "kernels.h"

namespace thcm {
namespace {
constexpr int DT = 192;            // threads per block: 32 cells x 6 rows, thread t always sees unknown t % 6
constexpr int DCELLS = DT / 6;

// centre-cell entries of one row; the rows the C++ glue overwrote come from the side buffer of the last assembly
__device__ __forceinline__ void load_diag_row(const DiagArgs& a, long long row, double (&d)[6]) {
  const double* side = nullptr;
  if (row == a.sp.rowintcon) side = a.sp.side;
  else if (row == a.sp.rowPfix1) side = a.sp.side + 6;
  else if (row == a.sp.rowPfix2) side = a.sp.side + 12;
  if (side && !a.stored_rows) {
#pragma unroll
    for (int b = 0; b < 6; b++) d[b] = side[b];
    return;
  }
  const unsigned di = __ldg(a.dinfo + row);
  const unsigned mask = di >> 26;
  const double* p = a.val + (size_t)__ldg(a.rowptr + row) + (di & 0x3ffffffu);
#pragma unroll
  for (int b = 0; b < 6; b++) {
    d[b] = 0.0;
    if ((mask >> b) & 1u) d[b] = __ldg(p++);
  }
}

// partial[blk*37 + ii*6 + jj] = sum over the block's OCEAN cells of J(6c+ii, 6c+jj); partial[blk*37+36] = their number.
// Fixed summation order (thread-strided rows, then threads of equal unknown in order, then blocks in order).
__global__ void __launch_bounds__(DT) average_block_kernel(const DiagArgs a, double* __restrict__ partial, unsigned int* ticket,
                                                            double* __restrict__ out) {
  __shared__ double sm[DT * 7];
  __shared__ bool last;
  double acc[6] = {0, 0, 0, 0, 0, 0};
  double cnt = 0.0;
  const long long nrows = 6 * a.ncell, stride = (long long)gridDim.x * DT;
  for (long long row = (long long)blockIdx.x * DT + threadIdx.x; row < nrows; row += stride) {
    const uint64_t g = __ldg(a.sig + row / 6);
    if (sbit(g, SIG_NOTOCEAN)) continue;
    double d[6];
    load_diag_row(a, row, d);
#pragma unroll
    for (int b = 0; b < 6; b++) acc[b] += d[b];
    cnt += 1.0;
  }
#pragma unroll
  for (int b = 0; b < 6; b++) sm[threadIdx.x * 7 + b] = acc[b];
  sm[threadIdx.x * 7 + 6] = cnt;
  __syncthreads();
  if (threadIdx.x < 37) {
    const int e = threadIdx.x;
    const int ri = e < 36 ? e / 6 : 3 /* PP rows count the cells (scaling.F90:48) */, cj = e < 36 ? e % 6 : 6;
    double s = 0.0;
    for (int t = ri; t < DT; t += 6) s += sm[t * 7 + cj];
    partial[(size_t)blockIdx.x * 37 + e] = s;
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) last = atomicAdd(ticket, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!last) return;
  __threadfence();
  if (threadIdx.x < 37) {
    double s = 0.0;
    for (unsigned b = 0; b < gridDim.x; b++) s += partial[(size_t)b * 37 + threadIdx.x];
    out[threadIdx.x] = s;
  }
  if (threadIdx.x == 0) *ticket = 0u;
}

__global__ void __launch_bounds__(256) scaling_fill_kernel(const uint64_t* __restrict__ sig, long long nrows, ScalTab t,
                                                            double* __restrict__ rowscal, double* __restrict__ colscal) {
  for (long long r = (long long)blockIdx.x * 256 + threadIdx.x; r < nrows; r += (long long)gridDim.x * 256) {
    const bool ocean = !sbit(__ldg(sig + r / 6), SIG_NOTOCEAN);
    const int ii = (int)(r % 6);
    if (rowscal) rowscal[r] = ocean ? t.r[ii] : 1.0;
    if (colscal) colscal[r] = ocean ? t.c[ii] : 1.0;
  }
}

// In-place Gauss-Jordan inverse with partial pivoting of a 6x6 block stored row-major at M (stride 1 between entries).
// Returns false if a pivot column is exactly zero (structurally singular block).
__device__ __forceinline__ bool invert6(double* M) {
  int piv[6];
#pragma unroll 1
  for (int k = 0; k < 6; k++) {
    int p = k;
    double best = fabs(M[k * 6 + k]);
    for (int i = k + 1; i < 6; i++) {
      const double v = fabs(M[i * 6 + k]);
      if (v > best) { best = v; p = i; }
    }
    piv[k] = p;
    if (best == 0.0) return false;
    if (p != k)
      for (int c = 0; c < 6; c++) { const double t = M[k * 6 + c]; M[k * 6 + c] = M[p * 6 + c]; M[p * 6 + c] = t; }
    const double d = 1.0 / M[k * 6 + k];
    M[k * 6 + k] = 1.0;
    for (int c = 0; c < 6; c++) M[k * 6 + c] *= d;
    for (int i = 0; i < 6; i++) {
      if (i == k) continue;
      const double f = M[i * 6 + k];
      M[i * 6 + k] = 0.0;
      for (int c = 0; c < 6; c++) M[i * 6 + c] -= f * M[k * 6 + c];
    }
  }
#pragma unroll 1
  for (int k = 5; k >= 0; k--) {
    const int p = piv[k];
    if (p != k)
      for (int r = 0; r < 6; r++) { const double t = M[r * 6 + k]; M[r * 6 + k] = M[r * 6 + p]; M[r * 6 + p] = t; }
  }
  return true;
}

// One block = 32 cells.  Thread t loads the centre entries of row t of the group (coalesced over the CSR runs), one
// thread per cell inverts its block in shared memory (stride 37: conflict-free), all threads write the 36-double
// inverses back as one contiguous run.  A singular block becomes the identity and is counted.
__global__ void __launch_bounds__(DT) block_inverse_kernel(const DiagArgs a, double* __restrict__ binv, int* __restrict__ nsingular) {
  __shared__ double sm[DCELLS * 37];
  const long long cell0 = (long long)blockIdx.x * DCELLS;
  const long long row = cell0 * 6 + threadIdx.x;
  if (row < 6 * a.ncell) {
    double d[6];
    load_diag_row(a, row, d);
    double* dst = sm + (threadIdx.x / 6) * 37 + (threadIdx.x % 6) * 6;
#pragma unroll
    for (int b = 0; b < 6; b++) dst[b] = d[b];
  }
  __syncthreads();
  if (threadIdx.x < DCELLS && cell0 + threadIdx.x < a.ncell) {
    double* M = sm + threadIdx.x * 37;
    if (!invert6(M)) {
      for (int e = 0; e < 36; e++) M[e] = (e % 7 == 0) ? 1.0 : 0.0;
      atomicAdd(nsingular, 1);
    }
  }
  __syncthreads();
  const long long ncl = min((long long)DCELLS, a.ncell - cell0);
  for (int t = threadIdx.x; t < ncl * 36; t += DT) __stcs(binv + cell0 * 36 + t, sm[(t / 36) * 37 + t % 36]);
}

// val[p] *= s[row(p)].  One block per run of LS_ROWS rows: the row boundaries sit in shared memory and every thread finds
// the row of its non-zero by bisection; loads and stores of the values are contiguous.
constexpr int LS_ROWS = 192;
__global__ void __launch_bounds__(256) left_scale_kernel(const int* __restrict__ rowptr, double* __restrict__ val,
                                                          const double* __restrict__ s, long long nrows) {
  __shared__ int rp[LS_ROWS + 1];
  __shared__ double sc[LS_ROWS];
  const long long r0 = (long long)blockIdx.x * LS_ROWS;
  const int nr = (int)min((long long)LS_ROWS, nrows - r0);
  for (int t = threadIdx.x; t <= nr; t += 256) rp[t] = __ldg(rowptr + r0 + t);
  for (int t = threadIdx.x; t < nr; t += 256) sc[t] = __ldg(s + r0 + t);
  __syncthreads();
  for (long long p = (long long)rp[0] + threadIdx.x; p < rp[nr]; p += 256) {
    int lo = 0, hi = nr; // rp[lo] <= p < rp[hi]
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (rp[mid] <= p) lo = mid; else hi = mid;
    }
    val[p] = val[p] * sc[lo];
  }
}

// this = a * A .* B + b * this  (Epetra_MultiVector::Multiply); b == 0 never reads `y`
__global__ void __launch_bounds__(256) multiply_kernel(double a, const double* __restrict__ A, const double* __restrict__ B, double b,
                                                        double* y, long long n) {
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long long)gridDim.x * 256) {
    const double ab = a * (A[i] * B[i]);
    y[i] = (b == 0.0) ? ab : ab + b * y[i];
  }
}
__global__ void __launch_bounds__(256) reciprocal_kernel(const double* __restrict__ A, double* y, long long n) {
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long long)gridDim.x * 256) y[i] = 1.0 / A[i];
}

int stream_grid(long long n, int threads) {
  const long long need = (n + threads - 1) / threads;
  const long long cap = (long long)red_blocks(); // one full wave of 256-thread blocks
  return (int)(need < 1 ? 1 : (need < cap ? need : cap));
}
} // namespace

int average_block_partials() { return 37 * red_blocks(); }
void launch_average_block(const DiagArgs& a, double* partial, unsigned int* ticket, double* out37, cudaStream_t st) {
  average_block_kernel<<<stream_grid(6 * a.ncell, DT), DT, 0, st>>>(a, partial, ticket, out37);
}
void launch_scaling_fill(const uint64_t* sig, long long nrows, const ScalTab& t, double* rowscal, double* colscal, cudaStream_t st) {
  scaling_fill_kernel<<<stream_grid(nrows, 256), 256, 0, st>>>(sig, nrows, t, rowscal, colscal);
}
void launch_block_inverse(const DiagArgs& a, double* binv, int* nsingular, cudaStream_t st) {
  const long long nb = (a.ncell + DCELLS - 1) / DCELLS;
  if (nb > 0) block_inverse_kernel<<<(unsigned)nb, DT, 0, st>>>(a, binv, nsingular);
}
void launch_left_scale(const int* rowptr, double* val, const double* s, long long nrows, cudaStream_t st) {
  const long long nb = (nrows + LS_ROWS - 1) / LS_ROWS;
  if (nb > 0) left_scale_kernel<<<(unsigned)nb, 256, 0, st>>>(rowptr, val, s, nrows);
}
void launch_multiply(double a, const double* A, const double* B, double b, double* y, long long n, cudaStream_t st) {
  multiply_kernel<<<stream_grid(n, 256), 256, 0, st>>>(a, A, B, b, y, n);
}
void launch_reciprocal(const double* A, double* y, long long n, cudaStream_t st) {
  reciprocal_kernel<<<stream_grid(n, 256), 256, 0, st>>>(A, y, n);
}

} // namespace thcm
