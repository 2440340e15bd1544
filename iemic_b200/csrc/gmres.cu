// gmres.cu -- vector kernels of the Krylov solver for sm_100a.
//
// Replaces the vector concept GMRESSolver relies on (dot / norm / update / scale, src/gmressolver/GMRESSolverDecl.H:7-16)
// and the modified Gram-Schmidt loop of GMRESSolver::solve (src/gmressolver/GMRESSolver.H:177-187) by fused kernels:
//
//   multi_dot   h[k] = <V_k, w> for all k in ONE pass over the basis (w is re-read once per 8 basis vectors),
//   multi_axpy  w -= sum_k h[k] V_k in one pass, with ||w||^2 of the result accumulated in the epilogue,
//   axpy_dot    the MGS step "w -= h_{k-1} V_{k-1}; h_k = <w, V_k>" in one pass (reproduces the reference's ordering),
//   givens      Hessenberg column, plane rotations and residual estimate on the device, so a whole Arnoldi
//               step needs no host synchronisation.
//
// Reductions are two-stage and deterministic: warp shuffles -> per-block partials in global memory -> the last
// block to arrive (atomic ticket) adds the partials in block order.  All loads are 16-byte vectorised; basis
// vectors are streamed with evict-first hints, w stays cacheable.
// HBM traffic per call: multi_dot (nv + ceil(nv/8)) * 8n, multi_axpy (nv + 2) * 8n, axpy_dot 4 * 8n.
#include <algorithm>
#include <cstdlib>
#include "kernels.h"
#include "gmres_givens.cuh"

namespace thcm {
namespace {
constexpr int RT = 256;          // threads per block
constexpr int KT = 8;            // basis vectors per register tile in multi_dot
constexpr int MAXV = 512;        // largest basis handled by one call
constexpr unsigned FULL = 0xffffffffu;

__device__ __forceinline__ double warp_sum(double s) {
  for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(FULL, s, o);
  return s;
}
// Block-level sum of NV values per thread -> partial[blockIdx.x * stride + k]; returns true in the last block
// (all threads) once every block has published its partials.
template <int NV>
__device__ __forceinline__ void block_publish(const double (&acc)[NV], int nvalid, double* partial, int stride, int k0) {
  __shared__ double red[NV][RT / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < NV; k++) {
    double s = warp_sum(acc[k]);
    if (lane == 0) red[k][warp] = s;
  }
  __syncthreads();
  if (threadIdx.x < NV && threadIdx.x < nvalid) {
    double s = 0.0;
#pragma unroll
    for (int w = 0; w < RT / 32; w++) s += red[threadIdx.x][w];
    partial[(size_t)blockIdx.x * stride + k0 + threadIdx.x] = s;
  }
  __syncthreads();
}
__device__ __forceinline__ bool last_block(unsigned int* ticket) {
  __shared__ bool is_last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned int t = atomicAdd(ticket, 1u);
    is_last = (t == gridDim.x - 1);
    if (is_last) *ticket = 0u; // ready for the next launch on this stream
  }
  __syncthreads();
  if (is_last) __threadfence();
  return is_last;
}
// out[k] = sum over blocks of partial[b*stride + k], fixed order; one warp per k
__device__ __forceinline__ void final_sum(const double* partial, int stride, int nv, double* out) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int k = warp; k < nv; k += (int)(blockDim.x >> 5)) {
    double s = 0.0;
    for (int b = lane; b < (int)gridDim.x; b += 32) s += __ldcg(partial + (size_t)b * stride + k);
    s = warp_sum(s);
    if (lane == 0) out[k] = s;
  }
}

// With gv.H set, the finishing block also performs the Arnoldi bookkeeping of this step (fused CGS2: the second
// multi-dot is the last reduction of the step), which saves the launch of the single-thread givens kernel.
__global__ void __launch_bounds__(RT) multi_dot_kernel(const double* __restrict__ V, long long ldv, int nv, const double* __restrict__ w,
                                                       long long n, double* out, double* partial, unsigned int* ticket, const PeerRed pr,
                                                       const GivensArgs gv) {
  const long long n2 = n >> 1; // double2 elements (n is even: 6 unknowns per cell)
  const double2* w2 = reinterpret_cast<const double2*>(w);
  for (int k0 = 0; k0 < nv; k0 += KT) {
    double acc[KT];
#pragma unroll
    for (int k = 0; k < KT; k++) acc[k] = 0.0;
    const int nk = min(KT, nv - k0);
    for (long long i = (long long)blockIdx.x * RT + threadIdx.x; i < n2; i += (long long)gridDim.x * RT) {
      const double2 wv = w2[i];
      double2 v[KT];
#pragma unroll
      for (int k = 0; k < KT; k++)
        if (k < nk) v[k] = __ldcs(reinterpret_cast<const double2*>(V + (size_t)(k0 + k) * ldv) + i);
#pragma unroll
      for (int k = 0; k < KT; k++)
        if (k < nk) acc[k] = fma(v[k].y, wv.y, fma(v[k].x, wv.x, acc[k]));
    }
    block_publish<KT>(acc, nk, partial, MAXV, k0);
  }
  if (last_block(ticket)) {
    final_sum(partial, MAXV, nv, out);
    if (pr.nranks) peer_allreduce_block(pr, out, nv);
    if (gv.H) {
      __syncthreads(); // out[] complete and visible to thread 0
      if (threadIdx.x == 0) givens_step(gv);
    }
  }
}

__global__ void __launch_bounds__(RT) multi_axpy_kernel(const double* __restrict__ V, long long ldv, int nv, const double* __restrict__ h,
                                                        double* __restrict__ w, long long n, double* nrm2_out, double* partial,
                                                        unsigned int* ticket, const double* inv_scale, const PeerRed pr) {
  __shared__ double hs[MAXV];
  for (int k = threadIdx.x; k < nv; k += RT) hs[k] = h[k];
  __syncthreads();
  const double sc = inv_scale ? 1.0 / *inv_scale : 1.0; // fused normalisation v_{j+1} = w / H(j+1,j)
  const long long n2 = n >> 1;
  double2* w2 = reinterpret_cast<double2*>(w);
  double acc[1] = {0.0};
  for (long long i = (long long)blockIdx.x * RT + threadIdx.x; i < n2; i += (long long)gridDim.x * RT) {
    double2 wv = w2[i];
    for (int k0 = 0; k0 < nv; k0 += KT) {
      double2 v[KT];
#pragma unroll
      for (int k = 0; k < KT; k++)
        if (k0 + k < nv) v[k] = __ldcs(reinterpret_cast<const double2*>(V + (size_t)(k0 + k) * ldv) + i);
#pragma unroll
      for (int k = 0; k < KT; k++)
        if (k0 + k < nv) { wv.x = fma(-hs[k0 + k], v[k].x, wv.x); wv.y = fma(-hs[k0 + k], v[k].y, wv.y); }
    }
    if (inv_scale) { wv.x = sc * wv.x; wv.y = sc * wv.y; }
    w2[i] = wv;
    acc[0] = fma(wv.y, wv.y, fma(wv.x, wv.x, acc[0]));
  }
  if (nrm2_out) {
    block_publish<1>(acc, 1, partial, MAXV, 0);
    if (last_block(ticket)) {
      final_sum(partial, MAXV, 1, nrm2_out);
      if (pr.nranks) peer_allreduce_block(pr, nrm2_out, 1);
    }
  }
}

// Fused second and third pass of a CGS2 step: w <- w - sum_k h[k] V_k and, from the SAME pass over the basis,
// out[k] = <V_k, w_new> (k < nv) and out[nv] = <w_new, w_new>.  A block keeps its 256-row tile of the basis in shared memory
// ((nv+1) x 256 doubles, row nv = the updated w) between the update and the dots, so the basis is read from HBM once
// instead of twice: (nv + 2) * 8n bytes for both passes.  Warp q owns the dots k = q, q+8, ...; fixed summation order.
constexpr int FUSE_MAXNV = 55;                 // largest basis the fused kernel takes ((nv+1) * FT * 8 B of shared memory per buffer)
// FT = rows per tile = threads per block, NB = shared-memory tiles (2: a warp may load the next tile while others still
// reduce the current one, one barrier per tile instead of two), MINB = resident blocks per SM the registers are capped for
// FK = loads in flight per thread in the update phase
template <int FT, int NB, int MINB, int FK>
__global__ void __launch_bounds__(FT, MINB) multi_axpy_dot_kernel(const double* __restrict__ V, long long ldv, int nv, const double* __restrict__ h,
                                                                  double* __restrict__ w, long long n, double* out, double* partial,
                                                                  unsigned int* ticket, const PeerRed pr, const GivensArgs gv) {
  constexpr int NW = FT / 32;
  constexpr int KPW = (FUSE_MAXNV + 1 + NW - 1) / NW;
  extern __shared__ double tile_all[];
  __shared__ double hs[FUSE_MAXNV + 1];
  for (int k = threadIdx.x; k < nv; k += FT) hs[k] = h[k];
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  double acc[KPW];
#pragma unroll
  for (int q = 0; q < KPW; q++) acc[q] = 0.0;
  int buf = 0;
  for (long long t0 = (long long)blockIdx.x * FT; t0 < n; t0 += (long long)gridDim.x * FT) {
    double* tile = tile_all + (size_t)buf * (nv + 1) * FT;
    double* wrow = tile + (size_t)nv * FT;
    const long long i = t0 + threadIdx.x;
    const bool in = i < n;
    const long long ic = in ? i : n - 1;      // rows past the end read a valid address and are weighted with w = 0 below
    double wv = w[ic];
    const double* vp = V + ic;
    double* tp = tile + threadIdx.x;
    int k0 = 0;
    for (; k0 + FK <= nv; k0 += FK) {         // full chunks: no predicates, FK loads in flight
      double v[FK];
#pragma unroll
      for (int k = 0; k < FK; k++) v[k] = __ldcs(vp + (size_t)k * ldv);
#pragma unroll
      for (int k = 0; k < FK; k++) { tp[k * FT] = v[k]; wv = fma(-hs[k0 + k], v[k], wv); }
      vp += (size_t)FK * ldv; tp += FK * FT;
    }
    if (k0 < nv) {
      const int rem = nv - k0;
      double v[FK - 1];
#pragma unroll
      for (int k = 0; k < FK - 1; k++) if (k < rem) v[k] = __ldcs(vp + (size_t)k * ldv);
#pragma unroll
      for (int k = 0; k < FK - 1; k++) if (k < rem) { tp[k * FT] = v[k]; wv = fma(-hs[k0 + k], v[k], wv); }
    }
    if (in) w[i] = wv; else wv = 0.0;
    wrow[threadIdx.x] = wv;
    __syncthreads();
    double wr[NW];
#pragma unroll
    for (int q = 0; q < NW; q++) wr[q] = wrow[lane + 32 * q];
#pragma unroll
    for (int q = 0; q < KPW; q++) {
      const int k = warp + q * NW;
      if (k > nv) break;
      const double* row = tile + (size_t)k * FT + lane;
      double sacc = acc[q];
#pragma unroll
      for (int r = 0; r < NW; r++) sacc = fma(row[32 * r], wr[r], sacc);
      acc[q] = sacc;
    }
    if (NB == 1) __syncthreads(); else buf ^= 1;
  }
  // block partials (RT/32 slots are not needed here: one value per dot and block)
#pragma unroll
  for (int q = 0; q < KPW; q++) {
    const int k = warp + q * NW;
    const double sacc = warp_sum(acc[q]);
    if (k <= nv && lane == 0) partial[(size_t)blockIdx.x * MAXV + k] = sacc;
  }
  if (last_block(ticket)) {
    final_sum(partial, MAXV, nv + 1, out);
    if (pr.nranks) peer_allreduce_block(pr, out, nv + 1);
    if (gv.H) {
      __syncthreads();
      if (threadIdx.x == 0) givens_step(gv);
    }
  }
}

__global__ void __launch_bounds__(RT) axpy_dot_kernel(const double* __restrict__ hprev, const double* __restrict__ vprev,
                                                      const double* __restrict__ vnext, double* __restrict__ w, long long n, double* out,
                                                      double* partial, unsigned int* ticket, const PeerRed pr) {
  const long long n2 = n >> 1;
  double2* w2 = reinterpret_cast<double2*>(w);
  const double hp = hprev ? *hprev : 0.0;
  double acc[1] = {0.0};
  for (long long i = (long long)blockIdx.x * RT + threadIdx.x; i < n2; i += (long long)gridDim.x * RT) {
    double2 wv = w2[i];
    if (hprev) {
      const double2 vp = __ldcs(reinterpret_cast<const double2*>(vprev) + i);
      wv.x = fma(-hp, vp.x, wv.x); wv.y = fma(-hp, vp.y, wv.y);
      w2[i] = wv;
    }
    const double2 vn = vnext ? __ldcs(reinterpret_cast<const double2*>(vnext) + i) : wv;
    acc[0] = fma(wv.y, vn.y, fma(wv.x, vn.x, acc[0]));
  }
  block_publish<1>(acc, 1, partial, MAXV, 0);
  if (last_block(ticket)) {
    final_sum(partial, MAXV, 1, out);
    if (pr.nranks) peer_allreduce_block(pr, out, 1);
  }
}

__global__ void __launch_bounds__(RT) dot_kernel(const double* __restrict__ a, const double* __restrict__ b, long long n, double* out,
                                                 double* partial, unsigned int* ticket, const PeerRed pr) {
  const long long n2 = n >> 1;
  double acc[1] = {0.0};
  for (long long i = (long long)blockIdx.x * RT + threadIdx.x; i < n2; i += (long long)gridDim.x * RT) {
    const double2 x = reinterpret_cast<const double2*>(a)[i], y = reinterpret_cast<const double2*>(b)[i];
    acc[0] = fma(x.y, y.y, fma(x.x, y.x, acc[0]));
  }
  block_publish<1>(acc, 1, partial, MAXV, 0);
  if (last_block(ticket)) {
    final_sum(partial, MAXV, 1, out);
    if (pr.nranks) peer_allreduce_block(pr, out, 1);
  }
}

__global__ void __launch_bounds__(RT) scale_copy_kernel(const double* __restrict__ x, double* __restrict__ y, long long n, const double* s, int inv, int sq) {
  double a = *s;
  if (sq) a = sqrt(a);
  if (inv) a = 1.0 / a;
  const long long n2 = n >> 1;
  for (long long i = (long long)blockIdx.x * RT + threadIdx.x; i < n2; i += (long long)gridDim.x * RT) {
    double2 v = reinterpret_cast<const double2*>(x)[i];
    v.x = a * v.x; v.y = a * v.y;
    reinterpret_cast<double2*>(y)[i] = v;
  }
}
__global__ void __launch_bounds__(RT) update_kernel(double a, const double* __restrict__ A, double b, double* __restrict__ y, long long n) {
  const long long n2 = n >> 1;
  for (long long i = (long long)blockIdx.x * RT + threadIdx.x; i < n2; i += (long long)gridDim.x * RT) {
    const double2 av = reinterpret_cast<const double2*>(A)[i];
    double2 yv = reinterpret_cast<double2*>(y)[i];
    if (b == 0.0) { yv.x = a * av.x; yv.y = a * av.y; } // Epetra semantics: b==0 ignores (possibly NaN) old content
    else { yv.x = a * av.x + b * yv.x; yv.y = a * av.y + b * yv.y; }
    reinterpret_cast<double2*>(y)[i] = yv;
  }
}
__global__ void __launch_bounds__(RT) fill_kernel(double a, double* __restrict__ y, long long n) {
  for (long long i = (long long)blockIdx.x * RT + threadIdx.x; i < n; i += (long long)gridDim.x * RT) y[i] = a;
}
__global__ void __launch_bounds__(RT) multi_axpy_plus_kernel(const double* __restrict__ Z, long long ldz, int nv, const double* __restrict__ yk,
                                                             double* __restrict__ x, long long n) {
  __shared__ double hs[MAXV];
  for (int k = threadIdx.x; k < nv; k += RT) hs[k] = yk[k];
  __syncthreads();
  const long long n2 = n >> 1;
  for (long long i = (long long)blockIdx.x * RT + threadIdx.x; i < n2; i += (long long)gridDim.x * RT) {
    double2 xv = reinterpret_cast<double2*>(x)[i];
    for (int k = 0; k < nv; k++) {
      const double2 v = __ldcs(reinterpret_cast<const double2*>(Z + (size_t)k * ldz) + i);
      xv.x = fma(hs[k], v.x, xv.x); xv.y = fma(hs[k], v.y, xv.y);
    }
    reinterpret_cast<double2*>(x)[i] = xv;
  }
}
__global__ void __launch_bounds__(RT) block_jacobi_kernel(const double* __restrict__ binv, const double* __restrict__ v, double* __restrict__ z, long long nb) {
  // one thread per output row: 6 consecutive threads share a block; row-major 6x6 blocks
  for (long long r = (long long)blockIdx.x * RT + threadIdx.x; r < 6 * nb; r += (long long)gridDim.x * RT) {
    const long long b = r / 6;
    const double* B = binv + r * 6;
    const double* vb = v + b * 6;
    double s = 0.0;
#pragma unroll
    for (int c = 0; c < 6; c++) s += B[c] * vb[c];
    z[r] = s;
  }
}
__global__ void __launch_bounds__(RT) residual_kernel(const double* __restrict__ b, double* __restrict__ r, long long n) {
  for (long long i = (long long)blockIdx.x * RT + threadIdx.x; i < n; i += (long long)gridDim.x * RT) r[i] = 1.0 * b[i] + -1.0 * r[i];
}

__global__ void givens_kernel(const GivensArgs g) {
  if (threadIdx.x == 0) givens_step(g);
}
__global__ void backsolve_kernel(const double* H, int ldh, const double* s, int k, double* y) {
  if (threadIdx.x != 0) return;
  for (int i = 0; i <= k; i++) y[i] = s[i];
  for (int i = k; i >= 0; i--) { // GMRESSolver.H:400-410
    y[i] /= H[(size_t)i * ldh + i];
    for (int jj = i - 1; jj >= 0; jj--) y[jj] -= H[(size_t)i * ldh + jj] * y[i];
  }
}
__global__ void cycle_init_kernel(const double* nrm2, double* s, int m, double* beta_out) {
  if (threadIdx.x != 0) return;
  double beta = sqrt(*nrm2);
  s[0] = beta;
  for (int i = 1; i <= m; i++) s[i] = 0.0;
  *beta_out = beta;
}
__global__ void sqrt_store_kernel(const double* nrm2, double* out) {
  if (threadIdx.x == 0) *out = sqrt(*nrm2);
}

// max |a_i| over the owned part; one block (used for log lines and divergence tests, never inside the solver loop)
__global__ void __launch_bounds__(1024, 1) absmax_kernel(const double* __restrict__ a, long long n, double* out) {
  __shared__ double red[32];
  double m = 0.0;
  for (long long i = threadIdx.x; i < n; i += blockDim.x) m = fmax(m, fabs(a[i]));
  for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_down_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
  __syncthreads();
  if (threadIdx.x < 32) {
    m = red[threadIdx.x];
    for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_down_sync(0xffffffffu, m, o));
    if (threadIdx.x == 0) *out = m;
  }
}

// ---- dot product without any allocation (vec_b200_dot_noalloc; no counterpart in the reference) ----
// Block partials and the arrival ticket live in a file-scope device array that every GPU loading the library holds until
// exit (NOALLOC_SLOTS * 8 bytes = 16.6 KB).  Every call shares it: calls that overlap on one GPU race.
constexpr int NOALLOC_MAXBLK = 2048;               // block partials [0, NOALLOC_MAXBLK)
constexpr int NOALLOC_SLOTS = NOALLOC_MAXBLK + 32;  // + one 256-byte aligned line for the ticket
static __device__ __align__(256) double g_noalloc_scratch[NOALLOC_SLOTS];

__global__ void __launch_bounds__(RT) dot_noalloc_kernel(const double* __restrict__ a, const double* __restrict__ b, long long n, double* out,
                                                         double* scratch) {
  double* partial = scratch;
  unsigned int* ticket = reinterpret_cast<unsigned int*>(scratch + NOALLOC_MAXBLK);
  const bool vec = (((size_t)a | (size_t)b) & 15) == 0;
  const long long n2 = vec ? (n >> 1) : 0;
  double acc[1] = {0.0};
  for (long long i = (long long)blockIdx.x * RT + threadIdx.x; i < n2; i += (long long)gridDim.x * RT) {
    const double2 x = reinterpret_cast<const double2*>(a)[i], y = reinterpret_cast<const double2*>(b)[i];
    acc[0] = fma(x.y, y.y, fma(x.x, y.x, acc[0]));
  }
  for (long long i = 2 * n2 + (long long)blockIdx.x * RT + threadIdx.x; i < n; i += (long long)gridDim.x * RT) acc[0] = fma(a[i], b[i], acc[0]); // odd tail / unaligned vectors
  block_publish<1>(acc, 1, partial, 1, 0);
  if (last_block(ticket)) final_sum(partial, 1, 1, out);
}

int g_red_blocks[64] = {0}; // per device
int grid_for(long long n) {
  long long need = ((n >> 1) + RT - 1) / RT;
  int cap = red_blocks();
  return (int)(need < 1 ? 1 : (need < cap ? need : cap));
}
} // namespace

int red_blocks() {
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  int& cached = g_red_blocks[dev & 63];
  if (cached == 0) {
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cached = sms * 8; // 8 resident blocks of 256 threads per SM: one full wave
  }
  return cached;
}

void launch_absmax(const double* a, long long n, double* out, cudaStream_t st) { absmax_kernel<<<1, 1024, 0, st>>>(a, n, out); }
long long dot_noalloc_max_n() { return 1ll << 30; }
// Returns 0, or a cudaError_t.  In host code the array's name is only a stand-in: the device address comes from
// cudaGetSymbolAddress, asked once per device with that device current.
int launch_dot_noalloc(const double* a, const double* b, long long n, double* out, cudaStream_t st) {
  static double* dev_ptr[64] = {nullptr};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev < 0 || dev >= 64) return (int)cudaErrorInvalidDevice;
  if (!dev_ptr[dev]) {
    double* p = nullptr;
    e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
    if (e != cudaSuccess) return (int)e;
    dev_ptr[dev] = p;
  }
  double* scratch = dev_ptr[dev];
  // the array keeps the last call's contents: the ticket must start at zero (an aborted launch may have left it counting)
  e = cudaMemsetAsync(scratch + NOALLOC_MAXBLK, 0, 32 * sizeof(double), st);
  if (e != cudaSuccess) return (int)e;
  int grid = grid_for(n);
  if (grid > NOALLOC_MAXBLK) grid = NOALLOC_MAXBLK;
  dot_noalloc_kernel<<<grid, RT, 0, st>>>(a, b, n, out, scratch);
  return (int)cudaGetLastError();
}

void launch_multi_dot(const double* V, long long ldv, int nv, const double* w, long long n, double* out, RedBuf rb, cudaStream_t st, const PeerRed& pr,
                      const GivensArgs* gv) {
  GivensArgs g;
  if (gv) g = *gv; else g.H = nullptr;
  multi_dot_kernel<<<grid_for(n), RT, 0, st>>>(V, ldv, nv, w, n, out, rb.partial, rb.ticket, pr, g);
}
void launch_multi_axpy(const double* V, long long ldv, int nv, const double* h, double* w, long long n, double* nrm2_out, RedBuf rb, cudaStream_t st,
                       const double* inv_scale, const PeerRed& pr) {
  multi_axpy_kernel<<<grid_for(n), RT, 0, st>>>(V, ldv, nv, h, w, n, nrm2_out, rb.partial, rb.ticket, inv_scale, pr);
}
int multi_axpy_dot_max_nv() { return FUSE_MAXNV; }
template <int FT, int NB, int MINB, int FK = 8>
static void launch_mad(const double* V, long long ldv, int nv, const double* h, double* w, long long n, double* out, RedBuf rb, cudaStream_t st,
                       const PeerRed& pr, const GivensArgs& g) {
  const int smem = NB * (nv + 1) * FT * (int)sizeof(double);
  static unsigned long long attr_done = 0; // one bit per device: a function attribute belongs to the device's copy of the kernel
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  if (!(attr_done >> (dev & 63) & 1ull)) {
    cudaFuncSetAttribute(multi_axpy_dot_kernel<FT, NB, MINB, FK>, cudaFuncAttributeMaxDynamicSharedMemorySize, NB * (FUSE_MAXNV + 1) * FT * (int)sizeof(double));
    attr_done |= 1ull << (dev & 63);
  }
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int per_sm = std::max(1, std::min(MINB, (220 * 1024) / (smem + 1024)));
  const long long need = (n + FT - 1) / FT, cap = std::min<long long>((long long)sms * per_sm, rb.nblk_max);
  multi_axpy_dot_kernel<FT, NB, MINB, FK><<<(int)(need < cap ? need : cap), FT, smem, st>>>(V, ldv, nv, h, w, n, out, rb.partial, rb.ticket, pr, g);
}
void launch_multi_axpy_dot(const double* V, long long ldv, int nv, const double* h, double* w, long long n, double* out, RedBuf rb, cudaStream_t st,
                           const PeerRed& pr, const GivensArgs* gv) {
  GivensArgs g;
  if (gv) g = *gv; else g.H = nullptr;
  // measured on B200 at 1 degree, GMRES(30) cycle, orthogonalisation per step: tiles of 256 rows x 3 blocks per SM 0.786 ms,
  // 128 x 6: 0.800, 160 x 5: 0.809, 16 instead of 8 loads in flight: 0.871 (registers), two shared-memory tiles per block: no gain
  static const int variant = getenv("IEMIC_GMRES_FUSE_VARIANT") ? atoi(getenv("IEMIC_GMRES_FUSE_VARIANT")) : 0;
  switch (variant) {
    case 1: launch_mad<128, 1, 6, 8>(V, ldv, nv, h, w, n, out, rb, st, pr, g); break;
    case 2: launch_mad<160, 1, 5, 8>(V, ldv, nv, h, w, n, out, rb, st, pr, g); break;
    default: launch_mad<256, 1, 3, 8>(V, ldv, nv, h, w, n, out, rb, st, pr, g); break;
  }
}
void launch_axpy_dot(const double* hprev, const double* vprev, const double* vnext, double* w, long long n, double* out, RedBuf rb, cudaStream_t st, const PeerRed& pr) {
  axpy_dot_kernel<<<grid_for(n), RT, 0, st>>>(hprev, vprev, vnext, w, n, out, rb.partial, rb.ticket, pr);
}
void launch_dot(const double* a, const double* b, long long n, double* out, RedBuf rb, cudaStream_t st, const PeerRed& pr) {
  dot_kernel<<<grid_for(n), RT, 0, st>>>(a, b, n, out, rb.partial, rb.ticket, pr);
}
void launch_scale_copy(const double* x, double* y, long long n, const double* s, int inv, int sq, cudaStream_t st) {
  scale_copy_kernel<<<grid_for(n), RT, 0, st>>>(x, y, n, s, inv, sq);
}
void launch_update(double a, const double* A, double b, double* y, long long n, cudaStream_t st) {
  update_kernel<<<grid_for(n), RT, 0, st>>>(a, A, b, y, n);
}
void launch_fill(double a, double* y, long long n, cudaStream_t st) { fill_kernel<<<grid_for(n), RT, 0, st>>>(a, y, n); }
void launch_multi_axpy_plus(const double* Z, long long ldz, int nv, const double* yk, double* x, long long n, cudaStream_t st) {
  multi_axpy_plus_kernel<<<grid_for(n), RT, 0, st>>>(Z, ldz, nv, yk, x, n);
}
void launch_block_jacobi(const double* binv, const double* v, double* z, long long nb, cudaStream_t st) {
  block_jacobi_kernel<<<grid_for(6 * nb * 2), RT, 0, st>>>(binv, v, z, nb);
}
void launch_residual(const double* b, double* r, long long n, cudaStream_t st) { residual_kernel<<<grid_for(n * 2), RT, 0, st>>>(b, r, n); }
void launch_givens(const GivensArgs& g, cudaStream_t st) { givens_kernel<<<1, 32, 0, st>>>(g); }
void launch_backsolve(const double* H, int ldh, const double* s, int k, double* y, cudaStream_t st) { backsolve_kernel<<<1, 32, 0, st>>>(H, ldh, s, k, y); }
void launch_cycle_init(const double* nrm2, double* s, int m, double* beta_out, cudaStream_t st) { cycle_init_kernel<<<1, 32, 0, st>>>(nrm2, s, m, beta_out); }
void launch_sqrt_store(const double* nrm2, double* out, cudaStream_t st) { sqrt_store_kernel<<<1, 32, 0, st>>>(nrm2, out); }

} // namespace thcm
