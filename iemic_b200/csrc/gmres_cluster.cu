// gmres_cluster.cu -- one GMRES(m) cycle of a TINY problem inside ONE thread-block cluster, the Krylov basis in distributed
// shared memory.
//
// The 2-D MOC configuration of the reference (test/2dmoc, src/tests/intt_2dmoc.C: 3x6x6 cells = 648 unknowns, GMRES(500)) and
// its 8x8x4 North Atlantic box are latency problems: a whole Krylov basis is 2.6 MB, one Arnoldi step a few hundred thousand
// multiply-adds.  With one kernel per operation a step costs 150 us at depth 250 (launches, L2 round trips of every reduction);
// the seven legs of intt_2dmoc spend 22 s on 148 000 such steps.
//
// Here the rows are dealt out to the CTAs of one cluster (up to 16, 6*k rows each so that a cell's 6x6 block stays in one CTA).
// Every CTA keeps ITS rows of all m+1 basis vectors in shared memory ((m+1) x R doubles), so
//   * the projections h = V^T w are partial sums over shared memory, finished by reading the other CTAs' partials through
//     DSMEM in rank order (every CTA ends up with bit-identical coefficients),
//   * the updates w -= V h never leave the SM,
//   * the matrix apply reads its input vector from a full copy in every CTA, which the owners refresh with remote stores,
//   * barriers are barrier.cluster (hardware), 3 per step (4 with the explicit norm of plain CGS2).
// CTA 0 keeps the rotations in shared memory and applies them with the running pair in registers.
// Arithmetic: GMRESSolver.H:142-213 with CGS2 (two projection passes, H = h1 + h2) and either the explicit norm (CGS2) or the
// Pythagoras norm of gmres.cu (CGS2F).  Deterministic: every sum has a fixed order.  The matrix rows are summed in 8
// consecutive segments (a different association than matAvec's left-to-right loop; the explicit residual that accepts a cycle
// is computed by the driver with the bit-exact product).
#include <cooperative_groups.h>
#include <cstdlib>
#include "kernels.h"

namespace cg = cooperative_groups;

namespace thcm {
namespace {
constexpr int TPB = 512;
constexpr unsigned FULL = 0xffffffffu;
constexpr int SMEM_LIMIT = 227 * 1024 - 1024; // dynamic part: the kernel has 1 KB of static shared memory

__device__ __forceinline__ double warp_tree(double v) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(FULL, v, o);
  return v;
}

__global__ void __launch_bounds__(TPB, 1) gmres_cluster_kernel(const ClusterArgs a) {
  cg::cluster_group cluster = cg::this_cluster();
  const int rank = (int)cluster.block_rank(), NC = (int)cluster.num_blocks();
  extern __shared__ __align__(16) double sm[];
  const int m2 = a.m + 2;
  double* Vs = sm;                                   // (m+1) x RP : this CTA's rows of the basis
  double* xf = Vs + (size_t)(a.m + 1) * a.RP;        // npad      : full input vector of the matrix apply
  double* hp = xf + a.npad;                          // 2 x m2    : this CTA's partial projections (pass 1, pass 2)
  double* hf = hp + 2 * m2;                          // m2        : summed coefficients of the current pass
  double* h1s = hf + m2;                             // m2        : pass-1 coefficients
  double* upd = h1s + m2;                            // KG x RP   : partial updates per group of basis vectors
  double* small = upd + (size_t)a.KG * a.RP;         // 8         : [0] norm partial, [1] norm, [2] stop flag (CTA 0)
  double* gcs = small + 8;                           // CTA 0: rotations, right-hand side, Hessenberg column
  double* gsn = gcs + m2;
  double* gs = gsn + m2;
  double* ghc = gs + m2;
  const int t = threadIdx.x, lane = t & 31;
  const int r0 = rank * a.R;
  int nr = (int)(a.n - r0 < a.R ? a.n - r0 : a.R);
  if (nr < 0) nr = 0;

  for (int r = t; r < nr; r += TPB) Vs[r] = a.V[r0 + r];
  for (int k = t; k < m2; k += TPB) { gs[k] = 0.0; gcs[k] = 0.0; gsn[k] = 0.0; }
  if (t == 0) { small[0] = 0.0; small[1] = 0.0; small[2] = 0.0; }
  __syncthreads();
  if (rank == 0 && t == 0) gs[0] = a.s[0];
  cluster.sync(); // every CTA of the cluster is resident: remote shared memory may be touched from here on

  // x rows of column `colv` (the vector itself or its block-Jacobi image, GMRESSolver.H:153-172) into every CTA's full copy
  auto push_x = [&](int colv) {
    for (int r = t; r < nr; r += TPB) {
      double xv;
      if (a.binv) {
        const long long gr = r0 + r, c = gr / 6;
        const int q = (int)(gr - 6 * c);
        const double* B = a.binv + 36 * c + 6 * q;
        const double* v = Vs + (size_t)colv * a.RP + (r - q);
        xv = 0.0;
#pragma unroll
        for (int k = 0; k < 6; k++) xv = fma(B[k], v[k], xv);
        a.Z[(size_t)colv * a.ld + gr] = xv;
      } else xv = Vs[(size_t)colv * a.RP + r];
      for (int c = 0; c < NC; c++) cluster.map_shared_rank(xf, c)[r0 + r] = xv;
    }
  };
  // hp[pass][k] = <V_k, w> over this CTA's rows, k < nk (vector kw is w itself)
  auto dots = [&](int pass, const double* w, int nk, int kw) {
    for (int k = t; k < nk; k += TPB) {
      const double* vk = (k == kw) ? w : Vs + (size_t)k * a.RP;
      double acc = 0.0;
      for (int r = 0; r < nr; r++) acc = fma(vk[r], w[r], acc);
      hp[pass * m2 + k] = acc;
    }
  };
  // hf[k] = sum over the CTAs in rank order
  auto gather = [&](int pass, int nk) {
    for (int k = t; k < nk; k += TPB) {
      double s = 0.0;
      for (int c = 0; c < NC; c++) s += cluster.map_shared_rank(hp, c)[pass * m2 + k];
      hf[k] = s;
    }
  };
  // w -= sum_{k < nk} hf[k] V_k on this CTA's rows
  auto update = [&](double* w, int nk) {
    const int g = t / a.R, rr = t - g * a.R;
    if (g < a.KG && rr < nr) {
      double p = 0.0;
      for (int k = g; k < nk; k += a.KG) p = fma(hf[k], Vs[(size_t)k * a.RP + rr], p);
      upd[g * a.RP + rr] = p;
    }
    __syncthreads();
    if (t < nr) {
      double x = w[t];
      for (int q = 0; q < a.KG; q++) x -= upd[q * a.RP + t];
      w[t] = x;
    }
    __syncthreads();
  };

  push_x(0);
  cluster.sync();

  for (int i = 0; i < a.maxcols; i++) {
    double* w = Vs + (size_t)(i + 1) * a.RP;
    // ---- w = A x : 8 lanes per row, consecutive segments, combined in lane order
    {
      const int slot = t >> 3, seg = t & 7;
      for (int rb = 0; rb < nr; rb += TPB / 8) {
        const int r = rb + slot;
        double s = 0.0;
        if (r < nr) {
          const int p0 = a.rowptr[r0 + r], p1 = a.rowptr[r0 + r + 1];
          const int per = (p1 - p0 + 7) >> 3;
          int q0 = p0 + seg * per, q1 = q0 + per;
          if (q1 > p1) q1 = p1;
          for (int p = q0; p < q1; p++) s = __dadd_rn(s, __dmul_rn(__ldg(a.val + p), xf[__ldg(a.col + p)]));
        }
        double tot = s;
        for (int d = 1; d < 8; d++) { const double o = __shfl_sync(FULL, s, d, 8); tot = __dadd_rn(tot, o); }
        if (seg == 0 && r < nr) w[r] = tot;
      }
    }
    __syncthreads();
    // ---- first projection
    dots(0, w, i + 1, -1);
    cluster.sync();
    gather(0, i + 1);
    __syncthreads();
    for (int k = t; k <= i; k += TPB) h1s[k] = hf[k];
    update(w, i + 1); // (its barriers also order h1s)
    // ---- second projection (fused mode: <w,w> rides along as entry i+1)
    const int nk2 = a.fused ? i + 2 : i + 1;
    dots(1, w, nk2, a.fused ? i + 1 : -1);
    cluster.sync();
    gather(1, nk2);
    __syncthreads();
    for (int k = t; k <= i; k += TPB) ghc[k] = h1s[k] + hf[k]; // Hessenberg column before the rotations (GMRESSolver.H:182-188)
    if (a.fused && t < 32) { // ||w - V h2||^2 = <w,w> - |h2|^2 (gmres.cu), the same number in every CTA
      double hh = 0.0;
      for (int k = lane; k <= i; k += 32) hh = fma(hf[k], hf[k], hh);
      hh = warp_tree(hh);
      if (lane == 0) {
        const double ww = hf[i + 1], d2 = ww - hh;
        small[1] = (d2 > 1.0e-28 * ww) ? sqrt(d2) : 0.0;
      }
    }
    update(w, i + 1);
    if (!a.fused) { // explicit norm of the twice-projected vector
      if (t < 32) {
        double p = 0.0;
        for (int r = lane; r < nr; r += 32) p = fma(w[r], w[r], p);
        p = warp_tree(p);
        if (lane == 0) small[0] = p;
      }
      cluster.sync();
      if (t == 0) {
        double s2 = 0.0;
        for (int c = 0; c < NC; c++) s2 += cluster.map_shared_rank(small, c)[0];
        small[1] = sqrt(s2);
      }
      __syncthreads();
    }
    const double nrm = small[1];
    const double scale = (nrm > 0.0) ? nrm : 1.0; // happy breakdown: H[i+1][i] = 0, divide by 1
    for (int r = t; r < nr; r += TPB) {
      const double v = w[r] / scale;
      w[r] = v;
      a.V[(size_t)(i + 1) * a.ld + r0 + r] = v;
    }
    __syncthreads();
    // ---- CTA 0, last warp: rotations and residual estimate, while the low warps hand out the next input vector
    if (rank == 0 && t == TPB - 32) {
      double* Hc = a.H + (size_t)i * a.ldh;
      double x = ghc[0];
      for (int k = 0; k < i; k++) { // GMRESSolver.H:190-193, the running pair stays in registers
        const double y = ghc[k + 1], c = gcs[k], s = gsn[k];
        Hc[k] = c * x + s * y;
        x = -s * x + c * y;
      }
      double y = nrm, c, s;
      if (y == 0.0) { c = 1.0; s = 0.0; }                       // GMRESSolver.H:258-289
      else if (fabs(y) > fabs(x)) { const double tmp = x / y; s = 1.0 / sqrt(1.0 + tmp * tmp); c = tmp * s; }
      else { const double tmp = y / x; c = 1.0 / sqrt(1.0 + tmp * tmp); s = tmp * c; }
      gcs[i] = c; gsn[i] = s;
      a.cs[i] = c; a.sn[i] = s;
      Hc[i] = c * x + s * y;
      Hc[i + 1] = -s * x + c * y;
      const double s0 = gs[i], s1 = gs[i + 1];
      gs[i] = c * s0 + s * s1;
      gs[i + 1] = -s * s0 + c * s1;
      a.s[i] = gs[i]; a.s[i + 1] = gs[i + 1];
      const double res = fabs(gs[i + 1]) / *a.bnorm;
      a.hist[i] = res;
      *a.out_cols = i + 1;
      if (a.check && res < a.tol) { small[2] = 1.0; *a.stop = 1; }
    }
    push_x(i + 1);
    cluster.sync();
    if (cluster.map_shared_rank(small, 0)[2] != 0.0) break; // set before the barrier, the same answer in every thread
  }
  cluster.sync(); // nobody leaves while its shared memory may still be read
}

struct Plan { int ok = 0, NC = 0, R = 0, RP = 0, KG = 0, npad = 0; size_t smem = 0; };

Plan make_plan(long long n, int m, int NC) {
  Plan p;
  if (n % 6 != 0 || n < 6) return p;
  const long long ncell = n / 6;
  const long long cells = (ncell + NC - 1) / NC;
  const long long R = 6 * cells;
  if (R > 252) return p; // the update phase wants at least two groups of basis vectors per row
  p.NC = NC; p.R = (int)R; p.RP = (int)R | 1; p.KG = TPB / p.R; p.npad = NC * p.R;
  const size_t m2 = (size_t)m + 2;
  const size_t doubles = (size_t)(m + 1) * p.RP + p.npad + 2 * m2 + m2 + m2 + (size_t)p.KG * p.RP + 8 + 4 * m2;
  p.smem = doubles * sizeof(double);
  p.ok = p.smem <= (size_t)SMEM_LIMIT;
  return p;
}

// the largest cluster this device schedules for the kernel at that much shared memory (0: none); cached per device and size
int schedulable(const Plan& p) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(p.NC); cfg.blockDim = dim3(TPB); cfg.dynamicSmemBytes = p.smem;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = p.NC; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  int ncl = 0;
  if (cudaOccupancyMaxActiveClusters(&ncl, gmres_cluster_kernel, &cfg) != cudaSuccess) { cudaGetLastError(); return 0; }
  return ncl;
}

Plan plan_for(long long n, int m) {
  static Plan cache; static long long cn = -1; static int cm = -1, cdev = -1;
  int dev = 0;
  cudaGetDevice(&dev);
  if (cn == n && cm == m && cdev == dev) return cache;
  Plan best;
  int cc_major = 0;
  cudaDeviceGetAttribute(&cc_major, cudaDevAttrComputeCapabilityMajor, dev);
  if (cc_major >= 9) {
    for (int NC : {16, 8}) {
      Plan p = make_plan(n, m, NC);
      if (!p.ok) continue;
      if (cudaFuncSetAttribute(gmres_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_LIMIT) != cudaSuccess) { cudaGetLastError(); break; }
      if (NC > 8 && cudaFuncSetAttribute(gmres_cluster_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess) { cudaGetLastError(); continue; }
      if (schedulable(p) >= 1) { best = p; break; }
    }
  }
  cache = best; cn = n; cm = m; cdev = dev;
  return best;
}
} // namespace

// 1 when the problem (n rows, restart length m) fits one cluster of this device
int gmres_cluster_fits(long long n, int m) { return plan_for(n, m).ok; }

// 0, a cudaError_t, or -1 when the problem does not fit (the caller takes another path)
int launch_gmres_cluster(ClusterArgs a, cudaStream_t st) {
  const Plan p = plan_for(a.n, a.m);
  if (!p.ok) return -1;
  a.R = p.R; a.RP = p.RP; a.KG = p.KG; a.npad = p.npad;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(p.NC); cfg.blockDim = dim3(TPB); cfg.dynamicSmemBytes = p.smem; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = p.NC; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  return (int)cudaLaunchKernelEx(&cfg, gmres_cluster_kernel, a);
}

} // namespace thcm
