// thcm_kernels.cu -- fused THCM residual and Jacobian assembly for sm_100a.
//
// Replaces the Fortran pipeline An=Al; nlin_*; vmix_*; boundaries; assemble; matAvec of the reference
// (src/ocean/usrc.F90:432-586) by two kernels that never materialise the stencil tensor:
//
//   jac_kernel : one thread block per 32x4x2 tile of cells.  The ghosted fields u,v,w,T,S that usol produces
//                (usrc.F90:997-1104) are staged once per tile in shared memory (radius-1 halo); each warp then owns a
//                line of 32 cells in i and walks the six rows U,V,W,P,T,S.  A row's <=24 entries are built in
//                registers (thcm_rows.cuh), compacted into a per-warp shared-memory scratch and copied to the static
//                CSR with fully coalesced 8-byte stores: the six rows of a cell, and the cells of a line, are
//                contiguous in the CSR value array, so a warp writes one contiguous ~26 KB run.
//                HBM traffic per cell: 48 B state + 8 B signature + 28 B row offsets in, 832 B values + 48 B diag(B) out.
//   rhs_kernel : same tiling; additionally stages the raw (un-zeroed) u,v,w,p so that the assembled row can be
//                contracted with the state in the reference's storage order (matAvec), then F = A(x)x + mix - Frc.
//
// Compiled with -fmad=false: the reference's Fortran is built without FMA contraction for baseline x86-64, and
// keeping every product and sum separately rounded makes these kernels agree with the CPU oracle bit for bit
// everywhere except the tanh inside the convective-adjustment term.
#include "kernels.h"
#include "thcm_stage.cuh"

namespace thcm {

namespace {
#ifndef JAC_TK
#define JAC_TK 2
#endif
#ifndef JAC_TJ
#define JAC_TJ 4
#endif
constexpr int TI = 32, TJ = JAC_TJ, TK = JAC_TK;
constexpr int SX = TI + 2, SY = TJ + 2, SZ = TK + 2;
constexpr int TCELLS = SX * SY * SZ; // 816 staged cells per tile
constexpr int NWARP = TJ * TK;       // one warp per line of the tile
constexpr int NTHREADS = NWARP * 32;
constexpr int STRIDE = 25;           // scratch doubles per lane: odd, so 8-byte accesses are bank-conflict free
constexpr unsigned FULL = 0xffffffffu;
#ifndef JAC_MINB
#define JAC_MINB 2
#endif

struct TileFields { // accessor concept F of thcm_rows.cuh over the shared-memory tile
  const double *pu, *pv, *pw, *pt, *ps;
  int base;
  __device__ __forceinline__ double u(int di, int dj, int dk) const { return pu[base + di + dj * SX + dk * SX * SY]; }
  __device__ __forceinline__ double v(int di, int dj, int dk) const { return pv[base + di + dj * SX + dk * SX * SY]; }
  __device__ __forceinline__ double w(int di, int dj, int dk) const { return pw[base + di + dj * SX + dk * SX * SY]; }
  __device__ __forceinline__ double t(int di, int dj, int dk) const { return pt[base + di + dj * SX + dk * SX * SY]; }
  __device__ __forceinline__ double s(int di, int dj, int dk) const { return ps[base + di + dj * SX + dk * SX * SY]; }
};
struct TileRaw { // accessor concept X: raw neighbour state
  const double *ru, *rv, *rw, *rp, *rt, *rs;
  int base;
  __device__ __forceinline__ double raw(int di, int dj, int dk, int var) const {
    const int o = base + di + dj * SX + dk * SX * SY;
    switch (var) {
      case UU: return ru[o];
      case VV: return rv[o];
      case WW: return rw[o];
      case PP: return rp[o];
      case TT: return rt[o];
      default: return rs[o];
    }
  }
};

// Copy the compacted rows of 32 cells from the warp scratch (lane-major, STRIDE apart) to the CSR value array.
// Consecutive lanes write consecutive doubles wherever consecutive rows are contiguous in the CSR.
template <int NE>
__device__ __noinline__ void warp_copy_out(double* __restrict__ val, const double* scr, int base, int len, int lane) { // rare: kept out of the hot path's instruction footprint
#pragma unroll
  for (int it = 0; it < NE; it++) {
    const int t = it * 32 + lane;
    const int cell = t / NE, q = t - cell * NE;
    const int b = __shfl_sync(FULL, base, cell);
    const int ln = __shfl_sync(FULL, len, cell);
#ifdef JAC_NO_STORE
    if (q < ln && scr[cell * STRIDE + q] == 1.2345e300) __stcs(val + (size_t)b + q, 0.0);
#else
    if (q < ln) __stcs(val + (size_t)b + q, scr[cell * STRIDE + q]);
#endif
  }
}

// Fast flavour for a warp whose 32 cells all carry complete rows (no domain edge, no dense row): the cell blocks are
// CELL_NNZ apart, so addresses follow from the lane-0 offset alone -- no shuffles, no predicates.
constexpr int CELL_NNZ = RowDef<UU>::NE + RowDef<VV>::NE + RowDef<WW>::NE + RowDef<PP>::NE + RowDef<TT>::NE + RowDef<SS>::NE; // 104
template <int A> constexpr int row_offset() {
  int o = 0;
  if (A > UU) o += RowDef<UU>::NE;
  if (A > VV) o += RowDef<VV>::NE;
  if (A > WW) o += RowDef<WW>::NE;
  if (A > PP) o += RowDef<PP>::NE;
  if (A > TT) o += RowDef<TT>::NE;
  return o;
}
// Lanes 0..NE-1 move one cell's row per step: both addresses are "register + compile-time immediate", i.e. one LDS and
// one STG per cell and no index arithmetic at all.  Each store covers NE*8 contiguous bytes (56..192 B).
template <int NE, int OFF>
__device__ __forceinline__ void warp_copy_out_full(double* __restrict__ val, const double* scr, int lane) {
  if (lane < NE) {
    const double* src = scr + lane;
    double* dst = val + OFF + lane;
#pragma unroll
    for (int cell = 0; cell < 32; cell++) {
#ifdef JAC_NO_STORE
      if (src[cell * STRIDE] == 1.2345e300) __stcs(dst + cell * CELL_NNZ, 0.0);
#else
      __stcs(dst + cell * CELL_NNZ, src[cell * STRIDE]);
#endif
    }
  }
}

// Same idea with run-time strides, for lines whose cells all have the SAME (clipped) row lengths: domain edges in j and
// k, partially filled last tile in i.  `cs` = values per cell, `off`/`len` = this row's offset and length inside a cell.
__device__ __noinline__ void warp_copy_out_uniform(double* __restrict__ val, const double* scr, int lane, int nact, int cs, int off, int len) {
  if (lane < len) {
    const double* src = scr + lane;
    double* dst = val + off + lane;
#pragma unroll 4
    for (int cell = 0; cell < nact; cell++) {
      __stcs(dst, src[0]);
      dst += cs; src += STRIDE;
    }
  }
}
enum { COPY_GENERAL = 0, COPY_UNIFORM = 1, COPY_FULL = 2 };
struct CopyPlan { int mode, b0, nact, cs; };

template <int A>
__device__ __forceinline__ double do_jac_row(const AsmArgs& a, const Cell& c, const TileFields& f, bool active, int row0,
                                             const int (&rp)[7], double* scr, int lane, const CopyPlan& cp, const MixFD* fd = nullptr) {
  double b = 0.0;
  int len = 0;
  if (active) {
    b = jac_row<A>(a.S, a.T, c, f, a.sp, row0 + A - 1, scr + lane * STRIDE, fd);
    len = (row0 + A - 1 == a.sp.rowintcon) ? 0 : rp[A] - rp[A - 1];
  }
  __syncwarp();
  if (cp.mode == COPY_FULL) warp_copy_out_full<RowDef<A>::NE, row_offset<A>()>(a.val + (size_t)cp.b0, scr, lane);
  else if (cp.mode == COPY_UNIFORM) {
    const int off = __shfl_sync(FULL, rp[A - 1] - rp[0], 0), ln = __shfl_sync(FULL, rp[A] - rp[A - 1], 0);
    warp_copy_out_uniform(a.val + (size_t)cp.b0, scr, lane, cp.nact, cp.cs, off, ln);
  } else warp_copy_out<RowDef<A>::NE>(a.val, scr, rp[A - 1], len, lane);
  __syncwarp();
  return b;
}

// Fill the shared-memory tile.  (Issuing all of a thread's loads before consuming them -- stage_plan/stage_apply --
// was measured: it costs registers in the row code that follows and did not pay.)
template <bool RHS>
__device__ __forceinline__ void stage_tile(const AsmArgs& a, double* sm, int gi0, int gj0, int gk0) {
  const Scalars& S = a.S;
  const LdVec ld{a.x};
  for (int t = threadIdx.x; t < TCELLS; t += NTHREADS) {
    const int ti = t % SX, tj = (t / SX) % SY, tk = t / (SX * SY);
    const int fi = gi0 + ti - 1, fj = gj0 + tj - 1, fk = gk0 + tk - 1;
    Staged s; s.u = 0.0; s.v = 0.0; s.w = 0.0; s.t = 0.0; s.s = 0.0;
    X6 raw; raw.u = 0.0; raw.v = 0.0; raw.w = 0.0; raw.p = 0.0; raw.t = 0.0; raw.s = 0.0;
    if (fi <= S.io0 + S.nloc && fj <= S.jo0 + S.mloc && fk <= S.l + 1) s = stage_cell_t(ld, a.sig, S, fi, fj, fk, raw);
    sm[t] = s.u; sm[TCELLS + t] = s.v; sm[2 * TCELLS + t] = s.w; sm[3 * TCELLS + t] = s.t; sm[4 * TCELLS + t] = s.s;
    // raw T,S equal the staged ones wherever matAvec can address them, so only u,v,w,p need a raw copy
    if (RHS) { sm[5 * TCELLS + t] = raw.u; sm[6 * TCELLS + t] = raw.v; sm[7 * TCELLS + t] = raw.w; sm[8 * TCELLS + t] = raw.p; }
  }
}

// Tiles whose cells (halo layer included) are all owned, interior cells of the global grid -- most tiles -- need none
// of the ghost rules: branch-free addressing, and all of a thread's loads are in flight together.
__device__ __forceinline__ bool tile_is_plain(const Scalars& S, int gi0, int gj0, int gk0) {
  return gi0 - 1 >= S.io0 && gi0 + TI <= S.io0 + S.nloc - 1 && gj0 - 1 >= S.jo0 && gj0 + TJ <= S.jo0 + S.mloc - 1 && gk0 >= 2 && gk0 + TK <= S.l;
}
template <bool RHS>
__device__ __forceinline__ void stage_tile_plain(const AsmArgs& a, double* sm, int gi0, int gj0, int gk0) {
  const Scalars& S = a.S;
  constexpr int NIT = (TCELLS + NTHREADS - 1) / NTHREADS;
  const size_t base = ((size_t)(gk0 - 2) * S.mloc + (gj0 - 1 - S.jo0)) * S.nloc + (gi0 - 1 - S.io0); // tile cell (0,0,0)
  double2 q[NIT][3];
  uint64_t sg[NIT];
#pragma unroll
  for (int it = 0; it < NIT; it++) {
    const int t = min((int)threadIdx.x + it * NTHREADS, TCELLS - 1);
    const int ti = t % SX, tj = (t / SX) % SY, tk = t / (SX * SY);
    const size_t c = base + ((size_t)tk * S.mloc + tj) * S.nloc + ti;
    const double2* p = reinterpret_cast<const double2*>(a.x + 6 * c);
    q[it][0] = __ldg(p); q[it][1] = __ldg(p + 1); q[it][2] = __ldg(p + 2);
    sg[it] = __ldg(a.sig + c);
  }
#pragma unroll
  for (int it = 0; it < NIT; it++) {
    const int t = threadIdx.x + it * NTHREADS;
    if (t < TCELLS) {
      const bool live = sbit(sg[it], SIG_LIVE), top = gk0 + t / (SX * SY) - 1 == S.l;
      sm[t] = live ? q[it][0].x : 0.0; sm[TCELLS + t] = live ? q[it][0].y : 0.0; sm[2 * TCELLS + t] = top ? 0.0 : q[it][1].x;
      sm[3 * TCELLS + t] = q[it][2].x; sm[4 * TCELLS + t] = q[it][2].y;
      if (RHS) { sm[5 * TCELLS + t] = q[it][0].x; sm[6 * TCELLS + t] = q[it][0].y; sm[7 * TCELLS + t] = q[it][1].x; sm[8 * TCELLS + t] = q[it][1].y; }
    }
  }
}

__global__ void __launch_bounds__(NTHREADS, JAC_MINB) jac_kernel(const __grid_constant__ AsmArgs a) {
  extern __shared__ double sm[];
  const Scalars& S = a.S;
  const int bx = blockIdx.x % a.ntx, by = (blockIdx.x / a.ntx) % a.nty, bz = blockIdx.x / (a.ntx * a.nty) + a.kz0;
  if (a.part) { // overlap with the halo exchange: rim tiles wait for it, the others do not (block-uniform exit)
    const bool rim = bx == 0 || bx == a.ntx - 1 || by == 0 || by == a.nty - 1;
    if (rim != (a.part == 2)) return;
  }
  const int gi0 = S.io0 + bx * TI, gj0 = S.jo0 + by * TJ, gk0 = 1 + bz * TK;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int lj = warp % TJ, lk = warp / TJ;
  const int i = gi0 + lane, j = gj0 + lj, k = gk0 + lk;
  const bool line = j < S.jo0 + S.mloc && k <= S.l; // warp-uniform
  const bool active = line && i < S.io0 + S.nloc;
  // this thread's own cell: signature and row offsets are requested before the tile is staged so that their latency
  // hides behind the staging loads
  const size_t cidx = active ? cell_index(S, i, j, k) : 0;
  Cell c; c.i = i; c.j = j; c.k = k; c.sig = active ? __ldg(a.sig + cidx) : 0ull;
  const int row0 = (int)(6 * cidx);
  // Row offsets: a line whose 32 cells all carry complete rows (almost all lines) needs only the first offset of every cell --
  // the cells are 104 values apart and the rows of a cell follow from the compile-time row lengths -- so one load per lane
  // (plus the end of the last cell) replaces seven.  Anything else loads the remaining offsets after the tile is staged.
  int rp[7];
  rp[0] = active ? __ldg(a.rowptr + row0) : 0;
  rp[6] = (active && lane == 31) ? __ldg(a.rowptr + row0 + 6) : 0;

  if (tile_is_plain(S, gi0, gj0, gk0)) stage_tile_plain<false>(a, sm, gi0, gj0, gk0);
  else stage_tile<false>(a, sm, gi0, gj0, gk0);
  __syncthreads();
  if (!line) return;
  {
    const int nxt = __shfl_down_sync(FULL, rp[0], 1);
    if (lane < 31) rp[6] = nxt;                 // rows are contiguous: the next cell starts where this one ends
    const bool special = a.sp.rowintcon >= row0 && a.sp.rowintcon < row0 + 6;
    const bool full = __all_sync(FULL, active && rp[6] - rp[0] == CELL_NNZ && !special);
    if (full) {
      rp[1] = rp[0] + row_offset<VV>(); rp[2] = rp[0] + row_offset<WW>(); rp[3] = rp[0] + row_offset<PP>();
      rp[4] = rp[0] + row_offset<TT>(); rp[5] = rp[0] + row_offset<SS>();
    } else {
#pragma unroll
      for (int q = 1; q < 7; q++) rp[q] = active ? __ldg(a.rowptr + row0 + q) : 0;
    }
  }
  TileFields f; f.pu = sm; f.pv = sm + TCELLS; f.pw = sm + 2 * TCELLS; f.pt = sm + 3 * TCELLS; f.ps = sm + 4 * TCELLS;
  f.base = (lane + 1) + (lj + 1) * SX + (lk + 1) * SX * SY;
  double* scr = sm + 5 * TCELLS + warp * 32 * STRIDE;

  // how this warp's rows go out: all 32 cells complete (compile-time strides), all active cells alike (run-time strides),
  // or anything else (non-periodic i edge, dense integral row): per-element shuffles
  CopyPlan cp;
  cp.b0 = __shfl_sync(FULL, rp[0], 0);
  cp.cs = __shfl_sync(FULL, rp[6] - rp[0], 0);
  cp.nact = __popc(__ballot_sync(FULL, active));
  {
    bool same = rp[0] == cp.b0 + lane * cp.cs && !(a.sp.rowintcon >= row0 && a.sp.rowintcon < row0 + 6);
#pragma unroll
    for (int q = 1; q <= 6; q++) { // the shuffle must be executed by every lane: no short-circuit in front of it
      const int ref = __shfl_sync(FULL, rp[q] - rp[0], 0);
      same = (rp[q] - rp[0] == ref) && same;
    }
    const bool uniform = __all_sync(FULL, !active || same);
    cp.mode = !uniform ? COPY_GENERAL : ((cp.nact == 32 && cp.cs == CELL_NNZ) ? COPY_FULL : COPY_UNIFORM);
  }
  double b[6];
  b[0] = do_jac_row<UU>(a, c, f, active, row0, rp, scr, lane, cp);
  b[1] = do_jac_row<VV>(a, c, f, active, row0, rp, scr, lane, cp);
  b[2] = do_jac_row<WW>(a, c, f, active, row0, rp, scr, lane, cp);
  b[3] = do_jac_row<PP>(a, c, f, active, row0, rp, scr, lane, cp);
  // convective-adjustment finite differences: evaluated once, shared by the T and the S row
  MixFD fd;
  if (active && S.vmix_on && !sbit(c.sig, SIG_NOTOCEAN))
    thcm_mix_fd(S, a.T, c, f.t(0, 0, -1), f.s(0, 0, -1), f.t(0, 0, 0), f.s(0, 0, 0), f.t(0, 0, 1), f.s(0, 0, 1), fd);
  b[4] = do_jac_row<TT>(a, c, f, active, row0, rp, scr, lane, cp, &fd);
  b[5] = do_jac_row<SS>(a, c, f, active, row0, rp, scr, lane, cp, &fd);

  // diag(B): the active cells of a line are consecutive rows -> one contiguous, coalesced run
  if (active) {
#pragma unroll
    for (int q = 0; q < 6; q++) scr[lane * 6 + q] = b[q];
  }
  __syncwarp();
  const int nact = __popc(__ballot_sync(FULL, active));
  const int base = __shfl_sync(FULL, row0, 0);
#pragma unroll
  for (int it = 0; it < 6; it++) {
    const int t = it * 32 + lane;
    if (t < 6 * nact) __stcs(a.diagB + (size_t)base + t, scr[t]);
  }
}

#ifndef RHS_MINB
#define RHS_MINB 2
#endif
__global__ void __launch_bounds__(NTHREADS, RHS_MINB) rhs_kernel(const __grid_constant__ AsmArgs a) {
  extern __shared__ double sm[];
  const Scalars& S = a.S;
  const int bx = blockIdx.x % a.ntx, by = (blockIdx.x / a.ntx) % a.nty, bz = blockIdx.x / (a.ntx * a.nty) + a.kz0;
  if (a.part) { // overlap with the halo exchange: rim tiles wait for it, the others do not (block-uniform exit)
    const bool rim = bx == 0 || bx == a.ntx - 1 || by == 0 || by == a.nty - 1;
    if (rim != (a.part == 2)) return;
  }
  const int gi0 = S.io0 + bx * TI, gj0 = S.jo0 + by * TJ, gk0 = 1 + bz * TK;
  if (tile_is_plain(S, gi0, gj0, gk0)) stage_tile_plain<true>(a, sm, gi0, gj0, gk0);
  else stage_tile<true>(a, sm, gi0, gj0, gk0);
  __syncthreads();

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int lj = warp % TJ, lk = warp / TJ;
  const int i = gi0 + lane, j = gj0 + lj, k = gk0 + lk;
  if (j >= S.jo0 + S.mloc || k > S.l) return;
  const bool active = i < S.io0 + S.nloc;
  const size_t cidx = active ? cell_index(S, i, j, k) : 0;
  Cell c; c.i = i; c.j = j; c.k = k; c.sig = active ? a.sig[cidx] : 0ull;
  TileFields f; f.pu = sm; f.pv = sm + TCELLS; f.pw = sm + 2 * TCELLS; f.pt = sm + 3 * TCELLS; f.ps = sm + 4 * TCELLS;
  f.base = (lane + 1) + (lj + 1) * SX + (lk + 1) * SX * SY;
  TileRaw xr; xr.ru = sm + 5 * TCELLS; xr.rv = sm + 6 * TCELLS; xr.rw = sm + 7 * TCELLS; xr.rp = sm + 8 * TCELLS;
  xr.rt = f.pt; xr.rs = f.ps; xr.base = f.base;
  const size_t row0 = 6 * cidx;
  double* scr = sm + 9 * TCELLS + warp * 32 * 6;

  // forcing of this line: one coalesced run of 6*nact doubles, read through the scratch
  const int nact = __popc(__ballot_sync(FULL, active));
  const size_t base = __shfl_sync(FULL, (unsigned long long)row0, 0);
#pragma unroll
  for (int it = 0; it < 6; it++) {
    const int t = it * 32 + lane;
    if (t < 6 * nact) scr[t] = __ldcs(a.frc + base + t);
  }
  __syncwarp();
  double Fv[6];
  if (active) {
    double frc[6];
#pragma unroll
    for (int q = 0; q < 6; q++) frc[q] = scr[lane * 6 + q];
    double mt = 0.0, ms = 0.0;
    if (S.vmix_on) thcm_mix(S, a.T, c, f.t(0, 0, -1), f.s(0, 0, -1), f.t(0, 0, 0), f.s(0, 0, 0), f.t(0, 0, 1), f.s(0, 0, 1), mt, ms);
    Fv[0] = rhs_row<UU>(S, a.T, c, f, xr, frc[0], 0.0);
    Fv[1] = rhs_row<VV>(S, a.T, c, f, xr, frc[1], 0.0);
    Fv[2] = rhs_row<WW>(S, a.T, c, f, xr, frc[2], 0.0);
    Fv[3] = rhs_row<PP>(S, a.T, c, f, xr, frc[3], 0.0);
    Fv[4] = rhs_row<TT>(S, a.T, c, f, xr, frc[4], mt);
    Fv[5] = rhs_row<SS>(S, a.T, c, f, xr, frc[5], ms);
  }
  __syncwarp();
  if (active) {
#pragma unroll
    for (int q = 0; q < 6; q++) scr[lane * 6 + q] = Fv[q];
  }
  __syncwarp();
#pragma unroll
  for (int it = 0; it < 6; it++) {
    const int t = it * 32 + lane;
    if (t < 6 * nact) __stcs(a.F + base + t, scr[t]);
  }
}

// ---- integral condition: sum over ocean S dofs of coef*x (THCM.C:1006-1018) ----
__global__ void __launch_bounds__(1024, 1) intcond_dot_kernel(const double* __restrict__ coefS, const double* __restrict__ x, long long ncell, double* out) {
  // single block: the SRES=0 configurations are small (2DMOC, test grids); deterministic summation order
  __shared__ double red[32];
  double s = 0.0;
  for (long long c = threadIdx.x; c < ncell; c += blockDim.x) s += coefS[c] * x[6 * c + 5];
  for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(FULL, s, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x < 32) {
    s = (threadIdx.x < (blockDim.x >> 5)) ? red[threadIdx.x] : 0.0;
    for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(FULL, s, o);
    if (threadIdx.x == 0) *out = s;
  }
}
__global__ void intcond_finish_kernel(const double* dotval, double sign, double corr, double* F, int row, int p1, int p2) {
  if (row >= 0) F[row] = sign * (*dotval - corr);
  if (p1 >= 0) F[p1] = 0.0;
  if (p2 >= 0) F[p2] = 0.0;
}
} // namespace

int jac_smem_bytes() { return (5 * TCELLS + NWARP * 32 * STRIDE) * (int)sizeof(double); }
int rhs_smem_bytes() { return (9 * TCELLS + NWARP * 32 * 6) * (int)sizeof(double); }

static int grid_of(const AsmArgs& a) { return a.ntx * a.nty * (a.nkz > 0 ? a.nkz : a.ntz); }
int asm_tile_levels() { return TK; }

void launch_jac(const AsmArgs& a0, cudaStream_t st) {
  AsmArgs a = a0;
  a.ntx = (a.S.nloc + TI - 1) / TI; a.nty = (a.S.mloc + TJ - 1) / TJ; a.ntz = (a.S.l + TK - 1) / TK;
  if (a.nkz <= 0) a.kz0 = 0;
  static bool attr = false;
  if (!attr) { cudaFuncSetAttribute(jac_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, jac_smem_bytes()); attr = true; }
  jac_kernel<<<grid_of(a), NTHREADS, jac_smem_bytes(), st>>>(a);
}
void launch_rhs(const AsmArgs& a0, cudaStream_t st) {
  AsmArgs a = a0;
  a.ntx = (a.S.nloc + TI - 1) / TI; a.nty = (a.S.mloc + TJ - 1) / TJ; a.ntz = (a.S.l + TK - 1) / TK;
  if (a.nkz <= 0) a.kz0 = 0;
  static bool attr = false;
  if (!attr) { cudaFuncSetAttribute(rhs_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, rhs_smem_bytes()); attr = true; }
  rhs_kernel<<<grid_of(a), NTHREADS, rhs_smem_bytes(), st>>>(a);
}
void launch_intcond_dot(const double* coefS, const double* x, long long ncell, double* partial, cudaStream_t st) {
  intcond_dot_kernel<<<1, 1024, 0, st>>>(coefS, x, ncell, partial);
}
void launch_intcond_finish(const double* dotval, double sign, double corr, double* F, int row, int p1, int p2, cudaStream_t st) {
  intcond_finish_kernel<<<1, 1, 0, st>>>(dotval, sign, corr, F, row, p1, p2);
}

} // namespace thcm
