// spmv_tile.cu -- y = J x as a structured stencil product over the static THCM graph, for sm_100a.
//
// Replaces Epetra_CrsMatrix::Apply (call sites src/ocean/Ocean.C:1144, 1355) / matAvec (src/ocean/matetc.F90:147-166) for the
// Jacobian this library assembles itself.  The CSR-stream kernel of spmv.cu treats the matrix as an anonymous CSR: 321 M warp
// instructions at 1 degree, bound by the L1 data pipe (gathers of x + shared-memory products).  Here the graph is known -- it
// is the CreateMaximalGraph pattern generated from the RowDef tables of thcm_rows.cuh -- so the kernel is the mirror image of
// the assembly kernel (thcm_kernels.cu):
//   * one thread block per 32x4x2 tile of cells; the six unknowns of the tile's cells and of its radius-1 halo (x is a stored
//     vector: owned cells, then the halo ring) are staged once in shared memory, structure-of-arrays;
//   * one warp per line of 32 cells in i, one lane per cell; the 104 values of a cell are contiguous in the CSR and the cells
//     of a line follow each other, so the warp streams the values of a row piece (U 24, V 22, W+P 18, T 20, S 20 values per
//     cell) with coalesced 16-byte loads into a per-warp scratch, transposed so that every lane finds its row in its own
//     208-byte slot; no column index is ever loaded for a complete cell: the neighbour of every entry is a compile-time
//     offset into the tile;
//   * every row is summed left to right in STORAGE order with separately rounded products and sums -- matAvec's loop -- so
//     the result is bit-identical to the reference's serial product (cells on the periodic seam sum in their permuted order);
//   * cells whose rows are clipped by the domain edge read their values straight from global memory (a lane walks its own
//     contiguous values) with the same slot logic the graph builder uses; the few cells that are both clipped and on the seam,
//     and the cell that carries the dense integral row, fall back to the generic CSR loop with loaded column indices.
// HBM traffic: 8 B per non-zero + x once + 8 B y per row + 8 B of row offsets per cell.
#include "kernels.h"
#include "thcm_stage.cuh"

namespace thcm {
namespace {
constexpr int TI = 32, TJ = 4, TK = 2;
constexpr int SX = TI + 2, SY = TJ + 2, SZ = TK + 2;
constexpr int TCELLS = SX * SY * SZ;
constexpr int NWARP = TJ * TK, NTHREADS = NWARP * 32;
constexpr int STRIDE = 26; // scratch doubles per lane (see thcm_kernels.cu)
constexpr unsigned FULL = 0xffffffffu;
constexpr int CELL_NNZ = SPMV_CELL_NNZ;
constexpr int LONG_ROW = 1024; // rows longer than this (the dense integral row of a large grid) are left to spmv_longrow_kernel

template <int A> constexpr int row_offset() {
  int o = 0;
  if (A > UU) o += RowDef<UU>::NE;
  if (A > VV) o += RowDef<VV>::NE;
  if (A > WW) o += RowDef<WW>::NE;
  if (A > PP) o += RowDef<PP>::NE;
  if (A > TT) o += RowDef<TT>::NE;
  return o;
}
// entry of row A that sits at storage position q of a complete cell: SIDE 0 static order, 1 / 2 periodic seam at i = 1 / i = n
template <int A, int SIDE> constexpr int entry_at(int q) {
  if (SIDE == 0) return q;
  for (int e = 0; e < RowDef<A>::NE; e++)
    if (seam_pos<A>(e, SIDE - 1) == q) return e;
  return -1;
}

struct XTile { // x of the tile's cells, structure of arrays: xs[var][tile cell]
  const double* xs;
  int base;
  template <int VAR> __device__ __forceinline__ double get(int di, int dj, int dk) const {
    return xs[(VAR - 1) * TCELLS + base + di + dj * SX + dk * SX * SY];
  }
};

// one complete row: vals[q] is the value at storage position q (shared-memory slot or global memory)
template <int A, int SIDE>
__device__ __forceinline__ double row_complete(const double* vals, const XTile& xt) {
  double s = 0.0;
  static_for<0, RowDef<A>::NE>([&](auto qq) {
    constexpr int q = decltype(qq)::value;
    constexpr int e = entry_at<A, SIDE>(q);
    constexpr int slot = RowDef<A>::e[e].slot, var = RowDef<A>::e[e].var;
    s = __dadd_rn(s, __dmul_rn(vals[q], xt.template get<var>(slot_di(slot), slot_dj(slot), slot_dk(slot))));
  });
  return s;
}
// a row clipped by the domain edge, not on the seam: static order, absent slots skipped (emit_row's second branch).  The lane
// first copies its row's values into its own scratch slot with independent loads (one memory round trip per row, not one per
// entry), then walks the static entries.
template <int A>
__device__ __forceinline__ double row_clipped(const double* __restrict__ gvals, int len, double* slot, const XTile& xt, const Scalars& S, const Cell& c) {
  double r[RowDef<A>::NE];
#pragma unroll
  for (int q = 0; q < RowDef<A>::NE; q++) r[q] = q < len ? __ldg(gvals + q) : 0.0;
#pragma unroll
  for (int q = 0; q < RowDef<A>::NE; q++) slot[q] = r[q];
  double s = 0.0;
  int pos = 0;
  static_for<0, RowDef<A>::NE>([&](auto qq) {
    constexpr int e = decltype(qq)::value;
    constexpr int sl = RowDef<A>::e[e].slot, var = RowDef<A>::e[e].var;
    if (slot_present(sl, S, c)) {
      s = __dadd_rn(s, __dmul_rn(slot[pos], xt.template get<var>(slot_di(sl), slot_dj(sl), slot_dk(sl))));
      pos++;
    }
  });
  return s;
}
// anything else: the CSR loop with loaded column indices, x from global memory
__device__ __noinline__ double row_generic(const int* __restrict__ rowptr, const int* __restrict__ col, const double* __restrict__ val,
                                           const double* __restrict__ x, int row) {
  const int p0 = __ldg(rowptr + row), p1 = __ldg(rowptr + row + 1);
  if (p1 - p0 > LONG_ROW) return 0.0; // written by spmv_longrow_kernel
  double s = 0.0;
  for (int p = p0; p < p1; p++) s = __dadd_rn(s, __dmul_rn(__ldg(val + p), __ldg(x + __ldg(col + p))));
  return s;
}

__device__ __forceinline__ bool tile_is_plain(const Scalars& S, int gi0, int gj0, int gk0) {
  return gi0 - 1 >= S.io0 && gi0 + TI <= S.io0 + S.nloc - 1 && gj0 - 1 >= S.jo0 && gj0 + TJ <= S.jo0 + S.mloc - 1 && gk0 >= 2 && gk0 + TK <= S.l;
}
// x of the tile (+ halo) into shared memory.  Positions outside the global domain hold 0 and are never referenced: the graph
// has no entry for them.
__device__ __forceinline__ void stage_x(const Scalars& S, const double* __restrict__ x, double* sm, int gi0, int gj0, int gk0) {
  constexpr int NIT = (TCELLS + NTHREADS - 1) / NTHREADS;
  if (tile_is_plain(S, gi0, gj0, gk0)) {
    const size_t base = ((size_t)(gk0 - 2) * S.mloc + (gj0 - 1 - S.jo0)) * S.nloc + (gi0 - 1 - S.io0); // tile cell (0,0,0)
    double2 q[NIT][3];
#pragma unroll
    for (int it = 0; it < NIT; it++) {
      const int t = min((int)threadIdx.x + it * NTHREADS, TCELLS - 1);
      const int ti = t % SX, tj = (t / SX) % SY, tk = t / (SX * SY);
      const double2* p = reinterpret_cast<const double2*>(x + 6 * (base + ((size_t)tk * S.mloc + tj) * S.nloc + ti));
      q[it][0] = __ldg(p); q[it][1] = __ldg(p + 1); q[it][2] = __ldg(p + 2);
    }
#pragma unroll
    for (int it = 0; it < NIT; it++) {
      const int t = threadIdx.x + it * NTHREADS;
      if (t < TCELLS) {
        sm[t] = q[it][0].x; sm[TCELLS + t] = q[it][0].y; sm[2 * TCELLS + t] = q[it][1].x;
        sm[3 * TCELLS + t] = q[it][1].y; sm[4 * TCELLS + t] = q[it][2].x; sm[5 * TCELLS + t] = q[it][2].y;
      }
    }
    return;
  }
  for (int t = threadIdx.x; t < TCELLS; t += NTHREADS) {
    const int ti = t % SX, tj = (t / SX) % SY, tk = t / (SX * SY);
    int fi = gi0 + ti - 1;
    const int fj = gj0 + tj - 1, fk = gk0 + tk - 1;
    double2 a = make_double2(0.0, 0.0), b = a, c = a;
    bool in = fi <= S.io0 + S.nloc && fj <= S.jo0 + S.mloc && fj >= 1 && fj <= S.m && fk >= 1 && fk <= S.l;
    if (S.periodic) { if (fi == 0) fi = S.n; else if (fi == S.n + 1) fi = 1; }
    in = in && fi >= 1 && fi <= S.n;
    if (in) {
      const double2* p = reinterpret_cast<const double2*>(x + 6 * cell_index(S, fi, fj, fk));
      a = __ldg(p); b = __ldg(p + 1); c = __ldg(p + 2);
    }
    sm[t] = a.x; sm[TCELLS + t] = a.y; sm[2 * TCELLS + t] = b.x; sm[3 * TCELLS + t] = b.y; sm[4 * TCELLS + t] = c.x; sm[5 * TCELLS + t] = c.y;
  }
}

// Coalesced load of one row piece of a complete line: the mirror of warp_copy_vec (thcm_kernels.cu).  Split in two halves so
// that the loads of the NEXT piece are in flight while the lanes work on the current one: issue() fills registers, park()
// writes them into the warp scratch, transposed (lane = (group g, 16-byte chunk q); group g handles cells g, g+G, ...).
template <int NP, int OFF> struct Piece {
  static_assert(NP % 2 == 0 && OFF % 2 == 0 && NP <= STRIDE, "16-byte pieces");
  static constexpr int NC = NP / 2, G = 32 / NC, NIT = (32 + G - 1) / G;
  double2 v[NIT];
  __device__ __forceinline__ void issue(const double* __restrict__ val_line, int lane, int nact) {
    const int g = lane / NC, q = lane - g * NC;
    if (g < G) {
      const double2* src = reinterpret_cast<const double2*>(val_line + g * CELL_NNZ + OFF) + q;
#pragma unroll
      for (int it = 0; it < NIT; it++)
        if (it * G + g < nact) v[it] = __ldcs(src + it * G * (CELL_NNZ / 2));
    }
  }
  __device__ __forceinline__ void park(double* scr, int lane, int nact) const {
    const int g = lane / NC, q = lane - g * NC;
    if (g < G) {
      double2* dst = reinterpret_cast<double2*>(scr + g * STRIDE) + q;
#pragma unroll
      for (int it = 0; it < NIT; it++)
        if (it * G + g < nact) dst[it * G * (STRIDE / 2)] = v[it];
    }
  }
};

template <int A> __device__ __forceinline__ double complete_row_side(const double* vals, const XTile& xt, int side) {
  if (side == 0) return row_complete<A, 0>(vals, xt);
  if (side == 1) return row_complete<A, 1>(vals, xt);
  return row_complete<A, 2>(vals, xt);
}

__global__ void __launch_bounds__(NTHREADS, 2) spmv_tile_kernel(const Scalars S, const int* __restrict__ rowptr, const int* __restrict__ col,
                                                                const double* __restrict__ val, const double* __restrict__ x, double* __restrict__ y) {
  extern __shared__ __align__(16) double sm[];
  // the bottom and the top layer of tiles hold the clipped rows (slower path): they are scheduled first, not in the tail
  const int ntz = gridDim.z, bz = blockIdx.z == 0 ? 0 : (blockIdx.z == 1 ? ntz - 1 : (int)blockIdx.z - 1);
  const int gi0 = S.io0 + blockIdx.x * TI, gj0 = S.jo0 + blockIdx.y * TJ, gk0 = 1 + bz * TK;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int lj = warp % TJ, lk = warp / TJ;
  const int i = gi0 + lane, j = gj0 + lj, k = gk0 + lk;
  const bool line = j < S.jo0 + S.mloc && k <= S.l;
  const bool active = line && i < S.io0 + S.nloc;
  const size_t cidx = active ? cell_index(S, i, j, k) : 0;
  const int row0 = (int)(6 * cidx);
  int p0 = 0, pe = 0;
  if (active) { p0 = __ldg(rowptr + row0); pe = __ldg(rowptr + row0 + 6); }
  // Is the line one contiguous, 16-byte aligned run of complete cells?  Then its values (26.6 KB) are pulled into L2 by ONE TMA
  // prefetch right away, in a single stream and behind the staging of x: the piece loads below visit every 832-byte cell five
  // times, 144..192 bytes at a time -- poison for DRAM page locality if they had to go to DRAM themselves.
  const bool complete = pe - p0 == CELL_NNZ;
  const int nact = __popc(__ballot_sync(FULL, active));
  const int b0 = __shfl_sync(FULL, p0, 0);
  const bool fast = nact > 0 && __all_sync(FULL, !active || (complete && p0 == b0 + lane * CELL_NNZ)) && (b0 & 1) == 0 && ((size_t)val & 15) == 0;
#ifndef SPMV_NO_PREFETCH
  if (fast && lane == 0) asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(val + (size_t)b0), "r"((unsigned)(nact * CELL_NNZ * 8)) : "memory");
#endif
  stage_x(S, x, sm, gi0, gj0, gk0);
  __syncthreads();
  if (!line) return;

  XTile xt; xt.xs = sm; xt.base = (lane + 1) + (lj + 1) * SX + (lk + 1) * SX * SY;
  double* scr = sm + 6 * TCELLS + warp * 32 * STRIDE;
  Cell c; c.i = i; c.j = j; c.k = k; c.sig = 0;
  const bool seam = row_has_seam(S, c);
  const int side = !seam ? 0 : (i == 1 ? 1 : 2);
  double out[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};

  if (fast) {
    // stream the line's values through the scratch piece by piece
    const double* vl = val + (size_t)b0;
    const double* mine = scr + lane * STRIDE;
    Piece<RowDef<UU>::NE, row_offset<UU>()> pu;
    Piece<RowDef<VV>::NE, row_offset<VV>()> pv;
    Piece<RowDef<WW>::NE + RowDef<PP>::NE, row_offset<WW>()> pw;
    Piece<RowDef<TT>::NE, row_offset<TT>()> pt;
    Piece<RowDef<SS>::NE, row_offset<SS>()> ps;
    pu.issue(vl, lane, nact);
    pu.park(scr, lane, nact);
    pv.issue(vl, lane, nact);                       // in flight while the U rows are summed
    __syncwarp();
    if (active) out[0] = complete_row_side<UU>(mine, xt, side);
    __syncwarp();
    pv.park(scr, lane, nact);
    pw.issue(vl, lane, nact);
    __syncwarp();
    if (active) out[1] = complete_row_side<VV>(mine, xt, side);
    __syncwarp();
    pw.park(scr, lane, nact);
    pt.issue(vl, lane, nact);
    __syncwarp();
    if (active) { out[2] = complete_row_side<WW>(mine, xt, side); out[3] = complete_row_side<PP>(mine + RowDef<WW>::NE, xt, side); }
    __syncwarp();
    pt.park(scr, lane, nact);
    ps.issue(vl, lane, nact);
    __syncwarp();
    if (active) out[4] = complete_row_side<TT>(mine, xt, side);
    __syncwarp();
    ps.park(scr, lane, nact);
    __syncwarp();
    if (active) out[5] = complete_row_side<SS>(mine, xt, side);
  } else if (active) {
    if (complete) { // complete cell in a line that is not one aligned run: the lane walks its own 104 values
      const double* v = val + (size_t)p0;
      out[0] = complete_row_side<UU>(v + row_offset<UU>(), xt, side); out[1] = complete_row_side<VV>(v + row_offset<VV>(), xt, side);
      out[2] = complete_row_side<WW>(v + row_offset<WW>(), xt, side); out[3] = complete_row_side<PP>(v + row_offset<PP>(), xt, side);
      out[4] = complete_row_side<TT>(v + row_offset<TT>(), xt, side); out[5] = complete_row_side<SS>(v + row_offset<SS>(), xt, side);
    } else if (!seam && pe - p0 < CELL_NNZ) { // clipped by the domain edge
      double* slot = scr + lane * STRIDE;
      int rp[7];
#pragma unroll
      for (int q = 0; q < 7; q++) rp[q] = __ldg(rowptr + row0 + q);
      out[0] = row_clipped<UU>(val + rp[0], rp[1] - rp[0], slot, xt, S, c); out[1] = row_clipped<VV>(val + rp[1], rp[2] - rp[1], slot, xt, S, c);
      out[2] = row_clipped<WW>(val + rp[2], rp[3] - rp[2], slot, xt, S, c); out[3] = row_clipped<PP>(val + rp[3], rp[4] - rp[3], slot, xt, S, c);
      out[4] = row_clipped<TT>(val + rp[4], rp[5] - rp[4], slot, xt, S, c); out[5] = row_clipped<SS>(val + rp[5], rp[6] - rp[5], slot, xt, S, c);
    } else { // clipped AND on the seam, or the cell of the dense integral row
#pragma unroll 1
      for (int q = 0; q < 6; q++) out[q] = row_generic(rowptr, col, val, x, row0 + q);
    }
  }
  if (active) {
    if (((size_t)y & 15) == 0) {
      double2* py = reinterpret_cast<double2*>(y + row0);
      py[0] = make_double2(out[0], out[1]); py[1] = make_double2(out[2], out[3]); py[2] = make_double2(out[4], out[5]);
    } else {
#pragma unroll
      for (int q = 0; q < 6; q++) y[row0 + q] = out[q];
    }
  }
}

// the dense salinity-integral row of a large grid (src/ocean/THCM.C:2494-2513): one block, tree reduction
__global__ void __launch_bounds__(1024, 1) spmv_longrow_kernel(const int* __restrict__ rowptr, const int* __restrict__ col, const double* __restrict__ val,
                                                               const double* __restrict__ x, double* __restrict__ y, int row) {
  __shared__ double red[32];
  const int p0 = rowptr[row], p1 = rowptr[row + 1];
  double s = 0.0;
  for (int p = p0 + threadIdx.x; p < p1; p += blockDim.x) s += val[p] * x[col[p]];
  for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(FULL, s, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x < 32) {
    s = red[threadIdx.x];
    for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(FULL, s, o);
    if (threadIdx.x == 0) y[row] = s;
  }
}
} // namespace

int spmv_tile_long_row() { return LONG_ROW; }
// long_row: local index of a row with more than spmv_tile_long_row() entries (-1: none)
void launch_spmv_tile(const Scalars& S, const int* rowptr, const int* col, const double* val, const double* x, double* y, int long_row, cudaStream_t st) {
  constexpr int smem = (6 * TCELLS + NWARP * 32 * STRIDE) * (int)sizeof(double);
  static unsigned long long attr_done = 0;
  int dev = 0;
  cudaGetDevice(&dev);
  if (!(attr_done >> (dev & 63) & 1ull)) { cudaFuncSetAttribute(spmv_tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem); attr_done |= 1ull << (dev & 63); }
  const dim3 grid((S.nloc + TI - 1) / TI, (S.mloc + TJ - 1) / TJ, (S.l + TK - 1) / TK);
  spmv_tile_kernel<<<grid, NTHREADS, smem, st>>>(S, rowptr, col, val, x, y);
  if (long_row >= 0) spmv_longrow_kernel<<<1, 1024, 0, st>>>(rowptr, col, val, x, y, long_row);
}

} // namespace thcm
