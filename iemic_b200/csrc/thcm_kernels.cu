// thcm_kernels.cu -- fused THCM residual and Jacobian assembly for sm_100a.
//
// Replaces the Fortran pipeline An=Al; nlin_*; vmix_*; boundaries; assemble; matAvec of the reference
// (src/ocean/usrc.F90:432-586) by kernels that never materialise the stencil tensor:
//
//   asm_kernel<JAC,RHS> : one thread block per 32x4x2 tile of cells.  The ghosted fields u,v,w,T,S that usol produces
//                (usrc.F90:997-1104) are staged once per tile in shared memory (radius-1 halo); each warp then owns a
//                line of 32 cells in i and walks the six rows U,V,W,P,T,S.
//                JAC: a row's <=24 entries are built in registers (thcm_rows.cuh) and written, in their final CSR
//                order, into the lane's own 208-byte slot of a per-warp shared-memory scratch.  The six rows of a cell
//                are contiguous in the static CSR, so every lane hands its piece (U 192 B, V 176 B, W+P 144 B, T 160 B,
//                S 160 B) to the TMA engine with one cp.async.bulk.global.shared::cta: no transposition, no per-element
//                store instructions.  A land cell is one 832-byte bulk copy from a constant block, and a tile of land
//                cells does not stage anything.  Lines whose rows are clipped by the domain edge (or whose first value
//                is not 16-byte aligned) keep the lane-transposed copy-out with ordinary stores.
//                RHS: the residual operator A_rhs(x) (nlin_rhs, a different operator than the Jacobian,
//                usrc.F90:775-870) is built from the SAME staged tile right after the Jacobian row, contracted with the
//                raw state in the reference's storage order (matAvec), and F = A(x)x + mix - Frc leaves through 16-byte
//                stores.  HBM traffic per cell (fused): 48 B state + 8 B signature + 48 B forcing in, 832 B values +
//                48 B diag(B) + 48 B residual out.
//   <true,false> is the Jacobian alone, <false,true> the residual alone (line searches, computeDFDPar).
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
#ifndef ASM_BULK
#define ASM_BULK 1
#endif
constexpr int TI = 32, TJ = JAC_TJ, TK = JAC_TK;
constexpr int SX = TI + 2, SY = TJ + 2, SZ = TK + 2;
constexpr int TCELLS = SX * SY * SZ; // 816 staged cells per tile
constexpr int NWARP = TJ * TK;       // one warp per line of the tile
constexpr int NTHREADS = NWARP * 32;
constexpr int STRIDE = 26;           // scratch doubles per lane: 208 B, so every lane's slot is 16-byte aligned (bulk copies; 16-byte
                                     // shared stores free of bank conflicts: 52 words per lane -> 8 lanes cover the 32 banks).
                                     // A smaller even stride works too (emit_slot hands back what does not fit), as long as the
                                     // ordinary-store path below finds room for a whole row per lane.
static_assert(STRIDE >= 24 && STRIDE % 2 == 0, "a U row has 24 entries; slots must stay 16-byte aligned");
constexpr unsigned FULL = 0xffffffffu;
#ifndef JAC_MINB
#define JAC_MINB 2
#endif
#ifndef RHS_MINB
#define RHS_MINB 3   // residual alone: 80 registers (112 bytes of spills) and 24 warps per SM beat 94 registers and 16 (0.172 -> 0.154 ms at 1 degree)
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

// ---- TMA bulk copies shared -> global (cp.async.bulk, SASS UBLKCP): per-thread bulk groups ----
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void bulk_s2g(double* gdst, const double* ssrc, unsigned bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"((unsigned)__cvta_generic_to_shared(ssrc)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); } // sources may be rewritten

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
template <int A> constexpr int centre_pos() { return row_offset<A>() + find_ent<A>(5, A); } // the diagonal entry inside a cell's 104 values
template <int A, int SIDE> constexpr int centre_seam() { return row_offset<A>() + seam_pos<A>(find_ent<A>(5, A), SIDE); } // same, cell on the periodic seam

// ---- copy-out with ordinary stores, for the lines the bulk path does not take ----
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
    if (q < ln) __stcs(val + (size_t)b + q, scr[cell * STRIDE + q]);
  }
}
// Lines whose active cells all have the SAME row lengths (complete cells at an odd offset, domain edges in j and k):
// lanes 0..len-1 move one cell's row per step.  `cs` = values per cell, `off`/`len` = this row's offset and length inside a cell.
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
enum { COPY_GENERAL = 0, COPY_UNIFORM = 1, COPY_BULK = 2 };
struct CopyPlan { int mode, b0, nact, cs; };

// Lines of complete cells at a 16-byte aligned offset: NP doubles per cell (NP even) at offset OFF inside the cell's
// CELL_NNZ values.  Lane = (group g, chunk q); group g moves cells g, g+G, ... : one LDS.128 + one STG.128 per step.
template <int NP, int OFF>
__device__ __forceinline__ void warp_copy_vec(double* __restrict__ val_line, const double* scr, int lane, int nact) {
  static_assert(NP % 2 == 0 && OFF % 2 == 0 && NP <= STRIDE, "16-byte pieces");
  constexpr int NC = NP / 2, G = 32 / NC;
  const int g = lane / NC, q = lane - g * NC;
  if (g < G) {
    const double2* src = reinterpret_cast<const double2*>(scr + g * STRIDE) + q;
    double2* dst = reinterpret_cast<double2*>(val_line + g * CELL_NNZ + OFF) + q;
    if (nact == 32) {
#pragma unroll
      for (int cc = 0; cc < 32; cc += G) {
        if (cc + G <= 32) __stcs(dst + cc * (CELL_NNZ / 2), src[cc * (STRIDE / 2)]);
        else if (cc + g < 32) __stcs(dst + cc * (CELL_NNZ / 2), src[cc * (STRIDE / 2)]);
      }
    } else {
#pragma unroll 4
      for (int cc = 0; cc < 32; cc += G)
        if (cc + g < nact) __stcs(dst + cc * (CELL_NNZ / 2), src[cc * (STRIDE / 2)]);
    }
  }
}

// A complete row into the lane's scratch slot in final order (SIDE 0: static order, 1 / 2: cell on the periodic seam at
// i = 1 / i = n, compile-time permutation); entries that do not fit the slot are returned in `tail`.
template <int A, int SIDE, bool ALIGNED>
__device__ __forceinline__ void emit_slot(const Row<A>& r, double* out, double (&tail)[2]) {
  constexpr int NE = Row<A>::NE, NS = NE < STRIDE ? NE : STRIDE;
  if constexpr (SIDE == 0) {
    if constexpr (ALIGNED) {
      double2* o2 = reinterpret_cast<double2*>(out);
#pragma unroll
      for (int q = 0; q + 1 < NS; q += 2) o2[q / 2] = make_double2(r.a[q], r.a[q + 1]);
      if constexpr (NS & 1) out[NS - 1] = r.a[NS - 1];
    } else {
#pragma unroll
      for (int q = 0; q < NS; q++) out[q] = r.a[q];
    }
    if constexpr (NE > STRIDE) { tail[0] = r.a[STRIDE]; tail[1] = r.a[STRIDE + 1]; }
  } else {
    static_for<0, NE>([&](auto q) {
      constexpr int e = decltype(q)::value;
      constexpr int pos = seam_pos<A>(e, SIDE - 1);
      if constexpr (pos < STRIDE) out[pos] = r.a[e];
      else tail[pos - STRIDE] = r.a[e];
    });
  }
}

// Jacobian row A of the line: registers -> scratch -> CSR.
template <int A>
__device__ __forceinline__ void jac_row_out(const AsmArgs& a, const Cell& c, const TileFields& f, bool active, int row0, double* scr, int lane,
                                            const CopyPlan& cp, const MixFD* fd) {
  Row<A> r;
  bool store = false;
  if (active) store = jac_row_build<A, TileFields>(r, a.S, a.T, c, f, a.sp, row0 + A - 1, fd);
  double* mine = scr + lane * STRIDE;
  if (cp.mode == COPY_BULK) {
    // Complete cells at a 16-byte aligned offset: the row goes into the lane's slot in final order with 16-byte shared
    // stores, then the warp moves the piece of all its cells with 16-byte loads and streaming stores -- lanes are split in
    // groups of NE/2, a group moves one cell's piece per step, all offsets are immediates.
    constexpr int PRE = (A == PP) ? RowDef<WW>::NE : 0;        // the P row rides behind the W row (7 + 11 values = 144 B)
    if (store) {
      double tail[2];
      if (!row_has_seam(a.S, c)) emit_slot<A, 0, (PRE % 2 == 0)>(r, mine + PRE, tail);
      else if (c.i == 1) emit_slot<A, 1, false>(r, mine + PRE, tail);
      else emit_slot<A, 2, false>(r, mine + PRE, tail);
    }
    if constexpr (A != WW) {
      constexpr int OFF = (A == PP) ? row_offset<WW>() : row_offset<A>();
      __syncwarp();
      warp_copy_vec<PRE + RowDef<A>::NE, OFF>(a.val + (size_t)cp.b0, scr, lane, cp.nact);
      __syncwarp();
    }
  } else {
    // rare path: the row offsets are (re)loaded here instead of being kept in registers across the whole kernel
    int p0 = 0, len = 0;
    if (active) { p0 = __ldg(a.rowptr + row0 + A - 1); len = store ? __ldg(a.rowptr + row0 + A) - p0 : 0; }
    if (store) emit_row<A>(r, a.S, c, mine);
    __syncwarp();
    if (cp.mode == COPY_UNIFORM) {
      const int off = __shfl_sync(FULL, p0, 0) - cp.b0, ln = __shfl_sync(FULL, len, 0);
      warp_copy_out_uniform(a.val + (size_t)cp.b0, scr, lane, cp.nact, cp.cs, off, ln);
    } else warp_copy_out<RowDef<A>::NE>(a.val, scr, p0, len, lane);
    __syncwarp();
  }
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
struct PlainLoads { double2 q[(TCELLS + NTHREADS - 1) / NTHREADS][3]; uint64_t sg[(TCELLS + NTHREADS - 1) / NTHREADS]; };
__device__ __forceinline__ void stage_plain_issue(const AsmArgs& a, PlainLoads& L, int gi0, int gj0, int gk0) {
  const Scalars& S = a.S;
  constexpr int NIT = (TCELLS + NTHREADS - 1) / NTHREADS;
  const size_t base = ((size_t)(gk0 - 2) * S.mloc + (gj0 - 1 - S.jo0)) * S.nloc + (gi0 - 1 - S.io0); // tile cell (0,0,0)
#pragma unroll
  for (int it = 0; it < NIT; it++) {
    const int t = min((int)threadIdx.x + it * NTHREADS, TCELLS - 1);
    const int ti = t % SX, tj = (t / SX) % SY, tk = t / (SX * SY);
    const size_t c = base + ((size_t)tk * S.mloc + tj) * S.nloc + ti;
    const double2* p = reinterpret_cast<const double2*>(a.x + 6 * c);
    L.q[it][0] = __ldg(p); L.q[it][1] = __ldg(p + 1); L.q[it][2] = __ldg(p + 2);
    L.sg[it] = __ldg(a.sig + c);
  }
}
template <bool RHS>
__device__ __forceinline__ void stage_plain_apply(const AsmArgs& a, const PlainLoads& L, double* sm, int gk0) {
  const Scalars& S = a.S;
  constexpr int NIT = (TCELLS + NTHREADS - 1) / NTHREADS;
#pragma unroll
  for (int it = 0; it < NIT; it++) {
    const int t = threadIdx.x + it * NTHREADS;
    if (t < TCELLS) {
      const bool live = sbit(L.sg[it], SIG_LIVE), top = gk0 + t / (SX * SY) - 1 == S.l;
      sm[t] = live ? L.q[it][0].x : 0.0; sm[TCELLS + t] = live ? L.q[it][0].y : 0.0; sm[2 * TCELLS + t] = top ? 0.0 : L.q[it][1].x;
      sm[3 * TCELLS + t] = L.q[it][2].x; sm[4 * TCELLS + t] = L.q[it][2].y;
      if (RHS) { sm[5 * TCELLS + t] = L.q[it][0].x; sm[6 * TCELLS + t] = L.q[it][0].y; sm[7 * TCELLS + t] = L.q[it][1].x; sm[8 * TCELLS + t] = L.q[it][1].y; }
    }
  }
}

// Residual of a non-ocean cell: the row is the identity (boundary.F90:383-391), so Au = x(row) and there is no mixing
// term; what rhs_row returns for it, without the row machinery.
__device__ __forceinline__ double rhs_land(uint64_t sig, double xraw, double frc) {
  const double Au = 1.0 * xraw + 0.0;
  double B = -Au - 0.0 + frc - 0.0;
  B = B * (sbit(sig, SIG_LANDVAL) ? 0.0 : 1.0);
  return B * -1.0;
}
__device__ __forceinline__ void store6(double* dst, const double (&v)[6]) { // 48 B per cell: 16-byte aligned when the vector is
  if (((size_t)dst & 15) == 0) {
    double2* p = reinterpret_cast<double2*>(dst);
    __stcs(p, make_double2(v[0], v[1])); __stcs(p + 1, make_double2(v[2], v[3])); __stcs(p + 2, make_double2(v[4], v[5]));
  } else {
#pragma unroll
    for (int q = 0; q < 6; q++) __stcs(dst + q, v[q]);
  }
}
__device__ __forceinline__ void load6(const double* src, double (&v)[6]) {
  if (((size_t)src & 15) == 0) {
    const double2* p = reinterpret_cast<const double2*>(src);
    const double2 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
    v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y; v[4] = c.x; v[5] = c.y;
  } else {
#pragma unroll
    for (int q = 0; q < 6; q++) v[q] = __ldg(src + q);
  }
}
// A line of land cells: six identity rows per cell.  The 832 bytes of a cell come from a constant block in shared memory
// through the TMA engine (cp.async.bulk is a uniform-datapath instruction: ONE lane issues the copies of the whole line);
// diag(B) = 0, residual from the cell's own state.
template <bool JAC, bool RHS>
__device__ __forceinline__ void land_line(const AsmArgs& a, const Cell& c, bool active, int row0, int b0, int nact, int lane, const double* ident,
                                          const double (&xown)[6], const double (&frc)[6]) {
  if (JAC) {
    const bool seam = a.S.periodic && a.S.nloc == a.S.n; // the first / last cell of the row then has its own column order
    if (lane == 0) {
      for (int cc = 0; cc < nact; cc++) {
        const int gi = c.i + cc;
        const int side = !seam ? 0 : (gi == 1 ? 1 : (gi == a.S.n ? 2 : 0));
        bulk_s2g(a.val + (size_t)b0 + (size_t)cc * CELL_NNZ, ident + side * CELL_NNZ, CELL_NNZ * 8);
      }
      bulk_commit();
    }
    if (active) {
      const double z[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
      store6(a.diagB + row0, z);
    }
  }
  if (RHS && active) {
    double Fv[6];
#pragma unroll
    for (int q = 0; q < 6; q++) Fv[q] = rhs_land(c.sig, xown[q], frc[q]);
    store6(a.F + row0, Fv);
  }
  if (JAC && lane == 0) bulk_wait_read(); // the block's shared memory must outlive the copies that read it
}

template <bool JAC, bool RHS>
__global__ void __launch_bounds__(NTHREADS, JAC ? JAC_MINB : RHS_MINB) asm_kernel(const __grid_constant__ AsmArgs a) {
  extern __shared__ __align__(16) double sm[];
  constexpr int NF = RHS ? 9 : 5;                     // staged field arrays
  constexpr int SCRW = JAC ? 32 * STRIDE : 0;         // scratch doubles per warp
  const Scalars& S = a.S;
  // surface layers (ocean: the expensive tiles) are scheduled first, deep layers (mostly land: bulk copies) last, so that the tail
  // of the grid is cheap (-1 % at 1 degree, -2.5 % on the 90x80x32 block a rank owns on 8 GPUs)
  const int bx = blockIdx.x, by = blockIdx.y, bz = (int)(gridDim.z - 1 - blockIdx.z) + a.kz0;
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
  const size_t cidx = active ? cell_index(S, i, j, k) : 0;
  Cell c; c.i = i; c.j = j; c.k = k; c.sig = active ? __ldg(a.sig + cidx) : 0ull;
  const int row0 = (int)(6 * cidx);
  // first and one-past-last value of the cell: a cell whose six rows are complete (all but domain-edge cells) has CELL_NNZ
  // values and its rows follow from the compile-time row lengths
  int p0 = 0, pe = 0;
  if (JAC && active) { p0 = __ldg(a.rowptr + row0); pe = __ldg(a.rowptr + row0 + 6); }
  const bool special = a.sp.rowintcon >= row0 && a.sp.rowintcon < row0 + 6;
  const bool pfix = a.sp.rowPfix1 == row0 + 3 || a.sp.rowPfix2 == row0 + 3;
  const bool bulk_cell = ASM_BULK && pe - p0 == CELL_NNZ && (p0 & 1) == 0 && !special && ((size_t)a.val & 15) == 0;

  // the 104 values of a land cell (six identity rows) in the three column orders a cell can have: static, periodic seam at
  // i = 1 (west neighbours sort last), seam at i = n (east neighbours sort first)
  double* ident = sm + NF * TCELLS + NWARP * SCRW;
  if (JAC) {
    for (int t = threadIdx.x; t < 3 * CELL_NNZ; t += NTHREADS) {
      const int side = t / CELL_NNZ, q = t - side * CELL_NNZ;
      bool one;
      if (side == 0)
        one = q == centre_pos<UU>() || q == centre_pos<VV>() || q == centre_pos<WW>() || q == centre_pos<PP>() || q == centre_pos<TT>() || q == centre_pos<SS>();
      else if (side == 1)
        one = q == centre_seam<UU, 0>() || q == centre_seam<VV, 0>() || q == centre_seam<WW, 0>() || q == centre_seam<PP, 0>() || q == centre_seam<TT, 0>() ||
              q == centre_seam<SS, 0>();
      else
        one = q == centre_seam<UU, 1>() || q == centre_seam<VV, 1>() || q == centre_seam<WW, 1>() || q == centre_seam<PP, 1>() || q == centre_seam<TT, 1>() ||
              q == centre_seam<SS, 1>();
      ident[t] = one ? 1.0 : 0.0;
    }
    fence_proxy_async();
  }

  // Tiles whose cells (halo layer included) are all owned, interior cells of the global grid -- most tiles -- have their
  // staging loads issued BEFORE the vote below, so that the loads of the cell's own signature / row offsets (which the
  // vote needs) and the tile's loads are one round trip to memory, not two.
  const bool plain = tile_is_plain(S, gi0, gj0, gk0);
  PlainLoads L;
  if (plain) stage_plain_issue(a, L, gi0, gj0, gk0);

  // A tile of land cells (a third of the 1-degree grid) needs no neighbour: nothing is staged.
  const bool easy_land = !active || (sbit(c.sig, SIG_NOTOCEAN) && (!JAC || (bulk_cell && !pfix)));
  if (__syncthreads_and(easy_land)) {
    double xown[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0}, frc[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
    if (RHS && active) { load6(a.x + row0, xown); load6(a.frc + row0, frc); }
    land_line<JAC, RHS>(a, c, active, row0, __shfl_sync(FULL, p0, 0), __popc(__ballot_sync(FULL, active)), lane, ident, xown, frc);
    return;
  }

  if (plain) stage_plain_apply<RHS>(a, L, sm, gk0);
  else stage_tile<RHS>(a, sm, gi0, gj0, gk0);
  __syncthreads();
  if (!line) return;

  TileFields f; f.pu = sm; f.pv = sm + TCELLS; f.pw = sm + 2 * TCELLS; f.pt = sm + 3 * TCELLS; f.ps = sm + 4 * TCELLS;
  f.base = (lane + 1) + (lj + 1) * SX + (lk + 1) * SX * SY;
  TileRaw xr; xr.ru = sm + 5 * TCELLS; xr.rv = sm + 6 * TCELLS; xr.rw = sm + 7 * TCELLS; xr.rp = sm + 8 * TCELLS;
  xr.rt = f.pt; xr.rs = f.ps; xr.base = f.base;
  double* scr = sm + NF * TCELLS + warp * SCRW;

  // How this warp's rows go out: every active cell complete at a 16-byte aligned offset -> bulk copies; all active cells
  // alike (same clipped row lengths, evenly spaced) -> lanes walk the cells with run-time strides; anything else
  // (non-periodic i edge, dense integral row) -> per-element shuffles.
  CopyPlan cp; cp.mode = COPY_GENERAL; cp.b0 = 0; cp.nact = 0; cp.cs = 0;
  if (JAC) {
    cp.nact = __popc(__ballot_sync(FULL, active));
    cp.b0 = __shfl_sync(FULL, p0, 0);
    cp.cs = __shfl_sync(FULL, pe - p0, 0);
    if (__all_sync(FULL, !active || (bulk_cell && p0 == cp.b0 + lane * CELL_NNZ))) cp.mode = COPY_BULK;
    else {
      bool same = p0 == cp.b0 + lane * cp.cs && !special;
#pragma unroll
      for (int q = 1; q <= 6; q++) { // the shuffle must be executed by every lane: no short-circuit in front of it
        const int mine = active ? __ldg(a.rowptr + row0 + q) - p0 : 0;
        const int ref = __shfl_sync(FULL, mine, 0);
        same = (mine == ref) && same;
      }
      cp.mode = __all_sync(FULL, !active || same) ? COPY_UNIFORM : COPY_GENERAL;
    }
    // a line of land cells inside a tile that has ocean: nothing to build
    if (cp.mode == COPY_BULK && __all_sync(FULL, !active || (sbit(c.sig, SIG_NOTOCEAN) && !pfix))) {
      double xown[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0}, frc[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
      if (RHS && active) {
#pragma unroll
        for (int q = 0; q < 6; q++) xown[q] = xr.raw(0, 0, 0, q + 1);
        load6(a.frc + row0, frc);
      }
      land_line<JAC, RHS>(a, c, active, row0, cp.b0, cp.nact, lane, ident, xown, frc);
      return;
    }
  }

  // convective adjustment: the base evaluation is the residual's mixing term, the six perturbed ones give the
  // forward-difference Jacobian entries shared by the T and the S row
  MixFD fd; fd.bt = 0.0; fd.bs = 0.0;
  double Fv[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
  // the row's forcing is requested before the Jacobian row is built, which hides its latency
#define ASM_ROW(A, FDP, MIXV)                                                                      \
  {                                                                                                \
    double frc = 0.0;                                                                              \
    if (RHS && active) frc = __ldg(a.frc + row0 + A - 1);                                          \
    if (JAC) jac_row_out<A>(a, c, f, active, row0, scr, lane, cp, FDP);                            \
    if (RHS && active) Fv[A - 1] = rhs_row<A>(S, a.T, c, f, xr, frc, MIXV);                        \
  }
  ASM_ROW(UU, nullptr, 0.0)
  ASM_ROW(VV, nullptr, 0.0)
  ASM_ROW(WW, nullptr, 0.0)
  ASM_ROW(PP, nullptr, 0.0)
  if (JAC) {
    if (active && S.vmix_on && !sbit(c.sig, SIG_NOTOCEAN))
      thcm_mix_fd(S, a.T, c, f.t(0, 0, -1), f.s(0, 0, -1), f.t(0, 0, 0), f.s(0, 0, 0), f.t(0, 0, 1), f.s(0, 0, 1), fd);
  } else if (active && S.vmix_on) {
    thcm_mix(S, a.T, c, f.t(0, 0, -1), f.s(0, 0, -1), f.t(0, 0, 0), f.s(0, 0, 0), f.t(0, 0, 1), f.s(0, 0, 1), fd.bt, fd.bs);
  }
  ASM_ROW(TT, &fd, fd.bt)
  ASM_ROW(SS, &fd, fd.bs)
#undef ASM_ROW

  // diag(B) and the residual: 48 contiguous bytes per cell, three 16-byte stores per lane
  if (active) {
    if (JAC) {
      double b[6];
      b[0] = diagb_entry<UU>(S, c.sig, a.sp, row0); b[1] = diagb_entry<VV>(S, c.sig, a.sp, row0 + 1); b[2] = diagb_entry<WW>(S, c.sig, a.sp, row0 + 2);
      b[3] = diagb_entry<PP>(S, c.sig, a.sp, row0 + 3); b[4] = diagb_entry<TT>(S, c.sig, a.sp, row0 + 4); b[5] = diagb_entry<SS>(S, c.sig, a.sp, row0 + 5);
      store6(a.diagB + row0, b);
    }
    if (RHS) store6(a.F + row0, Fv);
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

// Surface rows of the forcing vector after a parameter change (forcing.F90:4-218 through Host::forcing_profiles): one thread
// per owned surface cell copies the per-latitude values of an ocean cell into its six rows and applies the Frc zeroing of
// `boundaries` -- the same selection Host::build_forcing makes, no arithmetic, hence the same bits.
__global__ void __launch_bounds__(256) forcing_surface_kernel(const Scalars S, const uint64_t* __restrict__ sig, const double* __restrict__ fprof,
                                                              double* __restrict__ frc) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= S.nloc * S.mloc) return;
  const int li = t % S.nloc, lj = t / S.nloc, j = S.jo0 + lj;
  const size_t c = ((size_t)(S.l - 1) * S.mloc + lj) * S.nloc + li;
  const uint64_t g = __ldg(sig + c);
  const int stride = S.m + 2;
  double v[6];
  v[0] = frc_masked<UU>(g) ? 0.0 : __ldg(fprof + j);
  v[1] = frc_masked<VV>(g) ? 0.0 : __ldg(fprof + stride + j);
  v[2] = 0.0;
  v[3] = 0.0;
  v[4] = frc_masked<TT>(g) ? 0.0 : __ldg(fprof + 2 * stride + j);
  v[5] = frc_masked<SS>(g) ? 0.0 : __ldg(fprof + 3 * stride + j);
  store6(frc + 6 * c, v);
}

// THCM::evaluateB (THCM.C:1195-1235): diag(B) alone, from the signatures -- fillcolB plus the zeroed P-fix / integral rows
__global__ void __launch_bounds__(256) diagb_kernel(const Scalars S, const uint64_t* __restrict__ sig, const SpecialRows sp, double* __restrict__ diagB) {
  const long long c = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= S.nown) return;
  const uint64_t g = __ldg(sig + c);
  const int row0 = (int)(6 * c);
  double b[6];
  b[0] = diagb_entry<UU>(S, g, sp, row0); b[1] = diagb_entry<VV>(S, g, sp, row0 + 1); b[2] = diagb_entry<WW>(S, g, sp, row0 + 2);
  b[3] = diagb_entry<PP>(S, g, sp, row0 + 3); b[4] = diagb_entry<TT>(S, g, sp, row0 + 4); b[5] = diagb_entry<SS>(S, g, sp, row0 + 5);
  store6(diagB + row0, b);
}

__global__ void intcond_add_kernel(const double* dotval, double sign, double* y, int row) { y[row] = y[row] + sign * (*dotval); }

template <bool JAC, bool RHS> constexpr int asm_smem() {
  return ((RHS ? 9 : 5) * TCELLS + (JAC ? NWARP * 32 * STRIDE + 3 * CELL_NNZ : 0)) * (int)sizeof(double);
}
template <bool JAC, bool RHS> void launch_asm_t(const AsmArgs& a0, cudaStream_t st) {
  AsmArgs a = a0;
  a.ntx = (a.S.nloc + TI - 1) / TI; a.nty = (a.S.mloc + TJ - 1) / TJ; a.ntz = (a.S.l + TK - 1) / TK;
  if (a.nkz <= 0) a.kz0 = 0;
  constexpr int smem = asm_smem<JAC, RHS>();
  static unsigned long long attr_done = 0; // one bit per device: the attribute belongs to the device's copy of the function
  int dev = 0;
  cudaGetDevice(&dev);
  if (!(attr_done >> (dev & 63) & 1ull)) {
    cudaFuncSetAttribute(asm_kernel<JAC, RHS>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    attr_done |= 1ull << (dev & 63);
  }
  const dim3 grid(a.ntx, a.nty, a.nkz > 0 ? a.nkz : a.ntz);
  asm_kernel<JAC, RHS><<<grid, NTHREADS, smem, st>>>(a);
}
} // namespace

int jac_smem_bytes() { return asm_smem<true, false>(); }
int rhs_smem_bytes() { return asm_smem<false, true>(); }
int asm_smem_bytes() { return asm_smem<true, true>(); }
int asm_tile_levels() { return TK; }

void launch_jac(const AsmArgs& a, cudaStream_t st) { launch_asm_t<true, false>(a, st); }
void launch_rhs(const AsmArgs& a, cudaStream_t st) { launch_asm_t<false, true>(a, st); }
void launch_asm(const AsmArgs& a, cudaStream_t st) { launch_asm_t<true, true>(a, st); }
void launch_intcond_dot(const double* coefS, const double* x, long long ncell, double* partial, cudaStream_t st) {
  intcond_dot_kernel<<<1, 1024, 0, st>>>(coefS, x, ncell, partial);
}
void launch_forcing_surface(const Scalars& S, const uint64_t* sig, const double* fprof, double* frc, cudaStream_t st) {
  const int cells = S.nloc * S.mloc;
  forcing_surface_kernel<<<(cells + 255) / 256, 256, 0, st>>>(S, sig, fprof, frc);
}
void launch_diagb(const Scalars& S, const uint64_t* sig, const SpecialRows& sp, double* diagB, cudaStream_t st) {
  diagb_kernel<<<(int)((S.nown + 255) / 256), 256, 0, st>>>(S, sig, sp, diagB);
}
void launch_intcond_add(const double* dotval, double sign, double* y, int row, cudaStream_t st) { intcond_add_kernel<<<1, 1, 0, st>>>(dotval, sign, y, row); }
void launch_intcond_finish(const double* dotval, double sign, double corr, double* F, int row, int p1, int p2, cudaStream_t st) {
  intcond_finish_kernel<<<1, 1, 0, st>>>(dotval, sign, corr, F, row, p1, p2);
}

} // namespace thcm
