// thcm_stage.cuh -- the ghost-cell and land rules of usol (src/ocean/usrc.F90:997-1104) as a pure function.
//
// stage_cell() returns the value the reference's ghosted fields u,v,w,t,s hold at Fortran index
// (fi,fj,fk), fi in 0..n+1, fj in 0..m+1, fk in 0..l+1, computed directly from the interleaved state
// vector.  The kernels call it while filling their shared-memory tile, so the tile holds exactly what
// nlin_rhs/nlin_jac/vmix_fun see.  Quirks that are reproduced on purpose:
//   * periodic: u(0,j,k)=u(N,j,k) is copied BEFORE u(:,M,:)=0 and before the land zeroing (:1031-1032),
//     so it has its own liveness bit (SIG_LIVEWRAP);
//   * w(i,j,l)=0 only for i=1..n (:1072), the periodic copies w(0,j,l), w(N+1,j,l) keep the raw value;
//   * ghosts with two out-of-range indices are never assigned and stay 0.
#pragma once
#include "thcm_rows.cuh"

namespace thcm {

struct Staged { double u, v, w, t, s; };

// Storage index of global cell (fi,fj,fk), 1<=fi<=n etc.  Owned cells come first (i fastest, then j, then k:
// the local ordering of TRIOS::Domain::CreateMap, TRIOS_Domain.C:378-436), then per level a ring of
// 2*(nloc+2)+2*mloc halo cells: south row, north row, west column, east column.
THCM_HD int ring_size(const Scalars& S) { return 2 * (S.nloc + 2) + 2 * S.mloc; }
THCM_HD size_t cell_index(const Scalars& S, int fi, int fj, int fk) {
  int li = fi - S.io0, lj = fj - S.jo0;
  if (li < -1) li += S.n; else if (li > S.nloc) li -= S.n;   // periodic seam between two ranks
  if (li >= 0 && li < S.nloc && lj >= 0 && lj < S.mloc) return ((size_t)(fk - 1) * S.mloc + lj) * S.nloc + li;
  int r;
  if (lj < 0) r = li + 1;
  else if (lj >= S.mloc) r = (S.nloc + 2) + li + 1;
  else if (li < 0) r = 2 * (S.nloc + 2) + lj;
  else r = 2 * (S.nloc + 2) + S.mloc + lj;
  return (size_t)S.nown + (size_t)(fk - 1) * ring_size(S) + r;
}

// One cell of the interleaved state vector.
struct X6 { double u, v, w, p, t, s; };
struct LdPlain {
  const double* x;
  THCM_HD X6 operator()(size_t c) const { const double* q = x + 6 * c; X6 o; o.u = q[0]; o.v = q[1]; o.w = q[2]; o.p = q[3]; o.t = q[4]; o.s = q[5]; return o; }
};
#ifdef __CUDACC__
struct LdVec { // three 16-byte read-only loads per cell (the cell stride is 48 B, so every cell is 16-byte aligned)
  const double* x;
  __device__ __forceinline__ X6 operator()(size_t c) const {
    const double2* q = reinterpret_cast<const double2*>(x + 6 * c);
    double2 a = __ldg(q), b = __ldg(q + 1), d = __ldg(q + 2);
    X6 o; o.u = a.x; o.v = a.y; o.w = b.x; o.p = b.y; o.t = d.x; o.s = d.y; return o;
  }
};
#endif

// The ghost rule is split in two so that a kernel can issue the loads of several cells before consuming any of them:
// stage_plan() says which stored cell a (possibly ghost) position reads and by which rule, stage_apply() turns the loaded
// cell and its signature into the staged values.
enum { ST_ZERO = 0, ST_IN = 1, ST_GK = 2, ST_GJ = 3, ST_GI_W = 4, ST_GI_E = 5, ST_GI_WALL = 6 };
struct StagePlan { size_t c; int rule; bool edge, top; };
THCM_HD StagePlan stage_plan(const Scalars& S, int fi, int fj, int fk) {
  const int n = S.n, m = S.m, l = S.l;
  StagePlan p; p.c = 0; p.rule = ST_ZERO; p.edge = false; p.top = false;
  const bool gi = fi < 1 || fi > n, gj = fj < 1 || fj > m, gk = fk < 1 || fk > l;
  const int ng = (gi ? 1 : 0) + (gj ? 1 : 0) + (gk ? 1 : 0);
  if (ng > 1) return p;
  if (ng == 0) { p.rule = ST_IN; p.c = cell_index(S, fi, fj, fk); p.top = (fk == l); return p; }
  if (gk) { // usrc.F90:1066-1081: copies of the first / last layer taken before the land zeroing
    p.rule = ST_GK; p.c = cell_index(S, fi, fj, (fk < 1) ? 1 : l);
    p.edge = (fj == m) || (!S.periodic && fi == n);
    return p;
  }
  if (gj) { p.rule = ST_GJ; p.c = cell_index(S, fi, (fj < 1) ? 1 : m, fk); return p; } // usrc.F90:1053-1064
  // gi, usrc.F90:1029-1051
  if (S.periodic) { p.rule = (fi < 1) ? ST_GI_W : ST_GI_E; p.c = cell_index(S, (fi < 1) ? n : 1, fj, fk); }
  else { p.rule = ST_GI_WALL; p.c = cell_index(S, (fi < 1) ? 1 : n, fj, fk); }
  return p;
}
// `raw` receives the unmodified state of the cell the ghost rule reads from (the cell itself when it is inside the
// domain, the periodic image for an x-ghost): what matAvec multiplies the assembled row with (matetc.F90:147-166).
THCM_HD Staged stage_apply(const StagePlan& p, const X6& xc, uint64_t sg, X6& raw) {
  Staged o; o.u = 0.0; o.v = 0.0; o.w = 0.0; o.t = 0.0; o.s = 0.0;
  raw.u = 0.0; raw.v = 0.0; raw.w = 0.0; raw.p = 0.0; raw.t = 0.0; raw.s = 0.0;
  switch (p.rule) {
    case ST_IN: {
      raw = xc;
      const bool live = sbit(sg, SIG_LIVE);
      o.u = live ? xc.u : 0.0; o.v = live ? xc.v : 0.0;
      o.w = p.top ? 0.0 : xc.w;
      o.t = xc.t; o.s = xc.s;
    } break;
    case ST_GK:
      o.u = p.edge ? 0.0 : xc.u; o.v = p.edge ? 0.0 : xc.v;
      o.t = xc.t; o.s = xc.s;
      break;
    case ST_GJ: case ST_GI_WALL:
      o.t = xc.t; o.s = xc.s;
      break;
    case ST_GI_W: {
      raw = xc;
      const bool live = sbit(sg, SIG_LIVEWRAP);
      o.u = live ? xc.u : 0.0; o.v = live ? xc.v : 0.0;
      o.w = xc.w; o.t = xc.t; o.s = xc.s;
    } break;
    case ST_GI_E:
      raw = xc;
      o.w = xc.w; o.t = xc.t; o.s = xc.s;
      break;
    default: break;
  }
  return o;
}
template <class Ld>
THCM_HD Staged stage_cell_t(const Ld& ld, const uint64_t* __restrict__ sig, const Scalars& S, int fi, int fj, int fk, X6& raw) {
  const StagePlan p = stage_plan(S, fi, fj, fk);
  if (p.rule == ST_ZERO) { X6 z; z.u = 0.0; z.v = 0.0; z.w = 0.0; z.p = 0.0; z.t = 0.0; z.s = 0.0; return stage_apply(p, z, 0ull, raw); }
  return stage_apply(p, ld(p.c), sig[p.c], raw);
}

THCM_HD Staged stage_cell(const double* __restrict__ x, const uint64_t* __restrict__ sig, const Scalars& S, int fi, int fj, int fk) {
  X6 raw;
  return stage_cell_t(LdPlain{x}, sig, S, fi, fj, fk, raw);
}

} // namespace thcm
