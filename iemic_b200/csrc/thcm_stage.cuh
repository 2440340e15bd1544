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

// `raw` receives the unmodified state of the cell the ghost rule reads from (the cell itself when it is inside the
// domain, the periodic image for an x-ghost): what matAvec multiplies the assembled row with (matetc.F90:147-166).
template <class Ld>
THCM_HD Staged stage_cell_t(const Ld& ld, const uint64_t* __restrict__ sig, const Scalars& S, int fi, int fj, int fk, X6& raw) {
  const int n = S.n, m = S.m, l = S.l;
  Staged o; o.u = 0.0; o.v = 0.0; o.w = 0.0; o.t = 0.0; o.s = 0.0;
  raw.u = 0.0; raw.v = 0.0; raw.w = 0.0; raw.p = 0.0; raw.t = 0.0; raw.s = 0.0;
  const bool gi = fi < 1 || fi > n, gj = fj < 1 || fj > m, gk = fk < 1 || fk > l;
  const int ng = (gi ? 1 : 0) + (gj ? 1 : 0) + (gk ? 1 : 0);
  if (ng > 1) return o;
  if (ng == 0) {
    const size_t c = cell_index(S, fi, fj, fk);
    const X6 xc = ld(c);
    raw = xc;
    const bool live = sbit(sig[c], SIG_LIVE);
    o.u = live ? xc.u : 0.0; o.v = live ? xc.v : 0.0;
    o.w = (fk == l) ? 0.0 : xc.w;
    o.t = xc.t; o.s = xc.s;
    return o;
  }
  if (gk) { // usrc.F90:1066-1081: copies of the first / last layer taken before the land zeroing
    const int sk = (fk < 1) ? 1 : l;
    const X6 xc = ld(cell_index(S, fi, fj, sk));
    const bool edge = (fj == m) || (!S.periodic && fi == n);
    o.u = edge ? 0.0 : xc.u; o.v = edge ? 0.0 : xc.v;
    o.w = 0.0;
    o.t = xc.t; o.s = xc.s;
    return o;
  }
  if (gj) { // usrc.F90:1053-1064
    const int sj = (fj < 1) ? 1 : m;
    const X6 xc = ld(cell_index(S, fi, sj, fk));
    o.t = xc.t; o.s = xc.s;
    return o;
  }
  // gi, usrc.F90:1029-1051
  if (S.periodic) {
    const int si = (fi < 1) ? n : 1;
    const size_t c = cell_index(S, si, fj, fk);
    const X6 xc = ld(c);
    raw = xc;
    if (fi < 1) {
      const bool live = sbit(sig[c], SIG_LIVEWRAP);
      o.u = live ? xc.u : 0.0; o.v = live ? xc.v : 0.0;
    }
    o.w = xc.w;
    o.t = xc.t; o.s = xc.s;
  } else {
    const int si = (fi < 1) ? 1 : n;
    const X6 xc = ld(cell_index(S, si, fj, fk));
    o.t = xc.t; o.s = xc.s;
  }
  return o;
}

THCM_HD Staged stage_cell(const double* __restrict__ x, const uint64_t* __restrict__ sig, const Scalars& S, int fi, int fj, int fk) {
  X6 raw;
  return stage_cell_t(LdPlain{x}, sig, S, fi, fj, fk, raw);
}

} // namespace thcm
