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

THCM_HD size_t cell_index(const Scalars& S, int fi, int fj, int fk) {
  return ((size_t)(fk - 1) * S.m + (fj - 1)) * S.n + (fi - 1);
}

THCM_HD Staged stage_cell(const double* __restrict__ x, const uint64_t* __restrict__ sig, const Scalars& S, int fi, int fj, int fk) {
  const int n = S.n, m = S.m, l = S.l;
  Staged o; o.u = 0.0; o.v = 0.0; o.w = 0.0; o.t = 0.0; o.s = 0.0;
  const bool gi = fi < 1 || fi > n, gj = fj < 1 || fj > m, gk = fk < 1 || fk > l;
  const int ng = (gi ? 1 : 0) + (gj ? 1 : 0) + (gk ? 1 : 0);
  if (ng > 1) return o;
  if (ng == 0) {
    const size_t c = cell_index(S, fi, fj, fk);
    const double* xc = x + 6 * c;
    const bool live = sbit(sig[c], SIG_LIVE);
    o.u = live ? xc[0] : 0.0; o.v = live ? xc[1] : 0.0;
    o.w = (fk == l) ? 0.0 : xc[2];
    o.t = xc[4]; o.s = xc[5];
    return o;
  }
  if (gk) { // usrc.F90:1066-1081: copies of the first / last layer taken before the land zeroing
    const int sk = (fk < 1) ? 1 : l;
    const double* xc = x + 6 * cell_index(S, fi, fj, sk);
    const bool edge = (fj == m) || (!S.periodic && fi == n);
    o.u = edge ? 0.0 : xc[0]; o.v = edge ? 0.0 : xc[1];
    o.w = 0.0;
    o.t = xc[4]; o.s = xc[5];
    return o;
  }
  if (gj) { // usrc.F90:1053-1064
    const int sj = (fj < 1) ? 1 : m;
    const double* xc = x + 6 * cell_index(S, fi, sj, fk);
    o.t = xc[4]; o.s = xc[5];
    return o;
  }
  // gi, usrc.F90:1029-1051
  if (S.periodic) {
    const int si = (fi < 1) ? n : 1;
    const size_t c = cell_index(S, si, fj, fk);
    const double* xc = x + 6 * c;
    if (fi < 1) {
      const bool live = sbit(sig[c], SIG_LIVEWRAP);
      o.u = live ? xc[0] : 0.0; o.v = live ? xc[1] : 0.0;
    }
    o.w = xc[2];
    o.t = xc[4]; o.s = xc[5];
  } else {
    const int si = (fi < 1) ? 1 : n;
    const double* xc = x + 6 * cell_index(S, si, fj, fk);
    o.t = xc[4]; o.s = xc[5];
  }
  return o;
}

} // namespace thcm
