// gmres_givens.cuh -- Arnoldi bookkeeping of one step on the device (one thread): Hessenberg column, plane rotations
// (GMRESSolver.H:190-200, 258-289) and the residual estimate; shared by gmres.cu and gmres_cycle.cu.
#pragma once
#include "kernels.h"

namespace thcm {
// GMRESSolver.H:258-289
__device__ __forceinline__ void gen_rot(double& dx, double& dy, double& cs, double& sn) {
  if (dy == 0.0) { cs = 1.0; sn = 0.0; }
  else if (fabs(dy) > fabs(dx)) { double temp = dx / dy; sn = 1.0 / sqrt(1.0 + temp * temp); cs = temp * sn; }
  else { double temp = dy / dx; cs = 1.0 / sqrt(1.0 + temp * temp); sn = temp * cs; }
}
__device__ __forceinline__ void app_rot(double& dx, double& dy, double cs, double sn) {
  double temp = cs * dx + sn * dy;
  dy = -sn * dx + cs * dy;
  dx = temp;
}
__device__ __forceinline__ void givens_step(const GivensArgs& g) { // one thread
  const int j = g.j;
  double* Hc = g.H + (size_t)j * g.ldh;
  for (int k = 0; k <= j; k++) Hc[k] = g.h1[k] + (g.h2 ? g.h2[k] : 0.0);
  double nrm;
  if (g.pyth) { // ||w - V h2||^2 = ||w||^2 - |h2|^2 for orthonormal V: h2[j+1] carries <w,w> from the same reduction
    double hh = 0.0;
    for (int k = 0; k <= j; k++) hh = fma(g.h2[k], g.h2[k], hh);
    const double ww = g.h2[j + 1], d2 = ww - hh;
    // (happy) breakdown: the new direction lies in span(V) to working precision.  The difference is then pure cancellation
    // noise; H[j+1][j] = 0 ends the Krylov space here and the normalisation that rides on the next kernel divides by 1, not 0.
    nrm = (d2 > 1.0e-28 * ww) ? sqrt(d2) : 0.0;
  } else nrm = g.sq ? sqrt(*g.nrm2) : *g.nrm2;
  Hc[j + 1] = nrm;
  *g.hnext = (nrm > 0.0) ? nrm : 1.0;
  // the running pair stays in registers: going through Hc[k + 1] in memory makes every rotation wait for a store -> load
  // round trip (the compiler cannot prove that Hc does not alias cs / sn)
  double hx = Hc[0];
  for (int k = 0; k < j; k++) {
    double hy = Hc[k + 1];
    app_rot(hx, hy, g.cs[k], g.sn[k]);
    Hc[k] = hx;
    hx = hy;
  }
  double hy = nrm, c, s;
  gen_rot(hx, hy, c, s);
  app_rot(hx, hy, c, s);
  Hc[j] = hx; Hc[j + 1] = hy;
  g.cs[j] = c; g.sn[j] = s;
  app_rot(g.s[j], g.s[j + 1], g.cs[j], g.sn[j]);
  g.hist[j] = fabs(g.s[j + 1]) / *g.bnorm;
}
} // namespace thcm
