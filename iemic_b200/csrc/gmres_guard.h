// Host-side accuracy guard of the GMRES driver (capi.cu, gmres_b200_solve).
//
// GMRESSolver.H:213-246 accepts a cycle on the recurrence residual |s[i+1]| / ||b|| and only prints the explicit one.  On a
// singular Jacobian (THCM without fixed pressure points has the constant-pressure null vector) and a right-hand side at the
// rounding floor of F, b has a component outside the range of J; the least-squares coefficients y = R^-1 s then grow until
// eps ||H|| ||y|| exceeds ||b||, the recurrence keeps falling and the explicit residual is O(1): a "converged" solve returns a
// multiple of the null vector.  The reference's modified Gram-Schmidt loses orthogonality at that point and stagnates instead;
// CGS2 keeps the basis orthogonal and walks into the breakdown.  When the explicit residual contradicts the recurrence, the
// driver calls gmres_truncate: it estimates the attainable residual of every leading sub-problem,
//     est_j = hist_j + eps ||R_j||_F ||y_j|| / ||b||,        R_j y_j = s_j,
// and returns the earliest j whose estimate is within a factor two of the best one.
#pragma once
#include <cmath>
#include <vector>

namespace thcm {

// explicit residual that contradicts the recurrence residual (both relative to ||b||); `floor`: what rounding alone explains
inline bool gmres_lost_accuracy(double recurrence, double explicit_resid, double tol, double floor) {
  const double lim = std::fmax(std::fmax(10.0 * tol, 100.0 * recurrence), floor);
  return !(explicit_resid <= lim); // NaN counts as lost
}

// min(100 eps ||R||_F ||y|| / ||b||, 1e-8) of the full least-squares problem of this cycle (k columns): the size of the
// disagreement between recurrence and explicit residual that ordinary rounding produces
inline double gmres_rounding_floor(const double* R, int ldr, const double* s, int k, double normb) {
  const double eps = 2.220446049250313e-16;
  std::vector<double> y((size_t)k);
  for (int i = 0; i < k; i++) y[i] = s[i];
  double fro2 = 0.0;
  for (int i = k - 1; i >= 0; i--) {
    const double* col = R + (size_t)i * ldr;
    y[i] /= col[i];
    for (int q = 0; q < i; q++) y[q] -= col[q] * y[i];
    for (int q = 0; q <= i; q++) fro2 += col[q] * col[q];
  }
  double y2 = 0.0;
  for (int i = 0; i < k; i++) y2 += y[i] * y[i];
  const double f = 100.0 * eps * std::sqrt(fro2) * std::sqrt(y2) / normb;
  // capped: in the breakdown this guard exists for, ||y|| itself has exploded (1e13 ||b||) and would "explain" any residual
  return f >= 0.0 ? std::fmin(f, 1.0e-8) : 0.0; // NaN (zero pivot): no floor, the guard decides on the residuals alone
}

// R: rotated Hessenberg matrix, column j at R + j * ldr (upper triangular); s: rotated right-hand side; hist[j]: recurrence
// residual after j + 1 columns, relative to normb; k: columns of this cycle.  Returns the number of columns to keep and their
// coefficients in y.
inline int gmres_truncate(const double* R, int ldr, const double* s, const double* hist, int k, double normb, std::vector<double>& y) {
  const double eps = 2.220446049250313e-16;
  std::vector<double> est((size_t)k), yj((size_t)k);
  auto solve = [&](int j) { // leading j x j block
    for (int i = 0; i < j; i++) yj[i] = s[i];
    for (int i = j - 1; i >= 0; i--) {
      yj[i] /= R[(size_t)i * ldr + i];
      const double* col = R + (size_t)i * ldr;
      for (int q = 0; q < i; q++) yj[q] -= col[q] * yj[i];
    }
  };
  double fro2 = 0.0, best = INFINITY;
  for (int j = 1; j <= k; j++) {
    const double* col = R + (size_t)(j - 1) * ldr;
    for (int q = 0; q < j; q++) fro2 += col[q] * col[q];
    solve(j);
    double y2 = 0.0;
    for (int q = 0; q < j; q++) y2 += yj[q] * yj[q];
    est[j - 1] = hist[j - 1] + eps * std::sqrt(fro2) * std::sqrt(y2) / normb;
    if (!(est[j - 1] >= 0.0)) est[j - 1] = INFINITY; // NaN / Inf from a zero pivot
    if (est[j - 1] < best) best = est[j - 1];
  }
  if (!(best < INFINITY)) { y.clear(); return 0; } // no usable sub-problem (zero pivots throughout): keep the start vector
  int keep = k;
  for (int j = 1; j <= k; j++) if (est[j - 1] <= 2.0 * best) { keep = j; break; }
  solve(keep);
  y.assign(yj.begin(), yj.begin() + keep);
  return keep;
}

} // namespace thcm
