// gmres_oracle.cpp -- CPU restatement of src/gmressolver/GMRESSolver.H (TEST INFRASTRUCTURE ONLY).
//
// Follows GMRESSolver<Model,VectorPointer>::solve() (GMRESSolver.H:81-255) statement by statement
// for plain std::vector<double> vectors: single-pass modified Gram-Schmidt (:177-181), Givens
// rotations (:190-200, :258-289), back substitution (:400-410), Update (:292-313).
// Vector concept (GMRESSolverDecl.H:7-16): dot/norm are sequential sums, update(a,A,b) is
// this = a*A + b*this.
//
// PARITY: pinned against the reference's own code.  No reference test exercises src/gmressolver, but the
// header compiles here by itself: tests/gmres_reference.cpp instantiates the UNMODIFIED GMRESSolver.H (in place, from the
// reference tree) on plain host vectors and tests/test_gmres_reference_pin.py requires return code, iteration count, final
// recurrence residual and solution of this restatement to equal the reference's bit for bit (with and without the point-block
// preconditioner, one cycle and restarted, stopped by tolerance and by the iteration limit).
#include "thcm_oracle.h"
#include <cmath>
#include <vector>
#include <cstdio>

namespace {
typedef std::vector<double> Vec;
double dot(const Vec& a, const Vec& b) { double s = 0.0; for (size_t i = 0; i < a.size(); i++) s += a[i] * b[i]; return s; }
double norm(const Vec& a) { return std::sqrt(dot(a, a)); }
void update(Vec& self, double a, const Vec& A, double b) { for (size_t i = 0; i < self.size(); i++) self[i] = a * A[i] + b * self[i]; }
void scale(Vec& self, double a) { for (size_t i = 0; i < self.size(); i++) self[i] = a * self[i]; }

struct Model {
  int nrows; const int* rowptr; const int* col; const double* val;
  int prec; const double* blockinv;
  void applyMatrix(const Vec& v, Vec& out) const { thcm_oracle_spmv(nrows, rowptr, col, val, v.data(), out.data()); }
  void applyPrecon(const Vec& v, Vec& out) const { // point-block Jacobi: out_b = Binv_b * v_b
    for (int b = 0; b < nrows / 6; b++)
      for (int r = 0; r < 6; r++) {
        double s = 0.0;
        for (int c = 0; c < 6; c++) s += blockinv[(size_t)b * 36 + r * 6 + c] * v[(size_t)b * 6 + c];
        out[(size_t)b * 6 + r] = s;
      }
  }
};
// GMRESSolver.H:258-289
void GeneratePlaneRotation(double& dx, double& dy, double& cs, double& sn) {
  if (dy == 0.0) { cs = 1.0; sn = 0.0; }
  else if (std::fabs(dy) > std::fabs(dx)) { double temp = dx / dy; sn = 1.0 / std::sqrt(1.0 + temp * temp); cs = temp * sn; }
  else { double temp = dy / dx; cs = 1.0 / std::sqrt(1.0 + temp * temp); sn = temp * cs; }
}
void ApplyPlaneRotation(double& dx, double& dy, double& cs, double& sn) {
  double temp = cs * dx + sn * dy;
  dy = -sn * dx + cs * dy;
  dx = temp;
}
} // namespace

extern "C" int thcm_oracle_gmres(int nrows, const int* rowptr, const int* col, const double* val,
                                 const double* b_in, double* x_io, int m_, int maxit_, double tol_,
                                 int prec, const double* blockinv, double* hist, int* iters) {
  Model model{nrows, rowptr, col, val, prec, blockinv};
  bool prec_ = prec != 0, flexible_ = true; // right-preconditioned FGMRES (GMRESSolver.H:24,29-30)
  Vec x_(x_io, x_io + nrows), b_(b_in, b_in + nrows);
  int i, k, iter_ = 0;
  Vec s(m_ + 1, 0.0), cs(m_ + 1, 0.0), sn(m_ + 1, 0.0);
  std::vector<Vec> H(m_ + 1, Vec(m_, 0.0));
  Vec tmp(x_), r(x_), w(x_);
  double normb = norm(b_);
  model.applyMatrix(x_, r);
  update(r, 1.0, b_, -1.0);
  double beta = norm(r);
  if (normb == 0.0) normb = 1;
  double resid_ = beta / normb;
  int nh = 0;
  if (resid_ <= tol_) { *iters = 0; return 0; }
  std::vector<Vec> Z(m_ + 1), V(m_ + 1);
  int spaceSize = 0;
  int ret = 1;
  while (iter_ <= maxit_) {
    beta = norm(r);
    scale(r, 1.0 / beta);
    V[0] = r;
    s.assign(m_ + 1, 0.0);
    s[0] = beta;
    for (i = 0; i < m_ && iter_ <= maxit_; i++, iter_++) {
      if (prec_ && flexible_) { model.applyPrecon(V[i], tmp); Z[i] = tmp; model.applyMatrix(tmp, w); }
      else model.applyMatrix(V[i], w);
      for (k = 0; k <= i; k++) { H[k][i] = dot(w, V[k]); update(w, -H[k][i], V[k], 1.0); }
      H[i + 1][i] = norm(w);
      scale(w, 1.0 / H[i + 1][i]);
      V[i + 1] = w;
      spaceSize = i;
      for (k = 0; k < i; k++) ApplyPlaneRotation(H[k][i], H[k + 1][i], cs[k], sn[k]);
      GeneratePlaneRotation(H[i][i], H[i + 1][i], cs[i], sn[i]);
      ApplyPlaneRotation(H[i][i], H[i + 1][i], cs[i], sn[i]);
      ApplyPlaneRotation(s[i], s[i + 1], cs[i], sn[i]);
      resid_ = std::fabs(s[i + 1]) / normb;
      if (hist) hist[nh++] = resid_;
      if (resid_ < tol_) break;
    }
    // Update (GMRESSolver.H:292-313) with backSolve (:400-410)
    Vec y_(s);
    for (int ii = spaceSize; ii >= 0; ii--) { y_[ii] /= H[ii][ii]; for (int jj = ii - 1; jj >= 0; jj--) y_[jj] -= H[jj][ii] * y_[ii]; }
    for (int j = 0; j <= spaceSize; j++) update(x_, y_[j], (prec_ && flexible_) ? Z[j] : V[j], 1.0);
    model.applyMatrix(x_, r);
    update(r, 1.0, b_, -1.0);
    beta = norm(r);
    if (resid_ < tol_) { ret = 0; break; }
  }
  for (int q = 0; q < nrows; q++) x_io[q] = x_[q];
  *iters = iter_;
  if (hist) hist[nh] = beta / normb; // explicit residual after the last cycle (GMRESSolver.H:241-242)
  return ret;
}
