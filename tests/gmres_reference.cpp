// gmres_reference.cpp -- TEST INFRASTRUCTURE: the reference's UNMODIFIED GMRESSolver<Model, VectorPointer>
// (src/gmressolver/GMRESSolver.H, compiled in place from the reference tree, never copied) run on plain host vectors, to pin
// oracle/gmres_oracle.cpp (the restatement the GPU path is checked against) to the reference's own code.
//
// usage: gmres_reference <problem.bin> <out.bin>
//   problem.bin: int32 n, nnz, m, maxit, prec; double tol; int32 rowptr[n+1], col[nnz]; double val[nnz], b[n], blockinv[6n] (if prec)
//   out.bin    : int32 rc, iters; double resid; double x[n]
// The model applies the CSR matrix in row order and, as preconditioner, the 6x6 block inverses (right, flexible: the
// defaults of the reference's constructor, GMRESSolver.H:12-30).
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <memory>
#include <stdexcept>
#include <vector>
#include "Utils.H" // tests/compat: Teuchos::ParameterList stand-in

// the LAPACK routines of the 'Q' minimiser (GMRESSolver.H:71-78), never reached with the default scheme 'B'
extern "C" void dgels_(char*, int*, int*, int*, double*, int*, double*, int*, double*, int*, int*) { std::abort(); }
extern "C" void dgesv_(int*, int*, double*, int*, int*, double*, int*, int*) { std::abort(); }

struct Vec { // the Vector concept of GMRESSolverDecl.H:12-16 (+ what solve() calls)
  std::vector<double> v;
  Vec() {}
  explicit Vec(size_t n) : v(n, 0.0) {}
  void update(double a, const Vec& A, double b) { for (size_t i = 0; i < v.size(); i++) v[i] = a * A.v[i] + b * v[i]; }
  double norm() const { double s = 0; for (double x : v) s += x * x; return std::sqrt(s); }
  double dot(const Vec& o) const { double s = 0; for (size_t i = 0; i < v.size(); i++) s += v[i] * o.v[i]; return s; }
  void scale(double a) { for (double& x : v) x *= a; }
  void zero() { for (double& x : v) x = 0.0; }
};

struct CsrModel {
  int n = 0, prec = 0;
  std::vector<int> rp, col;
  std::vector<double> val, binv;
  void applyMatrix(const Vec& x, Vec& y) {
    for (int r = 0; r < n; r++) { double s = 0; for (int p = rp[r]; p < rp[r + 1]; p++) s += val[p] * x.v[col[p]]; y.v[r] = s; }
  }
  void applyPrecon(const Vec& x, Vec& y) {
    if (!prec) { y = x; return; }
    for (int c = 0; c < n / 6; c++) for (int i = 0; i < 6; i++) {
      double s = 0;
      for (int j = 0; j < 6; j++) s += binv[(size_t)c * 36 + i * 6 + j] * x.v[6 * c + j];
      y.v[6 * c + i] = s;
    }
  }
};

#include "GMRESSolver.H"

template <typename T> static void rd(FILE* f, T* p, size_t k) { if (std::fread(p, sizeof(T), k, f) != k) throw std::runtime_error("short read"); }

int main(int argc, char** argv) {
  if (argc < 3) return 2;
  try {
    FILE* f = std::fopen(argv[1], "rb");
    if (!f) throw std::runtime_error("cannot open problem");
    int hdr[5]; double tol;
    rd(f, hdr, 5); rd(f, &tol, 1);
    const int n = hdr[0], nnz = hdr[1], m = hdr[2], maxit = hdr[3], prec = hdr[4];
    CsrModel model; model.n = n; model.prec = prec;
    model.rp.resize(n + 1); model.col.resize(nnz); model.val.resize(nnz);
    auto b = std::make_shared<Vec>(n), x = std::make_shared<Vec>(n);
    rd(f, model.rp.data(), n + 1); rd(f, model.col.data(), nnz); rd(f, model.val.data(), nnz); rd(f, b->v.data(), n);
    if (prec) { model.binv.resize((size_t)6 * n); rd(f, model.binv.data(), (size_t)6 * n); }
    std::fclose(f);
    GMRESSolver<CsrModel, std::shared_ptr<Vec>> solver(model, x, b);
    solver.setSolution(x); solver.setRHS(b);
    auto pars = std::make_shared<Teuchos::ParameterList>("GMRES");
    pars->set("GMRES tolerance", tol); pars->set("GMRES iterations", maxit); pars->set("GMRES restart", m);
    pars->set("GMRES preconditioning", prec != 0); pars->set("GMRES verbosity", 0);
    // flexible_ defaults to true and Update() then reads Z, which is only filled when a preconditioner is applied (:158-163)
    pars->set("GMRES flexible", prec != 0);
    solver.setParameters(pars);
    const int rc = solver.solve();
    const int it = solver.getNumIters();
    const double res = solver.residual();
    FILE* g = std::fopen(argv[2], "wb");
    if (!g) throw std::runtime_error("cannot write result");
    std::fwrite(&rc, sizeof(int), 1, g); std::fwrite(&it, sizeof(int), 1, g); std::fwrite(&res, sizeof(double), 1, g);
    std::fwrite(x->v.data(), sizeof(double), n, g);
    std::fclose(g);
  } catch (const std::exception& e) {
    std::printf("gmres_reference failed: %s\n", e.what());
    return 1;
  }
  return 0;
}
