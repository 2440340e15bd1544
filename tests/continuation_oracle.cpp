// continuation_oracle.cpp -- TEST INFRASTRUCTURE: the reference's UNMODIFIED Continuation<Model> (src/continuation/
// Continuation.H, read in place from the reference tree) driving the CPU oracle (oracle/thcm_oracle.h) through the chain of
// continuations of the reference's integration test src/tests/intt_2dmoc.C:28-120 on the 3x6x6 2-D MOC box.  Writes the
// two end states (state1.bin, state2.bin) whose overturning streamfunction the reference asserts (psi_max = 14.7 Sv).
// Same stand-ins for Teuchos / Utils / JDQZ as tests/continuation_dropin.cpp (tests/compat).  Linear solves: the oracle's
// restatement of GMRESSolver.H with the point-block Jacobi of the current Jacobian, or (argv "lu") a dense LU with partial
// pivoting on the matrix regularised in its pressure null space, or (argv "cgs2") a host walk-through of the product's GMRES
// driver (capi.cu gmres_b200_solve: flexible right preconditioning, CGS2, acceptance on the recurrence residual, the accuracy
// guard of iemic_b200/csrc/gmres_guard.h; ORACLE_GUARD=0 switches the guard off and shows what it is for).
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>
#include "../oracle/thcm_oracle.h"
#include "ComplexVector.H"
#include "../iemic_b200/csrc/gmres_guard.h"

class HostVector { // the Epetra_Vector subset Continuation.H uses
 public:
  std::vector<double> v;
  HostVector() {}
  explicit HostVector(size_t n) : v(n, 0.0) {}
  int Update(double a, const HostVector& A, double b) { for (size_t i = 0; i < v.size(); i++) v[i] = a * A.v[i] + b * v[i]; return 0; }
  int Scale(double a) { for (double& x : v) x *= a; return 0; }
  int Dot(const HostVector& o, double* r) const { double s = 0; for (size_t i = 0; i < v.size(); i++) s += v[i] * o.v[i]; *r = s; return 0; }
  int Norm2(double* r) const { double s = 0; for (double x : v) s += x * x; *r = std::sqrt(s); return 0; }
  int NormInf(double* r) const { double s = 0; for (double x : v) s = std::max(s, std::fabs(x)); *r = s; return 0; }
  int PutScalar(double a) { std::fill(v.begin(), v.end(), a); return 0; }
  int Random() { unsigned long long s = 88172645463325252ull; for (double& x : v) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; x = (double)(s >> 11) / 4503599627370496.0 - 1.0; } return 0; }
  long long GlobalLength() const { return (long long)v.size(); }
};

class OracleModel {
 public:
  typedef std::shared_ptr<HostVector> VectorPtr;
  OracleModel(const thcm_oracle_config& c, const int* mask, const std::string& mode) : lu_(mode == "lu" || mode == "guardcheck"), cgs2_(mode == "cgs2") {
    o_ = thcm_oracle_create(&c, mask);
    if (!o_) throw std::runtime_error("thcm_oracle_create failed");
    n_ = (size_t)thcm_oracle_ndim(o_);
    state_ = std::make_shared<HostVector>(n_); rhs_ = std::make_shared<HostVector>(n_); sol_ = std::make_shared<HostVector>(n_);
    rp_.resize(n_ + 1); col_.resize((size_t)thcm_oracle_graph_nnz(o_)); val_.resize(col_.size()); diagB_.resize(n_);
    thcm_oracle_graph(o_, rp_.data(), col_.data());
  }
  ~OracleModel() { thcm_oracle_destroy(o_); }
  VectorPtr getState(char m = 'C') { return get(state_, m); }
  VectorPtr getRHS(char m = 'C') { return get(rhs_, m); }
  VectorPtr getSolution(char m = 'C') { return get(sol_, m); }
  void computeRHS() { thcm_oracle_eval_rhs(o_, state_->v.data(), rhs_->v.data()); }
  void computeJacobian() { thcm_oracle_eval_jac(o_, state_->v.data(), val_.data(), diagB_.data()); }
  void applyMatrix(const HostVector& x, HostVector& y) { thcm_oracle_spmv((int)n_, rp_.data(), col_.data(), val_.data(), x.v.data(), y.v.data()); }
  void solve(const VectorPtr& b) {
    std::fill(sol_->v.begin(), sol_->v.end(), 0.0);
    if (lu_) { solve_lu(b->v); return; }
    std::vector<double> binv = block_inverses();
    static const double tol = std::getenv("ORACLE_GMRES_TOL") ? std::atof(std::getenv("ORACLE_GMRES_TOL")) : 1e-10;
    static const bool guard = !(std::getenv("ORACLE_GUARD") && std::getenv("ORACLE_GUARD")[0] == '0');
    if (cgs2_) { solve_cgs2(b->v, binv, tol, guard); return; }
    std::vector<double> hist(20000, 0.0);
    int iters = 0;
    const int m = (int)std::min<size_t>(n_, 500);
    int rc = thcm_oracle_gmres((int)n_, rp_.data(), col_.data(), val_.data(), b->v.data(), sol_->v.data(), m, 4000, tol, 1, binv.data(), hist.data(), &iters);
    if (std::getenv("IEMIC_COMPAT_VERBOSE")) std::printf("OracleModel: GMRES rc = %d, i = %d\n", rc, iters);
  }
  double getPar(const std::string& name) { return thcm_oracle_get_par(o_, par2int(name)); }
  void setPar(const std::string& name, double v) { thcm_oracle_set_par(o_, par2int(name), v); }
  void preProcess() {}
  void postProcess() {}
  bool monitor() { return false; }
  std::string writeData(bool = false) { return ""; }
  void dumpBlocks() {}

 private:
  static int par2int(const std::string& s) { // THCM::par2int, src/ocean/THCM.C:1754-1813
    static const char* names[] = {"AL_T", "Rayleigh-Number", "Vertical Ekman-Number", "Horizontal Ekman-Number", "Rossby-Number", "MIXP", "RESC",
                                  "SPL1", "Salinity Homotopy", "Solar Forcing", "Horizontal Peclet-Number", "Vertical Peclet-Number", "P_VC", "LAMB",
                                  "Salinity Forcing", "Wind Forcing", "Temperature Forcing", "Nonlinear Factor", "Combined Forcing", "ARCL", "NLES",
                                  "IFRICB", "CONT", "Energy", "ALPC", "CMPR", "Flux Perturbation", "Salinity Perturbation", "MKAP", "SPL2"};
    for (int i = 0; i < 30; i++) if (s == names[i]) return i + 1;
    throw std::runtime_error("par2int: " + s);
  }
  VectorPtr get(const VectorPtr& v, char m) { return m == 'V' ? v : std::make_shared<HostVector>(*v); }
  std::vector<double> dense() const {
    std::vector<double> A(n_ * n_, 0.0);
    for (size_t r = 0; r < n_; r++) for (int p = rp_[r]; p < rp_[r + 1]; p++) A[r * n_ + col_[p]] += val_[p];
    return A;
  }
  std::vector<double> block_inverses() const {
    const size_t nb = n_ / 6;
    std::vector<double> out(nb * 36, 0.0);
    for (size_t b = 0; b < nb; b++) {
      double M[6][12];
      for (int i = 0; i < 6; i++) for (int j = 0; j < 12; j++) M[i][j] = (j - 6 == i) ? 1.0 : 0.0;
      for (int i = 0; i < 6; i++) { const size_t r = 6 * b + i; for (int p = rp_[r]; p < rp_[r + 1]; p++) if ((size_t)col_[p] / 6 == b) M[i][col_[p] % 6] = val_[p]; }
      bool ok = true;
      for (int k = 0; k < 6 && ok; k++) {
        int piv = k;
        for (int i = k + 1; i < 6; i++) if (std::fabs(M[i][k]) > std::fabs(M[piv][k])) piv = i;
        if (M[piv][k] == 0.0) { ok = false; break; }
        for (int j = 0; j < 12; j++) std::swap(M[k][j], M[piv][j]);
        const double d = 1.0 / M[k][k];
        for (int j = 0; j < 12; j++) M[k][j] *= d;
        for (int i = 0; i < 6; i++) if (i != k) { const double f = M[i][k]; for (int j = 0; j < 12; j++) M[i][j] -= f * M[k][j]; }
      }
      for (int i = 0; i < 6; i++) for (int j = 0; j < 6; j++) out[b * 36 + i * 6 + j] = ok ? M[i][j + 6] : (i == j ? 1.0 : 0.0);
    }
    return out;
  }
  void solve_lu(const std::vector<double>& b) { // dense LU with partial pivoting; tiny pivots (null space) are skipped
    std::vector<double> A = dense(), x = b;
    const size_t n = n_;
    std::vector<size_t> perm(n);
    double amax = 0.0;
    for (double a : A) amax = std::max(amax, std::fabs(a));
    std::vector<char> dead(n, 0);
    for (size_t k = 0; k < n; k++) {
      size_t piv = k;
      for (size_t i = k + 1; i < n; i++) if (std::fabs(A[i * n + k]) > std::fabs(A[piv * n + k])) piv = i;
      if (piv != k) { for (size_t j = 0; j < n; j++) std::swap(A[k * n + j], A[piv * n + j]); std::swap(x[k], x[piv]); }
      if (std::fabs(A[k * n + k]) < 1e-11 * amax) { dead[k] = 1; continue; }
      for (size_t i = k + 1; i < n; i++) {
        const double f = A[i * n + k] / A[k * n + k];
        if (f == 0.0) continue;
        A[i * n + k] = 0.0;
        for (size_t j = k + 1; j < n; j++) A[i * n + j] -= f * A[k * n + j];
        x[i] -= f * x[k];
      }
    }
    for (size_t kk = n; kk-- > 0;) {
      if (dead[kk]) { x[kk] = 0.0; continue; }
      double s = x[kk];
      for (size_t j = kk + 1; j < n; j++) s -= A[kk * n + j] * x[j];
      x[kk] = s / A[kk * n + kk];
    }
    sol_->v = x;
  }
  // the product's driver (iemic_b200/csrc/capi.cu, gmres_b200_solve) on the host: same order of operations, plain loops
  int solve_cgs2(const std::vector<double>& b, const std::vector<double>& binv, double tol, bool guard) {
    const size_t n = n_;
    const int m = (int)std::min<size_t>(n, 500), maxit = 4000, ldh = m + 1;
    std::vector<double>& x = sol_->v;
    auto dot = [&](const double* a, const double* c) { double t = 0; for (size_t i = 0; i < n; i++) t += a[i] * c[i]; return t; };
    auto amul = [&](const double* in, double* out) { thcm_oracle_spmv((int)n, rp_.data(), col_.data(), val_.data(), in, out); };
    auto prec = [&](const double* in, double* out) {
      for (size_t c = 0; c < n / 6; c++) for (int i = 0; i < 6; i++) { double t = 0; for (int j = 0; j < 6; j++) t += binv[c * 36 + i * 6 + j] * in[6 * c + j]; out[6 * c + i] = t; }
    };
    double normb = std::sqrt(dot(b.data(), b.data()));
    if (normb == 0.0) normb = 1.0;
    std::vector<double> r(n), w(n), tmp(n), V((size_t)(m + 1) * n), Z((size_t)m * n), H((size_t)ldh * m), cs(m), sn(m), s(m + 1), hist(m), y;
    auto residual = [&](const std::vector<double>& xx) { amul(xx.data(), r.data()); for (size_t i = 0; i < n; i++) r[i] = b[i] - r[i]; return std::sqrt(dot(r.data(), r.data())); };
    double beta = residual(x), resid = beta / normb;
    int iter = 0, rc = 1, trips = 0;
    double cycle_start = resid;
    if (resid <= tol) rc = 0;
    while (rc == 1 && iter <= maxit) {
      for (size_t i = 0; i < n; i++) V[i] = r[i] / beta;
      std::fill(s.begin(), s.end(), 0.0); s[0] = beta;
      int i, done = 0; bool conv = false;
      for (i = 0; i < m && iter <= maxit; i++, iter++) {
        double* vi = &V[(size_t)i * n];
        prec(vi, &Z[(size_t)i * n]); amul(&Z[(size_t)i * n], w.data());
        double* hc = &H[(size_t)i * ldh];
        std::vector<double> h1(i + 1), h2(i + 1);
        for (int k = 0; k <= i; k++) h1[k] = dot(&V[(size_t)k * n], w.data());
        for (int k = 0; k <= i; k++) for (size_t q = 0; q < n; q++) w[q] -= h1[k] * V[(size_t)k * n + q];
        for (int k = 0; k <= i; k++) h2[k] = dot(&V[(size_t)k * n], w.data());
        for (int k = 0; k <= i; k++) { for (size_t q = 0; q < n; q++) w[q] -= h2[k] * V[(size_t)k * n + q]; hc[k] = h1[k] + h2[k]; }
        double hn = std::sqrt(dot(w.data(), w.data()));
        for (size_t q = 0; q < n; q++) V[(size_t)(i + 1) * n + q] = w[q] / hn;
        for (int k = 0; k < i; k++) { const double t = cs[k] * hc[k] + sn[k] * hc[k + 1]; hc[k + 1] = -sn[k] * hc[k] + cs[k] * hc[k + 1]; hc[k] = t; }
        const double d = std::hypot(hc[i], hn); cs[i] = hc[i] / d; sn[i] = hn / d; hc[i] = d;
        s[i + 1] = -sn[i] * s[i]; s[i] = cs[i] * s[i];
        hist[i] = std::fabs(s[i + 1]) / normb; done = i + 1;
        if (hist[i] < tol) { conv = true; break; }
      }
      resid = hist[done - 1];
      auto candidate = [&](const std::vector<double>& yy) { tmp = x; for (size_t k = 0; k < yy.size(); k++) for (size_t q = 0; q < n; q++) tmp[q] += yy[k] * Z[k * n + q]; return residual(tmp) / normb; };
      y.assign(s.begin(), s.begin() + done);
      for (int a = done - 1; a >= 0; a--) { y[a] /= H[(size_t)a * ldh + a]; for (int q = 0; q < a; q++) y[q] -= H[(size_t)a * ldh + q] * y[a]; }
      double expl = candidate(y);
      if (guard && thcm::gmres_lost_accuracy(resid, expl, tol, 0.0) &&
          thcm::gmres_lost_accuracy(resid, expl, tol, thcm::gmres_rounding_floor(H.data(), ldh, s.data(), done, normb))) {
        // same walk as gmres_b200_solve: better of the full and the truncated update, one restart, then stop with code 2
        const std::vector<double> yfull = y;
        const double expl_full = expl;
        const int keep = thcm::gmres_truncate(H.data(), ldh, s.data(), hist.data(), done, normb, y);
        expl = candidate(y);                       // the truncated update is the trustworthy one, unless it is worse than the cycle's start
        if (!(expl <= cycle_start)) expl = (expl_full <= cycle_start) ? candidate(yfull) : candidate(std::vector<double>());
        x = tmp;
        if (std::getenv("IEMIC_COMPAT_VERBOSE")) std::printf("OracleModel: accuracy guard kept %d of %d columns, explicit residual %.3e\n", keep, done, expl);
        const bool progress = expl < 0.999 * cycle_start;
        if (expl <= tol) { rc = 0; break; }
        if (trips >= 1 || !progress || iter > maxit) { rc = 2; break; }
        trips++; cycle_start = expl; beta = expl * normb;
        continue;
      }
      trips = 0; cycle_start = expl;
      x = tmp; beta = expl * normb;
      if (conv) rc = 0;
    }
    if (std::getenv("IEMIC_COMPAT_VERBOSE")) std::printf("OracleModel: CGS2 GMRES rc = %d, i = %d\n", rc, iter);
    return rc;
  }
 public:
  // one linear solve at the rounding floor of F, with and without the accuracy guard: J of the current state (singular:
  // constant-pressure null vector), b = F(state) (noise with a component outside the range of J)
  void guardcheck() {
    computeRHS(); computeJacobian();
    const std::vector<double> b = rhs_->v, binv = block_inverses();
    double nb = 0; for (double v : b) nb += v * v; nb = std::sqrt(nb);
    for (int guard = 1; guard >= 0; guard--) {
      std::fill(sol_->v.begin(), sol_->v.end(), 0.0);
      const int rc = solve_cgs2(b, binv, 1e-10, guard != 0);
      double nx = 0; for (double v : sol_->v) nx += v * v; nx = std::sqrt(nx);
      std::printf("guardcheck: guard %d, rc %d, ||b|| = %.3e, ||x|| = %.3e\n", guard, rc, nb, nx);
    }
  }
 private:
  thcm_oracle* o_ = nullptr;
  size_t n_ = 0;
  bool lu_, cgs2_;
  VectorPtr state_, rhs_, sol_;
  std::vector<int> rp_, col_;
  std::vector<double> val_, diagB_;
};

#include "Continuation.H"
#include "continuation_cases.H"

static void write_state(OracleModel& m, const std::string& path) {
  auto x = m.getState('V');
  FILE* f = std::fopen(path.c_str(), "wb");
  if (!f) throw std::runtime_error("cannot write " + path);
  std::fwrite(x->v.data(), sizeof(double), x->v.size(), f);
  std::fclose(f);
}

int main(int argc, char** argv) {
  if (argc > 4 && std::string(argv[1]) == "xml") { // xml <ocean_params.xml> <continuation_params.xml> <mask dir> [mode]
    try {
      const std::string mode = argc > 5 ? argv[5] : "gmres";
      auto make = [&](const thcm_oracle_config& c, const std::vector<int>& mask) { return std::make_shared<OracleModel>(c, mask.data(), mode); };
      return run_ocean_from_files<Continuation<std::shared_ptr<OracleModel>>, thcm_oracle_config>(argv[2], argv[3], argv[4], make) == 0 ? 0 : 1;
    } catch (const std::exception& e) {
      std::printf("continuation on the oracle failed: %s\n", e.what());
      return 1;
    }
  }
  if (argc > 1 && std::string(argv[1]) == "cases") { print_cases<thcm_oracle_config>(); return 0; } // for tests/test_params.py
  if (argc > 2 && std::string(argv[1]) == "ocean") { // ocean <maskfile> [mode]: TEST(Ocean, Continuation), test_ocean.C:178-223
    try {
      const thcm_oracle_config c = test_ocean_config<thcm_oracle_config>();
      const std::vector<int> mask = read_mask_file(argv[2], (size_t)(c.n + 2) * (c.m + 2) * (c.l + 2));
      auto ocean = std::make_shared<OracleModel>(c, mask.data(), argc > 3 ? argv[3] : "gmres");
      return run_test_ocean_continuation<Continuation<std::shared_ptr<OracleModel>>>(ocean) == 0 ? 0 : 1;
    } catch (const std::exception& e) {
      std::printf("continuation on the oracle failed: %s\n", e.what());
      return 1;
    }
  }
  if (argc > 3 && std::string(argv[1]) == "reft") { // reft <maskfile> <state out> [mode]: src/tests/reft_ocean.C:31-80
    try {
      const thcm_oracle_config c = reft_ocean_config<thcm_oracle_config>();
      const std::vector<int> mask = read_mask_file(argv[2], (size_t)(c.n + 2) * (c.m + 2) * (c.l + 2));
      auto ocean = std::make_shared<OracleModel>(c, mask.data(), argc > 4 ? argv[4] : "gmres");
      const std::string out = argv[3];
      return run_reft_ocean_continuation<Continuation<std::shared_ptr<OracleModel>>>(ocean, [&] { write_state(*ocean, out); }) == 0 ? 0 : 1;
    } catch (const std::exception& e) {
      std::printf("continuation on the oracle failed: %s\n", e.what());
      return 1;
    }
  }
  const std::string outdir = argc > 1 ? argv[1] : ".";
  const std::string mode = argc > 2 ? argv[2] : "gmres";
  const int nlegs = argc > 3 ? std::atoi(argv[3]) : 7;
  const thcm_oracle_config c = moc2d_config<thcm_oracle_config>(); // test/2dmoc/ocean_params.xml
  const std::vector<int> mask = moc2d_mask(c.n, c.m, c.l);
  try {
    auto ocean = std::make_shared<OracleModel>(c, mask.data(), mode);
    moc2d_starting_parameters(ocean); // "Starting Parameters" of test/2dmoc/ocean_params.xml

    Teuchos::ParameterList pars = moc2d_continuation_params(); // test/2dmoc/continuation_params.xml

    struct Leg { const char* par; double dest, ds; const char* save; };
    const Leg legs[] = {{"Combined Forcing", 1.0, 0.1, nullptr}, {"CMPR", -0.2, -0.5, nullptr}, {"Salinity Forcing", 0.02, 0.5, nullptr},
                        {"CMPR", 0.0, 0.5, nullptr}, {"Salinity Forcing", 0.04, 0.5, "state1.bin"},
                        {"Salinity Forcing", 0.03, -0.5, nullptr}, {"Salinity Forcing", 0.04, -0.5, "state2.bin"}};
    int q = 0;
    for (const Leg& g : legs) {
      if (q++ >= nlegs) break;
      pars.set("continuation parameter", g.par);
      pars.set("destination 0", g.dest);
      pars.set("initial step size", g.ds);
      Continuation<std::shared_ptr<OracleModel>> leg(ocean, pars);
      const int st = leg.run();
      ocean->computeRHS();
      double nx, nf;
      ocean->getState('V')->Norm2(&nx); ocean->getRHS('V')->Norm2(&nf);
      std::printf("leg %s -> %g: status %d, value %.9f, ||x|| = %.6e, ||F|| = %.3e\n", g.par, g.dest, st, ocean->getPar(g.par), nx, nf);
      std::fflush(stdout);
      if (st != 0 || std::fabs(ocean->getPar(g.par) - g.dest) > 1e-5) return 1;
      if (g.save) write_state(*ocean, outdir + "/" + g.save);
    }
    if (mode == "guardcheck") ocean->guardcheck();
    std::printf("continuation on the oracle ok\n");
  } catch (const std::exception& e) {
    std::printf("continuation on the oracle failed: %s\n", e.what());
    return 1;
  }
  return 0;
}
