// fortran_abi.cpp -- seam S1 (include/thcm_fortran_abi.h): the Fortran symbols THCM.C binds (src/ocean/THCM.C:47-174),
// served by the B200 path through the native C ABI (include/iemic_b200.h).  Host-buffer adaptor, one process-global
// model like the Fortran module state.  No arithmetic happens here: rhs_ / matrix_ run the CUDA kernels and this file
// only converts between the reference's calling conventions and the device-resident static CSR.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <dlfcn.h>
#include <algorithm>
#include <vector>
#include "../../include/iemic_b200.h"
#include "../../include/thcm_fortran_abi.h"

namespace {

struct Global { // what m_global::initialize keeps (global.F90:66-170)
  bool set = false;
  double hdim = 0, qz = 0;
  int tres = 1, sres = 1, iza = 2, ite = 1, its = 1, coupled_T = 0, coupled_S = 0, forcing_type = 0;
} G;

struct Model {
  thcm_b200_t* h = nullptr;
  thcm_b200_config cfg;
  std::vector<int> landm;
  int vmix_fix = 0;
  // caller-owned CSR (m_mat::set_pointers, mat.F90:73-100)
  int *begA = nullptr, *jcoA = nullptr, *begF = nullptr, *jcoF = nullptr;
  double *coA = nullptr, *coB = nullptr, *coF = nullptr;
  int cap_rows = 0, cap_nnz = 0;
  // static graph of the device path and, per row, its entries in fillcolA's order (slot-major, then unknown)
  std::vector<int> rowptr, col, order;
  std::vector<double> val, diagB, F;
  bool have_B = false;
} M;

[[noreturn]] void die(const char* where) {
  std::fprintf(stderr, "thcm_b200 (Fortran ABI) %s: %s\n", where, iemic_b200_last_error());
  typedef void (*throw_fn)(const char*);
  if (throw_fn f = (throw_fn)dlsym(RTLD_DEFAULT, "thcm_throw_error_")) f(where); // THCM.C:2741, when the host program has it
  std::abort();
}
void need_model(const char* where) {
  if (!M.h) { std::fprintf(stderr, "thcm_b200 (Fortran ABI) %s called before init_\n", where); std::abort(); }
}

// fillcolA's order inside a row (assemble.F90:57-139): slot kk = 1..27, then unknown jj = 1..6.  The device CSR is sorted
// by column; the slot of an entry follows from the neighbour offset (shift, assemble.F90:142-179).
void build_order() {
  const int n = M.cfg.n, m = M.cfg.m;
  const size_t nrows = M.rowptr.size() - 1;
  M.order.resize(M.col.size());
  std::vector<std::pair<int, int>> key;
  for (size_t r = 0; r < nrows; r++) {
    const int c = (int)(r / 6), i = c % n, j = (c / n) % m, k = c / (n * m);
    key.clear();
    for (int p = M.rowptr[r]; p < M.rowptr[r + 1]; p++) {
      const int c2 = M.col[p] / 6, jj = M.col[p] % 6, i2 = c2 % n, j2 = (c2 / n) % m, k2 = c2 / (n * m);
      int di = i2 - i;
      if (di > 1) di -= n; else if (di < -1) di += n; // periodic image
      const int dj = j2 - j, dk = k2 - k;
      if (di < -1 || di > 1 || dj < -1 || dj > 1 || dk < -1 || dk > 1) { key.push_back({1000, p}); continue; } // dense integral row
      const int kk = 1 + (dj + 1) + 3 * (di + 1) + (dk == 0 ? 0 : (dk < 0 ? 9 : 18));
      key.push_back({kk * 8 + jj, p});
    }
    std::sort(key.begin(), key.end());
    for (size_t q = 0; q < key.size(); q++) M.order[M.rowptr[r] + q] = key[q].second;
  }
}

void create_model() {
  if (M.h) { thcm_b200_destroy(M.h); M.h = nullptr; }
  if (thcm_b200_create(&M.cfg, M.landm.data(), &M.h) != 0) die("init_");
  int64_t nr = 0, nv = 0, nnz = 0, ng = 0;
  if (thcm_b200_sizes(M.h, &nr, &nv, &nnz, &ng) != 0) die("init_");
  M.rowptr.resize(nr + 1); M.col.resize(nnz); M.val.resize(nnz); M.diagB.resize(nr); M.F.resize(nr);
  if (thcm_b200_graph_host(M.h, M.rowptr.data(), M.col.data()) != 0) die("init_");
  build_order();
  M.have_B = false;
}

void run_matrix(const double* un) {
  if (thcm_b200_jacobian_host(M.h, un, M.val.data(), M.diagB.data()) != 0) die("matrix_");
  M.have_B = true;
}

} // namespace

extern "C" {
#define EXPORT __attribute__((visibility("default")))

EXPORT void __m_global_MOD_initialize(int*, int*, int*, double*, double*, double*, double*, double* hdim, double* qz, int*,
                                      int*, int*, int* rd_mask, int* TRES, int* SRES, int* iza, int* ite, int* its, int*,
                                      int* coupled_T, int* coupled_S, int* forcing_type, const char*, const char*, const char*,
                                      const char*, const char*) {
  G.set = true; G.hdim = *hdim; G.qz = *qz; G.tres = *TRES; G.sres = *SRES; G.iza = *iza; G.ite = *ite; G.its = *its;
  G.coupled_T = *coupled_T; G.coupled_S = *coupled_S; G.forcing_type = *forcing_type;
  (void)rd_mask; // mask files are not read here: the mask arrives through init_ / set_landmask_
}
EXPORT void __m_global_MOD_finalize(void) { G = Global(); }

EXPORT void init_(int* n, int* m, int* l, int*, double* xmin, double* xmax, double* ymin, double* ymax, double* alphaT,
                  double* alphaS, int* ih, int* vmix, int* tap, int* rho_mixing, int* coriolis_on, int* periodic, int* landm,
                  double*, double*, double*, double*, double*) {
  if (!G.set) { std::fprintf(stderr, "thcm_b200 (Fortran ABI) init_ called before m_global::initialize\n"); std::abort(); }
  if (G.iza != 2 || G.ite != 1 || G.its != 1) {
    std::fprintf(stderr, "thcm_b200 (Fortran ABI) init_: only the idealised forcing (iza=2, ite=1, its=1) is implemented, got %d %d %d\n",
                 G.iza, G.ite, G.its);
    std::abort();
  }
  std::memset(&M.cfg, 0, sizeof(M.cfg));
  M.cfg.n = *n; M.cfg.m = *m; M.cfg.l = *l;
  M.cfg.xmin = *xmin; M.cfg.xmax = *xmax; M.cfg.ymin = *ymin; M.cfg.ymax = *ymax;
  M.cfg.periodic = *periodic; M.cfg.hdim = G.hdim; M.cfg.qz = G.qz; M.cfg.alphaT = *alphaT; M.cfg.alphaS = *alphaS;
  M.cfg.ih = *ih; M.cfg.vmix = *vmix; M.cfg.tap = *tap; M.cfg.rho_mixing = *rho_mixing; M.cfg.coriolis_on = *coriolis_on;
  M.cfg.forcing_type = G.forcing_type; M.cfg.tres = G.tres; M.cfg.sres = G.sres; M.cfg.iza = G.iza; M.cfg.ite = G.ite; M.cfg.its = G.its;
  M.cfg.coupled_T = G.coupled_T; M.cfg.coupled_S = G.coupled_S;
  M.cfg.int_sign = 1; M.cfg.intcond_i = -2; M.cfg.intcond_j = -1; // the integral row is C++ glue (THCM::intcond_S), not Fortran
  M.cfg.fix_pressure_points = 0;                                   // likewise (THCM::fixPressurePoints)
  M.landm.assign(landm, landm + (size_t)(*n + 2) * (*m + 2) * (*l + 2));
  create_model();
}
EXPORT void finalize_(void) {
  if (M.h) thcm_b200_destroy(M.h);
  M = Model();
}

EXPORT void setparcs_(int* param, double* value) { need_model("setparcs_"); if (thcm_b200_set_par(M.h, *param, *value) != 0) die("setparcs_"); }
EXPORT void getparcs_(int* param, double* value) { need_model("getparcs_"); if (thcm_b200_get_par(M.h, *param, value) != 0) die("getparcs_"); }

EXPORT void setsres_(int* sres) { // usrc.F90:417-429: switches the salinity surface condition, then forcing + lin
  need_model("setsres_");
  if (M.cfg.sres == *sres) return;
  double par[31];
  for (int p = 1; p <= 30; p++) if (thcm_b200_get_par(M.h, p, &par[p]) != 0) die("setsres_");
  M.cfg.sres = *sres;
  create_model();
  for (int p = 1; p <= 30; p++) if (thcm_b200_set_par(M.h, p, par[p]) != 0) die("setsres_");
}
EXPORT void set_landmask_(int* landm, int* periodic, int* reinit) {
  need_model("set_landmask_");
  (void)reinit;
  M.landm.assign(landm, landm + M.landm.size());
  if ((*periodic != 0) != (M.cfg.periodic != 0)) { // a different topology: new graph
    double par[31];
    for (int p = 1; p <= 30; p++) if (thcm_b200_get_par(M.h, p, &par[p]) != 0) die("set_landmask_");
    M.cfg.periodic = *periodic;
    create_model();
    for (int p = 1; p <= 30; p++) if (thcm_b200_set_par(M.h, p, par[p]) != 0) die("set_landmask_");
    return;
  }
  if (thcm_b200_set_landmask(M.h, M.landm.data()) != 0) die("set_landmask_");
  M.have_B = false;
}
EXPORT void __m_mix_MOD_set_vmix_fix(int* vmix_fix) { M.vmix_fix = *vmix_fix; }

EXPORT void __m_mat_MOD_get_array_sizes(int* nrows, int* nnz) { // mat.F90:56-70
  need_model("m_mat::get_array_sizes");
  *nrows = 6 * M.cfg.n * M.cfg.m * M.cfg.l;
  *nnz = *nrows * (6 * 27 + 1);
}
EXPORT void __m_mat_MOD_set_pointers(int* nrows, int* nnz, int* begA, int* jcoA, double* coA, double* coB, int* begF, int* jcoF,
                                     double* coF) {
  M.cap_rows = *nrows; M.cap_nnz = *nnz;
  M.begA = begA; M.jcoA = jcoA; M.coA = coA; M.coB = coB; M.begF = begF; M.jcoF = jcoF; M.coF = coF;
}

EXPORT void rhs_(double* un, double* b) { // usrc.F90:506-586; THCM::evaluate scales by -1 afterwards (THCM.C:1003)
  need_model("rhs_");
  if (thcm_b200_rhs_host(M.h, un, M.F.data()) != 0) die("rhs_");
  for (size_t r = 0; r < M.F.size(); r++) b[r] = -M.F[r];
}

EXPORT void matrix_(double* un) { // usrc.F90:432-504: coB by fillcolB, coA/jcoA/begA by fillcolA
  need_model("matrix_");
  if (!M.begA || !M.jcoA || !M.coA || !M.coB) { std::fprintf(stderr, "thcm_b200 (Fortran ABI) matrix_ called before m_mat::set_pointers\n"); std::abort(); }
  run_matrix(un);
  const size_t nrows = M.rowptr.size() - 1;
  if ((int)nrows > M.cap_rows) { std::fprintf(stderr, "thcm_b200 (Fortran ABI) matrix_: row arrays too small\n"); std::abort(); }
  int v = 1;
  for (size_t r = 0; r < nrows; r++) {
    M.begA[r] = v;
    for (int q = M.rowptr[r]; q < M.rowptr[r + 1]; q++) {
      const int p = M.order[q];
      const double a = M.val[p];
      if (std::fabs(a) > 1.0e-10) { // assemble.F90:115
        if (v > M.cap_nnz) { std::fprintf(stderr, "thcm_b200 (Fortran ABI) matrix_: value arrays too small\n"); std::abort(); }
        M.coA[v - 1] = a; M.jcoA[v - 1] = M.col[p] + 1;
        v++;
      }
    }
    M.coB[r] = M.diagB[r];
  }
  M.begA[nrows] = v;
}
EXPORT void fillcolb_(void) { // assemble.F90:18-54: depends on the mask and on Ro only
  need_model("fillcolb_");
  if (!M.coB) { std::fprintf(stderr, "thcm_b200 (Fortran ABI) fillcolb_ called before m_mat::set_pointers\n"); std::abort(); }
  if (!M.have_B) { std::vector<double> zero(M.F.size(), 0.0); run_matrix(zero.data()); }
  std::copy(M.diagB.begin(), M.diagB.end(), M.coB);
}

EXPORT void __m_thcm_utils_MOD_intcond_scaling(double* values, int* indices, int* len) { // thcm_utils.F90:285-323
  need_model("m_thcm_utils::intcond_scaling");
  std::vector<double> coeff(M.F.size());
  int row = -1;
  if (thcm_b200_intcond_host(M.h, coeff.data(), &row) != 0) die("m_thcm_utils::intcond_scaling");
  const int n = M.cfg.n, m = M.cfg.m, l = M.cfg.l;
  int v = 0;
  for (int k = 1; k <= l; k++) for (int j = 1; j <= m; j++) for (int i = 1; i <= n; i++)
    if (M.landm[((size_t)k * (m + 2) + j) * (n + 2) + i] == 0) { // OCEAN (par.F90:78)
      const size_t r = 6 * (((size_t)(k - 1) * m + (j - 1)) * n + (i - 1)) + 5;
      values[v] = coeff[r]; indices[v] = (int)r + 1;
      v++;
    }
  *len = v;
}


// m_scaling (src/ocean/scaling.F90), as THCM::RecomputeScaling calls it (THCM.C:1707-1721): average_block on the matrix of the
// last matrix_ call, then -- after the caller's SumAll / NumProc -- compute() with the globally averaged block.  The
// block average runs on the device; compute() hands out THCM-convention factors (the caller takes the reciprocals).
EXPORT void __m_scaling_MOD_average_block(double* db) {
  need_model("m_scaling::average_block");
  if (thcm_b200_average_block(M.h, db) != 0) die("m_scaling::average_block");
}
EXPORT void __m_scaling_MOD_compute(double* db, double* row_scaling, double* col_scaling) {
  need_model("m_scaling::compute");
  double dr[6], dc[6];
  if (thcm_b200_scal(db, dr, dc) != 0) { // scaling.F90:233-236: message, and the vectors stay as they are
    std::fprintf(stderr, " thcm scaling: diagonal block is singular up to working precision\n");
    return;
  }
  const int n = M.cfg.n, m = M.cfg.m, l = M.cfg.l;
  for (int k = 1; k <= l; k++) for (int j = 1; j <= m; j++) for (int i = 1; i <= n; i++) {
    const bool ocean = M.landm[((size_t)k * (m + 2) + j) * (n + 2) + i] == 0;
    const size_t r = 6 * (((size_t)(k - 1) * m + (j - 1)) * n + (i - 1));
    for (int ii = 0; ii < 6; ii++) { row_scaling[r + ii] = ocean ? dr[ii] : 1.0; col_scaling[r + ii] = ocean ? dc[ii] : 1.0; }
  }
}

} // extern "C"
