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
  int N = 0, M = 0, L = 0;
  std::vector<int> landm; // the global mask, when the host program hands one over with m_global::set_landm
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

EXPORT void __m_global_MOD_initialize(int* N, int* M, int* L, double*, double*, double*, double*, double* hdim, double* qz, int*,
                                      int*, int*, int* rd_mask, int* TRES, int* SRES, int* iza, int* ite, int* its, int*,
                                      int* coupled_T, int* coupled_S, int* forcing_type, const char*, const char*, const char*,
                                      const char*, const char*) {
  G.set = true; G.N = *N; G.M = *M; G.L = *L; G.hdim = *hdim; G.qz = *qz; G.tres = *TRES; G.sres = *SRES; G.iza = *iza; G.ite = *ite; G.its = *its;
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

// ---- the rest of the symbol surface THCM.C:47-174 binds ----
// Served where the hot path has an answer (global mask hand-over, idealised forcing fields = none, parameter print-out);
// everything that belongs to data-file forcing, time-dependent forcing or the coupled model (SURVEY rows out of scope / N2)
// refuses loudly through the same route as every other error here: message, thcm_throw_error_ if the host has it, abort.
namespace {
[[noreturn]] void refuse(const char* sym, const char* why) {
  std::fprintf(stderr, "thcm_b200 (Fortran ABI) %s: not served by the B200 path (%s)\n", sym, why);
  typedef void (*throw_fn)(const char*);
  if (throw_fn f = (throw_fn)dlsym(RTLD_DEFAULT, "thcm_throw_error_")) f(sym);
  std::abort();
}
size_t global_mask_len() { return (size_t)(G.N + 2) * (G.M + 2) * (G.L + 2); }
void need_global(const char* where) {
  if (!G.set) { std::fprintf(stderr, "thcm_b200 (Fortran ABI) %s called before m_global::initialize\n", where); std::abort(); }
}
void zero_surface(double* a) { if (a) std::fill(a, a + (size_t)G.N * G.M, 0.0); }
} // namespace

extern "C" {
// usrc.F90
EXPORT void writeparams_(void) { // usrc.F90:199-234: the parameter table
  need_model("writeparams_");
  for (int q = 1; q <= 30; q++) { double v = 0.0; thcm_b200_get_par(M.h, q, &v); std::printf(" par(%2d) = %.15e\n", q, v); }
}
EXPORT void stochastic_forcing_(void) { refuse("stochastic_forcing_", "stochastic forcing is not on the hot path"); }
EXPORT void write_levitus_(const char*) { refuse("write_levitus_", "Levitus data files are not read"); }
// global.F90: the global land mask.  Mask FILES are not read here; a host program that has the mask hands it over with set_landm
EXPORT void __m_global_MOD_set_maskfile(const char*) { refuse("m_global::set_maskfile", "mask files are not read; pass the mask through m_global::set_landm / init_"); }
EXPORT void __m_global_MOD_set_landm(int* landm) { need_global("m_global::set_landm"); G.landm.assign(landm, landm + global_mask_len()); }
EXPORT void __m_global_MOD_get_landm(int* landm) {
  need_global("m_global::get_landm");
  if (G.landm.size() != global_mask_len()) refuse("m_global::get_landm", "no global mask has been set: mask files are not read, call m_global::set_landm first");
  std::copy(G.landm.begin(), G.landm.end(), landm);
}
EXPORT void __m_global_MOD_get_current_landm(int* landm) { __m_global_MOD_get_landm(landm); }
EXPORT void __m_thcm_utils_MOD_get_landm(int* landm) { // the SUB-DOMAIN mask the Fortran layer works with (thcm_utils.F90)
  need_model("m_thcm_utils::get_landm");
  std::copy(M.landm.begin(), M.landm.end(), landm);
}
// idealised forcing (iza = 2, ite = its = 1): there are no data fields; the arrays THCM's constructor fetches and hands back to
// init_ (which ignores them) are zero
EXPORT void __m_global_MOD_get_windfield(double* taux, double* tauy) { need_global("m_global::get_windfield"); zero_surface(taux); zero_surface(tauy); }
EXPORT void __m_global_MOD_get_temforcing(double* tatm) { need_global("m_global::get_temforcing"); zero_surface(tatm); }
EXPORT void __m_global_MOD_get_salforcing(double* emip) { need_global("m_global::get_salforcing"); zero_surface(emip); }
EXPORT void __m_global_MOD_get_spert(double* spert) { need_global("m_global::get_spert"); zero_surface(spert); }
EXPORT void __m_global_MOD_get_internal_temforcing(double*) { refuse("m_global::get_internal_temforcing", "internal (data) forcing"); }
EXPORT void __m_global_MOD_get_internal_salforcing(double*) { refuse("m_global::get_internal_salforcing", "internal (data) forcing"); }
EXPORT void __m_global_MOD_get_monthly_forcing(double*, double*, double*, double*, int*) { refuse("m_global::get_monthly_forcing", "time-dependent forcing"); }
EXPORT void __m_global_MOD_get_monthly_internal_forcing(double*, double*, int*) { refuse("m_global::get_monthly_internal_forcing", "time-dependent forcing"); }
EXPORT void __m_global_MOD_get_land_temp(double*) { refuse("m_global::get_land_temp", "coupled model (SURVEY row N2)"); }
EXPORT void __m_monthly_MOD_set_forcing(double*, double*, double*, double*, int*) { refuse("m_monthly::set_forcing", "time-dependent forcing"); }
EXPORT void __m_monthly_MOD_set_internal_forcing(double*, double*, int*) { refuse("m_monthly::set_internal_forcing", "time-dependent forcing"); }
EXPORT void __m_monthly_MOD_update_forcing(double*, double*, double*, double*) { refuse("m_monthly::update_forcing", "time-dependent forcing"); }
EXPORT void __m_monthly_MOD_update_internal_forcing(double*, double*, double*) { refuse("m_monthly::update_internal_forcing", "time-dependent forcing"); }
EXPORT void __m_usr_MOD_set_internal_forcing(double*, double*) { refuse("m_usr::set_internal_forcing", "internal (data) forcing"); }
EXPORT void __m_thcm_utils_MOD_loadbal_weights(double*, double*, double*, double*) { refuse("m_thcm_utils::loadbal_weights", "load balancing is disabled in the reference too (TRIOS_Domain.H:271)"); }
// I-EMIC couplings (m_inserts / m_probe / m_integrals): the coupled model's surface fields, SURVEY row N2
#define REFUSE_N2_1(name) EXPORT void name(double*) { refuse(#name, "coupled model (SURVEY row N2)"); }
// served for Coupled Temperature = 1 (the fields Coupled Salinity = 1 needs are ignored by the Fortran itself while that flag is off;
// with it on the model is refused at creation): inserts.F90:12-166 -> thcm_b200_insert_field
#define SERVE_INSERT(name, which) EXPORT void name(double* f) { need_model(#name); if (thcm_b200_insert_field(M.h, which, f) != 0) die(#name); }
SERVE_INSERT(__m_inserts_MOD_insert_atmosphere_t, THCM_B200_FIELD_ATMOS_T) SERVE_INSERT(__m_inserts_MOD_insert_atmosphere_q, THCM_B200_FIELD_ATMOS_Q)
SERVE_INSERT(__m_inserts_MOD_insert_atmosphere_a, THCM_B200_FIELD_ATMOS_A) SERVE_INSERT(__m_inserts_MOD_insert_atmosphere_p, THCM_B200_FIELD_ATMOS_P)
SERVE_INSERT(__m_inserts_MOD_insert_seaice_q, THCM_B200_FIELD_SEAICE_Q) SERVE_INSERT(__m_inserts_MOD_insert_seaice_m, THCM_B200_FIELD_SEAICE_M)
SERVE_INSERT(__m_inserts_MOD_insert_seaice_g, THCM_B200_FIELD_SEAICE_G)
#undef SERVE_INSERT
// Ocean::synchronize's parameter calls and getdeps (Ocean.C:1469, 1490, 1545; usrc.F90:201-333): not in THCM.C's extern block, bound by Ocean.C
EXPORT void set_atmos_parameters_(void* pars) { need_model("set_atmos_parameters_"); if (thcm_b200_set_atmos_parameters(M.h, (const thcm_b200_atmos_pars*)pars) != 0) die("set_atmos_parameters_"); }
EXPORT void set_seaice_parameters_(void* pars) { need_model("set_seaice_parameters_"); if (thcm_b200_set_seaice_parameters(M.h, (const thcm_b200_seaice_pars*)pars) != 0) die("set_seaice_parameters_"); }
EXPORT void getdeps_(double* Ooa, double* Os, double* nus, double* eta, double* lvsc, double* qdim, double* pQSnd) {
  need_model("getdeps_");
  double o[7];
  if (thcm_b200_getdeps(M.h, o) != 0) die("getdeps_");
  *Ooa = o[0]; *Os = o[1]; *nus = o[2]; *eta = o[3]; *lvsc = o[4]; *qdim = o[5]; *pQSnd = o[6];
}
EXPORT void __m_probe_MOD_get_suno(double* suno) { need_model("m_probe::get_suno"); if (thcm_b200_get_suno(M.h, suno) != 0) die("m_probe::get_suno"); }
REFUSE_N2_1(__m_inserts_MOD_insert_emip) REFUSE_N2_1(__m_inserts_MOD_insert_adapted_emip)
REFUSE_N2_1(__m_inserts_MOD_insert_emip_pert) REFUSE_N2_1(__m_inserts_MOD_insert_tatm)
REFUSE_N2_1(__m_probe_MOD_get_emip) REFUSE_N2_1(__m_probe_MOD_get_adapted_emip) REFUSE_N2_1(__m_probe_MOD_get_emip_pert)
REFUSE_N2_1(__m_probe_MOD_get_atmosphere_t) REFUSE_N2_1(__m_probe_MOD_get_atmosphere_q) REFUSE_N2_1(__m_probe_MOD_get_atmosphere_p)
#undef REFUSE_N2_1
EXPORT void __m_integrals_MOD_salt_advection(double*, double*) { refuse("m_integrals::salt_advection", "diagnostic integrals are not on the hot path"); }
EXPORT void __m_integrals_MOD_salt_diffusion(double*, double*) { refuse("m_integrals::salt_diffusion", "diagnostic integrals are not on the hot path"); }
// probe.F90:381-471: dF_T/dM from the device; the three salinity derivatives are zero while Coupled Salinity = 0 (:446)
EXPORT void __m_probe_MOD_get_derivatives(double* un, double* dftdm, double* dfsdq, double* dfsdm, double* dfsdg) {
  need_model("m_probe::get_derivatives");
  const size_t nm = (size_t)M.cfg.n * M.cfg.m, nvec = M.F.size();
  void* d_x = nullptr;
  if (iemic_b200_malloc(&d_x, (int64_t)(nvec * sizeof(double))) != 0 || iemic_b200_memcpy_h2d(d_x, un, (int64_t)(nvec * sizeof(double))) != 0) die("m_probe::get_derivatives");
  const int rc = thcm_b200_get_dftdm(M.h, (const double*)d_x, dftdm);
  iemic_b200_free(d_x);
  if (rc != 0) die("m_probe::get_derivatives");
  for (size_t q = 0; q < nm; q++) { dfsdq[q] = 0.0; dfsdm[q] = 0.0; dfsdg[q] = 0.0; }
}
EXPORT void __m_probe_MOD_get_salflux(double*, double*, double*, double*, double*) { refuse("m_probe::get_salflux", "coupled model (SURVEY row N2)"); }
EXPORT void __m_probe_MOD_get_temflux(double*, double*, double*, double*, double*, double*, double*) { refuse("m_probe::get_temflux", "coupled model (SURVEY row N2)"); }
EXPORT void __m_probe_MOD_compute_evap(double*, double*) { refuse("m_probe::compute_evap", "coupled model (SURVEY row N2)"); }
} // extern "C"
