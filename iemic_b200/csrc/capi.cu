// capi.cu -- the C ABI of include/iemic_b200.h: handle management, device residency of the static data,
// kernel sequencing for assembly / SpMV / GMRES.  No arithmetic happens on the host in here except the
// per-parameter-change setup delegated to thcm::Host (thcm_host.cpp).
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <cstring>
#include <cmath>
#include <string>
#include <vector>
#include <stdexcept>
#include "../../include/iemic_b200.h"
#include "thcm_host.h"
#include "kernels.h"
#include "thcm_stage.cuh"
#include "gmres_guard.h"
#include "dist.h"

using namespace thcm;

namespace {
thread_local std::string g_err;
int fail(int code, const std::string& msg) { g_err = msg; return code; }
#define CU(call)                                                                                   \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess) return fail(IEMIC_B200_ECUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); \
  } while (0)
#define CHECK_H(h) if (!(h)) return fail(IEMIC_B200_EINVAL, "null handle")

template <class T> int dev_upload(T** d, const std::vector<T>& v, size_t pad = 0) {
  if (*d) { cudaFree(*d); *d = nullptr; }
  size_t bytes = (v.size() + pad) * sizeof(T);
  if (bytes == 0) bytes = sizeof(T);
  CU(cudaMalloc((void**)d, bytes));
  CU(cudaMemset(*d, 0, bytes));
  if (!v.empty()) CU(cudaMemcpy(*d, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
  return 0;
}
} // namespace

struct GmresWs {
  int m = 0; long long n = 0, ld = 0; bool haveZ = false;
  double *V = nullptr, *Z = nullptr, *w = nullptr, *r = nullptr, *tmp = nullptr;
  double *H = nullptr, *cs = nullptr, *sn = nullptr, *s = nullptr, *y = nullptr, *hist = nullptr;
  double* cyc_part = nullptr;  // block partials of the cooperative one-launch cycle (gmres_cycle.cu)
  double* cyc_flags = nullptr; // two ints: columns built, stopped on the tolerance (gmres_cycle.cu, gmres_cluster.cu)
};

struct TimedEv { int which; cudaEvent_t a, b; };

struct thcm_b200 {
  Host* host = nullptr;
  Dist* dist = nullptr;          // null on a single GPU
  cudaStream_t stream = 0;
  cudaStream_t copy_stream = nullptr; cudaEvent_t ev_rhs = nullptr, ev_copy = nullptr; // thcm_b200_assemble_host overlap
  cudaStream_t up_stream = nullptr; std::vector<cudaEvent_t> ev_slab;
  // static / per-parameter device data
  double* d_tab = nullptr; uint64_t* d_sig = nullptr; int* d_rowptr = nullptr; int* d_col = nullptr;
  double* d_val = nullptr; double* d_diagB = nullptr; double* d_frc = nullptr; double* d_coefS = nullptr;
  int* d_blkrow = nullptr; int nblk = 0; size_t d_frc_len = 0, d_tab_len = 0;
  int* d_blkp0 = nullptr; unsigned char* d_blkreg = nullptr; SpmvTab spmv_tab; size_t spmv_regular_cells = 0;
  uint64_t dev_version = 0, dev_mask_version = 0;
  // row N2: device copies of the coupling fields (whole surface) and of the per-cell surface T forcing
  double *d_msi = nullptr, *d_ftcell = nullptr, *d_cplf = nullptr; uint64_t dev_msi_version = 0;
  unsigned int* d_dinfo = nullptr; uint64_t dinfo_mask_version = 0; // diagonal-block positions (scaling.cu), built on first use
  // scratch vectors for the host-buffer entry points
  double *d_xs = nullptr, *d_ys = nullptr;
  // reductions and small device scalars
  RedBuf rb{nullptr, nullptr, 0};
  double* d_small = nullptr; // [0,512) h1, [512,1024) h2, 1024 nrm2, 1025 hnext, 1026 bnorm, 1027 beta, 1028 dot, 1029 intcond,
                             // [1030,1067) average block + count, [1070,1088) Fortran-level centre entries of the special rows, 1090 int counter
  GmresWs g;
  gmres_b200_precond_fn precond = nullptr; void* precond_user = nullptr; // right preconditioner supplied by the caller (GMRES_B200_PREC_CALLBACK)
  long long launches = 0;
  bool timing = false;
  std::vector<TimedEv> evs;
  double t_total[6] = {0, 0, 0, 0, 0, 0}; long long t_count[6] = {0, 0, 0, 0, 0, 0}; // 0 rhs, 1 jac, 2 spmv, 3 orthogonalisation, 4 fused jac+rhs
  long long nrows() const { return (long long)host->nrows_local(); }
  long long nvec() const { return (long long)host->nvec_local(); }
};

namespace {
struct Timed { // RAII bracket: records a pair of events around a kernel group when timing is on
  thcm_b200* h; int which; TimedEv ev{}; bool on; cudaStream_t st;
  Timed(thcm_b200* h_, int w, cudaStream_t st_ = nullptr) : h(h_), which(w), on(h_->timing), st(st_ ? st_ : h_->stream) {
    if (on) { ev.which = w; cudaEventCreate(&ev.a); cudaEventCreate(&ev.b); cudaEventRecord(ev.a, st); }
  }
  ~Timed() { if (on) { cudaEventRecord(ev.b, st); h->evs.push_back(ev); } }
};
int flush_timing(thcm_b200* h) {
  if (h->evs.empty()) return 0;
  CU(cudaStreamSynchronize(h->stream));
  for (auto& e : h->evs) {
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e.a, e.b);
    h->t_total[e.which] += ms; h->t_count[e.which]++;
    cudaEventDestroy(e.a); cudaEventDestroy(e.b);
  }
  h->evs.clear();
  return 0;
}

// Build the plan of the SpMV kernel (spmv.cu): blocks of whole regular cells (columns follow from the cell index, checked
// here against the real column array) and, for everything else, runs of <= rows_cap rows and <= nnz_cap non-zeros.
struct SpmvHostPlan { std::vector<int> blk_row, blk_p0; std::vector<unsigned char> blk_reg; SpmvTab tab; size_t regular_cells = 0; };
template <int A> void spmv_tab_rows(SpmvTab& T, int& pos, int nloc, long long nlm) {
  for (int e = 0; e < RowDef<A>::NE; e++) {
    const int slot = RowDef<A>::e[e].slot;
    T.off[pos++] = (int)(6 * (slot_di(slot) + (long long)slot_dj(slot) * nloc + slot_dk(slot) * nlm)) + RowDef<A>::e[e].var - 1;
  }
}
SpmvHostPlan spmv_plan(const Host& H) {
  SpmvHostPlan P;
  const std::vector<int>& rowptr = H.rowptr;
  const std::vector<int>& col = H.col;
  const int nrows = (int)rowptr.size() - 1, ncap = spmv_nnz_cap(), rcap = spmv_rows_cap(), ncell = nrows / 6;
  int pos = 0;
  const long long nlm = (long long)H.nloc * H.mloc;
  std::memset(&P.tab, 0, sizeof(P.tab));
  spmv_tab_rows<UU>(P.tab, pos, H.nloc, nlm); spmv_tab_rows<VV>(P.tab, pos, H.nloc, nlm); spmv_tab_rows<WW>(P.tab, pos, H.nloc, nlm);
  spmv_tab_rows<PP>(P.tab, pos, H.nloc, nlm); spmv_tab_rows<TT>(P.tab, pos, H.nloc, nlm); spmv_tab_rows<SS>(P.tab, pos, H.nloc, nlm);
  const bool use_regular = !(getenv("IEMIC_SPMV_REGULAR") && getenv("IEMIC_SPMV_REGULAR")[0] == '0') && pos == SPMV_CELL_NNZ;
  const int lens[6] = {RowDef<UU>::NE, RowDef<VV>::NE, RowDef<WW>::NE, RowDef<PP>::NE, RowDef<TT>::NE, RowDef<SS>::NE};
  std::vector<unsigned char> regular((size_t)ncell, 0);
  if (use_regular)
    for (int c = 0; c < ncell; c++) {
      bool ok = true;
      for (int a = 0; a < 6 && ok; a++) ok = rowptr[6 * c + a + 1] - rowptr[6 * c + a] == lens[a];
      const int p0 = rowptr[6 * c];
      for (int q = 0; q < SPMV_CELL_NNZ && ok; q++) ok = col[p0 + q] == 6 * c + P.tab.off[q];
      regular[c] = ok;
      P.regular_cells += ok;
    }
  const int G = spmv_regular_cells();
  if (G <= 0) std::fill(regular.begin(), regular.end(), 0);
  int r = 0;
  P.blk_row.push_back(0);
  while (r < nrows) {
    int r1;
    const int c = r / 6;
    if (r % 6 == 0 && c < ncell && regular[c]) {
      int g = 1;
      while (g < G && c + g < ncell && regular[c + g]) g++;
      r1 = r + 6 * g;
      P.blk_reg.push_back(1);
    } else {
      r1 = r + 1; // always take one row (a long row becomes a block of its own)
      auto stop = [&](int rr) { return rr % 6 == 0 && rr / 6 < ncell && regular[rr / 6]; }; // regular cells start their own block
      if (rowptr[r1] - rowptr[r] <= ncap)
        while (r1 < nrows && !stop(r1) && r1 - r < rcap && rowptr[r1 + 1] - rowptr[r] <= ncap) r1++;
      P.blk_reg.push_back(0);
    }
    P.blk_p0.push_back(rowptr[r]);
    P.blk_row.push_back(r1);
    r = r1;
  }
  return P;
}

int check_launch(thcm_b200* h, const char* what);
int refresh_device(thcm_b200* h) {
  Host* H = h->host;
  if (h->dev_mask_version != H->mask_version) {
    if (int rc = dev_upload(&h->d_sig, H->sig)) return rc;
    if (int rc = dev_upload(&h->d_rowptr, H->rowptr)) return rc;
    if (int rc = dev_upload(&h->d_col, H->col, 8)) return rc;
    std::vector<double> zeros;
    if (h->d_val) { cudaFree(h->d_val); h->d_val = nullptr; }
    CU(cudaMalloc((void**)&h->d_val, (H->col.size() + 8) * sizeof(double)));
    CU(cudaMemset(h->d_val, 0, (H->col.size() + 8) * sizeof(double)));
    if (h->d_diagB) { cudaFree(h->d_diagB); h->d_diagB = nullptr; }
    CU(cudaMalloc((void**)&h->d_diagB, (size_t)h->nrows() * sizeof(double)));
    CU(cudaMemset(h->d_diagB, 0, (size_t)h->nrows() * sizeof(double)));
    SpmvHostPlan sp = spmv_plan(*H);
    h->nblk = (int)sp.blk_row.size() - 1;
    h->spmv_tab = sp.tab; h->spmv_regular_cells = sp.regular_cells;
    if (getenv("IEMIC_DEBUG")) fprintf(stderr, "[iemic_b200] SpMV plan: %d blocks, %zu of %lld cells regular\n", h->nblk, sp.regular_cells, h->nrows() / 6);
    if (int rc = dev_upload(&h->d_blkrow, sp.blk_row)) return rc;
    if (int rc = dev_upload(&h->d_blkp0, sp.blk_p0)) return rc;
    if (int rc = dev_upload(&h->d_blkreg, sp.blk_reg)) return rc;
    // integral condition: compact per-cell coefficients, and the static dense Jacobian row (THCM.C:2121-2198)
    std::vector<double> coefS((size_t)h->nrows() / 6);
    for (size_t c = 0; c < coefS.size(); c++) coefS[c] = H->intcond[6 * c + 5];
    if (int rc = dev_upload(&h->d_coefS, coefS)) return rc;
    if (H->rowintcon >= 0 && H->owns_rowintcon) {
      const int r = H->rowintcon_local;
      std::vector<double> rowv((size_t)(H->rowptr[r + 1] - H->rowptr[r]));
      for (size_t q = 0; q < rowv.size(); q++) rowv[q] = H->cfg.int_sign * H->intcond_col(H->col[H->rowptr[r] + q]);
      CU(cudaMemcpy(h->d_val + H->rowptr[r], rowv.data(), rowv.size() * sizeof(double), cudaMemcpyHostToDevice));
    }
    if (!h->d_xs) {
      CU(cudaMalloc((void**)&h->d_xs, (size_t)h->nvec() * sizeof(double)));
      CU(cudaMemset(h->d_xs, 0, (size_t)h->nvec() * sizeof(double)));
      CU(cudaMalloc((void**)&h->d_ys, (size_t)h->nvec() * sizeof(double)));
      CU(cudaMemset(h->d_ys, 0, (size_t)h->nvec() * sizeof(double)));
    }
    h->dev_mask_version = H->mask_version;
    h->dev_version = 0;
  }
  if (h->dev_version != H->version) {
    // coefficient tables and per-latitude forcing profiles: one small buffer, reused (no allocation on a parameter change),
    // copied on the handle's stream so that kernels still reading the old values finish first
    const size_t ntab = H->tab.size(), nprof = H->fprof.size(), need = ntab + nprof + 8;
    if (!h->d_tab || h->d_tab_len < need) {
      if (h->d_tab) cudaFree(h->d_tab);
      CU(cudaMalloc((void**)&h->d_tab, need * sizeof(double)));
      h->d_tab_len = need;
    }
    CU(cudaMemcpyAsync(h->d_tab, H->tab.data(), ntab * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    if (nprof) CU(cudaMemcpyAsync(h->d_tab + ntab, H->fprof.data(), nprof * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    if (H->frc_host_valid && !(h->d_frc && h->d_frc_len == H->frc.size() && H->frc_dirty_from >= H->frc.size())) {
      // the host copy is current (creation, new land mask): upload what changed
      if (h->d_frc && h->d_frc_len == H->frc.size() && H->frc_dirty_from > 0) {
        const size_t a = std::min(H->frc_dirty_from, H->frc.size());
        if (a < H->frc.size()) CU(cudaMemcpyAsync(h->d_frc + a, H->frc.data() + a, (H->frc.size() - a) * sizeof(double), cudaMemcpyHostToDevice, h->stream));
      } else {
        CU(cudaStreamSynchronize(h->stream));
        if (int rc = dev_upload(&h->d_frc, H->frc)) return rc;
        h->d_frc_len = H->frc.size();
      }
      H->frc_dirty_from = H->frc.size();
    } else if (!H->frc_host_valid) {
      // a parameter changed: only the surface rows depend on it; a kernel rebuilds them from the profiles (row N1)
      launch_forcing_surface(H->S, h->d_sig, h->d_tab + ntab, h->d_frc, h->stream);
      h->launches++;
      if (H->cfg.coupled_T == 1) { // coupled to an atmosphere: the T rows are per-cell values (forcing.F90:73-96), row N2
        const size_t nc = H->ftcell.size();
        if (!h->d_ftcell) CU(cudaMalloc((void**)&h->d_ftcell, nc * sizeof(double)));
        CU(cudaMemcpyAsync(h->d_ftcell, H->ftcell.data(), nc * sizeof(double), cudaMemcpyHostToDevice, h->stream));
        launch_forcing_cells(H->S, h->d_sig, h->d_ftcell, h->d_frc, h->stream);
        h->launches++;
      }
      if (int rc = check_launch(h, "forcing_surface")) return rc;
    }
    if (H->cfg.coupled_T == 1 && h->dev_msi_version != H->cpl.msi_version) { // sea-ice mask the coefficients were built with
      const size_t nm = (size_t)H->n * H->m;
      if (!h->d_msi) CU(cudaMalloc((void**)&h->d_msi, nm * sizeof(double)));
      if (H->cpl.msi_lin.size() == nm) CU(cudaMemcpyAsync(h->d_msi, H->cpl.msi_lin.data(), nm * sizeof(double), cudaMemcpyHostToDevice, h->stream));
      h->dev_msi_version = H->cpl.msi_version;
    }
    h->dev_version = H->version;
    H->forcing_on_device = true; // from now on set_par leaves the rows to the kernel
  }
  return 0;
}

AsmArgs asm_args(thcm_b200* h, const double* d_x) {
  AsmArgs a;
  std::memset(&a, 0, sizeof(a));
  Host* H = h->host;
  a.S = H->S;
  a.T = H->tables(h->d_tab);
  a.T.msi = (H->cfg.coupled_T == 1 && H->cpl.have_msi) ? h->d_msi : nullptr;
  a.x = d_x; a.sig = h->d_sig; a.rowptr = h->d_rowptr; a.frc = h->d_frc; a.val = h->d_val; a.diagB = h->d_diagB;
  a.sp.rowintcon = H->owns_rowintcon ? H->rowintcon_local : -1;
  a.sp.rowPfix1 = H->rowPfix1_local; a.sp.rowPfix2 = H->rowPfix2_local;
  a.sp.side = (a.sp.rowintcon >= 0 || a.sp.rowPfix1 >= 0 || a.sp.rowPfix2 >= 0) ? h->d_small + 1070 : nullptr;
  return a;
}
int check_launch(thcm_b200* h, const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(IEMIC_B200_ECUDA, std::string(what) + ": " + cudaGetErrorString(e));
  (void)h;
  return 0;
}
int gmres_alloc(thcm_b200* h, int m, bool needZ) {
  GmresWs& g = h->g;
  const long long n = h->nrows();
  const long long ld = (h->nvec() + 31) / 32 * 32; // basis vectors double as SpMV inputs: room for the halo
  if (g.m >= m && g.n == n && (!needZ || g.haveZ)) return 0;
  double** ptrs[] = {&g.V, &g.Z, &g.w, &g.r, &g.tmp, &g.H, &g.cs, &g.sn, &g.s, &g.y, &g.hist, &g.cyc_part, &g.cyc_flags};
  for (auto p : ptrs) if (*p) { cudaFree(*p); *p = nullptr; }
  g.m = m; g.n = n; g.ld = ld; g.haveZ = needZ;
  CU(cudaMalloc((void**)&g.V, (size_t)(m + 1) * ld * sizeof(double)));
  CU(cudaMemset(g.V, 0, (size_t)(m + 1) * ld * sizeof(double)));
  if (needZ) { CU(cudaMalloc((void**)&g.Z, (size_t)(m + 1) * ld * sizeof(double))); CU(cudaMemset(g.Z, 0, (size_t)(m + 1) * ld * sizeof(double))); }
  CU(cudaMalloc((void**)&g.w, ld * sizeof(double)));
  CU(cudaMalloc((void**)&g.r, ld * sizeof(double)));
  CU(cudaMalloc((void**)&g.tmp, ld * sizeof(double)));
  CU(cudaMemset(g.w, 0, ld * sizeof(double))); CU(cudaMemset(g.r, 0, ld * sizeof(double))); CU(cudaMemset(g.tmp, 0, ld * sizeof(double)));
  CU(cudaMalloc((void**)&g.H, (size_t)(m + 1) * m * sizeof(double)));
  CU(cudaMalloc((void**)&g.cs, (m + 1) * sizeof(double)));
  CU(cudaMalloc((void**)&g.sn, (m + 1) * sizeof(double)));
  CU(cudaMalloc((void**)&g.s, (m + 2) * sizeof(double)));
  CU(cudaMalloc((void**)&g.y, (m + 2) * sizeof(double)));
  CU(cudaMalloc((void**)&g.hist, (m + 2) * sizeof(double)));
  if (!h->dist && n <= gmres_cycle_max_rows()) CU(cudaMalloc((void**)&g.cyc_part, (size_t)gmres_cycle_partials(m) * sizeof(double)));
  CU(cudaMalloc((void**)&g.cyc_flags, 2 * sizeof(double)));
  return 0;
}
void gmres_free(GmresWs& g) {
  double** ptrs[] = {&g.V, &g.Z, &g.w, &g.r, &g.tmp, &g.H, &g.cs, &g.sn, &g.s, &g.y, &g.hist, &g.cyc_part, &g.cyc_flags};
  for (auto p : ptrs) if (*p) { cudaFree(*p); *p = nullptr; }
  g.m = 0;
}
// allreduce helper: no-op on one GPU
// Reductions across ranks: with the peer-memory path the reducing kernel finishes the sum itself (argument block from
// red()); otherwise ncclAllReduce follows the kernel.
PeerRed red(thcm_b200* h) { return dist_next_red(h->dist); }
int allreduce(thcm_b200* h, double* d, int count, const PeerRed& pr = PeerRed{}) {
  if (!h->dist || pr.nranks) return 0;
  h->launches++;
  return dist_allreduce_sum(h->dist, d, count, h->stream) == 0 ? 0 : fail(IEMIC_B200_ENCCL, dist_last_error());
}
int spmv_impl(thcm_b200* h, double* d_x, double* d_y) {
  if (h->dist) {
    if (dist_halo_exchange(h->dist, d_x, h->stream) != 0) return fail(IEMIC_B200_ENCCL, dist_last_error());
    h->launches += dist_halo_launches(h->dist);
  }
  {
    Timed t(h, 2);
    static const bool tile = !(getenv("IEMIC_SPMV_TILE") && getenv("IEMIC_SPMV_TILE")[0] == '0');
    if (tile) { // the handle's matrix always has the static THCM pattern: structured product (spmv_tile.cu)
      Host* H = h->host;
      int long_row = -1;
      if (H->owns_rowintcon && H->rowintcon_local >= 0 &&
          H->rowptr[H->rowintcon_local + 1] - H->rowptr[H->rowintcon_local] > spmv_tile_long_row()) long_row = H->rowintcon_local;
      launch_spmv_tile(H->S, h->d_rowptr, h->d_col, h->d_val, d_x, d_y, long_row, h->stream);
      h->launches += long_row >= 0 ? 2 : 1;
    } else {
      SpmvPlan p{h->nblk, h->d_blkrow, spmv_nnz_cap(), h->d_blkp0, h->d_blkreg, h->spmv_tab};
      launch_spmv(p, h->d_rowptr, h->d_col, h->d_val, d_x, d_y, h->stream);
      h->launches++;
    }
  }
  if (int rc = check_launch(h, "spmv")) return rc;
  // Dense salinity-integral row on several GPUs (THCM.C:2494-2513): the owner's CSR row holds the columns it owns,
  // every other rank contributes sum coef*x over its own S unknowns through one all-reduced double.
  Host* H = h->host;
  if (h->dist && H->cfg.sres == 0 && H->rowintcon >= 0) {
    double* part = h->d_small + 1091;
    if (H->owns_rowintcon) { CU(cudaMemsetAsync(part, 0, sizeof(double), h->stream)); }
    else { launch_intcond_dot(h->d_coefS, d_x, h->nrows() / 6, part, h->stream); h->launches++; }
    if (int rc = allreduce(h, part, 1)) return rc;
    if (H->owns_rowintcon) { launch_intcond_add(part, (double)H->cfg.int_sign, d_y, H->rowintcon_local, h->stream); h->launches++; }
    return check_launch(h, "spmv integral row");
  }
  return 0;
}
} // namespace

extern "C" {

const char* iemic_b200_last_error(void) { return g_err.c_str(); }
const char* iemic_b200_version(void) { return "iemic_b200 0.1 (sm_100a)"; }
int iemic_b200_device_count(void) { int n = 0; if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; } return n; }

int thcm_b200_create_dist(const thcm_b200_config* cfg, const int* landm, int rank, int nranks, const void* nccl_id, thcm_b200_t** out) {
  if (!cfg || !landm || !out) return fail(IEMIC_B200_EINVAL, "null argument");
  if (nranks < 1 || rank < 0 || rank >= nranks) return fail(IEMIC_B200_EINVAL, "bad rank / nranks");
  if (iemic_b200_device_count() < 1) return fail(IEMIC_B200_ECUDA, "no CUDA device: the B200 path has no CPU fallback");
  thcm_b200* h = new thcm_b200();
  try {
    static_assert(sizeof(thcm_b200_config) == sizeof(Config), "config structs must stay layout compatible");
    Config c;
    std::memcpy(&c, cfg, sizeof(c));
    h->host = new Host(c, landm, rank, nranks);
  } catch (const std::exception& e) {
    delete h;
    return fail(IEMIC_B200_ENOTSUP, e.what());
  }
  if (nranks > 1) {
    h->dist = dist_create(*h->host, rank, nranks, nccl_id);
    if (!h->dist) { std::string m = dist_last_error(); delete h->host; delete h; return fail(IEMIC_B200_ENCCL, m); }
  }
  h->rb.nblk_max = red_blocks();
  cudaError_t e = cudaMalloc((void**)&h->rb.partial, (size_t)h->rb.nblk_max * 512 * sizeof(double));
  if (e == cudaSuccess) e = cudaMalloc((void**)&h->rb.ticket, 64);
  if (e == cudaSuccess) e = cudaMemset(h->rb.ticket, 0, 64);
  if (e == cudaSuccess) e = cudaMalloc((void**)&h->d_small, 1100 * sizeof(double));
  if (e == cudaSuccess) e = cudaMemset(h->d_small, 0, 1100 * sizeof(double));
  if (e != cudaSuccess) { delete h->host; delete h; return fail(IEMIC_B200_ECUDA, cudaGetErrorString(e)); }
  if (int rc = refresh_device(h)) { thcm_b200_destroy(h); return rc; }
  *out = h;
  return 0;
}
int thcm_b200_create(const thcm_b200_config* cfg, const int* landm, thcm_b200_t** out) {
  return thcm_b200_create_dist(cfg, landm, 0, 1, nullptr, out);
}
int thcm_b200_comm_unique_id(void* id) { return dist_unique_id(id) == 0 ? 0 : fail(IEMIC_B200_ENCCL, dist_last_error()); }

int thcm_b200_destroy(thcm_b200_t* h) {
  CHECK_H(h);
  cudaStreamSynchronize(h->stream);
  flush_timing(h);
  void* ptrs[] = {h->d_tab, h->d_sig, h->d_rowptr, h->d_col, h->d_val, h->d_diagB, h->d_frc, h->d_coefS, h->d_blkrow, h->d_blkp0, h->d_blkreg, h->d_xs, h->d_ys, h->d_dinfo,
                  h->d_msi, h->d_ftcell, h->d_cplf, h->rb.partial, h->rb.ticket, h->d_small};
  for (void* p : ptrs) if (p) cudaFree(p);
  gmres_free(h->g);
  if (h->copy_stream) { cudaStreamDestroy(h->copy_stream); cudaEventDestroy(h->ev_rhs); cudaEventDestroy(h->ev_copy); }
  if (h->up_stream) cudaStreamDestroy(h->up_stream);
  for (cudaEvent_t e : h->ev_slab) cudaEventDestroy(e);
  if (h->dist) dist_destroy(h->dist);
  delete h->host;
  delete h;
  return 0;
}
int thcm_b200_set_stream(thcm_b200_t* h, void* s) { CHECK_H(h); h->stream = (cudaStream_t)s; return 0; }
int thcm_b200_sync(thcm_b200_t* h) {
  CHECK_H(h);
  CU(cudaStreamSynchronize(h->stream));
  if (dist_peer_error(h->dist) != 0) return fail(IEMIC_B200_ENCCL, dist_last_error());
  return 0;
}

int thcm_b200_set_par(thcm_b200_t* h, int idx1, double value) {
  CHECK_H(h);
  if (idx1 < 1 || idx1 > NPAR) return fail(IEMIC_B200_EINVAL, "parameter index out of range");
  try { h->host->set_par(idx1, value); } catch (const std::exception& e) { return fail(IEMIC_B200_ENOTSUP, e.what()); }
  return 0;
}
// ---- row N2: coupling to an atmosphere / sea-ice model (Coupled Temperature = 1) ----
int thcm_b200_insert_field(thcm_b200_t* h, int which, const double* field) {
  CHECK_H(h);
  if (!field) return fail(IEMIC_B200_EINVAL, "insert_field: null field");
  if (which < 0 || which > THCM_B200_FIELD_SEAICE_G) return fail(IEMIC_B200_EINVAL, "insert_field: unknown field");
  try { h->host->insert_field(which, field); } catch (const std::exception& e) { return fail(IEMIC_B200_ENOTSUP, e.what()); }
  return 0;
}
int thcm_b200_set_atmos_parameters(thcm_b200_t* h, const thcm_b200_atmos_pars* pars) {
  CHECK_H(h);
  if (!pars) return fail(IEMIC_B200_EINVAL, "set_atmos_parameters: null argument");
  static_assert(sizeof(thcm_b200_atmos_pars) == 18 * sizeof(double), "atmos_pars must stay 18 doubles (usrc.F90:250-269)");
  try { h->host->set_atmos_parameters(reinterpret_cast<const double*>(pars)); } catch (const std::exception& e) { return fail(IEMIC_B200_ENOTSUP, e.what()); }
  return 0;
}
int thcm_b200_set_seaice_parameters(thcm_b200_t* h, const thcm_b200_seaice_pars* pars) {
  CHECK_H(h);
  if (!pars) return fail(IEMIC_B200_EINVAL, "set_seaice_parameters: null argument");
  static_assert(sizeof(thcm_b200_seaice_pars) == 7 * sizeof(double), "seaice_pars must stay 7 doubles (usrc.F90:303-311)");
  try { h->host->set_seaice_parameters(reinterpret_cast<const double*>(pars)); } catch (const std::exception& e) { return fail(IEMIC_B200_ENOTSUP, e.what()); }
  return 0;
}
int thcm_b200_getdeps(thcm_b200_t* h, double* out7) {
  CHECK_H(h);
  if (!out7) return fail(IEMIC_B200_EINVAL, "getdeps: null argument");
  h->host->getdeps(out7);
  return 0;
}
int thcm_b200_get_suno(thcm_b200_t* h, double* suno) { // THCM::getSunO: the shortwave profile on every surface point
  CHECK_H(h);
  if (!suno) return fail(IEMIC_B200_EINVAL, "get_suno: null argument");
  Host* H = h->host;
  for (int j = 1; j <= H->m; j++) for (int i = 1; i <= H->n; i++) suno[(size_t)(j - 1) * H->n + (i - 1)] = H->cpl.suno[j];
  return 0;
}
int thcm_b200_get_dftdm(thcm_b200_t* h, const double* d_x, double* dftdm) {
  CHECK_H(h);
  if (!d_x || !dftdm) return fail(IEMIC_B200_EINVAL, "get_dftdm: null argument");
  Host* H = h->host;
  const size_t nc = (size_t)H->nloc * H->mloc, nm = (size_t)H->n * H->m;
  if (H->cfg.coupled_T != 1) { std::fill(dftdm, dftdm + nc, 0.0); return 0; } // probe.F90:429
  if (int rc = refresh_device(h)) return rc;
  // tatm | albe | qatm | suno | out
  const size_t need = 3 * nm + (size_t)H->m + 2 + nc;
  if (!h->d_cplf) CU(cudaMalloc((void**)&h->d_cplf, need * sizeof(double)));
  const Host::Coupling& c = H->cpl;
  const std::vector<double>* f[3] = {&c.tatm, &c.albe, &c.qatm};
  const double* dev[3] = {nullptr, nullptr, nullptr};
  for (int q = 0; q < 3; q++)
    if (f[q]->size() == nm) {
      CU(cudaMemcpyAsync(h->d_cplf + q * nm, f[q]->data(), nm * sizeof(double), cudaMemcpyHostToDevice, h->stream));
      dev[q] = h->d_cplf + q * nm;
    }
  CU(cudaMemcpyAsync(h->d_cplf + 3 * nm, c.suno.data(), ((size_t)H->m + 2) * sizeof(double), cudaMemcpyHostToDevice, h->stream));
  DftdmArgs a;
  a.S = H->S; a.sig = h->d_sig; a.x = d_x; a.tatm = dev[0]; a.albe = dev[1]; a.qatm = dev[2]; a.suno = h->d_cplf + 3 * nm;
  a.c = H->dftdm_scalars(); // products in the order of the Fortran expression, formed by the FMA-free host code
  a.out = h->d_cplf + 3 * nm + H->m + 2;
  launch_dftdm(a, h->stream);
  h->launches++;
  if (int rc = check_launch(h, "dftdm")) return rc;
  CU(cudaMemcpyAsync(dftdm, a.out, nc * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  return 0;
}
// Ocean::getBlock(Atmosphere) (Ocean.C:1538-1664), T rows (Coupled Salinity = 1 is not on this path): for every ocean surface
// point three entries -dF_T/dT_atm, -dF_T/dalbedo, -dF_T/dq_atm in the reference's push order, rows and columns global
int thcm_b200_atmos_block(thcm_b200_t* h, int* rows, int* beg, int* jco, double* co, int* nrows_out) {
  CHECK_H(h);
  if (!rows || !beg || !jco || !co || !nrows_out) return fail(IEMIC_B200_EINVAL, "atmos_block: null argument");
  *nrows_out = h->host->atmos_block(rows, beg, jco, co);
  return 0;
}
// Ocean::getBlock(SeaIce) (Ocean.C:1667-1746), T rows: -dF_T/dM at the sea-ice mask unknown; dF_T/dM from the device (get_dftdm)
int thcm_b200_seaice_block(thcm_b200_t* h, const double* d_x, int* rows, int* beg, int* jco, double* co, int* nrows_out) {
  CHECK_H(h);
  if (!d_x || !rows || !beg || !jco || !co || !nrows_out) return fail(IEMIC_B200_EINVAL, "seaice_block: null argument");
  Host* H = h->host;
  if (H->nranks > 1) return fail(IEMIC_B200_ENOTSUP, "seaice_block: the derivative field is not gathered over the ranks yet (one GPU only)");
  const int N = H->n, M = H->m, L = H->l;
  std::vector<double> d((size_t)N * M);
  if (int rc = thcm_b200_get_dftdm(h, d_x, d.data())) return rc;
  *nrows_out = H->seaice_block(d.data(), rows, beg, jco, co);
  return 0;
}

// CouplingBlock (CouplingBlock.H): rows of the ocean that a neighbouring model's unknowns enter
struct coupling_b200 { thcm_b200* h = nullptr; int nrows = 0; int *d_rows = nullptr, *d_beg = nullptr, *d_jco = nullptr; double* d_co = nullptr; int64_t ncols = 0; };
int coupling_b200_create(thcm_b200_t* h, int nrows, const int* rows, const int* beg, const int* jco, const double* co, int64_t ncols, coupling_b200_t** out) {
  CHECK_H(h);
  if (!out || nrows < 0 || (nrows > 0 && (!rows || !beg || !jco || !co))) return fail(IEMIC_B200_EINVAL, "coupling_create: bad argument");
  Host* H = h->host;
  const int N = H->n, M = H->m;
  // keep the rows this rank owns, local row numbers; entries of a row sorted by column as Epetra's FillComplete leaves them
  std::vector<int> lr, lb(1, 0), lj; std::vector<double> lc;
  for (int r = 0; r < nrows; r++) {
    const int g = rows[r], X = g % 6, cell = g / 6, i = cell % N, j = (cell / N) % M, k = cell / (N * M);
    if (g < 0 || (size_t)g >= H->ndim) return fail(IEMIC_B200_EINVAL, "coupling_create: row outside the ocean");
    if (i + 1 < H->io0 || i + 1 >= H->io0 + H->nloc || j + 1 < H->jo0 || j + 1 >= H->jo0 + H->mloc) continue;
    std::vector<std::pair<int, double>> e;
    for (int p = beg[r]; p < beg[r + 1]; p++) {
      if (jco[p] < 0 || jco[p] >= ncols) return fail(IEMIC_B200_EINVAL, "coupling_create: column outside the neighbouring model");
      e.emplace_back(jco[p], co[p]);
    }
    std::stable_sort(e.begin(), e.end(), [](const std::pair<int, double>& a, const std::pair<int, double>& b) { return a.first < b.first; });
    lr.push_back((int)(6 * cell_index(H->S, i + 1, j + 1, k + 1)) + X);
    for (auto& q : e) { lj.push_back(q.first); lc.push_back(q.second); }
    lb.push_back((int)lj.size());
  }
  coupling_b200* b = new coupling_b200();
  b->h = h; b->nrows = (int)lr.size(); b->ncols = ncols;
  int rc = 0;
  if (b->nrows > 0) {
    if (!rc) rc = dev_upload(&b->d_rows, lr);
    if (!rc) rc = dev_upload(&b->d_beg, lb);
    if (!rc) rc = dev_upload(&b->d_jco, lj);
    if (!rc) rc = dev_upload(&b->d_co, lc);
  }
  if (rc) { coupling_b200_destroy(b); return rc; }
  *out = b;
  return 0;
}
int coupling_b200_apply(coupling_b200_t* b, const double* d_in, double* d_out, int accumulate) {
  if (!b || !b->h) return fail(IEMIC_B200_EINVAL, "coupling_apply: null block");
  if (!d_in || !d_out) return fail(IEMIC_B200_EINVAL, "coupling_apply: null vector");
  thcm_b200* h = b->h;
  if (!accumulate) CU(cudaMemsetAsync(d_out, 0, (size_t)h->nrows() * sizeof(double), h->stream)); // out = C in: zero outside the block's rows
  launch_coupling_apply(b->nrows, b->d_rows, b->d_beg, b->d_jco, b->d_co, d_in, d_out, accumulate, h->stream);
  h->launches++;
  return check_launch(h, "coupling_apply");
}
void coupling_b200_destroy(coupling_b200_t* b) {
  if (!b) return;
  void* ptrs[] = {b->d_rows, b->d_beg, b->d_jco, b->d_co};
  for (void* p : ptrs) if (p) cudaFree(p);
  delete b;
}
int thcm_b200_get_par(thcm_b200_t* h, int idx1, double* value) {
  CHECK_H(h);
  if (idx1 < 1 || idx1 > NPAR || !value) return fail(IEMIC_B200_EINVAL, "parameter index out of range");
  *value = h->host->get_par(idx1);
  return 0;
}
int thcm_b200_set_landmask(thcm_b200_t* h, const int* landm) {
  CHECK_H(h);
  try { h->host->set_landmask(landm); } catch (const std::exception& e) { return fail(IEMIC_B200_ENOTSUP, e.what()); }
  CU(cudaStreamSynchronize(h->stream));
  gmres_free(h->g);
  if (h->d_xs) { cudaFree(h->d_xs); cudaFree(h->d_ys); h->d_xs = h->d_ys = nullptr; }
  return refresh_device(h);
}
int thcm_b200_set_intcond_correction(thcm_b200_t* h, double corr) { CHECK_H(h); h->host->intCorrection = corr; return 0; }

int thcm_b200_sizes(thcm_b200_t* h, int64_t* nrows, int64_t* nvec, int64_t* nnz, int64_t* nglob) {
  CHECK_H(h);
  if (nrows) *nrows = h->nrows();
  if (nvec) *nvec = h->nvec();
  if (nnz) *nnz = (int64_t)h->host->col.size();
  if (nglob) *nglob = (int64_t)h->host->ndim;
  return 0;
}
int thcm_b200_block(thcm_b200_t* h, int* i0, int* j0, int* nloc, int* mloc, int* npN, int* npM) {
  CHECK_H(h);
  Host* H = h->host;
  if (i0) *i0 = H->S.io0 - 1;
  if (j0) *j0 = H->S.jo0 - 1;
  if (nloc) *nloc = H->S.nloc;
  if (mloc) *mloc = H->S.mloc;
  if (npN) *npN = H->npN;
  if (npM) *npM = H->npM;
  return 0;
}
int thcm_b200_global_ids(thcm_b200_t* h, int64_t* gid) {
  CHECK_H(h);
  const std::vector<long long>& g = h->host->gid;
  for (size_t q = 0; q < g.size(); q++) gid[q] = g[q];
  return 0;
}

int thcm_b200_halo_exchange(thcm_b200_t* h, double* d_x) {
  CHECK_H(h);
  if (!h->dist) return 0;
  if (dist_halo_exchange(h->dist, d_x, h->stream) != 0) return fail(IEMIC_B200_ENCCL, dist_last_error());
  h->launches += dist_halo_launches(h->dist);
  return 0;
}

} // extern "C" (re-opened below)

namespace {
int rhs_part(thcm_b200* h, const double* d_x, double* d_F, int part, cudaStream_t st = nullptr, int kz0 = 0, int nkz = 0) {
  AsmArgs a = asm_args(h, d_x);
  a.F = d_F; a.part = part; a.kz0 = kz0; a.nkz = nkz;
  if (!st) st = h->stream;
  {
    Timed t(h, 0, st);
    launch_rhs(a, st);
    h->launches++;
  }
  return check_launch(h, "asm_kernel<false,true> (residual)");
}
int rhs_epilogue(thcm_b200* h, const double* d_x, double* d_F) { // THCM.C:1006-1033
  Host* H = h->host;
  const bool intc = H->cfg.sres == 0 && H->rowintcon >= 0;
  if (intc) {
    launch_intcond_dot(h->d_coefS, d_x, h->nrows() / 6, h->d_small + 1029, h->stream);
    h->launches++;
    if (int rc = allreduce(h, h->d_small + 1029, 1)) return rc;
  }
  const int r = (intc && H->owns_rowintcon) ? H->rowintcon_local : -1;
  if (r >= 0 || H->rowPfix1_local >= 0 || H->rowPfix2_local >= 0) {
    launch_intcond_finish(h->d_small + 1029, (double)H->cfg.int_sign, H->intCorrection, d_F, r, H->rowPfix1_local, H->rowPfix2_local, h->stream);
    h->launches++;
  }
  return check_launch(h, "rhs epilogue");
}
int jac_part(thcm_b200* h, const double* d_x, int part, int kz0 = 0, int nkz = 0) {
  AsmArgs a = asm_args(h, d_x);
  a.part = part; a.kz0 = kz0; a.nkz = nkz;
  {
    Timed t(h, 1);
    launch_jac(a, h->stream);
    h->launches++;
  }
  return check_launch(h, "asm_kernel<true,false> (Jacobian)");
}
// Jacobian and residual of one state in ONE kernel: the residual operator is built from the tile the Jacobian pass staged.
int asm_part(thcm_b200* h, const double* d_x, double* d_F, int part, int kz0 = 0, int nkz = 0) {
  AsmArgs a = asm_args(h, d_x);
  a.F = d_F; a.part = part; a.kz0 = kz0; a.nkz = nkz;
  {
    Timed t(h, 4);
    launch_asm(a, h->stream);
    h->launches++;
  }
  return check_launch(h, "asm_kernel");
}
} // namespace

extern "C" {
int thcm_b200_rhs(thcm_b200_t* h, const double* d_x, double* d_F) {
  CHECK_H(h);
  if (int rc = refresh_device(h)) return rc;
  if (int rc = rhs_part(h, d_x, d_F, 0)) return rc;
  return rhs_epilogue(h, d_x, d_F);
}

int thcm_b200_jacobian(thcm_b200_t* h, const double* d_x) {
  CHECK_H(h);
  if (int rc = refresh_device(h)) return rc;
  return jac_part(h, d_x, 0);
}

// THCM::evaluateB (THCM.C:1195-1235): the mass-matrix diagonal alone (time steppers ask for it without a Jacobian)
int thcm_b200_evaluate_b(thcm_b200_t* h, double* d_diagB) {
  CHECK_H(h);
  if (int rc = refresh_device(h)) return rc;
  AsmArgs a = asm_args(h, nullptr);
  launch_diagb(a.S, h->d_sig, a.sp, d_diagB ? d_diagB : h->d_diagB, h->stream);
  h->launches++;
  return check_launch(h, "diagb");
}

int thcm_b200_exchange_mode(thcm_b200_t* h) {
  if (!h) return -1;
  return !h->dist ? 0 : (dist_p2p(h->dist) ? 2 : 1);
}

// Residual and/or Jacobian on one state with the halo exchange hidden behind the tiles that do not need it.
int thcm_b200_assemble(thcm_b200_t* h, double* d_x, double* d_F, int jacobian) {
  CHECK_H(h);
  if (int rc = refresh_device(h)) return rc;
  static const bool fused = !(getenv("IEMIC_ASM_FUSED") && getenv("IEMIC_ASM_FUSED")[0] == '0');
  if (!h->dist) {
    if (d_F && jacobian && fused) { if (int rc = asm_part(h, d_x, d_F, 0)) return rc; return rhs_epilogue(h, d_x, d_F); }
    if (d_F) { if (int rc = rhs_part(h, d_x, d_F, 0)) return rc; if (int rc = rhs_epilogue(h, d_x, d_F)) return rc; }
    if (jacobian) if (int rc = jac_part(h, d_x, 0)) return rc;
    return 0;
  }
  if (d_F && jacobian && fused && !getenv("IEMIC_ASM_PLAN")) { // halo exchange, then the fused kernel
    h->launches += dist_halo_launches(h->dist);
    if (dist_halo_exchange(h->dist, d_x, h->stream) != 0) return fail(IEMIC_B200_ENCCL, dist_last_error());
    if (int rc = asm_part(h, d_x, d_F, 0)) return rc;
    return rhs_epilogue(h, d_x, d_F);
  }
  // Plans (IEMIC_ASM_PLAN, default 1; measured on 8 B200 at 1 degree: 0.141 / 0.128 / 0.137 ms per assembly):
  //   0  halo exchange, Jacobian, residual, one after the other on the handle's stream
  //   1  halo exchange, then the residual kernel on a second stream beside the Jacobian kernel (fills its partial waves)
  //   2  halo exchange started, Jacobian tiles that read no halo cell, exchange finished, rim tiles beside the residual
  const int plan_env = getenv("IEMIC_ASM_PLAN") ? atoi(getenv("IEMIC_ASM_PLAN")) : 1;
  const int plan = (d_F && jacobian) ? plan_env : (plan_env == 2 && jacobian ? 2 : 0);
  if (plan != 0 && !h->copy_stream) {
    CU(cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
    CU(cudaEventCreateWithFlags(&h->ev_rhs, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&h->ev_copy, cudaEventDisableTiming));
  }
  h->launches += dist_halo_launches(h->dist);
  if (plan == 2) {
    if (dist_halo_begin(h->dist, d_x, h->stream) != 0) return fail(IEMIC_B200_ENCCL, dist_last_error());
    if (int rc = jac_part(h, d_x, 1)) return rc;
    if (dist_halo_end(h->dist, h->stream) != 0) return fail(IEMIC_B200_ENCCL, dist_last_error());
  } else if (dist_halo_exchange(h->dist, d_x, h->stream) != 0) return fail(IEMIC_B200_ENCCL, dist_last_error());
  if (plan != 0 && d_F) {
    CU(cudaEventRecord(h->ev_rhs, h->stream));
    CU(cudaStreamWaitEvent(h->copy_stream, h->ev_rhs, 0));
    if (int rc = rhs_part(h, d_x, d_F, 0, h->copy_stream)) return rc;
    CU(cudaEventRecord(h->ev_copy, h->copy_stream));
    if (int rc = jac_part(h, d_x, plan == 2 ? 2 : 0)) return rc;
    CU(cudaStreamWaitEvent(h->stream, h->ev_copy, 0));
  } else {
    if (jacobian) if (int rc = jac_part(h, d_x, plan == 2 ? 2 : 0)) return rc;
    if (d_F) if (int rc = rhs_part(h, d_x, d_F, 0)) return rc;
  }
  if (d_F) if (int rc = rhs_epilogue(h, d_x, d_F)) return rc;
  return 0;
}

int thcm_b200_csr(thcm_b200_t* h, const int** rp, const int** col, double** val, const double** diagB) {
  CHECK_H(h);
  if (int rc = refresh_device(h)) return rc;
  if (rp) *rp = h->d_rowptr;
  if (col) *col = h->d_col;
  if (val) *val = h->d_val;
  if (diagB) *diagB = h->d_diagB;
  return 0;
}
int thcm_b200_graph_host(thcm_b200_t* h, int* rowptr, int* col) {
  CHECK_H(h);
  Host* H = h->host;
  if (rowptr) std::memcpy(rowptr, H->rowptr.data(), H->rowptr.size() * sizeof(int));
  if (col) std::memcpy(col, H->col.data(), H->col.size() * sizeof(int));
  return 0;
}
int thcm_b200_values_host(thcm_b200_t* h, double* val, double* diagB) {
  CHECK_H(h);
  CU(cudaStreamSynchronize(h->stream));
  if (val) CU(cudaMemcpy(val, h->d_val, h->host->col.size() * sizeof(double), cudaMemcpyDeviceToHost));
  if (diagB) CU(cudaMemcpy(diagB, h->d_diagB, (size_t)h->nrows() * sizeof(double), cudaMemcpyDeviceToHost));
  return 0;
}
int thcm_b200_set_values_host(thcm_b200_t* h, const double* val) {
  CHECK_H(h);
  if (int rc = refresh_device(h)) return rc;
  CU(cudaStreamSynchronize(h->stream));
  CU(cudaMemcpy(h->d_val, val, h->host->col.size() * sizeof(double), cudaMemcpyHostToDevice));
  return 0;
}
int thcm_b200_forcing_host(thcm_b200_t* h, double* frc) {
  CHECK_H(h);
  h->host->refresh_host_forcing(); // after a parameter change only the device copy is current: rebuild the host one on demand
  std::memcpy(frc, h->host->frc.data(), h->host->frc.size() * sizeof(double));
  return 0;
}
int thcm_b200_intcond_host(thcm_b200_t* h, double* coeff, int* rowintcon) {
  CHECK_H(h);
  if (coeff) std::memcpy(coeff, h->host->intcond.data(), h->host->intcond.size() * sizeof(double));
  if (rowintcon) *rowintcon = h->host->rowintcon;
  return 0;
}

int thcm_b200_rhs_host(thcm_b200_t* h, const double* x, double* F) {
  CHECK_H(h);
  if (int rc = refresh_device(h)) return rc;
  const size_t nb = (size_t)h->nrows() * sizeof(double);
  CU(cudaMemcpyAsync(h->d_xs, x, nb, cudaMemcpyHostToDevice, h->stream));
  if (int rc = thcm_b200_halo_exchange(h, h->d_xs)) return rc;
  if (int rc = thcm_b200_rhs(h, h->d_xs, h->d_ys)) return rc;
  CU(cudaMemcpyAsync(F, h->d_ys, nb, cudaMemcpyDeviceToHost, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  return 0;
}
int thcm_b200_jacobian_host(thcm_b200_t* h, const double* x, double* val, double* diagB) {
  CHECK_H(h);
  if (int rc = refresh_device(h)) return rc;
  const size_t nb = (size_t)h->nrows() * sizeof(double);
  CU(cudaMemcpyAsync(h->d_xs, x, nb, cudaMemcpyHostToDevice, h->stream));
  if (int rc = thcm_b200_halo_exchange(h, h->d_xs)) return rc;
  if (int rc = thcm_b200_jacobian(h, h->d_xs)) return rc;
  if (val) CU(cudaMemcpyAsync(val, h->d_val, h->host->col.size() * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  if (diagB) CU(cudaMemcpyAsync(diagB, h->d_diagB, nb, cudaMemcpyDeviceToHost, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  return 0;
}

int thcm_b200_assemble_host(thcm_b200_t* h, const double* x, double* F, int jacobian) {
  CHECK_H(h);
  if (int rc = refresh_device(h)) return rc;
  const size_t nb = (size_t)h->nrows() * sizeof(double);
  if (!h->copy_stream) {
    CU(cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
    CU(cudaEventCreateWithFlags(&h->ev_rhs, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&h->ev_copy, cudaEventDisableTiming));
  }
  if (h->dist) { // several GPUs: the halo exchange needs the whole state on board
    CU(cudaMemcpyAsync(h->d_xs, x, nb, cudaMemcpyHostToDevice, h->stream));
    if (int rc = thcm_b200_assemble(h, h->d_xs, F ? h->d_ys : nullptr, jacobian)) return rc;
    if (F) CU(cudaMemcpyAsync(F, h->d_ys, nb, cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    return 0;
  }
  // One GPU: the state goes up in slabs of levels, the kernels of a slab start as soon as the slab above it has arrived,
  // and the residual of a slab goes down while later slabs are still going up (PCIe is full duplex): the call costs about
  // one upload instead of upload + kernels + download.
  Host* H = h->host;
  const int tk = asm_tile_levels(), ntz = (H->l + tk - 1) / tk;
  const int want = getenv("IEMIC_SLABS") ? std::max(1, atoi(getenv("IEMIC_SLABS"))) : 8;
  const int nslab = std::min(ntz, want), per = (ntz + nslab - 1) / nslab; // tile layers per slab
  const size_t lev = (size_t)6 * H->n * H->m;                          // unknowns per level
  if (!h->up_stream) CU(cudaStreamCreateWithFlags(&h->up_stream, cudaStreamNonBlocking));
  while ((int)h->ev_slab.size() < 2 * nslab) { cudaEvent_t e; CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming)); h->ev_slab.push_back(e); }
  CU(cudaEventRecord(h->ev_rhs, h->stream));             // earlier work on the handle's stream (e.g. a solve reading d_xs)
  CU(cudaStreamWaitEvent(h->up_stream, h->ev_rhs, 0));
  CU(cudaStreamWaitEvent(h->copy_stream, h->ev_rhs, 0));
  auto lev0 = [&](int s) { return std::min(H->l, s * per * tk); };      // first level (0-based) of slab s
  auto work = [&](int s) -> int {                                       // kernels of slab s; needs the levels of slab s+1 on board
    const int kz0 = s * per, nkz = std::min(per, ntz - kz0);
    if (nkz <= 0) return 0;
    if (jacobian && F) { if (int rc = asm_part(h, h->d_xs, h->d_ys, 0, kz0, nkz)) return rc; }
    else if (jacobian) { if (int rc = jac_part(h, h->d_xs, 0, kz0, nkz)) return rc; }
    if (F) {
      if (!jacobian) if (int rc = rhs_part(h, h->d_xs, h->d_ys, 0, nullptr, kz0, nkz)) return rc;
      const bool last = kz0 + nkz >= ntz;
      if (last) if (int rc = rhs_epilogue(h, h->d_xs, h->d_ys)) return rc; // integral condition / pressure rows live in the top level
      CU(cudaEventRecord(h->ev_slab[nslab + s], h->stream));
      CU(cudaStreamWaitEvent(h->copy_stream, h->ev_slab[nslab + s], 0));
      const size_t a = lev * lev0(s), b = lev * lev0(s + 1);
      if (b > a) CU(cudaMemcpyAsync(F + a, h->d_ys + a, (b - a) * sizeof(double), cudaMemcpyDeviceToHost, h->copy_stream));
    }
    return 0;
  };
  for (int s = 0; s < nslab; s++) {
    const size_t a = lev * lev0(s), b = lev * lev0(s + 1);
    if (b > a) CU(cudaMemcpyAsync(h->d_xs + a, x + a, (b - a) * sizeof(double), cudaMemcpyHostToDevice, h->up_stream));
    CU(cudaEventRecord(h->ev_slab[s], h->up_stream));
    CU(cudaStreamWaitEvent(h->stream, h->ev_slab[s], 0));
    if (s >= 1) if (int rc = work(s - 1)) return rc;
  }
  if (int rc = work(nslab - 1)) return rc;
  CU(cudaEventRecord(h->ev_copy, h->copy_stream));
  CU(cudaStreamWaitEvent(h->stream, h->ev_copy, 0));
  CU(cudaStreamSynchronize(h->stream));
  return 0;
}

int thcm_b200_spmv(thcm_b200_t* h, double* d_x, double* d_y) {
  CHECK_H(h);
  if (int rc = refresh_device(h)) return rc;
  return spmv_impl(h, d_x, d_y);
}
int thcm_b200_spmv_host(thcm_b200_t* h, const double* x, double* y) {
  CHECK_H(h);
  if (int rc = refresh_device(h)) return rc;
  const size_t nb = (size_t)h->nrows() * sizeof(double);
  CU(cudaMemcpyAsync(h->d_xs, x, nb, cudaMemcpyHostToDevice, h->stream));
  if (int rc = spmv_impl(h, h->d_xs, h->d_ys)) return rc;
  CU(cudaMemcpyAsync(y, h->d_ys, nb, cudaMemcpyDeviceToHost, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  return 0;
}

// ---- diagonal blocks: THCM scaling (THCM::RecomputeScaling, THCM.C:1694-1750), point-block Jacobi, row scaling ----
namespace {
int diag_args(thcm_b200* h, DiagArgs& a, int stored_rows) {
  if (int rc = refresh_device(h)) return rc;
  Host* H = h->host;
  if (!h->d_dinfo || h->dinfo_mask_version != H->mask_version) {
    std::vector<unsigned> info;
    try { info = H->diag_info(); } catch (const std::exception& e) { return fail(IEMIC_B200_EINVAL, e.what()); }
    if (int rc = dev_upload(&h->d_dinfo, info)) return rc;
    h->dinfo_mask_version = H->mask_version;
  }
  a.rowptr = h->d_rowptr; a.dinfo = h->d_dinfo; a.val = h->d_val; a.sig = h->d_sig; a.ncell = h->nrows() / 6;
  a.sp.rowintcon = H->owns_rowintcon ? H->rowintcon_local : -1;
  a.sp.rowPfix1 = H->rowPfix1_local; a.sp.rowPfix2 = H->rowPfix2_local;
  a.sp.side = h->d_small + 1070;
  a.stored_rows = stored_rows;
  return 0;
}
} // namespace

int thcm_b200_average_block(thcm_b200_t* h, double* db) {
  CHECK_H(h);
  if (!db) return fail(IEMIC_B200_EINVAL, "average_block: null output");
  DiagArgs a;
  if (int rc = diag_args(h, a, 0)) return rc;
  double* out = h->d_small + 1030;
  launch_average_block(a, h->rb.partial, h->rb.ticket + 8, out, h->stream);
  h->launches++;
  if (int rc = check_launch(h, "average_block")) return rc;
  if (int rc = allreduce(h, out, 37)) return rc;
  double s[37];
  CU(cudaMemcpyAsync(s, out, sizeof(s), cudaMemcpyDeviceToHost, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  const double nl = s[36];
  for (int ii = 0; ii < 6; ii++) for (int jj = 0; jj < 6; jj++) db[ii + 6 * jj] = nl > 0.0 ? s[ii * 6 + jj] / nl : s[ii * 6 + jj];
  return 0;
}

int thcm_b200_scal(const double* db, double* dr, double* dc) {
  if (!db || !dr || !dc) return fail(IEMIC_B200_EINVAL, "scal: null argument");
  return scal(db, dr, dc) ? 0 : fail(IEMIC_B200_EINVAL, "thcm scaling: diagonal block is singular up to working precision");
}

int thcm_b200_recompute_scaling(thcm_b200_t* h, double* d_rowscal, double* d_colscal, double* rs6, double* cs6) {
  CHECK_H(h);
  double db[36], dr[6], dc[6];
  if (int rc = thcm_b200_average_block(h, db)) return rc;
  if (!scal(db, dr, dc)) return fail(IEMIC_B200_EINVAL, "thcm scaling: diagonal block is singular up to working precision");
  ScalTab t;
  for (int q = 0; q < 6; q++) { t.r[q] = 1.0 / dr[q]; t.c[q] = 1.0 / dc[q]; } // Trilinos' are the inverse of THCM's, THCM.C:1726-1730
  double mean = 0.5 * (t.r[4] + t.r[5]); t.r[4] = mean; t.r[5] = mean;       // T and S scaled alike, THCM.C:1736-1745
  mean = 0.5 * (t.c[4] + t.c[5]); t.c[4] = mean; t.c[5] = mean;
  if (rs6) std::memcpy(rs6, t.r, sizeof(t.r));
  if (cs6) std::memcpy(cs6, t.c, sizeof(t.c));
  if (d_rowscal || d_colscal) {
    launch_scaling_fill(h->d_sig, h->nrows(), t, d_rowscal, d_colscal, h->stream);
    h->launches++;
    return check_launch(h, "scaling_fill");
  }
  return 0;
}

int thcm_b200_block_jacobi(thcm_b200_t* h, double* d_blockinv, int* nsingular) {
  CHECK_H(h);
  if (!d_blockinv) return fail(IEMIC_B200_EINVAL, "block_jacobi: null output");
  DiagArgs a;
  if (int rc = diag_args(h, a, 1)) return rc;
  int* d_cnt = reinterpret_cast<int*>(h->d_small + 1090);
  CU(cudaMemsetAsync(d_cnt, 0, sizeof(int), h->stream));
  launch_block_inverse(a, d_blockinv, d_cnt, h->stream);
  h->launches++;
  if (int rc = check_launch(h, "block_inverse")) return rc;
  if (nsingular) {
    CU(cudaMemcpyAsync(nsingular, d_cnt, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
  }
  return 0;
}

int thcm_b200_left_scale(thcm_b200_t* h, const double* d_scal) {
  CHECK_H(h);
  if (int rc = refresh_device(h)) return rc;
  launch_left_scale(h->d_rowptr, h->d_val, d_scal, h->nrows(), h->stream);
  h->launches++;
  return check_launch(h, "left_scale");
}

int vec_b200_multiply(thcm_b200_t* h, double a, const double* A, const double* B, double b, double* y) {
  CHECK_H(h);
  launch_multiply(a, A, B, b, y, h->nrows(), h->stream);
  h->launches++;
  return check_launch(h, "multiply");
}
int vec_b200_reciprocal(thcm_b200_t* h, const double* A, double* y) {
  CHECK_H(h);
  launch_reciprocal(A, y, h->nrows(), h->stream);
  h->launches++;
  return check_launch(h, "reciprocal");
}

// ---- vectors ----
int vec_b200_dot(thcm_b200_t* h, const double* a, const double* b, double* result) {
  CHECK_H(h);
  const PeerRed pr = red(h);
  launch_dot(a, b, h->nrows(), h->d_small + 1028, h->rb, h->stream, pr);
  h->launches++;
  if (int rc = allreduce(h, h->d_small + 1028, 1, pr)) return rc;
  CU(cudaMemcpyAsync(result, h->d_small + 1028, sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  if (dist_peer_error(h->dist) != 0) return fail(IEMIC_B200_ENCCL, dist_last_error());
  return 0;
}
// Epetra_Vector::NormInf: max over the owned parts of all ranks (slot [rank] of a zeroed array, summed across ranks)
int vec_b200_norminf(thcm_b200_t* h, const double* d_a, double* result) {
  CHECK_H(h);
  const int nr = h->host->nranks, me = h->host->rank;
  if (nr > 8) return fail(IEMIC_B200_ENOTSUP, "vec_b200_norminf: more than 8 ranks");
  double* slots = h->d_small + 1092; // 8 doubles
  CU(cudaMemsetAsync(slots, 0, 8 * sizeof(double), h->stream));
  launch_absmax(d_a, h->nrows(), slots + me, h->stream);
  h->launches++;
  if (int rc = check_launch(h, "absmax")) return rc;
  if (int rc = allreduce(h, slots, nr)) return rc;
  double v[8];
  CU(cudaMemcpyAsync(v, slots, nr * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  double m = 0.0;
  for (int q = 0; q < nr; q++) m = std::fmax(m, v[q]);
  *result = m;
  return 0;
}
// host <-> device copies of vector data ordered on the handle's stream (a user stream set with thcm_b200_set_stream does not
// synchronise with the legacy stream plain cudaMemcpy runs on)
int thcm_b200_upload(thcm_b200_t* h, double* d_dst, const double* src, int64_t n) {
  CHECK_H(h);
  CU(cudaMemcpyAsync(d_dst, src, (size_t)n * sizeof(double), cudaMemcpyHostToDevice, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  return 0;
}
int thcm_b200_download(thcm_b200_t* h, double* dst, const double* d_src, int64_t n) {
  CHECK_H(h);
  CU(cudaMemcpyAsync(dst, d_src, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  return 0;
}
int vec_b200_dot_noalloc(const double* d_a, const double* d_b, int64_t n, double* d_out, void* stream) {
  if (!d_a || !d_b || !d_out) return fail(IEMIC_B200_EINVAL, "vec_b200_dot_noalloc: null pointer");
  if (n < 0 || n > dot_noalloc_max_n())
    return fail(IEMIC_B200_EINVAL, "vec_b200_dot_noalloc: n = " + std::to_string((long long)n) + " is outside 0.." + std::to_string(dot_noalloc_max_n()) +
                                       " (the static reduction scratch of this entry point is sized for that)");
  const int e = launch_dot_noalloc(d_a, d_b, n, d_out, (cudaStream_t)stream);
  if (e != 0) return fail(IEMIC_B200_ECUDA, std::string("vec_b200_dot_noalloc: ") + cudaGetErrorString((cudaError_t)e));
  return 0;
}
int vec_b200_norm2(thcm_b200_t* h, const double* a, double* result) {
  int rc = vec_b200_dot(h, a, a, result);
  if (rc == 0) *result = std::sqrt(*result);
  return rc;
}
int vec_b200_update(thcm_b200_t* h, double a, const double* A, double b, double* y) {
  CHECK_H(h);
  launch_update(a, A, b, y, h->nrows(), h->stream);
  h->launches++;
  return check_launch(h, "update");
}
int vec_b200_scale(thcm_b200_t* h, double a, double* y) { return vec_b200_update(h, a, y, 0.0, y); }
int vec_b200_putscalar(thcm_b200_t* h, double a, double* y) {
  CHECK_H(h);
  launch_fill(a, y, h->nrows(), h->stream);
  h->launches++;
  return check_launch(h, "fill");
}

// ---- GMRES (GMRESSolver.H:81-255) ----
int gmres_b200_solve(thcm_b200_t* h, const double* d_b, double* d_x, int m, int maxit, double tol, int ortho, int prec,
                     const double* d_blockinv, int check_every, double* hist_out, int* iters_out) {
  CHECK_H(h);
  if (m < 1 || m > 500) return fail(IEMIC_B200_EINVAL, "GMRES restart length must be in 1..500");
  if (ortho != GMRES_B200_MGS && ortho != GMRES_B200_CGS2 && ortho != GMRES_B200_CGS2F) return fail(IEMIC_B200_EINVAL, "unknown orthogonalisation");
  if (prec == GMRES_B200_PREC_BLOCKJACOBI && !d_blockinv) return fail(IEMIC_B200_EINVAL, "block Jacobi needs d_blockinv");
  if (prec != GMRES_B200_PREC_NONE && prec != GMRES_B200_PREC_BLOCKJACOBI && prec != GMRES_B200_PREC_CALLBACK) return fail(IEMIC_B200_EINVAL, "unknown preconditioner");
  if (prec == GMRES_B200_PREC_CALLBACK && !h->precond) return fail(IEMIC_B200_EINVAL, "GMRES_B200_PREC_CALLBACK without gmres_b200_set_precond");
  if (int rc = refresh_device(h)) return rc;
  const bool flex = prec != GMRES_B200_PREC_NONE;
  if (int rc = gmres_alloc(h, m, flex)) return rc;
  GmresWs& g = h->g;
  const long long n = h->nrows(), ld = g.ld;
  cudaStream_t st = h->stream;
  double* sm = h->d_small;
  double *h1 = sm, *h2 = sm + 512, *nrm2 = sm + 1024, *hnext = sm + 1025, *bnorm = sm + 1026, *beta = sm + 1027;
  auto L = [&](int k = 1) { h->launches += k; };

  // normb
  { const PeerRed pr = red(h); launch_dot(d_b, d_b, n, nrm2, h->rb, st, pr); L();
  if (int rc = allreduce(h, nrm2, 1, pr)) return rc; }
  launch_sqrt_store(nrm2, bnorm, st); L();
  double normb = 0.0;
  CU(cudaMemcpyAsync(&normb, bnorm, sizeof(double), cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  if (normb == 0.0) { normb = 1.0; CU(cudaMemcpyAsync(bnorm, &normb, sizeof(double), cudaMemcpyHostToDevice, st)); }
  // r = b - A x  (x is copied into a halo-capable work vector first)
  CU(cudaMemcpyAsync(g.tmp, d_x, n * sizeof(double), cudaMemcpyDeviceToDevice, st));
  if (int rc = spmv_impl(h, g.tmp, g.r)) return rc;
  launch_residual(d_b, g.r, n, st); L();
  { const PeerRed pr = red(h); launch_dot(g.r, g.r, n, nrm2, h->rb, st, pr); L();
  if (int rc = allreduce(h, nrm2, 1, pr)) return rc; }
  double r2 = 0.0;
  CU(cudaMemcpyAsync(&r2, nrm2, sizeof(double), cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  double resid = std::sqrt(r2) / normb;
  int iter = 0, nh = 0, ret = 1;
  if (resid <= tol) { if (iters_out) *iters_out = 0; return 0; }

  std::vector<double> hbuf(m + 2);
  int guard_trips = 0;
  double cycle_start_resid = resid;
  static const bool guard_on = !(getenv("IEMIC_GMRES_GUARD") && getenv("IEMIC_GMRES_GUARD")[0] == '0');
  while (iter <= maxit) {
    // beta = ||r|| ; V0 = r / beta ; s = beta e1   (nrm2 holds ||r||^2)
    launch_cycle_init(nrm2, g.s, m, beta, st); L();
    launch_scale_copy(g.r, g.V, n, beta, 1, 0, st); L();
    int i, space = 0, done_in_cycle = 0;
    bool conv = false;
    // Small single-GPU problems: the whole cycle is ONE launch -- same algorithm, no launch and no host round trip per
    // iteration.  Tiny ones (the basis fits the shared memory of one cluster: 2-D MOC, the 8x8x4 box) run in gmres_cluster.cu,
    // L2-resident ones in the cooperative kernel of gmres_cycle.cu.  IEMIC_GMRES_CYCLE=0 keeps the kernel-per-operation path
    // below, =coop skips the cluster kernel, =cluster skips the cooperative one.
    const char* cyc_env = getenv("IEMIC_GMRES_CYCLE");
    const bool cyc_any = !(cyc_env && cyc_env[0] == '0') && !h->dist && prec != GMRES_B200_PREC_CALLBACK && iter <= maxit;
    const bool use_cluster = cyc_any && !(cyc_env && !strcmp(cyc_env, "coop")) && ortho != GMRES_B200_MGS && gmres_cluster_fits(n, m);
    const bool use_coop = cyc_any && !use_cluster && !(cyc_env && !strcmp(cyc_env, "cluster")) && g.cyc_part && ortho == GMRES_B200_CGS2F &&
                          n <= gmres_cycle_max_rows();
    if (use_cluster || use_coop) {
      Host* HH = h->host;
      int* flags = reinterpret_cast<int*>(g.cyc_flags);
      const int maxcols = std::min(m, maxit + 1 - iter); // the loop below runs while i < m && iter <= maxit
      CU(cudaMemsetAsync(flags, 0, 2 * sizeof(int), st));
      if (use_cluster) {
        ClusterArgs ca;
        ca.rowptr = h->d_rowptr; ca.col = h->d_col; ca.val = h->d_val; ca.binv = flex ? d_blockinv : nullptr;
        ca.V = g.V; ca.Z = g.Z; ca.ld = ld; ca.n = n; ca.m = m; ca.maxcols = maxcols; ca.fused = ortho == GMRES_B200_CGS2F ? 1 : 0;
        ca.tol = tol; ca.check = check_every > 0 ? 1 : 0; ca.out_cols = flags; ca.stop = flags + 1;
        ca.H = g.H; ca.ldh = m + 1; ca.cs = g.cs; ca.sn = g.sn; ca.s = g.s; ca.hist = g.hist; ca.bnorm = bnorm;
        ca.R = ca.RP = ca.KG = ca.npad = 0;
        Timed t(h, 3);
        if (int e = launch_gmres_cluster(ca, st)) return fail(IEMIC_B200_ECUDA, std::string("gmres cluster kernel: ") + (e > 0 ? cudaGetErrorString((cudaError_t)e) : "does not fit"));
      } else {
        CycleArgs ca;
        ca.rowptr = h->d_rowptr; ca.col = h->d_col; ca.val = h->d_val; ca.binv = flex ? d_blockinv : nullptr;
        ca.V = g.V; ca.Z = g.Z; ca.ld = ld; ca.n = n;
        ca.long_row = (HH->owns_rowintcon && HH->rowintcon_local >= 0 &&
                       HH->rowptr[HH->rowintcon_local + 1] - HH->rowptr[HH->rowintcon_local] > 256) ? HH->rowintcon_local : -1;
        ca.part = g.cyc_part; ca.ldp = 0; ca.m = m; ca.maxcols = maxcols;
        ca.tol = tol; ca.check = check_every > 0 ? 1 : 0;
        ca.out_cols = flags; ca.stop = flags + 1;
        ca.h1 = h1; ca.h2 = h2;
        GivensArgs& gc = ca.gv;
        gc.H = g.H; gc.ldh = m + 1; gc.cs = g.cs; gc.sn = g.sn; gc.s = g.s; gc.h1 = h1; gc.h2 = h2; gc.nrm2 = nrm2; gc.sq = 1; gc.pyth = 1;
        gc.bnorm = bnorm; gc.hist = g.hist; gc.hnext = hnext; gc.j = 0;
        Timed t(h, 3);
        if (int e = launch_gmres_cycle(ca, st)) return fail(IEMIC_B200_ECUDA, std::string("gmres cycle kernel: ") + cudaGetErrorString((cudaError_t)e));
      }
      L();
      int fl[2] = {0, 0}; // columns built, stopped on the tolerance
      CU(cudaMemcpyAsync(fl, flags, 2 * sizeof(int), cudaMemcpyDeviceToHost, st));
      CU(cudaStreamSynchronize(st));
      const int cols = fl[0];
      if (cols < 1 || cols > maxcols) return fail(IEMIC_B200_ECUDA, "gmres cycle kernel returned no column");
      conv = fl[1] != 0;
      // GMRESSolver.H:142-213 leaves its loop BEFORE counting the converged iteration
      i = cols; iter += conv ? cols - 1 : cols; done_in_cycle = cols; space = cols - 1;
    } else
    for (i = 0; i < m && iter <= maxit; i++, iter++) {
      double* vi = g.V + (size_t)i * ld;
      const bool fused = ortho == GMRES_B200_CGS2F;
      double* w = fused ? g.V + (size_t)(i + 1) * ld : g.w; // fused mode builds v_{i+1} in place
      if (flex) {
        double* zi = g.Z + (size_t)i * ld;
        if (prec == GMRES_B200_PREC_CALLBACK) { // model_.applyPrecon(v, z) of GMRESSolver.H:153-172, on the solver's stream
          if (int rc = check_launch(h, "gmres (before the preconditioner callback)")) return rc;
          const int prc = h->precond(h->precond_user, vi, zi, (int64_t)n, (void*)st);
          if (prc != 0) return fail(IEMIC_B200_ECALLBACK, "preconditioner callback returned " + std::to_string(prc));
        } else { launch_block_jacobi(d_blockinv, vi, zi, n / 6, st); L(); }
        if (int rc = spmv_impl(h, zi, w)) return rc;
      } else if (int rc = spmv_impl(h, vi, w)) return rc;
      {
        Timed t(h, 3);
        bool givens_done = false;
        if (fused) {
          // two reductions per step: the second one also carries <w,w>, the norm of the twice-projected vector follows
          // from Pythagoras, and the normalisation rides on the second update.
          { const PeerRed pr = red(h); launch_multi_dot(g.V, ld, i + 1, w, n, h1, h->rb, st, pr); L();
          if (int rc = allreduce(h, h1, i + 1, pr)) return rc; }
          // first update and second reduction share one pass over the basis while its tile fits in shared memory
          static const bool fuse23_on = !(getenv("IEMIC_GMRES_FUSE23") && getenv("IEMIC_GMRES_FUSE23")[0] == '0');
          const bool fuse23 = fuse23_on && i + 1 <= multi_axpy_dot_max_nv();
          if (!fuse23) { launch_multi_axpy(g.V, ld, i + 1, h1, w, n, nullptr, h->rb, st); L(); }
          { const PeerRed pr = red(h);
            // the Arnoldi bookkeeping rides in the finishing block of this reduction unless NCCL still has to sum h2
            const bool fuse_gv = !h->dist || pr.nranks > 0;
            GivensArgs gf;
            gf.H = g.H; gf.ldh = m + 1; gf.cs = g.cs; gf.sn = g.sn; gf.s = g.s; gf.h1 = h1; gf.h2 = h2; gf.nrm2 = nrm2; gf.sq = 1; gf.pyth = 1;
            gf.bnorm = bnorm; gf.hist = g.hist; gf.hnext = hnext; gf.j = i;
            if (fuse23) launch_multi_axpy_dot(g.V, ld, i + 1, h1, w, n, h2, h->rb, st, pr, fuse_gv ? &gf : nullptr);
            else launch_multi_dot(g.V, ld, i + 2, w, n, h2, h->rb, st, pr, fuse_gv ? &gf : nullptr);
            L();
            givens_done = fuse_gv;
          if (int rc = allreduce(h, h2, i + 2, pr)) return rc; }
        } else if (ortho == GMRES_B200_CGS2) {
          { const PeerRed pr = red(h); launch_multi_dot(g.V, ld, i + 1, g.w, n, h1, h->rb, st, pr); L();
          if (int rc = allreduce(h, h1, i + 1, pr)) return rc; }
          launch_multi_axpy(g.V, ld, i + 1, h1, g.w, n, nullptr, h->rb, st); L();
          { const PeerRed pr = red(h); launch_multi_dot(g.V, ld, i + 1, g.w, n, h2, h->rb, st, pr); L();
          if (int rc = allreduce(h, h2, i + 1, pr)) return rc; }
          { const PeerRed pr = red(h); launch_multi_axpy(g.V, ld, i + 1, h2, g.w, n, nrm2, h->rb, st, nullptr, pr); L();
          if (int rc = allreduce(h, nrm2, 1, pr)) return rc; }
        } else { // MGS, dependent chain exactly as GMRESSolver.H:177-181
          for (int k = 0; k <= i; k++) {
            { const PeerRed pr = red(h); launch_axpy_dot(k ? h1 + k - 1 : nullptr, k ? g.V + (size_t)(k - 1) * ld : nullptr, g.V + (size_t)k * ld, g.w, n, h1 + k, h->rb, st, pr); L();
            if (int rc = allreduce(h, h1 + k, 1, pr)) return rc; }
          }
          { const PeerRed pr = red(h); launch_axpy_dot(h1 + i, g.V + (size_t)i * ld, nullptr, g.w, n, nrm2, h->rb, st, pr); L();
          if (int rc = allreduce(h, nrm2, 1, pr)) return rc; }
        }
        GivensArgs ga;
        ga.H = g.H; ga.ldh = m + 1; ga.cs = g.cs; ga.sn = g.sn; ga.s = g.s; ga.h1 = h1; ga.h2 = (ortho != GMRES_B200_MGS) ? h2 : nullptr;
        ga.nrm2 = nrm2; ga.sq = 1; ga.pyth = fused ? 1 : 0; ga.bnorm = bnorm; ga.hist = g.hist; ga.hnext = hnext; ga.j = i;
        if (!givens_done) { launch_givens(ga, st); L(); }
        if (fused) { launch_multi_axpy(g.V, ld, i + 1, h2, w, n, nullptr, h->rb, st, hnext); L(); }
        else { launch_scale_copy(g.w, g.V + (size_t)(i + 1) * ld, n, hnext, 1, 0, st); L(); }
      }
      space = i; done_in_cycle = i + 1;
      if (check_every > 0 && ((i + 1) % check_every == 0)) {
        // look at the residuals of the last check_every iterations in order: the FIRST one below the tolerance is where the
        // reference (which tests every iteration) stops; columns computed beyond it are simply not used
        const int q0 = i + 1 - check_every;
        CU(cudaMemcpyAsync(hbuf.data() + q0, g.hist + q0, check_every * sizeof(double), cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        for (int q = q0; q <= i; q++)
          if (hbuf[q] < tol) { conv = true; iter -= i - q; space = q; done_in_cycle = q + 1; break; }
        if (conv) break;
      }
    }
    if (int rc = check_launch(h, "gmres cycle")) return rc;
    // residual history of this cycle
    CU(cudaMemcpyAsync(hbuf.data(), g.hist, done_in_cycle * sizeof(double), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    for (int q = 0; q < done_in_cycle; q++) if (hist_out) hist_out[nh++] = hbuf[q];
    if (done_in_cycle > 0) resid = hbuf[done_in_cycle - 1];
    if (!conv && resid < tol) conv = true;
    // Update: y = H^-1 s ; x += Z y   (GMRESSolver.H:292-313), built in the work vector first: the explicit residual
    // (GMRESSolver.H:228-242) decides whether it is taken as it is
    auto candidate = [&](int ncols) -> int {
      CU(cudaMemcpyAsync(g.tmp, d_x, n * sizeof(double), cudaMemcpyDeviceToDevice, st));
      launch_multi_axpy_plus(flex ? g.Z : g.V, ld, ncols, g.y, g.tmp, n, st); L();
      if (int rc = spmv_impl(h, g.tmp, g.r)) return rc;
      launch_residual(d_b, g.r, n, st); L();
      { const PeerRed pr = red(h); launch_dot(g.r, g.r, n, nrm2, h->rb, st, pr); L();
      if (int rc = allreduce(h, nrm2, 1, pr)) return rc; }
      CU(cudaMemcpyAsync(&r2, nrm2, sizeof(double), cudaMemcpyDeviceToHost, st));
      CU(cudaStreamSynchronize(st));
      return 0;
    };
    launch_backsolve(g.H, m + 1, g.s, space, g.y, st); L();
    if (int rc = candidate(space + 1)) return rc;
    const double expl_full = std::sqrt(r2) / normb;
    bool lost = guard_on && ortho != GMRES_B200_MGS && done_in_cycle > 0 && gmres_lost_accuracy(resid, expl_full, tol, 0.0);
    std::vector<double> hH, hs;
    if (lost) {
      // the recurrence and the explicit residual disagree: is it more than the rounding of this least-squares problem,
      // eps ||R||_F ||y|| / ||b||, explains?  (a constant floor would cut legitimate solves with tol below it short)
      hH.resize((size_t)(m + 1) * done_in_cycle); hs.resize(m + 1);
      CU(cudaMemcpyAsync(hH.data(), g.H, hH.size() * sizeof(double), cudaMemcpyDeviceToHost, st));
      CU(cudaMemcpyAsync(hs.data(), g.s, hs.size() * sizeof(double), cudaMemcpyDeviceToHost, st));
      CU(cudaStreamSynchronize(st));
      lost = gmres_lost_accuracy(resid, expl_full, tol, gmres_rounding_floor(hH.data(), m + 1, hs.data(), done_in_cycle, normb));
    }
    if (lost) {
      // the least-squares coefficients outgrew what double precision carries (singular Jacobian, b outside its range):
      // candidate two keeps the leading columns the estimate of gmres_guard.h still trusts; the better of the two is taken
      std::vector<double> ytr;
      const int keep = gmres_truncate(hH.data(), m + 1, hs.data(), hbuf.data(), done_in_cycle, normb, ytr);
      CU(cudaMemcpyAsync(g.y, ytr.data(), keep * sizeof(double), cudaMemcpyHostToDevice, st));
      if (int rc = candidate(keep)) return rc;
      const double expl_trunc = std::sqrt(r2) / normb;
      // The truncated update is the trustworthy one (the full one carries the exploded coefficients: on a singular Jacobian a
      // multiple of the null vector, whatever its residual says).  It must not be worse than where the cycle started, though:
      // then the full update is taken if IT improved on the start, else the iterate stays where it was.
      double expl_best = expl_trunc;
      if (!(expl_trunc <= cycle_start_resid)) {
        if (expl_full <= cycle_start_resid) { launch_backsolve(g.H, m + 1, g.s, space, g.y, st); L(); if (int rc = candidate(space + 1)) return rc; }
        else if (int rc = candidate(0)) return rc; // zero columns: x itself, with its residual vector rebuilt for the restart
        expl_best = std::sqrt(r2) / normb;
      }
      CU(cudaMemcpyAsync(d_x, g.tmp, n * sizeof(double), cudaMemcpyDeviceToDevice, st));
      // GMRESSolver.H:213-246 restarts from the explicit residual, which repairs ordinary recurrence drift: do that once.
      // Two trips in a row, or a cycle that did not reduce the explicit residual, is the breakdown: stop with code 2.
      const bool progress = expl_best < 0.999 * cycle_start_resid;
      if (expl_best <= tol) { ret = 0; break; }
      if (guard_trips >= 1 || !progress || iter > maxit) { ret = 2; break; }
      guard_trips++;
      cycle_start_resid = expl_best;
      continue; // g.r and nrm2 hold the residual of the iterate just stored
    }
    guard_trips = 0;
    cycle_start_resid = expl_full;
    CU(cudaMemcpyAsync(d_x, g.tmp, n * sizeof(double), cudaMemcpyDeviceToDevice, st));
    if (conv) { ret = 0; break; }
  }
  CU(cudaMemcpyAsync(&r2, nrm2, sizeof(double), cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  if (dist_peer_error(h->dist) != 0) return fail(IEMIC_B200_ENCCL, dist_last_error());
  if (hist_out) hist_out[nh] = std::sqrt(r2) / normb; // explicit residual after the last cycle (:241-242)
  if (iters_out) *iters_out = iter;
  if (int rc = check_launch(h, "gmres")) return rc;
  return ret;
}

int vec_b200_block_jacobi_apply(thcm_b200_t* h, const double* d_blockinv, const double* d_v, double* d_z) {
  CHECK_H(h);
  if (!d_blockinv || !d_v || !d_z) return fail(IEMIC_B200_EINVAL, "block_jacobi_apply: null pointer");
  launch_block_jacobi(d_blockinv, d_v, d_z, h->nrows() / 6, h->stream);
  h->launches++;
  return check_launch(h, "block_jacobi_apply");
}

int gmres_b200_set_precond(thcm_b200_t* h, gmres_b200_precond_fn fn, void* user) {
  CHECK_H(h);
  h->precond = fn; h->precond_user = user;
  return 0;
}

int gmres_b200_solve_host(thcm_b200_t* h, const double* b, double* x, int m, int maxit, double tol, int ortho, int prec,
                          const double* blockinv, int check_every, double* hist, int* iters) {
  CHECK_H(h);
  if (int rc = refresh_device(h)) return rc;
  const long long n = h->nrows();
  double *d_b = nullptr, *d_x = nullptr, *d_bi = nullptr;
  CU(cudaMalloc((void**)&d_b, n * sizeof(double)));
  CU(cudaMalloc((void**)&d_x, n * sizeof(double)));
  CU(cudaMemcpy(d_b, b, n * sizeof(double), cudaMemcpyHostToDevice));
  CU(cudaMemcpy(d_x, x, n * sizeof(double), cudaMemcpyHostToDevice));
  if (prec == GMRES_B200_PREC_BLOCKJACOBI && blockinv) {
    CU(cudaMalloc((void**)&d_bi, n * 6 * sizeof(double)));
    CU(cudaMemcpy(d_bi, blockinv, n * 6 * sizeof(double), cudaMemcpyHostToDevice));
  }
  int rc = gmres_b200_solve(h, d_b, d_x, m, maxit, tol, ortho, prec, d_bi, check_every, hist, iters);
  if (rc >= 0) cudaMemcpy(x, d_x, n * sizeof(double), cudaMemcpyDeviceToHost);
  cudaFree(d_b); cudaFree(d_x); if (d_bi) cudaFree(d_bi);
  return rc;
}

int thcm_b200_launch_count(thcm_b200_t* h, int64_t* launches) { CHECK_H(h); *launches = h->launches; return 0; }
int thcm_b200_timing_enable(thcm_b200_t* h, int on) { CHECK_H(h); if (!on) flush_timing(h); h->timing = on != 0; return 0; }
int thcm_b200_timing_reset(thcm_b200_t* h) {
  CHECK_H(h);
  flush_timing(h);
  for (int q = 0; q < 6; q++) { h->t_total[q] = 0; h->t_count[q] = 0; }
  return 0;
}
int thcm_b200_timing_get(thcm_b200_t* h, int which, double* total_ms, int64_t* count) {
  CHECK_H(h);
  if (which < 0 || which > 5) return fail(IEMIC_B200_EINVAL, "which must be 0..5");
  if (int rc = flush_timing(h)) return rc;
  if (total_ms) *total_ms = h->t_total[which];
  if (count) *count = h->t_count[which];
  return 0;
}

int iemic_b200_malloc(void** p, int64_t bytes) { CU(cudaMalloc(p, (size_t)bytes)); return 0; }
int iemic_b200_free(void* p) { CU(cudaFree(p)); return 0; }
int iemic_b200_memcpy_h2d(void* d, const void* s, int64_t bytes) { CU(cudaMemcpy(d, s, (size_t)bytes, cudaMemcpyHostToDevice)); return 0; }
int iemic_b200_memcpy_d2h(void* d, const void* s, int64_t bytes) { CU(cudaMemcpy(d, s, (size_t)bytes, cudaMemcpyDeviceToHost)); return 0; }

} // extern "C"
