// host_rowcheck.cpp -- TEST HARNESS ONLY.  Compiles the __host__ __device__ row code of the product
// (iemic_b200/csrc/thcm_rows.cuh, thcm_stage.cuh, thcm_cell.cuh) with the host compiler and evaluates it
// cell by cell, so that the arithmetic the CUDA kernels execute can be compared against the oracle on a
// machine without a GPU.  Nothing in the product links this file; the product path is CUDA only.
#include <cstring>
#include <stdexcept>
#include <cstdio>
#include "../iemic_b200/csrc/thcm_host.h"
#include "../iemic_b200/csrc/thcm_stage.cuh"
#include "../iemic_b200/csrc/thcm_cell.cuh"

using namespace thcm;

namespace {
struct HostFields { // accessor F: direct evaluation of the staging rule
  const double* x; const uint64_t* sig; const Scalars* S; int i, j, k;
  double u(int di, int dj, int dk) const { return stage_cell(x, sig, *S, i + di, j + dj, k + dk).u; }
  double v(int di, int dj, int dk) const { return stage_cell(x, sig, *S, i + di, j + dj, k + dk).v; }
  double w(int di, int dj, int dk) const { return stage_cell(x, sig, *S, i + di, j + dj, k + dk).w; }
  double t(int di, int dj, int dk) const { return stage_cell(x, sig, *S, i + di, j + dj, k + dk).t; }
  double s(int di, int dj, int dk) const { return stage_cell(x, sig, *S, i + di, j + dj, k + dk).s; }
};
struct HostRaw { // accessor X: raw neighbour state through shift() (assemble.F90:142-179)
  const double* x; const Scalars* S; int i, j, k;
  double raw(int di, int dj, int dk, int var) const {
    int i2 = i + di, j2 = j + dj, k2 = k + dk;
    if (S->periodic) { if (i2 == 0) i2 = S->n; else if (i2 == S->n + 1) i2 = 1; }
    return x[6 * cell_index(*S, i2, j2, k2) + var - 1];
  }
};
} // namespace

extern "C" {
void* hostcheck_create_dist(const Config* cfg, const int* landm, int rank, int nranks, char* err, int errlen) {
  try { return new Host(*cfg, landm, rank, nranks); }
  catch (const std::exception& e) { if (err) { strncpy(err, e.what(), errlen - 1); err[errlen - 1] = 0; } return nullptr; }
}
long long hostcheck_nrows(void* h) { return (long long)((Host*)h)->nrows_local(); }
long long hostcheck_nvec(void* h) { return (long long)((Host*)h)->nvec_local(); }
void hostcheck_gid(void* h, long long* g) { Host* H = (Host*)h; memcpy(g, H->gid.data(), sizeof(long long) * H->gid.size()); }
void hostcheck_block(void* h, int* out) { Host* H = (Host*)h; out[0] = H->io0 - 1; out[1] = H->jo0 - 1; out[2] = H->nloc; out[3] = H->mloc; out[4] = H->npN; out[5] = H->npM; }
int hostcheck_halo_npeers(void* h) { return (int)((Host*)h)->halo_plan().size(); }
void hostcheck_halo_peer(void* h, int idx, int* info) {
  auto plan = ((Host*)h)->halo_plan();
  info[0] = plan[idx].rank; info[1] = (int)plan[idx].send_cells.size(); info[2] = (int)plan[idx].recv_cells.size();
}
void hostcheck_halo_lists(void* h, int idx, int* send, int* recv) {
  auto plan = ((Host*)h)->halo_plan();
  memcpy(send, plan[idx].send_cells.data(), sizeof(int) * plan[idx].send_cells.size());
  memcpy(recv, plan[idx].recv_cells.data(), sizeof(int) * plan[idx].recv_cells.size());
}
void* hostcheck_create(const Config* cfg, const int* landm, char* err, int errlen) {
  try { return new Host(*cfg, landm); }
  catch (const std::exception& e) { if (err) { strncpy(err, e.what(), errlen - 1); err[errlen - 1] = 0; } return nullptr; }
}
void hostcheck_destroy(void* h) { delete (Host*)h; }
int hostcheck_set_par(void* h, int idx, double v, char* err, int errlen) {
  try { ((Host*)h)->set_par(idx, v); return 0; }
  catch (const std::exception& e) { if (err) { strncpy(err, e.what(), errlen - 1); err[errlen - 1] = 0; } return -1; }
}
// coupling to an atmosphere / sea-ice model
int hostcheck_insert(void* h, int which, const double* field, char* err, int errlen) {
  try { ((Host*)h)->insert_field(which, field); return 0; } catch (std::exception& e) { snprintf(err, errlen, "%s", e.what()); return 1; }
}
void hostcheck_set_atmos_parameters(void* h, const double* pars) { ((Host*)h)->set_atmos_parameters(pars); }
void hostcheck_set_seaice_parameters(void* h, const double* pars) { ((Host*)h)->set_seaice_parameters(pars); }
void hostcheck_getdeps(void* h, double* out) { ((Host*)h)->getdeps(out); }
void hostcheck_suno(void* h, double* out) { Host* H = (Host*)h; for (int j = 1; j <= H->m; j++) out[j - 1] = H->cpl.suno[j]; }
// m_probe::get_derivatives, coupled_T part: the loop of dftdm_kernel (coupled.cu) around the shared expression thcm_dftdm
void hostcheck_dftdm(void* h, const double* x, double* out) {
  Host* H = (Host*)h;
  const Scalars& S = H->S;
  const DftdmScal a = H->dftdm_scalars();
  const Host::Coupling& c = H->cpl;
  const size_t nm = (size_t)H->n * H->m;
  for (int j = S.jo0; j < S.jo0 + S.mloc; j++) for (int i = S.io0; i < S.io0 + S.nloc; i++) {
    const size_t ci = cell_index(S, i, j, S.l), g = (size_t)(j - 1) * S.n + (i - 1), t = (size_t)(j - S.jo0) * S.nloc + (i - S.io0);
    out[t] = 0.0;
    if (H->cfg.coupled_T != 1 || sbit(H->sig[ci], SIG_NOTOCEAN)) continue;
    out[t] = thcm_dftdm(a, x[6 * ci + 4], x[6 * ci + 5], c.tatm.size() == nm ? c.tatm[g] : 0.0, c.albe.size() == nm ? c.albe[g] : 0.0,
                        c.qatm.size() == nm ? c.qatm[g] : 0.0, c.suno[j]);
  }
}
int hostcheck_atmos_block(void* h, int* rows, int* beg, int* jco, double* co) { return ((Host*)h)->atmos_block(rows, beg, jco, co); }
int hostcheck_seaice_block(void* h, const double* dftdm, int* rows, int* beg, int* jco, double* co) { return ((Host*)h)->seaice_block(dftdm, rows, beg, jco, co); }
double hostcheck_get_par(void* h, int idx) { return ((Host*)h)->get_par(idx); }
long long hostcheck_nnz(void* h) { return (long long)((Host*)h)->col.size(); }
int hostcheck_rowintcon(void* h) { return ((Host*)h)->rowintcon; }
void hostcheck_graph(void* h, int* rowptr, int* col) {
  Host* H = (Host*)h;
  memcpy(rowptr, H->rowptr.data(), sizeof(int) * H->rowptr.size());
  memcpy(col, H->col.data(), sizeof(int) * H->col.size());
}
void hostcheck_frc(void* h, double* frc) { Host* H = (Host*)h; memcpy(frc, H->frc.data(), sizeof(double) * H->frc.size()); }
void hostcheck_sig(void* h, unsigned long long* sig) { Host* H = (Host*)h; memcpy(sig, H->sig.data(), 8 * H->sig.size()); }
void hostcheck_intcond(void* h, double* c) { Host* H = (Host*)h; memcpy(c, H->intcond.data(), sizeof(double) * H->nrows_local()); }

static void jac_with_side(void* h, const double* x, double* val, double* diagB, double* side);
void hostcheck_jac(void* h, const double* x, double* val, double* diagB) { jac_with_side(h, x, val, diagB, nullptr); }

// The product's scaling path on the CPU: Jacobian rows with the side buffer of the special rows, the diagonal-block
// positions of Host::diag_info, the sum over OCEAN cells exactly as average_block_kernel forms it (other order), thcm::scal.
// db: column-major 6x6 like the C ABI; rs/cs: the Trilinos-convention factors of thcm_b200_recompute_scaling.  Returns 0 if singular.
int hostcheck_scaling(void* h, const double* x, double* db, double* rs, double* cs) {
  Host* H = (Host*)h;
  std::vector<double> val(H->col.size()), diagB(H->nrows_local());
  double side[18] = {0};
  jac_with_side(h, x, val.data(), diagB.data(), side);
  std::vector<unsigned> info = H->diag_info();
  double sum[36] = {0}; double nl = 0.0;
  const int sp[3] = {H->owns_rowintcon ? H->rowintcon_local : -1, H->rowPfix1_local, H->rowPfix2_local};
  for (size_t r = 0; r < H->nrows_local(); r++) {
    if (sbit(H->sig[r / 6], SIG_NOTOCEAN)) continue;
    double d[6] = {0, 0, 0, 0, 0, 0};
    int special = -1;
    for (int q = 0; q < 3; q++) if ((int)r == sp[q]) special = q;
    if (special >= 0) for (int b = 0; b < 6; b++) d[b] = side[6 * special + b];
    else {
      const double* p = val.data() + H->rowptr[r] + (info[r] & 0x3ffffffu);
      for (int b = 0; b < 6; b++) if ((info[r] >> (26 + b)) & 1u) d[b] = *p++;
    }
    for (int b = 0; b < 6; b++) sum[(r % 6) * 6 + b] += d[b];
    if (r % 6 == 3) nl += 1.0;
  }
  for (int ii = 0; ii < 6; ii++) for (int jj = 0; jj < 6; jj++) db[ii + 6 * jj] = nl > 0 ? sum[ii * 6 + jj] / nl : sum[ii * 6 + jj];
  double dr[6], dc[6];
  if (!scal(db, dr, dc)) return 0;
  for (int q = 0; q < 6; q++) { rs[q] = 1.0 / dr[q]; cs[q] = 1.0 / dc[q]; }
  double mean = 0.5 * (rs[4] + rs[5]); rs[4] = mean; rs[5] = mean;
  mean = 0.5 * (cs[4] + cs[5]); cs[4] = mean; cs[5] = mean;
  return 1;
}
int hostcheck_scal(const double* db, double* dr, double* dc) { return scal(db, dr, dc) ? 1 : 0; }

static void jac_with_side(void* h, const double* x, double* val, double* diagB, double* side) {
  Host* H = (Host*)h;
  const Scalars& S = H->S;
  Tables T = H->host_tables();
  SpecialRows sp{H->owns_rowintcon ? H->rowintcon_local : -1, H->rowPfix1_local, H->rowPfix2_local, side};
  for (size_t q = 0; q < H->col.size(); q++) val[q] = 0.0;
  for (int k = 1; k <= S.l; k++) for (int j = S.jo0; j < S.jo0 + S.mloc; j++) for (int i = S.io0; i < S.io0 + S.nloc; i++) {
    size_t ci = cell_index(S, i, j, k);
    Cell c{i, j, k, H->sig[ci]};
    HostFields f{x, H->sig.data(), &S, i, j, k};
    int row = (int)(6 * ci);
    diagB[row + 0] = jac_row<UU>(S, T, c, f, sp, row + 0, val + H->rowptr[row + 0]);
    diagB[row + 1] = jac_row<VV>(S, T, c, f, sp, row + 1, val + H->rowptr[row + 1]);
    diagB[row + 2] = jac_row<WW>(S, T, c, f, sp, row + 2, val + H->rowptr[row + 2]);
    diagB[row + 3] = jac_row<PP>(S, T, c, f, sp, row + 3, val + H->rowptr[row + 3]);
    diagB[row + 4] = jac_row<TT>(S, T, c, f, sp, row + 4, val + H->rowptr[row + 4]);
    diagB[row + 5] = jac_row<SS>(S, T, c, f, sp, row + 5, val + H->rowptr[row + 5]);
  }
  if (H->owns_rowintcon) { // intcond_S (THCM.C:2121-2198)
    int r = H->rowintcon_local;
    for (int p = H->rowptr[r]; p < H->rowptr[r + 1]; p++) val[p] = H->cfg.int_sign * H->intcond[H->col[p]];
  }
}

void hostcheck_rhs(void* h, const double* x, double* Fo) {
  Host* H = (Host*)h;
  const Scalars& S = H->S;
  Tables T = H->host_tables();
  for (int k = 1; k <= S.l; k++) for (int j = S.jo0; j < S.jo0 + S.mloc; j++) for (int i = S.io0; i < S.io0 + S.nloc; i++) {
    size_t ci = cell_index(S, i, j, k);
    Cell c{i, j, k, H->sig[ci]};
    HostFields f{x, H->sig.data(), &S, i, j, k};
    HostRaw xr{x, &S, i, j, k};
    size_t row = 6 * ci;
    double mt = 0.0, ms = 0.0;
    if (S.vmix_on) thcm_mix(S, T, c, f.t(0, 0, -1), f.s(0, 0, -1), f.t(0, 0, 0), f.s(0, 0, 0), f.t(0, 0, 1), f.s(0, 0, 1), mt, ms);
    Fo[row + 0] = rhs_row<UU>(S, T, c, f, xr, H->frc[row + 0], 0.0);
    Fo[row + 1] = rhs_row<VV>(S, T, c, f, xr, H->frc[row + 1], 0.0);
    Fo[row + 2] = rhs_row<WW>(S, T, c, f, xr, H->frc[row + 2], 0.0);
    Fo[row + 3] = rhs_row<PP>(S, T, c, f, xr, H->frc[row + 3], 0.0);
    Fo[row + 4] = rhs_row<TT>(S, T, c, f, xr, H->frc[row + 4], mt);
    Fo[row + 5] = rhs_row<SS>(S, T, c, f, xr, H->frc[row + 5], ms);
  }
  if (H->cfg.sres == 0 && H->owns_rowintcon) { // THCM.C:1006-1018 (single rank: the dot is local)
    double d = 0.0;
    for (size_t q = 0; q < H->nrows_local(); q++) d += H->intcond[q] * x[q];
    Fo[H->rowintcon_local] = H->cfg.int_sign * (d - H->intCorrection);
  }
  if (H->rowPfix1_local >= 0) Fo[H->rowPfix1_local] = 0.0;
  if (H->rowPfix2_local >= 0) Fo[H->rowPfix2_local] = 0.0;
}
}
