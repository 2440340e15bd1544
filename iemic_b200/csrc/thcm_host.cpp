// thcm_host.cpp -- see thcm_host.h.  Compiled by the host compiler with -ffp-contract=off so that the
// coefficient tables carry exactly the roundings of the reference's (FMA-free) Fortran expressions.
#include "thcm_host.h"
#include "thcm_stage.cuh"
#include <cmath>
#include <cstring>
#include <stdexcept>
#include <algorithm>

namespace thcm {
namespace {
const double pi = 3.14159265358979323846; // par.F90:14
// usr.F90:129-160
const double omegadim = 7.292e-05, r0dim = 6.37e+06, udim = 0.1e+00, gdim = 9.8e+00, rhodim = 1.024e+03;
const double s0 = 35.0, cp0 = 4.2e+03, deltas = 1.0, t0 = 15, deltat = 1.0;
// atm.F90:5-19
const double rhoa = 1.25, ce = 1.3e-03, ch = 0.94 * ce, cpa = 1000., uw = 8.5, c0 = 0.43, sun0 = 1360., lv = 2.5e+06;
const double alpt1 = 2.93, alpt2 = 8.3e-02, alpt3 = 6.6e-04;
const double ah = 2.5e+05, av = 1.0e-03, kappah = 1.0e+03, kappav = 1.0e-04;
const double zmin = -1.0, zmax = 0.0;
enum { OCEAN = 0, LAND = 1, PERIO = 3 };

// grid.F90:62-95
double fz(double z, double qz) {
  if (qz > 1.0) return -1 + std::tanh(qz * (z + 1)) / std::tanh(qz);
  return z + (1. - qz) * z * (1 - z);
}
double dfdz(double z, double qz) {
  if (qz > 1.0) { double c = std::cosh(qz * (z + 1)); return qz / (std::tanh(qz) * c * c); }
  return 1.0 + (1. - qz) * (1. - 2. * z);
}
// spf.F90:795-853
double amh(double y, int ih) { return ih == 0 ? 1.0 : 1. + 10.0 * std::exp(-5 * y * y); }
double bmh(double y, int ih) { return ih == 0 ? 1.0 : 1.0 + 10.0 * std::exp(-5 * y * y); }
double bmhy(double y, int ih) { return ih == 0 ? 0.0 : -10. * 10.0 * y * std::exp(-5 * y * y); }
} // namespace

Host::Host(const Config& c, const int* landm_in, int rank_, int nranks_) : cfg(c), n(c.n), m(c.m), l(c.l), rank(rank_), nranks(nranks_) {
  if (n < 3 || m < 2 || l < 2) throw std::runtime_error("thcm_b200: grid too small (need n>=3, m>=2, l>=2)");
  if (nranks < 1 || rank < 0 || rank >= nranks) throw std::runtime_error("thcm_b200: bad rank / nranks");
  ncell = (size_t)n * m * l;
  ndim = 6 * ncell;
  decompose();
  if (105ull * (size_t)nloc * mloc * l > 0x7fffffffull) throw std::runtime_error("thcm_b200: sub-domain too large for 32-bit CSR offsets");
  version = 1; mask_version = 0; intCorrection = 0.0;
  make_grid();
  stpnt();
  atmos_coef();
  check_supported();
  set_landmask(landm_in); // signatures, graph, tables, forcing
}

// TRIOS::Domain::Decomp2D (TRIOS_Domain.C:81-153): factor nranks = npN*npM, row-major rank grid, remainders to the first ranks
void Host::decompose() {
  int t1 = nranks, t2 = 1;
  npM = t1; npN = t2;
  double r_min = 100;
  while (t1 > 0) {
    t2 = nranks / t1;
    double r = std::abs(m / t1 - n / t2);
    if (t1 * t2 == nranks && r <= r_min) { r_min = r; npM = t1; npN = t2; }
    t1--;
  }
  pidN = rank % npN; pidM = (rank - pidN) / npN;
  mloc = m / npM; nloc = n / npN;
  int Moff = pidM * (m / npM), Noff = pidN * (n / npN);
  const int remM = m % npM, remN = n % npN;
  if (pidM < remM) mloc++;
  if (pidN < remN) nloc++;
  Moff += std::min(remM, pidM); Noff += std::min(remN, pidN);
  io0 = Noff + 1; jo0 = Moff + 1;
  if (nloc < 2 || mloc < 2) throw std::runtime_error("thcm_b200: sub-domain smaller than 2x2 columns");
  if (npN > 1 && nloc > n - 2) throw std::runtime_error("thcm_b200: too few columns per rank for the periodic halo");
}

namespace {
struct Block { int io0, jo0, nloc, mloc; };
// block of rank r (TRIOS_Domain.C:120-153)
Block block_of(int n, int m, int npN, int npM, int r) {
  const int pN = r % npN, pM = (r - pN) / npN;
  Block b;
  b.mloc = m / npM; b.nloc = n / npN;
  int Moff = pM * (m / npM), Noff = pN * (n / npN);
  const int remM = m % npM, remN = n % npN;
  if (pM < remM) b.mloc++;
  if (pN < remN) b.nloc++;
  Moff += std::min(remM, pM); Noff += std::min(remN, pN);
  b.io0 = Noff + 1; b.jo0 = Moff + 1;
  return b;
}
// global (fi,fj,fk) of stored cell c of block b; false for an unused ring slot
bool stored_cell_of(const Block& b, int n, int m, int l, bool wrap, size_t c, int& fi, int& fj, int& fk) {
  const size_t nown = (size_t)b.nloc * b.mloc * l;
  int li, lj;
  if (c < nown) {
    li = (int)(c % b.nloc); lj = (int)((c / b.nloc) % b.mloc); fk = (int)(c / ((size_t)b.nloc * b.mloc)) + 1;
  } else {
    const int R = 2 * (b.nloc + 2) + 2 * b.mloc;
    const size_t q = c - nown;
    fk = (int)(q / R) + 1;
    const int r = (int)(q % R);
    if (r < b.nloc + 2) { lj = -1; li = r - 1; }
    else if (r < 2 * (b.nloc + 2)) { lj = b.mloc; li = r - (b.nloc + 2) - 1; }
    else if (r < 2 * (b.nloc + 2) + b.mloc) { li = -1; lj = r - 2 * (b.nloc + 2); }
    else { li = b.nloc; lj = r - 2 * (b.nloc + 2) - b.mloc; }
  }
  fi = b.io0 + li; fj = b.jo0 + lj;
  if (fj < 1 || fj > m) return false;
  if (fi < 1 || fi > n) {
    if (!wrap) return false; // with one rank in x the kernels wrap by themselves
    fi = fi < 1 ? fi + n : fi - n;
  }
  return true;
}
} // namespace

bool Host::stored_cell(size_t c, int& fi, int& fj, int& fk) const {
  Block b{io0, jo0, nloc, mloc};
  return stored_cell_of(b, n, m, l, cfg.periodic && npN > 1, c, fi, fj, fk);
}

std::vector<Host::HaloPeer> Host::halo_plan() const {
  std::vector<HaloPeer> plan;
  if (nranks == 1) return plan;
  const bool wrap = cfg.periodic && npN > 1;
  const Block me{io0, jo0, nloc, mloc};
  auto owner = [&](int fi, int fj) {
    for (int r = 0; r < nranks; r++) {
      Block b = block_of(n, m, npN, npM, r);
      if (fi >= b.io0 && fi < b.io0 + b.nloc && fj >= b.jo0 && fj < b.jo0 + b.mloc) return r;
    }
    return -1;
  };
  std::vector<HaloPeer> byrank(nranks);
  for (int r = 0; r < nranks; r++) byrank[r].rank = r;
  Scalars Sm; std::memset(&Sm, 0, sizeof(Sm));
  Sm.n = n; Sm.m = m; Sm.l = l; Sm.periodic = cfg.periodic != 0; Sm.io0 = io0; Sm.jo0 = jo0; Sm.nloc = nloc; Sm.mloc = mloc; Sm.nown = (long long)nloc * mloc * l;
  // what I receive: my ring cells, in my ring order
  {
    const size_t nown = (size_t)nloc * mloc * l, R = 2 * (nloc + 2) + 2 * mloc;
    for (size_t c = nown; c < nown + R * l; c++) {
      int fi, fj, fk;
      if (!stored_cell_of(me, n, m, l, wrap, c, fi, fj, fk)) continue;
      int o = owner(fi, fj);
      if (o >= 0 && o != rank) byrank[o].recv_cells.push_back((int)c);
      else if (o == rank) byrank[o].recv_cells.push_back((int)c); // periodic image of my own cell (npN == 2 corner cases)
    }
  }
  // what I send: for every other rank, its ring cells that I own, in ITS ring order
  for (int r = 0; r < nranks; r++) {
    Block b = block_of(n, m, npN, npM, r);
    const size_t nown = (size_t)b.nloc * b.mloc * l, R = 2 * (b.nloc + 2) + 2 * b.mloc;
    for (size_t c = nown; c < nown + R * l; c++) {
      int fi, fj, fk;
      if (!stored_cell_of(b, n, m, l, wrap, c, fi, fj, fk)) continue;
      if (fi >= io0 && fi < io0 + nloc && fj >= jo0 && fj < jo0 + mloc) byrank[r].send_cells.push_back((int)cell_index(Sm, fi, fj, fk));
    }
  }
  for (int r = 0; r < nranks; r++)
    if (!byrank[r].send_cells.empty() || !byrank[r].recv_cells.empty()) plan.push_back(byrank[r]);
  return plan;
}

// grid.F90:2-60
void Host::make_grid() {
  x.assign(n + 1, 0); y.assign(m + 2, 0); z.assign(l + 1, 0); xu.assign(n + 1, 0); yv.assign(m + 1, 0);
  zw.assign(l + 1, 0); dfzT.assign(l + 1, 0); dfzW.assign(l + 1, 0);
  dx = (cfg.xmax - cfg.xmin) / n;
  dy = (cfg.ymax - cfg.ymin) / m;
  dz = (zmax - zmin) / l;
  for (int i = 1; i <= n; i++) { x[i] = ((double)i - 0.5) * dx + cfg.xmin; xu[i] = ((double)i) * dx + cfg.xmin; }
  xu[0] = cfg.xmin;
  for (int j = 1; j <= m; j++) { y[j] = ((double)j - 0.5) * dy + cfg.ymin; yv[j] = ((double)j) * dy + cfg.ymin; }
  y[0] = y[1] - dy; y[m + 1] = y[m] + dy; yv[0] = cfg.ymin;
  for (int k = 1; k <= l; k++) {
    double ze = ((double)k - 0.5) * dz + zmin, zwe = ((double)k) * dz + zmin;
    z[k] = fz(ze, cfg.qz); zw[k] = fz(zwe, cfg.qz);
    dfzT[k] = dfdz(ze, cfg.qz); dfzW[k] = dfdz(zwe, cfg.qz);
  }
  zw[0] = zmin; dfzW[0] = dfdz(zmin, cfg.qz);
}

// usrc.F90:1175-1213 (atmos_coef) with QTnd, QSnd of init (usrc.F90:117-120)
void Host::atmos_coef() {
  const double dzne = dz * dfzT[l];
  cpl.QTnd = r0dim / (udim * cp0 * rhodim * cfg.hdim * dzne);
  cpl.QSnd = s0 * r0dim / (deltas * udim * cfg.hdim * dzne);
  const double muoa = rhoa * ch * cpa * uw;
  cpl.Os = sun0 * c0 / 4 * cpl.QTnd;
  cpl.Ooa = muoa * cpl.QTnd;
  cpl.nus = 0.0; cpl.lvsc = 0.0;
  cpl.suno.assign(m + 2, 0.0);
  for (int j = 1; j <= m; j++) {
    const double sj = std::sin(y[j]);
    cpl.suno[j] = cpl.Os * (1 - .482 * (3 * (sj * sj) - 1.) / 2.);
  }
}

// inserts.F90:12-166
void Host::insert_field(int which, const double* field) {
  const bool cT = cfg.coupled_T == 1, cS = cfg.coupled_S == 1;
  std::vector<double>* dst = nullptr;
  switch (which) {
    case F_TATM: if (cT) dst = &cpl.tatm; break;
    case F_QATM: if (cT || cS) dst = &cpl.qatm; break;
    case F_ALBE: if (cT) dst = &cpl.albe; break;
    case F_MSI: dst = &cpl.msi; break;
    case F_PATM: case F_GSI: if (cS) throw std::runtime_error("thcm_b200: Coupled Salinity=1 is not on the B200 path yet"); return; // ignored by the Fortran too
    case F_QSA: return; // enters the salinity forcing only (forcing.F90:181-184), which needs Coupled Salinity=1
    default: throw std::runtime_error("thcm_b200: unknown coupling field");
  }
  if (!dst) return; // the Fortran ignores the insert as well when the flag is off
  dst->assign(field, field + (size_t)n * m);
}
void Host::coupled_changed() {
  build_tables();
  if (forcing_on_device) { forcing_profiles(); frc_host_valid = false; }
  else build_forcing(true);
  version++;
}
void Host::set_atmos_parameters(const double* p) {
  cpl.qdim = p[1]; cpl.nuq = p[2]; cpl.eta = p[3]; cpl.dqso = p[4]; cpl.eo0 = p[7]; cpl.albe0 = p[12]; cpl.albed = p[13];
  cpl.nus = par[P_COMB] * par[P_SALT] * cpl.eta * cpl.qdim * cpl.QSnd;
  cpl.lvsc = par[P_COMB] * par[P_TEMP] * rhodim * lv * cpl.QTnd;
  coupled_changed();
}
void Host::set_seaice_parameters(const double* p) {
  cpl.zeta = p[0]; cpl.a0 = p[1]; cpl.Lf = p[2]; cpl.Qvar = p[5]; cpl.Q0 = p[6];
  coupled_changed();
}
void Host::getdeps(double* o) const {
  o[0] = cpl.Ooa; o[1] = cpl.Os; o[2] = cpl.nus; o[3] = cpl.eta; o[4] = cpl.lvsc; o[5] = cpl.qdim;
  o[6] = par[P_COMB] * par[P_SALT] * cpl.QSnd;
}

DftdmScal Host::dftdm_scalars() {
  DftdmScal a;
  cpl.nus = par[P_COMB] * par[P_SALT] * cpl.eta * cpl.qdim * cpl.QSnd; // probe.F90:414: get_derivatives refreshes the module variable
  a.cs = par[P_COMB] * par[P_SUNP];
  a.albe0 = cpl.albe0; a.albed = cpl.albed; a.Ooa = cpl.Ooa;
  a.leq = cpl.lvsc * cpl.eta * cpl.qdim;
  a.dq = deltat / cpl.qdim * cpl.dqso;
  a.lveo = cpl.lvsc * cpl.eo0;
  a.qtz = cpl.QTnd * cpl.zeta;
  a.a0 = cpl.a0; a.s0 = s0; a.t0 = t0;
  return a;
}
int Host::atmos_block(int* rows, int* beg, int* jco, double* co) const {
  const int N = n, M = m, L = l;
  const double comb = par[P_COMB], sunp = par[P_SUNP];
  int nr = 0, el = 0;
  beg[0] = 0;
  if (cfg.coupled_T != 1) return 0;
  for (int j = 0; j != M; ++j) for (int i = 0; i != N; ++i) {
    if (lm(i + 1, j + 1, L) != 0) continue; // (*landmask_.global_surface)[sr] == 0
    const size_t sr = (size_t)j * N + i;
    const double Msi = cpl.msi.size() == (size_t)N * M ? cpl.msi[sr] : 0.0, Ssw = cpl.suno[j + 1];
    const double dTFT = cpl.Ooa * (1.0 - Msi);
    const double dAFT = -comb * sunp * Ssw * cpl.albed * (1.0 - Msi);
    const double dQFT = cpl.lvsc * cpl.eta * cpl.qdim * (1.0 - Msi);
    rows[nr] = 6 * ((L - 1) * N * M + N * j + i) + TT - 1;
    co[el] = -dTFT; jco[el++] = 3 * (N * j + i) + 0; // ATMOS_TT_ (Atmosphere::interface_row, dof 3)
    co[el] = -dAFT; jco[el++] = 3 * (N * j + i) + 2; // ATMOS_AA_
    co[el] = -dQFT; jco[el++] = 3 * (N * j + i) + 1; // ATMOS_QQ_
    beg[++nr] = el;
  }
  return nr;
}
int Host::seaice_block(const double* dftdm, int* rows, int* beg, int* jco, double* co) const {
  const int N = n, M = m, L = l;
  int nr = 0, el = 0;
  beg[0] = 0;
  if (cfg.coupled_T != 1) return 0;
  for (int j = 0; j != M; ++j) for (int i = 0; i != N; ++i) {
    if (lm(i + 1, j + 1, L) != 0) continue;
    rows[nr] = 6 * ((L - 1) * N * M + N * j + i) + TT - 1;
    co[el] = -dftdm[(size_t)j * N + i]; jco[el++] = 4 * (N * j + i) + 3 - 1; // SEAICE_MM_ = 3, dof 4 (SeaIce::find_row0)
    beg[++nr] = el;
  }
  return nr;
}

// usrc.F90:1136-1172 and mix_imp.f:92-107
void Host::stpnt() {
  for (int i = 0; i <= NPAR; i++) par[i] = 0.0;
  const double hdim = cfg.hdim;
  par[P_AL_T] = 0.1 / (2 * omegadim * rhodim * hdim * udim * dz * dfzT[l]);
  par[P_RAYL] = cfg.alphaT * gdim * hdim / (2 * omegadim * udim * r0dim);
  par[P_EK_V] = av / (2 * omegadim * hdim * hdim);
  par[P_EK_H] = ah / (2 * omegadim * r0dim * r0dim);
  par[P_ROSB] = udim / (2 * omegadim * r0dim);
  par[P_PE_H] = kappah / (udim * r0dim);
  par[P_PE_V] = kappav * r0dim / (udim * hdim * hdim);
  par[P_P_VC] = 2.5e+04 * par[P_PE_V];
  par[P_LAMB] = cfg.alphaS / cfg.alphaT;
  par[P_BIOT] = r0dim / (75. * 3600. * 24. * udim);
  par[P_ALPC] = 1.0; par[P_ENER] = 1.0e+02;
  par[P_SPL1] = 2.0e+03; par[P_SPL2] = 0.01;
  if (cfg.vmix == 0) { par[P_MIXP] = 0.0; par[P_P_VC] = 0.0; par[P_ALPC] = 1.0; par[P_ENER] = 1.0e+2; par[P_MKAP] = 0.0; }
}

// What the B200 path implements; everything else fails loudly instead of silently diverging.
void Host::check_supported() const {
  if (cfg.vmix != 0 && cfg.vmix != 1) throw std::runtime_error("thcm_b200: Mixing must be 0 or 1 (adaptive partitioning, vmix=2, is not on the B200 path)");
  if (cfg.iza != 2 || cfg.ite != 1 || cfg.its != 1) throw std::runtime_error("thcm_b200: only idealized forcing (Wind Forcing Type=2, Levitus T=1, Levitus S=1) is supported");
  if (cfg.coupled_S != 0) throw std::runtime_error("thcm_b200: Coupled Salinity=1 is not on the B200 path yet");
  if (std::abs(cfg.int_sign) != 1) throw std::runtime_error("thcm_b200: invalid integral sign"); // THCM.C:277-280
}

void Host::set_par(int idx1, double value) {
  if (idx1 >= 1 && idx1 <= NPAR) par[idx1] = value;
  // neutral physics / GM / consistent mixing would add 27-point T,S couplings outside the static graph
  if (cfg.vmix != 0 && (par[P_MIXP] * par[P_PE_H] != 0.0 || par[P_MKAP] * par[P_PE_H] != 0.0 || (1.0 - par[P_ALPC]) * par[P_ENER] * par[P_PE_V] != 0.0))
    throw std::runtime_error("thcm_b200: MIXP/MKAP/ALPC settings enable neutral-physics or consistent mixing, which the B200 path does not implement");
  build_tables();
  // forcing.F90 depends on the parameters through the surface terms only (the body term is an exact zero).  With a device
  // attached only the per-latitude profiles are rebuilt here; the rows are filled by a kernel and the host copy goes stale.
  if (forcing_on_device) { forcing_profiles(); frc_host_valid = false; }
  else build_forcing(true);
  version++;
}

// init (usrc.F90:79-107) and set_landmask (usrc.F90:336-395) border rules
void Host::set_landmask(const int* landm_in) {
  landm.assign((size_t)(n + 2) * (m + 2) * (l + 2), OCEAN);
  const bool per = cfg.periodic != 0;
  auto L = [&](int i, int j, int k) -> int& { return landm[(size_t)i + (size_t)(n + 2) * (j + (size_t)(m + 2) * k)]; };
  size_t pos = 0;
  for (int k = 0; k <= l + 1; k++) for (int j = 0; j <= m + 1; j++) for (int i = 0; i <= n + 1; i++) {
    int v = landm_in[pos++];
    if (!per && v == PERIO) v = OCEAN;
    L(i, j, k) = v;
  }
  for (int i = 1; i <= n; i++) for (int j = 1; j <= m; j++) for (int k = l; k >= 2; k--) // land inversion fix
    if (L(i, j, k) == LAND && L(i, j, k - 1) == OCEAN) L(i, j, k - 1) = LAND;
  for (int k = 0; k <= l + 1; k++) for (int j = 0; j <= m + 1; j++) if (!per) { L(0, j, k) = LAND; L(n + 1, j, k) = LAND; }
  for (int k = 0; k <= l + 1; k++) for (int i = 0; i <= n + 1; i++) { L(i, 0, k) = LAND; L(i, m + 1, k) = LAND; }
  for (int j = 0; j <= m + 1; j++) for (int i = 0; i <= n + 1; i++) { L(i, j, 0) = LAND; L(i, j, l + 1) = LAND; }
  for (int k = 1; k <= l; k++) for (int j = 1; j <= m; j++) for (int i = 1; i <= n; i++)
    if (L(i, j, k) != OCEAN && L(i, j, k) != LAND) throw std::runtime_error("thcm_b200: interior land mask values must be 0 (ocean) or 1 (land)");
  // storage descriptors first: everything below indexes through cell_index()
  std::memset(&S, 0, sizeof(S));
  S.n = n; S.m = m; S.l = l; S.periodic = cfg.periodic != 0;
  S.io0 = io0; S.jo0 = jo0; S.nloc = nloc; S.mloc = mloc; S.nown = (long long)nloc * mloc * l;
  const size_t nstored = (size_t)S.nown + (nranks > 1 ? (size_t)l * ring_size(S) : 0);
  auto owned = [&](int gi, int gj) { return gi >= io0 && gi < io0 + nloc && gj >= jo0 && gj < jo0 + mloc; };
  auto lrow = [&](int gi, int gj, int gk, int X) { return (int)(6 * cell_index(S, gi, gj, gk)) + X - 1; };
  // THCM.C:647-704 integral condition row
  rowintcon = -1; rowintcon_local = -1; owns_rowintcon = false;
  if (cfg.sres == 0 && cfg.intcond_i != -2) { // -2: no integral row, the S row stays what the Fortran layer alone makes it
    int Nic = cfg.intcond_i == -1 ? n - 1 : cfg.intcond_i, Mic = cfg.intcond_j == -1 ? m - 1 : cfg.intcond_j;
    if (Nic < 0 || Nic >= n || Mic < 0 || Mic >= m || L(Nic + 1, Mic + 1, l) != OCEAN)
      throw std::runtime_error("thcm_b200: integral row coordinates give a land point"); // THCM.C:678-682
    rowintcon = 6 * (((l - 1) * m + Mic) * n + Nic) + SS - 1;
    if (owned(Nic + 1, Mic + 1)) { owns_rowintcon = true; rowintcon_local = lrow(Nic + 1, Mic + 1, l, SS); }
  }
  rowPfix1 = -1; rowPfix2 = -1; rowPfix1_local = -1; rowPfix2_local = -1;
  if (cfg.fix_pressure_points) { // THCM.C:763-772
    rowPfix1 = 6 * (((l - 1) * m + (m - 1)) * n + (n - 1)) + PP - 1;
    rowPfix2 = 6 * (((l - 1) * m + (m - 1)) * n + (n - 2)) + PP - 1;
    if (owned(n, m)) rowPfix1_local = lrow(n, m, l, PP);
    if (owned(n - 1, m)) rowPfix2_local = lrow(n - 1, m, l, PP);
  }
  // thcm_utils.F90:285-323: coefficient of every stored S dof
  intcond.assign(6 * nstored, 0.0);
  gid.assign(6 * nstored, -1);
  for (size_t c = 0; c < nstored; c++) {
    int fi, fj, fk;
    if (!stored_cell(c, fi, fj, fk)) continue;
    const long long g = 6 * ((((long long)fk - 1) * m + (fj - 1)) * n + (fi - 1));
    for (int q = 0; q < 6; q++) gid[6 * c + q] = g + q;
    if (L(fi, fj, fk) == OCEAN) intcond[6 * c + SS - 1] = std::cos(y[fj]) * dfzT[fk];
  }
  build_signatures();
  build_tables();
  build_graph();
  build_forcing();
  version++; mask_version++;
}

void Host::build_signatures() {
  const size_t nstored = gid.size() / 6;
  sig.assign(nstored, 0);
  const bool per = cfg.periodic != 0;
  auto isLand = [&](int i, int j, int k) { return lm(i, j, k) == LAND; };
  auto isoc = [&](int i, int j, int k) { int v = lm(i, j, k); return v == OCEAN || v == PERIO; };
  for (size_t c = 0; c < nstored; c++) {
    int i, j, k;
    if (!stored_cell(c, i, j, k)) continue;
    uint64_t g = 0;
    for (int s = 1; s <= 27; s++)
      if (isLand(i + slot_di(s), j + slot_dj(s), k + slot_dk(s))) g |= 1ull << (s - 1);
    if (i < n) {
      if (isLand(i + 2, j - 1, k)) g |= 1ull << SIG_SOUTHEE;
      if (isLand(i + 2, j, k)) g |= 1ull << SIG_EASTEAST;
      if (isLand(i + 2, j + 1, k)) g |= 1ull << SIG_NORTHEE;
      if (j < m && isLand(i + 2, j + 2, k)) g |= 1ull << SIG_NNORTHEE;
    }
    if (j < m && isLand(i, j + 2, k)) g |= 1ull << SIG_NN;
    if (lm(i, j, k) != OCEAN) g |= 1ull << SIG_NOTOCEAN;
    if (lm(i, j, k) == 1) g |= 1ull << SIG_LANDVAL;
    // usol: u,v at corner (i,j) are zeroed by any land cell (i',j') in {i,i+1}x{j,j+1} with i'<=n, j'<=m (:1087-1102),
    // by the north wall j=M (:1055-1058) and, without periodicity, by the east wall i=N (:1044-1047).
    bool live = !(j == m) && !(!per && i == n);
    for (int di = 0; di <= 1 && live; di++) for (int dj = 0; dj <= 1; dj++) {
      int ii = i + di, jj = j + dj;
      if (ii <= n && jj <= m && lm(ii, jj, k) == 1) live = false;
    }
    if (live) g |= 1ull << SIG_LIVE;
    if (per && i == n) { // u(0,j,k): zeroed only by land cells (1,j,k) and (1,j+1,k)
      bool lw = !(lm(1, j, k) == 1) && !(j + 1 <= m && lm(1, j + 1, k) == 1);
      if (lw) g |= 1ull << SIG_LIVEWRAP;
    }
    if (isoc(i, j, k) && isoc(i, j, k + 1)) g |= 1ull << SIG_FACE_UP;
    if (isoc(i, j, k) && isoc(i, j, k - 1)) g |= 1ull << SIG_FACE_DN;
    if (!isLand(i + 1, j, k)) g |= 1ull << SIG_BU;
    if (!isLand(i, j + 1, k)) g |= 1ull << SIG_BV;
    sig[c] = g;
  }
}


Tables Host::tables(const double* base) const {
  Tables T;
  const size_t nj = (size_t)m + 2, nk = (size_t)l + 1;
  const double* p = base;
  const double** jt[NJTAB] = {&T.cdxu, &T.cdyu, &T.cyv, &T.tanv, &T.cdxt, &T.cdyt,
                              &T.uu_we, &T.uu_s, &T.uu_n, &T.uu_ch, &T.uv_c, &T.uv_w, &T.uv_e,
                              &T.vv_we, &T.vv_s, &T.vv_n, &T.vv_ch, &T.vu_c, &T.vu_w, &T.vu_e,
                              &T.pu, &T.pv_s, &T.pv_n, &T.tt_we, &T.tt_s, &T.tt_n, &T.tt_ch, nullptr, nullptr};
  for (int q = 0; q < NJTAB; q++) { if (jt[q]) *jt[q] = p; p += nj; }
  const double** kt[NKTAB] = {&T.uu_b, &T.uu_t, &T.uu_cv, &T.tt_b, &T.tt_t, &T.tt_cv, &T.wp, &T.pw, &T.tdzi8, &T.dfzT, &T.dzdfzT, &T.dzdfzW, &T.r_dfzT, &T.r_dzdfzT, &T.r_dzdfzW};
  for (int q = 0; q < NKTAB; q++) { *kt[q] = p; p += nk; }
  T.msi = nullptr; // set by whoever owns a copy of the sea-ice mask in the same memory space as `base`
  return T;
}

// Coefficients of the linear operator Al (lin, usrc.F90:588-772) and of the atoms in spf.F90, tabulated
// per latitude / per level with the reference's own expression order.
void Host::build_tables() {
  const size_t nj = (size_t)m + 2, nk = (size_t)l + 1;
  tab.assign(NJTAB * nj + NKTAB * nk, 0.0);
  Tables T = tables(tab.data());
  auto W = [](const double* p) { return const_cast<double*>(p); };
  const int ih = cfg.ih;
  const double EV = par[P_EK_V], EH = par[P_EK_H], ph = (1 - par[P_MIXP]) * par[P_PE_H], pv = par[P_PE_V];
  const double lambda = par[P_LAMB], xes = par[P_NLES], bi = par[P_BIOT], Ra = par[P_RAYL];
  for (int j = 0; j <= m; j++) {
    W(T.cdxu)[j] = 1.0 / (2 * std::cos(yv[j]) * dx);  // unlin/vnlin costdxi (spf.F90:563,...), gradp cosdxi (:331)
    W(T.cdyu)[j] = 1.0 / (2 * std::cos(yv[j]) * dy);
    W(T.cyv)[j] = std::cos(yv[j]);
    W(T.tanv)[j] = std::tan(yv[j]);
  }
  for (int j = 0; j <= m + 1; j++) {
    W(T.cdxt)[j] = 1.0 / (4 * std::cos(y[j]) * dx);   // tnlin costdxi (spf.F90:403,415)
    W(T.cdyt)[j] = 1.0 / (4 * std::cos(y[j]) * dy);
  }
  for (int j = 1; j <= m; j++) {
    const bool in = j <= m - 1; // uderiv/vderiv/coriolis/gradp loop j=1..m-1
    double c = 1.0 / (std::cos(yv[j]) * dx), cosdx2i = c * c;
    double r = 1.0 / dy, rdy2i = r * r;
    // uderiv (spf.F90:13-76)
    double uxx2 = in ? amh(yv[j], ih) * cosdx2i : 0.0, uxx5 = in ? -(uxx2 + uxx2) : 0.0;
    double uyy4 = in ? rdy2i * bmh(y[j], ih) * std::cos(y[j]) / std::cos(yv[j]) : 0.0;
    double uyy6 = in ? rdy2i * bmh(y[j + 1], ih) * std::cos(y[j + 1]) / std::cos(yv[j]) : 0.0;
    double uyy5 = in ? -(uyy4 + uyy6) : 0.0;
    double tand2 = 1 - std::tan(yv[j]) * std::tan(yv[j]);
    double ucsi5 = in ? bmh(yv[j], ih) * tand2 + std::tan(yv[j]) * bmhy(yv[j], ih) : 0.0;
    double tanv = std::tan(yv[j]), cosv = std::cos(yv[j]);
    double vxs2 = in ? (bmhy(yv[j], ih) - (amh(yv[j], ih) + bmh(yv[j], ih)) * tanv) / (dx * cosv) : 0.0;
    double vxs8 = in ? -(bmhy(yv[j], ih) - (amh(yv[j], ih) + bmh(yv[j], ih)) * tanv) / (dx * cosv) : 0.0;
    double corv = in ? std::sin(yv[j]) * cfg.coriolis_on : 0.0;
    W(T.uu_we)[j] = -EH * (uxx2 + 0.0 + 0.0) - EV * 0.0;
    W(T.uu_s)[j] = -EH * (0.0 + uyy4 + 0.0) - EV * 0.0;
    W(T.uu_n)[j] = -EH * (0.0 + uyy6 + 0.0) - EV * 0.0;
    W(T.uu_ch)[j] = -EH * (uxx5 + uyy5 + ucsi5);
    W(T.uv_c)[j] = -corv - EH * 0.0;
    W(T.uv_w)[j] = -0.0 - EH * vxs2;
    W(T.uv_e)[j] = -0.0 - EH * vxs8;
    // vderiv (spf.F90:78-141)
    double vxx2 = in ? bmh(yv[j], ih) * cosdx2i : 0.0, vxx5 = in ? -2 * bmh(yv[j], ih) * cosdx2i : 0.0;
    double vyy4 = in ? rdy2i * amh(y[j], ih) * std::cos(y[j]) / std::cos(yv[j]) : 0.0;
    double vyy6 = in ? rdy2i * amh(y[j + 1], ih) * std::cos(y[j + 1]) / std::cos(yv[j]) : 0.0;
    double vyy5 = in ? -(vyy4 + vyy6) : 0.0;
    double vcsi5 = in ? bmh(yv[j], ih) - amh(yv[j], ih) * std::tan(yv[j]) * std::tan(yv[j]) + bmhy(yv[j], ih) * std::tan(yv[j]) : 0.0;
    double uxs2 = in ? -((amh(yv[j], ih) + bmh(yv[j], ih)) * tanv - bmhy(yv[j], ih)) / (dx * cosv) : 0.0;
    double uxs8 = in ? ((amh(yv[j], ih) + bmh(yv[j], ih)) * tanv - bmhy(yv[j], ih)) / (dx * cosv) : 0.0;
    W(T.vv_we)[j] = -EH * (vxx2 + 0.0 + 0.0) - EV * 0.0;
    W(T.vv_s)[j] = -EH * (0.0 + vyy4 + 0.0) - EV * 0.0;
    W(T.vv_n)[j] = -EH * (0.0 + vyy6 + 0.0) - EV * 0.0;
    W(T.vv_ch)[j] = -EH * (vxx5 + vyy5 + vcsi5);
    W(T.vu_c)[j] = corv - EH * 0.0;
    W(T.vu_w)[j] = 0.0 - EH * uxs2;
    W(T.vu_e)[j] = 0.0 - EH * uxs8;
    // pderiv (spf.F90:143-196)
    W(T.pu)[j] = 1.0 / (2 * std::cos(y[j]) * dx);
    double cos2i = 1. / (2 * std::cos(y[j]) * dy);
    W(T.pv_s)[j] = -std::cos(yv[j - 1]) * cos2i;
    W(T.pv_n)[j] = std::cos(yv[j]) * cos2i;
    // tderiv 3,4 (spf.F90:216-236), surface-land factor (1-landm(i,j,l)) == 1 on every ocean row
    double ct = 1.0 / (std::cos(y[j]) * dx), ctdx2i = ct * ct;
    double txx2 = ctdx2i, txx5 = -2 * ctdx2i;
    double tyy4 = (rdy2i * std::cos(yv[j - 1]) / std::cos(y[j])), tyy6 = (rdy2i * std::cos(yv[j]) / std::cos(y[j]));
    double tyy5 = -(tyy4 + tyy6);
    W(T.tt_we)[j] = -ph * (txx2 + 0.0) - pv * 0.0;
    W(T.tt_s)[j] = -ph * (0.0 + tyy4) - pv * 0.0;
    W(T.tt_n)[j] = -ph * (0.0 + tyy6) - pv * 0.0;
    W(T.tt_ch)[j] = -ph * (txx5 + tyy5);
  }
  const double rz = 1.0 / dz, rdz2i = rz * rz, dzi = 1.0 / dz;
  for (int k = 1; k <= l; k++) {
    double h1 = 1. / (dfzT[k] * dfzW[k]), h2 = 1. / (dfzT[k] * dfzW[k - 1]);
    double uzz14 = h2 * rdz2i, uzz23 = h1 * rdz2i, uzz5 = -(uzz14 + uzz23); // uderiv(4)
    W(T.uu_b)[k] = -0.0 - EV * uzz14;
    W(T.uu_t)[k] = -0.0 - EV * uzz23;
    W(T.uu_cv)[k] = EV * uzz5;
    double tzz14 = h2 * rdz2i, tzz23 = (k <= l - 1) ? h1 * rdz2i : 0.0, tzz5 = -(tzz14 + tzz23); // tderiv(5)
    W(T.tt_b)[k] = -0.0 - pv * tzz14;
    W(T.tt_t)[k] = -0.0 - pv * tzz23;
    W(T.tt_cv)[k] = pv * tzz5;
    W(T.wp)[k] = dzi / dfzW[k];   // gradp(3)
    W(T.pw)[k] = dzi / dfzT[k];   // pderiv(3)
    W(T.tdzi8)[k] = 1.0 / (8 * dfzT[k] * dz);
    W(T.dfzT)[k] = dfzT[k];
    W(T.dzdfzT)[k] = dz * dfzT[k];
    W(T.r_dfzT)[k] = 1.0 / dfzT[k];
    W(T.r_dzdfzT)[k] = 1.0 / (dz * dfzT[k]);
  }
  for (int k = 0; k <= l; k++) { W(T.dzdfzW)[k] = dz * dfzW[k]; W(T.r_dzdfzW)[k] = 1.0 / (dz * dfzW[k]); }

  std::memset(&S, 0, sizeof(S));
  S.n = n; S.m = m; S.l = l; S.periodic = cfg.periodic != 0;
  S.io0 = io0; S.jo0 = jo0; S.nloc = nloc; S.mloc = mloc; S.nown = (long long)nloc * mloc * l;
  S.epsr = par[P_ROSB];
  S.w_tt = -Ra * (1. + xes * alpt1) * 1.0 / 2.;   // Al(:,WW,TT) = -Ra*(1.+xes*alpt1)*tbc/2.
  S.w_ss = lambda * Ra * 1.0 / 2.;                // Al(:,WW,SS) = lambda*Ra*tbc/2.
  S.ra_a2 = Ra * xes * alpt2; S.ra_a3 = Ra * xes * alpt3;
  S.tt_surf = cfg.tres * bi * 1.0; S.ss_surf = cfg.sres * bi * 1.0;
  if (cfg.coupled_T == 1) { // usrc.F90:725-741 with tc = sc = 1: Ooa*tc, then + dedt*sc + mc*(QTnd*zeta*tc - Ooa*tc - dedt*sc)
    const double dedt = cpl.lvsc * cpl.eta * cpl.qdim * (deltat / cpl.qdim) * cpl.dqso;
    S.coupled_T = 1;
    S.tt_surf = cpl.Ooa * 1.0;
    S.tt_dedt = dedt * 1.0;
    S.tt_mc = cpl.QTnd * cpl.zeta * 1.0 - cpl.Ooa * 1.0 - dedt * 1.0;
    S.ts_mc = -cpl.QTnd * cpl.zeta * cpl.a0;
    if (cpl.msi_lin != cpl.msi) { // masksi(mc, msi) of this `lin` call (usrc.F90:723)
      cpl.msi_lin = cpl.msi;
      cpl.have_msi = false;
      for (double v : cpl.msi_lin) if (v != 0.0) { cpl.have_msi = true; break; }
      cpl.msi_version++;
    }
  }
  S.dyi = 1. / (2 * dy); S.dzi = dzi; S.tdzi2 = 1.0 / (2 * dz);
  S.kvc = par[P_P_VC]; S.lambda = lambda; S.xes = xes;
  S.fac = cfg.alphaT * par[P_SPL1];
  S.vmix_on = cfg.vmix >= 1;
  S.rho_mixing = (cfg.rho_mixing != 0) && (xes == 0.0);
  S.dz = dz;
}

namespace {
// forcing.F90:489-534
double wfun(double yy, int v1) {
  if (v1 == 1) return 0.2 - 0.8 * std::sin(6 * std::fabs(yy)) - 0.5 * (1 - std::tanh(10 * std::fabs(yy))) - 0.5 * (1 - std::tanh(10 * (pi / 2 - std::fabs(yy))));
  return 0.0;
}
} // namespace

// forcing.F90:4-218 (idealized, uncoupled salinity): the surface forcing of an OCEAN cell depends on its latitude only, so it is
// evaluated once per latitude -- fprof = [fu | fv | ft | fs], each (m + 2) doubles indexed by the global Fortran j -- with the
// reference's expressions and summation order.  The host fill below and the device kernel (forcing_surface_kernel, capi.cu /
// thcm_kernels.cu) both only copy these numbers into the rows, so they cannot disagree.
void Host::forcing_profiles() {
  const int TRES = cfg.tres, SRES = cfg.sres;
  const double gymin = cfg.ymin, gymax = cfg.ymax;
  auto temfun = [&](double yy) {
    if (cfg.forcing_type == 2) return std::cos(pi * (yy - gymin) / (gymax - gymin));
    return std::cos(pi * yy / gymax) + par[P_CMPR] * std::sin(pi * yy / gymax);
  };
  auto salfun = [&](double yy) {
    if (cfg.forcing_type == 2) return std::cos(pi * (yy - gymin) / (gymax - gymin));
    if (cfg.forcing_type == 1) return (std::cos(pi * yy / gymax) + par[P_FPER] * yy / gymax) / std::cos(yy);
    return std::cos(pi * yy / gymax) + par[P_FPER] * yy / gymax;
  };
  // thcm_forcing_integral_ (THCM.C:2704-2740): a GLOBAL surface integral; every rank holds the global mask, so it is
  // evaluated redundantly in the reference's summation order instead of through an allreduce.  Only the flux-forced
  // configurations (TRES = 0 / SRES = 0) need it.
  std::vector<double> tj(m + 2, 0.0), sj(m + 2, 0.0), cosy(m + 2, 0.0);
  for (int j = 1; j <= m; j++) { cosy[j] = std::cos(y[j]); tj[j] = temfun(y[j]); sj[j] = salfun(y[j]); }
  auto qint = [&](const std::vector<double>* prof, bool times_ocean) { // prof == nullptr: the zero field
    double lf = 0.0, ls = 0.0;
    for (int j = 1; j <= m; j++) for (int i = 1; i <= n; i++) {
      const int ocean = 1 - lm(i, j, l);
      const double field = prof ? (times_ocean ? (*prof)[j] * ocean : (*prof)[j]) : 0.0;
      lf = field * cosy[j] * ocean + lf;
      ls = cosy[j] * ocean + ls;
    }
    return lf / ls;
  };
  fprof.assign(4 * (size_t)(m + 2), 0.0);
  double *fu = fprof.data(), *fv = fu + (m + 2), *ft = fv + (m + 2), *fs = ft + (m + 2);
  const double sigma = par[P_COMB] * par[P_WIND] * par[P_AL_T];
  for (int j = 1; j <= m - 1; j++) { fu[j] = sigma * wfun(yv[j], 1); fv[j] = sigma * wfun(yv[j], 2); }
  const double etabi = par[P_COMB] * par[P_TEMP] * (1 - TRES + TRES * par[P_BIOT]);
  double temcor = 0.0;
  if (cfg.coupled_T == 0) {
    if (TRES == 0) temcor = qint(&tj, false);
    for (int j = 1; j <= m; j++) ft[j] = etabi * (tj[j] - temcor);
  }
  if (cfg.coupled_T == 1) { // forcing.F90:73-96: per cell, from the inserted fields (zero until an atmosphere / sea ice is attached)
    ftcell.assign((size_t)nloc * mloc, 0.0);
    auto fld = [&](const std::vector<double>& f, int i, int j) { return f.empty() ? 0.0 : f[(size_t)(j - 1) * n + (i - 1)]; };
    for (int j = jo0; j < jo0 + mloc; j++) for (int i = io0; i < io0 + nloc; i++) {
      const double QToa = par[P_COMB] * par[P_SUNP] * cpl.suno[j] * (1 - cpl.albe0 - cpl.albed * fld(cpl.albe, i, j)) +
                          cpl.Ooa * fld(cpl.tatm, i, j) + cpl.lvsc * cpl.eta * cpl.qdim * fld(cpl.qatm, i, j) - cpl.lvsc * cpl.eo0;
      const double QTos = cpl.QTnd * cpl.zeta * (cpl.a0 * s0 - t0);
      ftcell[(size_t)(j - jo0) * nloc + (i - io0)] = (QToa + fld(cpl.msi, i, j) * (QTos - QToa)) * (1 - lm(i, j, l));
    }
  }
  const double gamma = par[P_COMB] * par[P_SALT] * (1 - SRES + SRES * par[P_BIOT]);
  double salcor = 0.0, adapted_salcor = 0.0, spertcor = 0.0;
  if (SRES == 0) { salcor = qint(&sj, true); adapted_salcor = qint(nullptr, false); spertcor = qint(nullptr, false); }
  for (int j = 1; j <= m; j++) // emip = salfun * (1 - landm) = salfun on an ocean cell
    fs[j] = gamma * (1 - par[P_HMTP]) * (sj[j] * 1 - salcor) + gamma * par[P_HMTP] * (0.0 - adapted_salcor) +
            par[P_SPER] * (1 - SRES + SRES * par[P_BIOT]) * (0.0 - spertcor);
}

// Frc on the host: the surface rows from fprof, the body rows (an exact zero), then the Frc zeroing of boundary.F90
void Host::build_forcing(bool surface_only) {
  const size_t nrows = 6 * (size_t)S.nown;
  const size_t top = 6 * (size_t)nloc * mloc * (l - 1); // first row of the surface level (levels are the slowest index)
  if (surface_only && frc.size() == nrows) {
    std::fill(frc.begin() + top, frc.end(), 0.0);
    frc_dirty_from = std::min(frc_dirty_from, top);
  } else {
    surface_only = false;
    frc.assign(nrows, 0.0);
    frc_dirty_from = 0;
  }
  forcing_profiles();
  frc_host_valid = true;
  const int i1 = io0, i2 = io0 + nloc - 1, j1 = jo0, j2 = jo0 + mloc - 1; // owned block, global Fortran indices
  auto row = [&](int i, int j, int k, int X) { return 6 * cell_index(S, i, j, k) + X - 1; };
  const double *fu = fprof.data(), *fv = fu + (m + 2), *ft = fv + (m + 2), *fs = ft + (m + 2);
  for (int j = j1; j <= j2; j++) for (int i = i1; i <= i2; i++) {
    frc[row(i, j, l, UU)] = fu[j]; frc[row(i, j, l, VV)] = fv[j];
    frc[row(i, j, l, TT)] = cfg.coupled_T == 1 ? ftcell[(size_t)(j - jo0) * nloc + (i - io0)] : ft[j];
    frc[row(i, j, l, SS)] = fs[j];
  }
  if (!surface_only)
  for (int k = 1; k <= l - 1; k++) for (int j = j1; j <= j2; j++) for (int i = i1; i <= i2; i++) // internal T/S forcing is zero (THCM.C:540-544)
    frc[row(i, j, k, WW)] = -par[P_COMB] * (1 - lm(i, j, k)) * par[P_RAYL] * (par[P_LAMB] * (0.0 + 0.0) / 2. - (0.0 + 0.0) / 2.);
  // rows whose forcing `boundaries` resets to zero on every call
  for (size_t c = surface_only ? top / 6 : 0; c < (size_t)S.nown; c++) {
    uint64_t g = sig[c];
    if (frc_masked<UU>(g)) frc[6 * c + 0] = 0.0;
    if (frc_masked<VV>(g)) frc[6 * c + 1] = 0.0;
    if (frc_masked<WW>(g)) frc[6 * c + 2] = 0.0;
    if (frc_masked<PP>(g)) frc[6 * c + 3] = 0.0;
    if (frc_masked<TT>(g)) frc[6 * c + 4] = 0.0;
    if (frc_masked<SS>(g)) frc[6 * c + 5] = 0.0;
  }
}

// Static pattern of THCM::CreateMaximalGraph (THCM.C:2241-2539): generated from the same RowDef tables the
// kernels use, so the kernels' emit order and the graph cannot disagree.  Column = stored index of the neighbour cell
// (cell_index) * 6 + var; entry order inside a row = emit_row's order.
template <int A> static void graph_row(const Scalars& S, const Cell& c, std::vector<int>& col) {
  int cols[Row<A>::NE], gcols[Row<A>::NE], cnt = 0;
  for (int q = 0; q < Row<A>::NE; q++) {
    int slot = RowDef<A>::e[q].slot;
    if (slot_present(slot, S, c)) {
      gcols[cnt] = slot_col(slot, RowDef<A>::e[q].var, S, c);
      int i2 = c.i + slot_di(slot), j2 = c.j + slot_dj(slot), k2 = c.k + slot_dk(slot);
      if (S.periodic) { if (i2 == 0) i2 = S.n; else if (i2 == S.n + 1) i2 = 1; }
      cols[cnt] = (int)(6 * cell_index(S, i2, j2, k2)) + RowDef<A>::e[q].var - 1;
      cnt++;
    }
  }
  if (row_has_seam(S, c)) { // single rank in x: order the row like the global graph (= the order the kernels' seam emit uses).
    // Sorted by GLOBAL column: with several ranks in y the north / south neighbours of a block's edge rows sit in the halo ring,
    // whose stored indices follow all owned cells.
    int ord[Row<A>::NE];
    for (int q = 0; q < cnt; q++) ord[q] = q;
    std::sort(ord, ord + cnt, [&](int a, int b) { return gcols[a] < gcols[b]; });
    int tmp[Row<A>::NE];
    for (int q = 0; q < cnt; q++) tmp[q] = cols[ord[q]];
    for (int q = 0; q < cnt; q++) cols[q] = tmp[q];
  }
  col.insert(col.end(), cols, cols + cnt);
}
void Host::build_graph() {
  const size_t nrows = 6 * (size_t)S.nown;
  rowptr.assign(nrows + 1, 0);
  col.clear();
  col.reserve(104 * (size_t)S.nown + (rowintcon_local >= 0 ? (size_t)S.nown : 0));
  size_t r = 0;
  for (int k = 1; k <= l; k++) for (int j = jo0; j < jo0 + mloc; j++) for (int i = io0; i < io0 + nloc; i++) {
    Cell c{i, j, k, 0};
    rowptr[r++] = (int)col.size(); graph_row<UU>(S, c, col);
    rowptr[r++] = (int)col.size(); graph_row<VV>(S, c, col);
    rowptr[r++] = (int)col.size(); graph_row<WW>(S, c, col);
    rowptr[r++] = (int)col.size(); graph_row<PP>(S, c, col);
    rowptr[r++] = (int)col.size(); graph_row<TT>(S, c, col);
    rowptr[r++] = (int)col.size();
    if ((int)(r - 1) == rowintcon_local) { // dense integral-condition row (THCM.C:2494-2513), sorted: all S dofs
      for (size_t cc = 0; cc < (size_t)S.nown; cc++) col.push_back((int)(6 * cc + SS - 1));
    } else graph_row<SS>(S, c, col);
  }
  rowptr[nrows] = (int)col.size();
}

// Where the 6x6 diagonal block sits inside every row of the static graph: offset of the first entry whose column belongs
// to the row's own cell (bits 0..25) and which unknowns are present (bits 26..31).  The centre-slot entries are adjacent
// in both the static (dk,dj,di,var) order and the sorted order of seam rows.
std::vector<unsigned> Host::diag_info() const {
  const size_t nrows = rowptr.size() - 1;
  std::vector<unsigned> info(nrows, 0u);
  for (size_t r = 0; r < nrows; r++) {
    const int cell = (int)(r / 6);
    int first = -1;
    unsigned mask = 0;
    for (int p = rowptr[r]; p < rowptr[r + 1]; p++)
      if (col[p] / 6 == cell) {
        if (first < 0) first = p - rowptr[r];
        mask |= 1u << (col[p] % 6);
      }
    if (first < 0) first = 0;
    if (first >= (1 << 26)) throw std::runtime_error("diag_info: row too long");
    info[r] = (unsigned)first | (mask << 26);
  }
  return info;
}

// m_scaling::scal (src/ocean/scaling.F90:198-284): invert the averaged 6x6 block and derive the row / column factors
// "special for the oceanography problem".  db is column-major like the Fortran array, db[ii + 6*jj].  The reference calls
// LAPACK dgetrf / dgecon / dgetri; here: LU with partial pivoting (the dgetf2 pivot rule), the inverse column by column,
// and the exact 1/(|A|_inf |A^-1|_inf) where dgecon estimates it (used only for the singularity test :233-236).
bool scal(const double* db, double* dr, double* dc) {
  const int N = 6;
  double A[N][N], LU[N][N], inv[N][N];
  for (int i = 0; i < N; i++) for (int j = 0; j < N; j++) { A[i][j] = db[i + N * j]; LU[i][j] = A[i][j]; }
  double anorm = 0.0;
  for (int i = 0; i < N; i++) { double s = 0.0; for (int j = 0; j < N; j++) s += std::fabs(A[i][j]); anorm = std::max(anorm, s); }
  int piv[N];
  for (int k = 0; k < N; k++) {
    int p = k;
    for (int i = k + 1; i < N; i++) if (std::fabs(LU[i][k]) > std::fabs(LU[p][k])) p = i;
    piv[k] = p;
    if (LU[p][k] == 0.0) return false;
    if (p != k) for (int j = 0; j < N; j++) std::swap(LU[k][j], LU[p][j]);
    for (int i = k + 1; i < N; i++) {
      LU[i][k] /= LU[k][k];
      for (int j = k + 1; j < N; j++) LU[i][j] -= LU[i][k] * LU[k][j];
    }
  }
  for (int c = 0; c < N; c++) {
    double b[N];
    for (int i = 0; i < N; i++) b[i] = (i == c) ? 1.0 : 0.0;
    for (int k = 0; k < N; k++) std::swap(b[k], b[piv[k]]);
    for (int i = 1; i < N; i++) for (int j = 0; j < i; j++) b[i] -= LU[i][j] * b[j];
    for (int i = N - 1; i >= 0; i--) { for (int j = i + 1; j < N; j++) b[i] -= LU[i][j] * b[j]; b[i] /= LU[i][i]; }
    for (int i = 0; i < N; i++) inv[i][c] = b[i];
  }
  double inorm = 0.0;
  for (int i = 0; i < N; i++) { double s = 0.0; for (int j = 0; j < N; j++) s += std::fabs(inv[i][j]); inorm = std::max(inorm, s); }
  const double rcond = (anorm > 0.0 && inorm > 0.0) ? (1.0 / anorm) / inorm : 0.0;
  if (1.0 + rcond == 1.0) return false; // "diagonal block is singular up to working precision"
  auto mat = [&](int i, int j) -> double& { return inv[i - 1][j - 1]; }; // 1-based like the Fortran
  double idc, idr;
  dr[0] = 1.0; dc[0] = 1.0;
  idc = std::sqrt(mat(1, 1) / mat(2, 2));
  dr[1] = 1 / idc; dc[1] = dr[1];
  idr = std::sqrt(std::fabs(mat(1, 1) / mat(4, 4)));
  dr[3] = 1 / idr; dc[3] = dr[3];
  if (std::fabs(mat(4, 3)) > std::fabs(mat(3, 3))) idr = mat(1, 1) / (idr * mat(4, 3));
  else idr = std::sqrt(std::fabs(mat(1, 1) / mat(3, 3)));
  dr[2] = 2 / idr; dc[2] = dr[2];
  if (std::fabs(mat(4, 5) * mat(5, 4)) < .01 * std::fabs(mat(4, 4) * mat(5, 5))) { mat(4, 5) = 1; mat(5, 4) = 1; }
  idc = std::sqrt(std::fabs(mat(1, 1) * mat(4, 5) / (mat(5, 4) * mat(5, 5))));
  idr = mat(1, 1) / (idc * mat(5, 5));
  dr[4] = 1 / idr; dc[4] = 1 / idc;
  if (std::fabs(mat(4, 6) * mat(6, 4)) < .01 * std::fabs(mat(4, 4) * mat(6, 6))) { mat(4, 6) = 1; mat(6, 4) = 1; }
  idc = std::sqrt(std::fabs(mat(1, 1) * mat(4, 6) / (mat(6, 4) * mat(6, 6))));
  idr = mat(1, 1) / (idc * mat(6, 6));
  dr[5] = 1 / idr; dc[5] = 1 / idc;
  return true;
}


} // namespace thcm
