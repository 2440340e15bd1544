// thcm_oracle.cpp -- CPU oracle: structure-faithful C++ restatement of the reference's
// Fortran THCM kernels and the THCM.C glue.  TEST INFRASTRUCTURE ONLY (see thcm_oracle.h).
//
// Every routine cites the reference file:line it follows (paths relative to /root/reference).
// The dense tensors Al/An(27,6,6,n,m,l) are materialised exactly as the reference does.
// Compile WITHOUT fp contraction (-ffp-contract=off): the reference is built for baseline
// x86-64 (src/CMakeLists.txt:24, no -march), i.e. without FMA.
//
// PARITY: the Fortran cannot be built here (no Fortran compiler).  Pinned by the golden state the reference
// wrote (test/ocean/ocean_reference.h5 is a root of this F: I7), by the known answer of the reference's integration test
// intt_2dmoc (I6: reached both by a solver of ours and by the reference's own unmodified Continuation.H driving this oracle,
// tests/continuation_oracle.cpp) and by the invariants I1-I5; details in thcm_oracle.h.
#include "thcm_oracle.h"
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <algorithm>

namespace {

// par.F90:37-66 -- parameter enumeration (1-based)
enum { AL_T = 1, RAYL, EK_V, EK_H, ROSB, MIXP, RESC, SPL1, HMTP, SUNP, PE_H, PE_V, P_VC, LAMB, SALT,
       WIND, TEMP, BIOT, COMB, ARCL, NLES, IFRICB, CONT, ENER, ALPC, CMPR, FPER, SPER, MKAP, SPL2 };
enum { UU = 1, VV, WW, PP, TT, SS };              // par.F90:69-74
enum { OCEAN = 0, LAND = 1, WATER = 2, PERIO = 3 }; // par.F90:77-80
const int NUN = 6, NP = 27, NPAR = 30;
const double pi = 3.14159265358979323846;         // par.F90:14

// usr.F90:129-160 -- fixed physical parameters
const double omegadim = 7.292e-05, r0dim = 6.37e+06, udim = 0.1e+00, gdim = 9.8e+00, rhodim = 1.024e+03;
const double t0 = 15, deltat = 1.0, deltas = 1.0, s0 = 35.0, cp0 = 4.2e+03;
const double alpt1 = 2.93, alpt2 = 8.3e-02, alpt3 = 6.6e-04;
const double ah = 2.5e+05, av = 1.0e-03, kappah = 1.0e+03, kappav = 1.0e-04;
const double zmin = -1.0, zmax = 0.0;
// atm.F90:5-19
const double hdima = 8400., rhoa = 1.25, uatm = 0.0, ce = 1.3e-03, ch = 0.94 * ce, cpa = 1000., uw = 8.5;
const double d0 = 3.1e+06, c0 = 0.43, arad = 216.0, brad = 1.5, sun0 = 1360., lv = 2.5e+06;

// grid.F90:62-95
double fz(double z, double qz) {
  double th = std::tanh(qz * (z + 1)), tth = std::tanh(qz);
  if (qz > 1.0) return -1 + th / tth;
  return z + (1. - qz) * z * (1 - z);
}
double dfdz(double z, double qz) {
  double chv = std::cosh(qz * (z + 1)), tth = std::tanh(qz);
  if (qz > 1.0) return qz / (tth * chv * chv);
  return 1.0 + (1. - qz) * (1. - 2. * z);
}
// spf.F90:795-853
double amh(double y, int ih) { return ih == 0 ? 1.0 : 1. + 10.0 * std::exp(-5 * y * y); }
double bmh(double y, int ih) { return ih == 0 ? 1.0 : 1.0 + 10.0 * std::exp(-5 * y * y); }
double bmhy(double y, int ih) { return ih == 0 ? 0.0 : -10. * 10.0 * y * std::exp(-5 * y * y); }

// Fortran-style array with arbitrary lower bounds (column major).
struct F3 {
  int lo[3], sz[3];
  std::vector<double> a;
  void alloc(int l0, int h0, int l1, int h1, int l2, int h2) {
    lo[0] = l0; lo[1] = l1; lo[2] = l2; sz[0] = h0 - l0 + 1; sz[1] = h1 - l1 + 1; sz[2] = h2 - l2 + 1;
    a.assign((size_t)sz[0] * sz[1] * sz[2], 0.0);
  }
  inline double& operator()(int i, int j, int k) {
    return a[(size_t)(i - lo[0]) + (size_t)sz[0] * ((j - lo[1]) + (size_t)sz[1] * (k - lo[2]))];
  }
  void zero() { std::fill(a.begin(), a.end(), 0.0); }
};

} // namespace

struct thcm_oracle {
  thcm_oracle_config cfg;
  int n, m, l, ndim;
  bool periodic;
  double par[NPAR + 1];
  // grid (grid.F90)
  double dx, dy, dz;
  std::vector<double> x, y, z, xu, yv, zw, ze, zwe, dfzT, dfzW; // indexed as in Fortran (0 or 1 based)
  std::vector<int> landm;  // (0:n+1,0:m+1,0:l+1)
  // forcing fields (usr.F90:104-113)
  std::vector<double> Frc, taux, tauy, tatm, emip, spert, adapted_emip, qatm, albe, patm, msi, gsi, qsa;
  std::vector<double> internal_temp, internal_salt;
  double QTnd, QSnd;
  // atm.F90 / ice.F90 module variables
  double qdim, nuq, nus, eta, dqso, eo0, albe0, albed, lvsc, Ooa, Os;
  std::vector<double> suno, suna, dat, davt;
  double zeta, a0, Lf, Qvar, Q0;
  // mix.F90
  int vmix_flag, vmix_temp, vmix_salt, vmix_fix;
  // m_mat
  std::vector<double> Al, An;   // (27,6,6,n,m,l)
  std::vector<int> begA, jcoA;  // 1-based contents
  std::vector<double> coA, coB;
  // THCM.C glue
  std::vector<int> g_rowptr, g_col;
  int rowintcon;                // 0-based global row, -1 if none
  int rowPfix1, rowPfix2;
  double intCorrection;
  std::vector<double> intcondCoeff;
  int64_t dropped;

  inline int& lm(int i, int j, int k) { return landm[(size_t)i + (size_t)(n + 2) * (j + (size_t)(m + 2) * k)]; }
  inline size_t aidx(int s, int A, int B, int i, int j, int k) const {
    return (size_t)(s - 1) + 27 * ((size_t)(A - 1) + 6 * ((size_t)(B - 1) + 6 * ((size_t)(i - 1) + (size_t)n * ((j - 1) + (size_t)m * (k - 1)))));
  }
  inline double& AN(int s, int A, int B, int i, int j, int k) { return An[aidx(s, A, B, i, j, k)]; }
  inline double& AL(int s, int A, int B, int i, int j, int k) { return Al[aidx(s, A, B, i, j, k)]; }
  // matetc.F90:123-144
  inline int find_row2(int i, int j, int k, int XX) const { return NUN * ((k - 1) * n * m + n * (j - 1) + i - 1) + XX; }
  inline double& s2(std::vector<double>& f, int i, int j) { return f[(size_t)(i - 1) + (size_t)n * (j - 1)]; }
  inline double& s3(std::vector<double>& f, int i, int j, int k) { return f[(size_t)(i - 1) + (size_t)n * ((j - 1) + (size_t)m * (k - 1))]; }
};

namespace {
typedef thcm_oracle T;

// atom(np,n,m,l)
struct Atom {
  int n, m, l;
  std::vector<double> a;
  Atom(const T& t) : n(t.n), m(t.m), l(t.l), a((size_t)27 * t.n * t.m * t.l, 0.0) {}
  inline double& operator()(int s, int i, int j, int k) { return a[(size_t)(s - 1) + 27 * ((size_t)(i - 1) + (size_t)n * ((j - 1) + (size_t)m * (k - 1)))]; }
  void zero() { std::fill(a.begin(), a.end(), 0.0); }
};

// grid.F90:2-60
void grid(T& t) {
  int n = t.n, m = t.m, l = t.l;
  double qz = t.cfg.qz;
  t.x.assign(n + 1, 0); t.y.assign(m + 2, 0); t.z.assign(l + 1, 0); t.xu.assign(n + 1, 0); t.yv.assign(m + 1, 0);
  t.zw.assign(l + 1, 0); t.ze.assign(l + 1, 0); t.zwe.assign(l + 1, 0); t.dfzT.assign(l + 1, 0); t.dfzW.assign(l + 1, 0);
  t.dx = (t.cfg.xmax - t.cfg.xmin) / n;
  t.dy = (t.cfg.ymax - t.cfg.ymin) / m;
  t.dz = (zmax - zmin) / l;
  for (int i = 1; i <= n; i++) { t.x[i] = ((double)i - 0.5) * t.dx + t.cfg.xmin; t.xu[i] = ((double)i) * t.dx + t.cfg.xmin; }
  t.xu[0] = t.cfg.xmin;
  for (int j = 1; j <= m; j++) { t.y[j] = ((double)j - 0.5) * t.dy + t.cfg.ymin; t.yv[j] = ((double)j) * t.dy + t.cfg.ymin; }
  t.y[0] = t.y[1] - t.dy;
  t.y[m + 1] = t.y[m] + t.dy;
  t.yv[0] = t.cfg.ymin;
  for (int k = 1; k <= l; k++) {
    t.ze[k] = ((double)k - 0.5) * t.dz + zmin;
    t.zwe[k] = ((double)k) * t.dz + zmin;
    t.z[k] = fz(t.ze[k], qz);
    t.zw[k] = fz(t.zwe[k], qz);
    t.dfzT[k] = dfdz(t.ze[k], qz);
    t.dfzW[k] = dfdz(t.zwe[k], qz);
  }
  t.zw[0] = zmin;
  t.dfzW[0] = dfdz(zmin, qz);
}

// usrc.F90:1136-1172 (stpnt) + mix_imp.f:92-107 (vmix_par)
void stpnt(T& t) {
  double* par = t.par;
  double hdim = t.cfg.hdim, alphaT = t.cfg.alphaT, alphaS = t.cfg.alphaS;
  for (int i = 0; i <= NPAR; i++) par[i] = 0.0;
  par[AL_T] = 0.1 / (2 * omegadim * rhodim * hdim * udim * t.dz * t.dfzT[t.l]);
  par[RAYL] = alphaT * gdim * hdim / (2 * omegadim * udim * r0dim);
  par[EK_V] = av / (2 * omegadim * hdim * hdim);
  par[EK_H] = ah / (2 * omegadim * r0dim * r0dim);
  par[ROSB] = udim / (2 * omegadim * r0dim);
  par[HMTP] = 0.0; par[SUNP] = 0.0;
  par[PE_H] = kappah / (udim * r0dim);
  par[PE_V] = kappav * r0dim / (udim * hdim * hdim);
  par[P_VC] = 2.5e+04 * par[PE_V];
  par[LAMB] = alphaS / alphaT;
  par[SALT] = 0.0; par[WIND] = 0.0; par[TEMP] = 0.0;
  par[BIOT] = r0dim / (75. * 3600. * 24. * udim);
  par[COMB] = 0.0; par[NLES] = 0.0; par[CMPR] = 0.0;
  par[ALPC] = 1.0; par[ENER] = 1.0e+02; par[MIXP] = 0.0; par[MKAP] = 0.0;
  par[SPL1] = 2.0e+03; par[SPL2] = 0.01;
  if (t.cfg.vmix == 0) { par[MIXP] = 0.0; par[P_VC] = 0.0; par[ALPC] = 1.0; par[ENER] = 1.0e+2; par[MKAP] = 0.0; }
}

// mix_imp.f:58-90 (vmix_init); the DSM colouring of vmix_part is replaced, see vmix_jac below.
void vmix_init(T& t) {
  int vmix = t.cfg.vmix;
  if (vmix == 0) { t.vmix_flag = 0; t.vmix_fix = 1; }
  else if (vmix == 1) { t.vmix_flag = 1; t.vmix_fix = 1; }
  else if (vmix == 2) { t.vmix_flag = 2; t.vmix_fix = 0; }
  else t.vmix_flag = -1;
  switch (t.vmix_flag) {
    case 0: t.vmix_temp = 0; t.vmix_salt = 0; break;
    case 1: t.vmix_temp = 1; t.vmix_salt = 1; break;
    case 2: t.vmix_temp = 0; t.vmix_salt = 0; break;
    default: t.vmix_temp = 0; t.vmix_salt = 0;
  }
}

// usrc.F90:1175-1213
void atmos_coef(T& t) {
  double muoa = rhoa * ch * cpa * uw;
  t.Os = sun0 * c0 / 4 * t.QTnd;
  t.Ooa = muoa * t.QTnd;
  double As = sun0 * (1 - c0) / (4 * muoa);
  t.nus = 0.0;
  t.lvsc = 0.0;
  int m = t.m;
  t.suno.assign(m + 1, 0); t.suna.assign(m + 1, 0); t.dat.assign(m + 1, 0); t.davt.assign(m + 1, 0);
  for (int j = 1; j <= m; j++) {
    t.dat[j] = 0.9 + 1.5 * std::exp(-12 * t.y[j] * t.y[j] / pi);
    double sj = std::sin(t.y[j]);
    t.suno[j] = t.Os * (1 - .482 * (3 * (sj * sj) - 1.) / 2.);
    t.suna[j] = As * (1 - .482 * (3 * (sj * sj) - 1.) / 2.);
  }
  for (int j = 0; j <= m; j++) t.davt[j] = 0.9 + 1.5 * std::exp(-12 * t.yv[j] * t.yv[j] / pi);
}

// forcing.F90:489-534
double wfun(double yy, int v1) {
  if (v1 == 1)
    return 0.2 - 0.8 * std::sin(6 * std::fabs(yy)) - 0.5 * (1 - std::tanh(10 * std::fabs(yy))) -
           0.5 * (1 - std::tanh(10 * (pi / 2 - std::fabs(yy))));
  return 0.0;
}
double temfun(T& t, double yy) {
  double ymin = t.cfg.ymin, ymax = t.cfg.ymax; // m_global values (single domain: identical)
  if (t.cfg.forcing_type == 2) return std::cos(pi * (yy - ymin) / (ymax - ymin));
  return std::cos(pi * yy / ymax) + t.par[CMPR] * std::sin(pi * yy / ymax);
}
double salfun(T& t, double yy) {
  double ymin = t.cfg.ymin, ymax = t.cfg.ymax;
  if (t.cfg.forcing_type == 2) return std::cos(pi * (yy - ymin) / (ymax - ymin));
  if (t.cfg.forcing_type == 1) return (std::cos(pi * yy / ymax) + t.par[FPER] * yy / ymax) / std::cos(yy);
  return std::cos(pi * yy / ymax) + t.par[FPER] * yy / ymax;
}
// forcing.F90:536-548 -> THCM.C:2704-2740 (thcm_forcing_integral_)
double qint(T& t, std::vector<double>& field) {
  double lfsint = 0.0, lsint = 0.0;
  for (int j = 1; j <= t.m; j++)
    for (int i = 1; i <= t.n; i++) {
      lfsint = t.s2(field, i, j) * std::cos(t.y[j]) * (1 - t.lm(i, j, t.l)) + lfsint;
      lsint = std::cos(t.y[j]) * (1 - t.lm(i, j, t.l)) + lsint;
    }
  return lfsint / lsint;
}

// forcing.F90:4-218
void forcing(T& t) {
  int n = t.n, m = t.m, l = t.l;
  double* par = t.par;
  int TRES = t.cfg.tres, SRES = t.cfg.sres;
  std::fill(t.Frc.begin(), t.Frc.end(), 0.0);
  double sigma = par[COMB] * par[WIND] * par[AL_T];
  if (t.cfg.iza == 2)
    for (int j = 1; j <= m; j++)
      for (int i = 1; i <= n; i++) { t.s2(t.taux, i, j) = wfun(t.yv[j], 1); t.s2(t.tauy, i, j) = wfun(t.yv[j], 2); }
  for (int j = 1; j <= m - 1; j++)
    for (int i = 1; i <= n; i++) {
      t.Frc[t.find_row2(i, j, l, UU) - 1] = sigma * t.s2(t.taux, i, j);
      t.Frc[t.find_row2(i, j, l, VV) - 1] = sigma * t.s2(t.tauy, i, j);
    }
  double etabi = par[COMB] * par[TEMP] * (1 - TRES + TRES * par[BIOT]);
  double temcor = 0.0;
  if (t.cfg.ite == 1 && t.cfg.coupled_T == 0) {
    for (int j = 1; j <= m; j++)
      for (int i = 1; i <= n; i++) t.s2(t.tatm, i, j) = temfun(t, t.y[j]);
    if (TRES == 0) temcor = qint(t, t.tatm);
  }
  for (int j = 1; j <= m; j++)
    for (int i = 1; i <= n; i++) {
      if (t.cfg.coupled_T == 1) {
        double QToa = par[COMB] * par[SUNP] * t.suno[j] * (1 - t.albe0 - t.albed * t.s2(t.albe, i, j)) +
                      t.Ooa * t.s2(t.tatm, i, j) + t.lvsc * t.eta * t.qdim * t.s2(t.qatm, i, j) - t.lvsc * t.eo0;
        double QTos = t.QTnd * t.zeta * (t.a0 * s0 - t0);
        t.Frc[t.find_row2(i, j, l, TT) - 1] = (QToa + t.s2(t.msi, i, j) * (QTos - QToa)) * (1 - t.lm(i, j, l));
      } else {
        t.Frc[t.find_row2(i, j, l, TT) - 1] = etabi * (t.s2(t.tatm, i, j) - temcor);
      }
    }
  double gamma;
  if (t.cfg.coupled_S == 1) gamma = par[COMB] * par[SALT];
  else gamma = par[COMB] * par[SALT] * (1 - SRES + SRES * par[BIOT]);
  double salcor = 0.0, adapted_salcor = 0.0, spertcor = 0.0;
  if (t.cfg.its == 1) {
    for (int j = 1; j <= m; j++)
      for (int i = 1; i <= n; i++) t.s2(t.emip, i, j) = salfun(t, t.y[j]) * (1 - t.lm(i, j, l));
    if (SRES == 0 && t.cfg.coupled_S == 0) salcor = qint(t, t.emip);
  }
  if (SRES == 0 && t.cfg.coupled_S == 0) {
    adapted_salcor = qint(t, t.adapted_emip);
    spertcor = qint(t, t.spert);
  }
  double pQSnd = par[COMB] * par[SALT] * t.QSnd;
  for (int j = 1; j <= m; j++)
    for (int i = 1; i <= n; i++) {
      if (t.cfg.coupled_S == 1) {
        double QSoa = pQSnd * (t.eo0 - t.eta * t.qdim * t.s2(t.qatm, i, j) - t.s2(t.patm, i, j));
        double QSos = pQSnd * (t.zeta * (t.a0 * s0 - t0) - t.Qvar * t.s2(t.qsa, i, j) - t.Q0) / (rhodim * t.Lf);
        t.Frc[t.find_row2(i, j, l, SS) - 1] = (QSoa + t.s2(t.msi, i, j) * (QSos - QSoa) - t.s2(t.gsi, i, j)) * (1 - t.lm(i, j, l));
      } else {
        t.Frc[t.find_row2(i, j, l, SS) - 1] = gamma * (1 - par[HMTP]) * (t.s2(t.emip, i, j) - salcor) +
                                             gamma * par[HMTP] * (t.s2(t.adapted_emip, i, j) - adapted_salcor) +
                                             par[SPER] * (1 - SRES + SRES * par[BIOT]) * (t.s2(t.spert, i, j) - spertcor);
      }
    }
  for (int k = 1; k <= l - 1; k++)
    for (int j = 1; j <= m; j++)
      for (int i = 1; i <= n; i++)
        t.Frc[t.find_row2(i, j, k, WW) - 1] =
            -par[COMB] * (1 - t.lm(i, j, k)) * par[RAYL] *
            (par[LAMB] * (t.s3(t.internal_salt, i, j, k) + t.s3(t.internal_salt, i, j, k + 1)) / 2. -
             (t.s3(t.internal_temp, i, j, k) + t.s3(t.internal_temp, i, j, k + 1)) / 2.);
}

// ---------------------------------------------------------------- spf.F90 linear atoms
// spf.F90:13-76
void uderiv(T& t, int type, Atom& atom) {
  int n = t.n, m = t.m, l = t.l, ih = t.cfg.ih;
  double dx = t.dx, dy = t.dy, dz = t.dz;
  atom.zero();
  switch (type) {
    case 1: for (int k = 1; k <= l; k++) for (int j = 1; j <= m; j++) for (int i = 1; i <= n; i++) atom(5, i, j, k) = 1.0; break;
    case 2:
      for (int j = 1; j <= m - 1; j++) {
        double c = 1.0 / (std::cos(t.yv[j]) * dx); double cosdx2i = c * c;
        for (int i = 1; i <= n; i++) for (int k = 1; k <= l; k++) {
          atom(2, i, j, k) = amh(t.yv[j], ih) * cosdx2i;
          atom(8, i, j, k) = amh(t.yv[j], ih) * cosdx2i;
          atom(5, i, j, k) = -(atom(2, i, j, k) + atom(8, i, j, k));
        }
      }
      break;
    case 3: {
      double r = 1.0 / dy, rdy2i = r * r;
      for (int i = 1; i <= n; i++) for (int j = 1; j <= m - 1; j++) for (int k = 1; k <= l; k++) {
        atom(4, i, j, k) = rdy2i * bmh(t.y[j], ih) * std::cos(t.y[j]) / std::cos(t.yv[j]);
        atom(6, i, j, k) = rdy2i * bmh(t.y[j + 1], ih) * std::cos(t.y[j + 1]) / std::cos(t.yv[j]);
        atom(5, i, j, k) = -(atom(4, i, j, k) + atom(6, i, j, k));
      }
    } break;
    case 4: {
      double r = 1.0 / dz, rdz2i = r * r;
      for (int k = 1; k <= l; k++) {
        double h1 = 1. / (t.dfzT[k] * t.dfzW[k]), h2 = 1. / (t.dfzT[k] * t.dfzW[k - 1]);
        for (int j = 1; j <= m; j++) for (int i = 1; i <= n; i++) {
          atom(14, i, j, k) = h2 * rdz2i;
          atom(23, i, j, k) = h1 * rdz2i;
          atom(5, i, j, k) = -(atom(14, i, j, k) + atom(23, i, j, k));
        }
      }
    } break;
    case 5:
      for (int j = 1; j <= m - 1; j++) {
        double tand2 = 1 - std::tan(t.yv[j]) * std::tan(t.yv[j]);
        double v = bmh(t.yv[j], ih) * tand2 + std::tan(t.yv[j]) * bmhy(t.yv[j], ih);
        for (int i = 1; i <= n; i++) for (int k = 1; k <= l; k++) atom(5, i, j, k) = v;
      }
      break;
    case 6:
      for (int j = 1; j <= m - 1; j++) {
        double tand2 = std::tan(t.yv[j]), cosd2 = std::cos(t.yv[j]);
        double v2 = (bmhy(t.yv[j], ih) - (amh(t.yv[j], ih) + bmh(t.yv[j], ih)) * tand2) / (dx * cosd2);
        double v8 = -(bmhy(t.yv[j], ih) - (amh(t.yv[j], ih) + bmh(t.yv[j], ih)) * tand2) / (dx * cosd2);
        for (int i = 1; i <= n; i++) for (int k = 1; k <= l; k++) { atom(2, i, j, k) = v2; atom(8, i, j, k) = v8; }
      }
      break;
  }
}
// spf.F90:78-141
void vderiv(T& t, int type, Atom& atom) {
  int n = t.n, m = t.m, l = t.l, ih = t.cfg.ih;
  double dx = t.dx, dy = t.dy, dz = t.dz;
  atom.zero();
  switch (type) {
    case 1: for (int k = 1; k <= l; k++) for (int j = 1; j <= m; j++) for (int i = 1; i <= n; i++) atom(5, i, j, k) = 1.0; break;
    case 2:
      for (int i = 1; i <= n; i++) for (int j = 1; j <= m - 1; j++) {
        double c = 1.0 / (std::cos(t.yv[j]) * dx); double cosdx2i = c * c;
        for (int k = 1; k <= l; k++) {
          atom(2, i, j, k) = bmh(t.yv[j], ih) * cosdx2i;
          atom(5, i, j, k) = -2 * bmh(t.yv[j], ih) * cosdx2i;
          atom(8, i, j, k) = bmh(t.yv[j], ih) * cosdx2i;
        }
      }
      break;
    case 3: {
      double r = 1.0 / dy, dy2i = r * r;
      for (int i = 1; i <= n; i++) for (int j = 1; j <= m - 1; j++) for (int k = 1; k <= l; k++) {
        atom(4, i, j, k) = dy2i * amh(t.y[j], ih) * std::cos(t.y[j]) / std::cos(t.yv[j]);
        atom(6, i, j, k) = dy2i * amh(t.y[j + 1], ih) * std::cos(t.y[j + 1]) / std::cos(t.yv[j]);
        atom(5, i, j, k) = -(atom(4, i, j, k) + atom(6, i, j, k));
      }
    } break;
    case 4: {
      double r = 1.0 / dz, rdz2i = r * r;
      for (int k = 1; k <= l; k++) {
        double h1 = 1. / (t.dfzT[k] * t.dfzW[k]), h2 = 1. / (t.dfzT[k] * t.dfzW[k - 1]);
        for (int j = 1; j <= m; j++) for (int i = 1; i <= n; i++) {
          atom(14, i, j, k) = h2 * rdz2i;
          atom(23, i, j, k) = h1 * rdz2i;
          atom(5, i, j, k) = -(atom(14, i, j, k) + atom(23, i, j, k));
        }
      }
    } break;
    case 5:
      for (int j = 1; j <= m - 1; j++) {
        double v = bmh(t.yv[j], ih) - amh(t.yv[j], ih) * std::tan(t.yv[j]) * std::tan(t.yv[j]) + bmhy(t.yv[j], ih) * std::tan(t.yv[j]);
        for (int i = 1; i <= n; i++) for (int k = 1; k <= l; k++) atom(5, i, j, k) = v;
      }
      break;
    case 6:
      for (int j = 1; j <= m - 1; j++) {
        double tand2 = std::tan(t.yv[j]), cosd2 = std::cos(t.yv[j]);
        double v2 = -((amh(t.yv[j], ih) + bmh(t.yv[j], ih)) * tand2 - bmhy(t.yv[j], ih)) / (dx * cosd2);
        double v8 = ((amh(t.yv[j], ih) + bmh(t.yv[j], ih)) * tand2 - bmhy(t.yv[j], ih)) / (dx * cosd2);
        for (int i = 1; i <= n; i++) for (int k = 1; k <= l; k++) { atom(2, i, j, k) = v2; atom(8, i, j, k) = v8; }
      }
      break;
  }
}
// spf.F90:143-196
void pderiv(T& t, int type, Atom& atom) {
  int n = t.n, m = t.m, l = t.l;
  atom.zero();
  switch (type) {
    case 1:
      for (int j = 1; j <= m; j++) {
        double cos2i = 1.0 / (2 * std::cos(t.y[j]) * t.dx);
        for (int i = 1; i <= n; i++) for (int k = 1; k <= l; k++) {
          atom(2, i, j, k) = -cos2i; atom(4, i, j, k) = cos2i; atom(1, i, j, k) = -cos2i; atom(5, i, j, k) = cos2i;
        }
      }
      break;
    case 2:
      for (int j = 1; j <= m; j++) {
        double cos2i = 1. / (2 * std::cos(t.y[j]) * t.dy);
        double cvm = std::cos(t.yv[j - 1]), cv = std::cos(t.yv[j]);
        for (int i = 1; i <= n; i++) for (int k = 1; k <= l; k++) {
          atom(4, i, j, k) = -cvm * cos2i; atom(2, i, j, k) = cv * cos2i;
          atom(1, i, j, k) = -cvm * cos2i; atom(5, i, j, k) = cv * cos2i;
        }
      }
      break;
    case 3: {
      double dzi = 1.0 / t.dz;
      for (int k = 1; k <= l; k++) for (int j = 1; j <= m; j++) for (int i = 1; i <= n; i++) {
        atom(5, i, j, k) = dzi / t.dfzT[k];
        atom(14, i, j, k) = -dzi / t.dfzT[k];
      }
    } break;
  }
}
// spf.F90:198-283
void tderiv(T& t, int type, Atom& atom) {
  int n = t.n, m = t.m, l = t.l;
  atom.zero();
  switch (type) {
    case 1: case 2:
      for (int j = 1; j <= m; j++) for (int i = 1; i <= n; i++) atom(5, i, j, l) = 1.0;
      break;
    case 3:
      for (int i = 1; i <= n; i++) for (int j = 1; j <= m; j++) {
        double c = 1.0 / (std::cos(t.y[j]) * t.dx); double cosdx2i = c * c;
        for (int k = 1; k <= l; k++) {
          atom(2, i, j, k) = cosdx2i * (1 - t.lm(i, j, l));
          atom(5, i, j, k) = -2 * cosdx2i * (1 - t.lm(i, j, l));
          atom(8, i, j, k) = cosdx2i * (1 - t.lm(i, j, l));
        }
      }
      break;
    case 4: {
      double r = 1.0 / t.dy, dy2i = r * r;
      for (int i = 1; i <= n; i++) for (int k = 1; k <= l; k++) for (int j = 1; j <= m; j++) {
        atom(4, i, j, k) = (dy2i * std::cos(t.yv[j - 1]) / std::cos(t.y[j])) * (1 - t.lm(i, j, l));
        atom(6, i, j, k) = (dy2i * std::cos(t.yv[j]) / std::cos(t.y[j])) * (1 - t.lm(i, j, l));
        atom(5, i, j, k) = -(atom(4, i, j, k) + atom(6, i, j, k));
      }
    } break;
    case 5: {
      double r = 1.0 / t.dz, dz2i = r * r;
      for (int k = 1; k <= l - 1; k++) {
        double h1 = 1. / (t.dfzT[k] * t.dfzW[k]), h2 = 1. / (t.dfzT[k] * t.dfzW[k - 1]);
        for (int j = 1; j <= m; j++) for (int i = 1; i <= n; i++) {
          atom(14, i, j, k) = h2 * dz2i * (1 - t.lm(i, j, l));
          atom(23, i, j, k) = h1 * dz2i * (1 - t.lm(i, j, l));
          atom(5, i, j, k) = -(atom(14, i, j, k) + atom(23, i, j, k));
        }
      }
      int k = l;
      double h2 = 1. / (t.dfzT[k] * t.dfzW[k - 1]);
      for (int i = 1; i <= n; i++) for (int j = 1; j <= m; j++) {
        atom(14, i, j, k) = h2 * dz2i * (1 - t.lm(i, j, l));
        atom(23, i, j, k) = 0.0;
        atom(5, i, j, k) = -(atom(14, i, j, k) + atom(23, i, j, k));
      }
    } break;
    case 6:
      for (int i = 1; i <= n; i++) for (int j = 1; j <= m; j++) for (int k = 1; k <= l; k++) {
        atom(23, i, j, k) = 1.0 * (1 - t.lm(i, j, l));
        atom(5, i, j, k) = 1.0 * (1 - t.lm(i, j, l));
      }
      break;
    case 7:
      for (int j = 1; j <= m; j++) for (int i = 1; i <= n; i++) atom(5, i, j, 1) = 1.0;
      break;
  }
}
// spf.F90:286-316
void coriolis(T& t, int type, Atom& atom) {
  atom.zero();
  if (type == 1 || type == 2)
    for (int i = 1; i <= t.n; i++) for (int j = 1; j <= t.m - 1; j++) {
      double corv = std::sin(t.yv[j]) * t.cfg.coriolis_on;
      for (int k = 1; k <= t.l; k++) atom(5, i, j, k) = corv;
    }
}
// spf.F90:319-359
void gradp(T& t, int type, Atom& atom) {
  int n = t.n, m = t.m, l = t.l;
  atom.zero();
  switch (type) {
    case 1:
      for (int i = 1; i <= n; i++) for (int j = 1; j <= m - 1; j++) {
        double cosdxi = 1. / (2 * std::cos(t.yv[j]) * t.dx);
        for (int k = 1; k <= l; k++) { atom(5, i, j, k) = -cosdxi; atom(6, i, j, k) = -cosdxi; atom(8, i, j, k) = cosdxi; atom(9, i, j, k) = cosdxi; }
      }
      break;
    case 2: {
      double dyi = 1. / (2 * t.dy);
      for (int i = 1; i <= n; i++) for (int j = 1; j <= m - 1; j++) for (int k = 1; k <= l; k++) {
        atom(5, i, j, k) = -dyi; atom(8, i, j, k) = -dyi; atom(6, i, j, k) = dyi; atom(9, i, j, k) = dyi;
      }
    } break;
    case 3: {
      double dzi = 1. / t.dz;
      for (int k = 1; k <= l; k++) for (int j = 1; j <= m; j++) for (int i = 1; i <= n; i++) {
        atom(5, i, j, k) = -dzi / t.dfzW[k];
        atom(23, i, j, k) = dzi / t.dfzW[k];
      }
    } break;
  }
}
// spf.F90:347-359
void masksi(T& t, Atom& atom, std::vector<double>& mask) {
  atom.zero();
  for (int j = 1; j <= t.m; j++) for (int i = 1; i <= t.n; i++) atom(5, i, j, t.l) = t.s2(mask, i, j);
}

// An/Al(:,A,B,:,:,1:l) (+)= c*atom helpers (Fortran array statements of usrc.F90)
template <class F> void for_all(T& t, F f) {
  for (int k = 1; k <= t.l; k++) for (int j = 1; j <= t.m; j++) for (int i = 1; i <= t.n; i++) for (int s = 1; s <= 27; s++) f(s, i, j, k);
}

// usrc.F90:588-772
void lin(T& t) {
  double* par = t.par;
  Atom ucsi(t), uxx(t), uyy(t), uzz(t), uxs(t), fu(t), px(t), sc(t), tcb(t), mc(t);
  double EV = par[EK_V], EH = par[EK_H], ph = (1 - par[MIXP]) * par[PE_H], pv = par[PE_V];
  double lambda = par[LAMB], xes = par[NLES], bi = par[BIOT], Ra = par[RAYL];
  int TRES = t.cfg.tres, SRES = t.cfg.sres;
  std::fill(t.Al.begin(), t.Al.end(), 0.0);
  // u-equation
  uderiv(t, 2, uxx); uderiv(t, 3, uyy); uderiv(t, 4, uzz); uderiv(t, 5, ucsi); uderiv(t, 6, uxs);
  coriolis(t, 1, fu); gradp(t, 1, px);
  for_all(t, [&](int s, int i, int j, int k) {
    t.AL(s, UU, UU, i, j, k) = -EH * (uxx(s, i, j, k) + uyy(s, i, j, k) + ucsi(s, i, j, k)) - EV * uzz(s, i, j, k);
    t.AL(s, UU, VV, i, j, k) = -fu(s, i, j, k) - EH * uxs(s, i, j, k);
    t.AL(s, UU, PP, i, j, k) = px(s, i, j, k);
  });
  // v-equation
  vderiv(t, 2, uxx); vderiv(t, 3, uyy); vderiv(t, 4, uzz); vderiv(t, 5, ucsi); vderiv(t, 6, uxs);
  coriolis(t, 2, fu); gradp(t, 2, px);
  for_all(t, [&](int s, int i, int j, int k) {
    t.AL(s, VV, UU, i, j, k) = fu(s, i, j, k) - EH * uxs(s, i, j, k);
    t.AL(s, VV, VV, i, j, k) = -EH * (uxx(s, i, j, k) + uyy(s, i, j, k) + ucsi(s, i, j, k)) - EV * uzz(s, i, j, k);
    t.AL(s, VV, PP, i, j, k) = px(s, i, j, k);
  });
  // w-equation
  gradp(t, 3, px); tderiv(t, 6, uxs);
  for_all(t, [&](int s, int i, int j, int k) {
    t.AL(s, WW, PP, i, j, k) = px(s, i, j, k);
    t.AL(s, WW, TT, i, j, k) = -Ra * (1. + xes * alpt1) * uxs(s, i, j, k) / 2.;
    t.AL(s, WW, SS, i, j, k) = lambda * Ra * uxs(s, i, j, k) / 2.;
  });
  // p-equation
  pderiv(t, 1, uxx); pderiv(t, 2, uyy); pderiv(t, 3, uzz);
  for_all(t, [&](int s, int i, int j, int k) {
    t.AL(s, PP, UU, i, j, k) = uxx(s, i, j, k);
    t.AL(s, PP, VV, i, j, k) = uyy(s, i, j, k);
    t.AL(s, PP, WW, i, j, k) = uzz(s, i, j, k);
  });
  // T-equation
  Atom& tc = fu; Atom& txx = uxx; Atom& tyy = uyy; Atom& tzz = uzz;
  tderiv(t, 1, tc); tderiv(t, 2, sc); tderiv(t, 3, txx); tderiv(t, 4, tyy); tderiv(t, 5, tzz); tderiv(t, 7, tcb);
  masksi(t, mc, t.msi);
  double dedt = t.lvsc * t.eta * t.qdim * (deltat / t.qdim) * t.dqso;
  if (t.cfg.coupled_T == 1) {
    for_all(t, [&](int s, int i, int j, int k) {
      t.AL(s, TT, TT, i, j, k) = -ph * (txx(s, i, j, k) + tyy(s, i, j, k)) - pv * tzz(s, i, j, k) + t.Ooa * tc(s, i, j, k) +
                                 dedt * sc(s, i, j, k) +
                                 mc(s, i, j, k) * (t.QTnd * t.zeta * tc(s, i, j, k) - t.Ooa * tc(s, i, j, k) - dedt * sc(s, i, j, k));
      t.AL(s, TT, SS, i, j, k) = -t.QTnd * t.zeta * t.a0 * mc(s, i, j, k);
    });
  } else {
    for_all(t, [&](int s, int i, int j, int k) {
      t.AL(s, TT, TT, i, j, k) = -ph * (txx(s, i, j, k) + tyy(s, i, j, k)) - pv * tzz(s, i, j, k) + TRES * bi * tc(s, i, j, k);
    });
  }
  // S-equation
  dedt = t.nus * (deltat / t.qdim) * t.dqso;
  double pQSnd = par[COMB] * par[SALT] * t.QSnd;
  if (t.cfg.coupled_S == 1) {
    for_all(t, [&](int s, int i, int j, int k) {
      t.AL(s, SS, SS, i, j, k) = -ph * (txx(s, i, j, k) + tyy(s, i, j, k)) - pv * tzz(s, i, j, k) -
                                 mc(s, i, j, k) * pQSnd * t.zeta * t.a0 / (rhodim * t.Lf);
      double QSoa = -dedt * sc(s, i, j, k);
      double QSos = pQSnd * t.zeta / (rhodim * t.Lf);
      t.AL(s, SS, TT, i, j, k) = QSoa + mc(s, i, j, k) * (QSos - QSoa);
    });
  } else {
    for_all(t, [&](int s, int i, int j, int k) {
      t.AL(s, SS, SS, i, j, k) = -ph * (txx(s, i, j, k) + tyy(s, i, j, k)) - pv * tzz(s, i, j, k) + SRES * bi * sc(s, i, j, k);
    });
  }
}

struct Fields { F3 u, v, w, p, t, s; };
void alloc_fields(T& t, Fields& f) {
  int n = t.n, m = t.m, l = t.l;
  f.u.alloc(0, n, 0, m, 0, l + 1); f.v.alloc(0, n, 0, m, 0, l + 1);
  f.w.alloc(0, n + 1, 0, m + 1, 0, l); f.p.alloc(0, n + 1, 0, m + 1, 0, l + 1);
  f.t.alloc(0, n + 1, 0, m + 1, 0, l + 1); f.s.alloc(0, n + 1, 0, m + 1, 0, l + 1);
}
// usrc.F90:997-1104
void usol(T& tt, const double* un, Fields& f) {
  int n = tt.n, m = tt.m, l = tt.l, N = n, M = m;
  F3 &u = f.u, &v = f.v, &w = f.w, &p = f.p, &t = f.t, &s = f.s;
  u.zero(); v.zero(); w.zero(); p.zero(); t.zero(); s.zero();
  for (int k = 1; k <= l; k++) for (int j = 1; j <= m; j++) for (int i = 1; i <= n; i++) {
    u(i, j, k) = un[tt.find_row2(i, j, k, UU) - 1];
    v(i, j, k) = un[tt.find_row2(i, j, k, VV) - 1];
    w(i, j, k) = un[tt.find_row2(i, j, k, WW) - 1];
    p(i, j, k) = un[tt.find_row2(i, j, k, PP) - 1];
    t(i, j, k) = un[tt.find_row2(i, j, k, TT) - 1];
    s(i, j, k) = un[tt.find_row2(i, j, k, SS) - 1];
  }
  for (int k = 1; k <= l; k++) for (int j = 1; j <= m; j++) {
    if (tt.periodic) {
      u(0, j, k) = u(N, j, k); v(0, j, k) = v(N, j, k);
      w(N + 1, j, k) = w(1, j, k); w(0, j, k) = w(N, j, k);
      p(N + 1, j, k) = p(1, j, k); p(0, j, k) = p(N, j, k);
      t(N + 1, j, k) = t(1, j, k); t(0, j, k) = t(N, j, k);
      s(N + 1, j, k) = s(1, j, k); s(0, j, k) = s(N, j, k);
    } else {
      u(0, j, k) = 0.0; u(N, j, k) = 0.0; v(0, j, k) = 0.0; v(N, j, k) = 0.0;
      p(0, j, k) = 0.0; p(N + 1, j, k) = 0.0;
      t(0, j, k) = t(1, j, k); t(N + 1, j, k) = t(N, j, k);
      s(0, j, k) = s(1, j, k); s(N + 1, j, k) = s(N, j, k);
    }
  }
  for (int k = 1; k <= l; k++) for (int i = 1; i <= n; i++) {
    u(i, 0, k) = 0.0; u(i, M, k) = 0.0; v(i, 0, k) = 0.0; v(i, M, k) = 0.0;
    p(i, 0, k) = 0.0; p(i, M + 1, k) = 0.0;
    t(i, 0, k) = t(i, 1, k); t(i, M + 1, k) = t(i, M, k);
    s(i, 0, k) = s(i, 1, k); s(i, M + 1, k) = s(i, M, k);
  }
  for (int j = 1; j <= m; j++) for (int i = 1; i <= n; i++) {
    u(i, j, 0) = u(i, j, 1); u(i, j, l + 1) = u(i, j, l);
    v(i, j, 0) = v(i, j, 1); v(i, j, l + 1) = v(i, j, l);
    w(i, j, l) = 0.0; w(i, j, 0) = 0.0;
    p(i, j, l + 1) = 0.0; p(i, j, 0) = 0.0;
    t(i, j, l + 1) = t(i, j, l); t(i, j, 0) = t(i, j, 1);
    s(i, j, l + 1) = s(i, j, l); s(i, j, 0) = s(i, j, 1);
  }
  for (int i = 1; i <= n; i++) for (int j = 1; j <= m; j++) for (int k = 1; k <= l; k++)
    if (tt.lm(i, j, k) == 1) {
      u(i, j, k) = 0.0; v(i, j, k) = 0.0; u(i - 1, j, k) = 0.0; v(i - 1, j, k) = 0.0;
      u(i, j - 1, k) = 0.0; v(i, j - 1, k) = 0.0; u(i - 1, j - 1, k) = 0.0; v(i - 1, j - 1, k) = 0.0;
    }
}

// spf.F90:362-478 ; c is T or S
void tnlin(T& tt, int type, Atom& atom, Fields& f, F3& c) {
  int n = tt.n, m = tt.m, l = tt.l;
  F3 &u = f.u, &v = f.v, &w = f.w;
  double dx = tt.dx, dy = tt.dy, dz = tt.dz;
  atom.zero();
#define LS(i, j) (1 - tt.lm(i, j, l))
  switch (type) {
    case 1: for (int k = 1; k <= l; k++) for (int j = 1; j <= m; j++) for (int i = 1; i <= n; i++) atom(5, i, j, k) = 1.0; break;
    case 2:
      for (int k = 1; k <= l; k++) for (int j = 1; j <= m; j++) {
        double costdxi = 1.0 / (4 * std::cos(tt.y[j]) * dx);
        for (int i = 1; i <= n; i++) {
          atom(2, i, j, k) = -(c(i, j, k) + c(i - 1, j, k)) * costdxi * LS(i, j);
          atom(4, i, j, k) = (c(i + 1, j, k) + c(i, j, k)) * costdxi * LS(i, j);
          atom(1, i, j, k) = -(c(i, j, k) + c(i - 1, j, k)) * costdxi * LS(i, j);
          atom(5, i, j, k) = (c(i + 1, j, k) + c(i, j, k)) * costdxi * LS(i, j);
        }
      }
      break;
    case 3:
      for (int k = 1; k <= l; k++) for (int j = 1; j <= m; j++) {
        double costdxi = 1.0 / (4 * std::cos(tt.y[j]) * dx);
        for (int i = 1; i <= n; i++) {
          atom(2, i, j, k) = -(u(i - 1, j, k) + u(i - 1, j - 1, k)) * costdxi * LS(i, j);
          atom(8, i, j, k) = (u(i, j, k) + u(i, j - 1, k)) * costdxi * LS(i, j);
          atom(5, i, j, k) = atom(2, i, j, k) + atom(8, i, j, k);
        }
      }
      break;
    case 4:
      for (int k = 1; k <= l; k++) for (int j = 1; j <= m; j++) {
        double costdxi = 1.0 / (4 * std::cos(tt.y[j]) * dy);
        for (int i = 1; i <= n; i++) {
          atom(4, i, j, k) = -costdxi * (c(i, j, k) + c(i, j - 1, k)) * std::cos(tt.yv[j - 1]) * LS(i, j);
          atom(1, i, j, k) = -costdxi * (c(i, j, k) + c(i, j - 1, k)) * std::cos(tt.yv[j - 1]) * LS(i, j);
          atom(5, i, j, k) = costdxi * (c(i, j + 1, k) + c(i, j, k)) * std::cos(tt.yv[j]) * LS(i, j);
          atom(2, i, j, k) = costdxi * (c(i, j + 1, k) + c(i, j, k)) * std::cos(tt.yv[j]) * LS(i, j);
        }
      }
      break;
    case 5:
      for (int k = 1; k <= l; k++) for (int j = 1; j <= m; j++) {
        double costdxi = 1.0 / (4 * std::cos(tt.y[j]) * dy);
        for (int i = 1; i <= n; i++) {
          atom(4, i, j, k) = -(v(i, j - 1, k) + v(i - 1, j - 1, k)) * costdxi * std::cos(tt.yv[j - 1]) * LS(i, j);
          atom(6, i, j, k) = (v(i, j, k) + v(i - 1, j, k)) * costdxi * std::cos(tt.yv[j]) * LS(i, j);
          atom(5, i, j, k) = atom(4, i, j, k) + atom(6, i, j, k);
        }
      }
      break;
    case 6: {
      double tdzi = 1.0 / (2 * dz);
      for (int j = 1; j <= m; j++) for (int i = 1; i <= n; i++) {
        for (int k = 1; k <= l - 1; k++) {
          atom(14, i, j, k) = -tdzi * LS(i, j) * (c(i, j, k) + c(i, j, k - 1)) / tt.dfzT[k];
          atom(5, i, j, k) = tdzi * LS(i, j) * (c(i, j, k + 1) + c(i, j, k)) / tt.dfzT[k];
        }
        int k = l;
        atom(14, i, j, k) = -tdzi * LS(i, j) * (c(i, j, k) + c(i, j, k - 1)) / tt.dfzT[k];
        atom(5, i, j, k) = 0.0;
      }
    } break;
    case 7: {
      double tdzi = 1.0 / (2 * dz);
      for (int k = 1; k <= l; k++) for (int j = 1; j <= m; j++) for (int i = 1; i <= n; i++) {
        atom(14, i, j, k) = -w(i, j, k - 1) * LS(i, j) * tdzi / tt.dfzT[k];
        atom(23, i, j, k) = w(i, j, k) * LS(i, j) * tdzi / tt.dfzT[k];
        atom(5, i, j, k) = atom(14, i, j, k) + atom(23, i, j, k);
      }
    } break;
  }
#undef LS
}
// spf.F90:480-541
void wnlin(T& tt, int type, Atom& atom, F3& t) {
  int n = tt.n, m = tt.m, l = tt.l;
  atom.zero();
  for (int k = 1; k <= l - 1; k++) for (int j = 1; j <= m; j++) for (int i = 1; i <= n; i++) {
    switch (type) {
      case 1:
        atom(23, i, j, k) = (t(i, j, k) + t(i, j, k + 1)) / 2.;
        atom(5, i, j, k) = (t(i, j, k) + t(i, j, k + 1)) / 2.;
        break;
      case 2:
        atom(23, i, j, k) = t(i, j, k + 1) / 4.;
        atom(5, i, j, k) = (t(i, j, k) + 2 * t(i, j, k + 1)) / 4.;
        break;
      case 3: {
        double q = t(i, j, k) + t(i, j, k + 1);
        atom(5, i, j, k) = 0.375 * (q * q);
        atom(23, i, j, k) = 0.375 * (q * q);
      } break;
      case 4:
        atom(5, i, j, k) = 0.125 * (t(i, j, k) * t(i, j, k) + 3 * t(i, j, k + 1) * t(i, j, k) + 3 * t(i, j, k + 1) * t(i, j, k + 1));
        atom(23, i, j, k) = 0.125 * t(i, j, k + 1) * t(i, j, k + 1);
        break;
    }
  }
}
// spf.F90:543-665
void unlin(T& tt, int type, Atom& atom, Fields& f) {
  int n = tt.n, m = tt.m, l = tt.l;
  F3 &u = f.u, &v = f.v, &w = f.w;
  double dx = tt.dx, dy = tt.dy, dz = tt.dz;
  std::vector<double>& yv = tt.yv;
  atom.zero();
  switch (type) {
    case 1: case 2: {
      double fac = (type == 2) ? 2.0 : 1.0;
      for (int j = 1; j <= m; j++) {
        double costdxi = 1.0 / (2 * std::cos(yv[j]) * dx);
        for (int k = 1; k <= l; k++) {
          for (int i = 1; i <= n - 1; i++) atom(8, i, j, k) = (type == 2 ? fac * u(i + 1, j, k) : u(i + 1, j, k)) * costdxi;
          for (int i = 2; i <= n; i++) atom(2, i, j, k) = -(type == 2 ? fac * u(i - 1, j, k) : u(i - 1, j, k)) * costdxi;
        }
      }
    } break;
    case 3: case 4: {
      F3& q = (type == 3) ? v : u;
      for (int k = 1; k <= l; k++) for (int i = 1; i <= n; i++) {
        for (int j = 2; j <= m; j++) atom(4, i, j, k) = -q(i, j - 1, k) * std::cos(yv[j - 1]) * (1.0 / (2 * std::cos(yv[j]) * dy));
        for (int j = 1; j <= m - 1; j++) atom(6, i, j, k) = q(i, j + 1, k) * std::cos(yv[j + 1]) * (1.0 / (2 * std::cos(yv[j]) * dy));
      }
    } break;
    case 5:
      for (int k = 1; k <= l; k++) {
        double tdzi = 1.0 / (8 * tt.dfzT[k] * dz);
        for (int j = 1; j <= m; j++) for (int i = 1; i <= n; i++) {
          atom(23, i, j, k) = (w(i, j, k) + w(i, j + 1, k) + w(i + 1, j, k) + w(i + 1, j + 1, k)) * tdzi;
          atom(14, i, j, k) = -(w(i, j, k - 1) + w(i, j + 1, k - 1) + w(i + 1, j, k - 1) + w(i + 1, j + 1, k - 1)) * tdzi;
          atom(5, i, j, k) = atom(14, i, j, k) + atom(23, i, j, k);
        }
      }
      break;
    case 6:
      for (int j = 1; j <= m; j++) for (int i = 1; i <= n; i++) for (int k = 1; k <= l; k++) {
        double tdzi = 1.0 / (8 * tt.dfzT[k] * dz);
        double up = (u(i, j, k) + u(i, j, k + 1)) * tdzi, dn = -(u(i, j, k) + u(i, j, k - 1)) * tdzi;
        atom(5, i, j, k) = up; atom(6, i, j, k) = up; atom(8, i, j, k) = up; atom(9, i, j, k) = up;
        atom(14, i, j, k) = dn; atom(15, i, j, k) = dn; atom(17, i, j, k) = dn; atom(18, i, j, k) = dn;
      }
      break;
    case 7: case 8: {
      F3& q = (type == 7) ? v : u;
      for (int k = 1; k <= l; k++) for (int j = 1; j <= m; j++) {
        double tanr = std::tan(yv[j]);
        for (int i = 1; i <= n; i++) atom(5, i, j, k) = q(i, j, k) * tanr;
      }
    } break;
  }
}
// spf.F90:667-793
void vnlin(T& tt, int type, Atom& atom, Fields& f) {
  int n = tt.n, m = tt.m, l = tt.l;
  F3 &u = f.u, &v = f.v, &w = f.w;
  double dx = tt.dx, dy = tt.dy, dz = tt.dz;
  std::vector<double>& yv = tt.yv;
  atom.zero();
  switch (type) {
    case 1: case 2: {
      F3& q = (type == 1) ? u : v;
      for (int k = 1; k <= l; k++) for (int j = 1; j <= m; j++) {
        double costdxi = 1.0 / (2 * std::cos(yv[j]) * dx);
        for (int i = 1; i <= n - 1; i++) atom(8, i, j, k) = q(i + 1, j, k) * costdxi;
        for (int i = 2; i <= n; i++) atom(2, i, j, k) = -q(i - 1, j, k) * costdxi;
      }
    } break;
    case 3: case 4:
      for (int k = 1; k <= l; k++) for (int i = 1; i <= n; i++) {
        for (int j = 1; j <= m - 1; j++) {
          double costdxi = 1.0 / (2 * std::cos(yv[j]) * dy);
          atom(6, i, j, k) = (type == 4 ? 2 * v(i, j + 1, k) : v(i, j + 1, k)) * std::cos(yv[j + 1]) * costdxi;
        }
        for (int j = 2; j <= m; j++) {
          double costdxi = 1.0 / (2 * std::cos(yv[j]) * dy);
          atom(4, i, j, k) = (type == 4 ? -2 * v(i, j - 1, k) : -v(i, j - 1, k)) * std::cos(yv[j - 1]) * costdxi;
        }
      }
      break;
    case 5:
      for (int k = 1; k <= l; k++) {
        double tdzi = 1.0 / (8 * tt.dfzT[k] * dz);
        for (int j = 1; j <= m; j++) for (int i = 1; i <= n; i++) {
          atom(23, i, j, k) = (w(i, j, k) + w(i, j + 1, k) + w(i + 1, j, k) + w(i + 1, j + 1, k)) * tdzi;
          atom(14, i, j, k) = -(w(i, j, k - 1) + w(i, j + 1, k - 1) + w(i + 1, j, k - 1) + w(i + 1, j + 1, k - 1)) * tdzi;
          atom(5, i, j, k) = atom(14, i, j, k) + atom(23, i, j, k);
        }
      }
      break;
    case 6:
      for (int j = 1; j <= m; j++) for (int i = 1; i <= n; i++) for (int k = 1; k <= l; k++) {
        double tdzi = 1.0 / (8 * tt.dfzT[k] * dz);
        double up = (v(i, j, k) + v(i, j, k + 1)) * tdzi, dn = -(v(i, j, k) + v(i, j, k - 1)) * tdzi;
        atom(5, i, j, k) = up; atom(6, i, j, k) = up; atom(8, i, j, k) = up; atom(9, i, j, k) = up;
        atom(14, i, j, k) = dn; atom(15, i, j, k) = dn; atom(17, i, j, k) = dn; atom(18, i, j, k) = dn;
      }
      break;
    case 7: case 8:
      for (int k = 1; k <= l; k++) for (int j = 1; j <= m; j++) {
        double tanr = std::tan(yv[j]);
        for (int i = 1; i <= n; i++) atom(5, i, j, k) = (type == 8 ? 2 * u(i, j, k) : u(i, j, k)) * tanr;
      }
      break;
  }
}

// An(:,A,B,:,:,1:l) += c * (sum of atoms)
void add_atoms(T& t, int A, int B, double c, bool scaled, std::initializer_list<Atom*> atoms) {
  for_all(t, [&](int s, int i, int j, int k) {
    double sum = 0.0; bool first = true;
    for (Atom* a : atoms) { if (first) { sum = (*a)(s, i, j, k); first = false; } else sum = sum + (*a)(s, i, j, k); }
    double& an = t.AN(s, A, B, i, j, k);
    an = scaled ? an + c * sum : an + sum;
  });
}

// usrc.F90:775-870
void nlin_rhs(T& t, const double* un) {
  double* par = t.par;
  Fields f; alloc_fields(t, f);
  double epsr = par[ROSB], Ra = par[RAYL], xes = par[NLES];
  usol(t, un, f);
  Atom a1(t), a2(t), a3(t), a4(t);
  // u-equation
  unlin(t, 1, a1, f); unlin(t, 3, a2, f); unlin(t, 5, a3, f); unlin(t, 7, a4, f);
  add_atoms(t, UU, UU, epsr, true, {&a1, &a2, &a3, &a4});
  // v-equation
  vnlin(t, 1, a1, f); vnlin(t, 3, a2, f); vnlin(t, 5, a3, f); vnlin(t, 7, a4, f);
  add_atoms(t, VV, UU, epsr, true, {&a4});
  add_atoms(t, VV, VV, epsr, true, {&a1, &a2, &a3});
  // w-equation
  wnlin(t, 2, a1, f.t); wnlin(t, 4, a2, f.t);
  for_all(t, [&](int s, int i, int j, int k) {
    double& an = t.AN(s, WW, TT, i, j, k);
    an = an - Ra * xes * alpt2 * a1(s, i, j, k) + Ra * xes * alpt3 * a2(s, i, j, k);
  });
  // T-equation
  tnlin(t, 3, a1, f, f.t); tnlin(t, 5, a2, f, f.t); tnlin(t, 7, a3, f, f.t);
  for_all(t, [&](int s, int i, int j, int k) {
    double& an = t.AN(s, TT, TT, i, j, k);
    an = an + a1(s, i, j, k) + a2(s, i, j, k) + a3(s, i, j, k);
  });
  // S-equation
  tnlin(t, 3, a1, f, f.s); tnlin(t, 5, a2, f, f.s); tnlin(t, 7, a3, f, f.s);
  for_all(t, [&](int s, int i, int j, int k) {
    double& an = t.AN(s, SS, SS, i, j, k);
    an = an + a1(s, i, j, k) + a2(s, i, j, k) + a3(s, i, j, k);
  });
}

// usrc.F90:873-995
void nlin_jac(T& t, const double* un) {
  double* par = t.par;
  Fields f; alloc_fields(t, f);
  double epsr = par[ROSB], Ra = par[RAYL], xes = par[NLES];
  usol(t, un, f);
  Atom Urux(t), uvy1(t), Urvy1(t), uwz(t), Urwz(t), uvy2(t), Urvy2(t);
  unlin(t, 2, Urux, f); unlin(t, 3, uvy1, f); unlin(t, 4, Urvy1, f); unlin(t, 5, uwz, f);
  unlin(t, 6, Urwz, f); unlin(t, 7, uvy2, f); unlin(t, 8, Urvy2, f);
  add_atoms(t, UU, UU, epsr, true, {&Urux, &uvy1, &uwz, &uvy2});
  add_atoms(t, UU, VV, epsr, true, {&Urvy1, &Urvy2});
  add_atoms(t, UU, WW, epsr, true, {&Urwz});
  Atom &uvx = Urux, &uVrx = uvy1, &Vrvy = Urvy1, &vwz = uwz, &Vrwz = Urwz, &Urt2 = uvy2;
  vnlin(t, 1, uvx, f); vnlin(t, 2, uVrx, f); vnlin(t, 4, Vrvy, f); vnlin(t, 5, vwz, f); vnlin(t, 6, Vrwz, f); vnlin(t, 8, Urt2, f);
  add_atoms(t, VV, UU, epsr, true, {&Urt2, &uVrx});
  add_atoms(t, VV, VV, epsr, true, {&uvx, &Vrvy, &vwz});
  add_atoms(t, VV, WW, epsr, true, {&Vrwz});
  Atom &t2r = Urux, &t3r = uvy1;
  wnlin(t, 1, t2r, f.t); wnlin(t, 3, t3r, f.t);
  for_all(t, [&](int s, int i, int j, int k) {
    double& an = t.AN(s, WW, TT, i, j, k);
    an = an - Ra * xes * alpt2 * t2r(s, i, j, k) + Ra * xes * alpt3 * t3r(s, i, j, k);
  });
  Atom &urTx = Urux, &Utrx = uvy1, &vrTy = Urvy1, &Vtry = uwz, &wrTz = Urwz, &Wtrz = uvy2;
  for (int pass = 0; pass < 2; pass++) {
    F3& c = pass == 0 ? f.t : f.s;
    int X = pass == 0 ? TT : SS;
    tnlin(t, 2, urTx, f, c); tnlin(t, 3, Utrx, f, c); tnlin(t, 4, vrTy, f, c);
    tnlin(t, 5, Vtry, f, c); tnlin(t, 6, wrTz, f, c); tnlin(t, 7, Wtrz, f, c);
    add_atoms(t, X, UU, 1.0, false, {&urTx});
    add_atoms(t, X, VV, 1.0, false, {&vrTy});
    add_atoms(t, X, WW, 1.0, false, {&wrTz});
    for_all(t, [&](int s, int i, int j, int k) {
      double& an = t.AN(s, X, X, i, j, k);
      an = an + Utrx(s, i, j, k) + Vtry(s, i, j, k) + Wtrz(s, i, j, k);
    });
  }
}

// boundary.F90:2-393 -- statement for statement
void boundaries(T& t) {
  int n = t.n, m = t.m, l = t.l;
  auto fold = [&](int dst, int src, int B, int i, int j, int k) { // An(dst,:,B) += An(src,:,B)
    for (int A = 1; A <= 6; A++) t.AN(dst, A, B, i, j, k) = t.AN(dst, A, B, i, j, k) + t.AN(src, A, B, i, j, k);
  };
  auto zc = [&](int s, int B, int i, int j, int k) { for (int A = 1; A <= 6; A++) t.AN(s, A, B, i, j, k) = 0.0; };       // An(s,:,B)=0
  auto zs = [&](int s, int i, int j, int k) { for (int B = 1; B <= 6; B++) for (int A = 1; A <= 6; A++) t.AN(s, A, B, i, j, k) = 0.0; }; // An(s,:,:)=0
  auto zrow = [&](int A, int i, int j, int k) { for (int B = 1; B <= 6; B++) for (int s = 1; s <= 27; s++) t.AN(s, A, B, i, j, k) = 0.0; }; // An(:,A,:)=0
  auto zuv = [&](int s, int i, int j, int k) { zc(s, UU, i, j, k); zc(s, VV, i, j, k); };
  for (int i = 1; i <= n; i++) for (int j = 1; j <= m; j++) for (int k = 1; k <= l; k++) {
    int southw = t.lm(i - 1, j - 1, k), west = t.lm(i - 1, j, k), nwest = t.lm(i - 1, j + 1, k);
    int south = t.lm(i, j - 1, k), center = t.lm(i, j, k), north = t.lm(i, j + 1, k);
    int southe = t.lm(i + 1, j - 1, k), east = t.lm(i + 1, j, k), neast = t.lm(i + 1, j + 1, k);
    int southwb = t.lm(i - 1, j - 1, k - 1), westb = t.lm(i - 1, j, k - 1), nwestb = t.lm(i - 1, j + 1, k - 1);
    int southb = t.lm(i, j - 1, k - 1), bottom = t.lm(i, j, k - 1), northb = t.lm(i, j + 1, k - 1);
    int southeb = t.lm(i + 1, j - 1, k - 1), eastb = t.lm(i + 1, j, k - 1), neastb = t.lm(i + 1, j + 1, k - 1);
    int southwt = t.lm(i - 1, j - 1, k + 1), westt = t.lm(i - 1, j, k + 1), nwestt = t.lm(i - 1, j + 1, k + 1);
    int southt = t.lm(i, j - 1, k + 1), top = t.lm(i, j, k + 1), northt = t.lm(i, j + 1, k + 1);
    int southet = t.lm(i + 1, j - 1, k + 1), eastt = t.lm(i + 1, j, k + 1), neastt = t.lm(i + 1, j + 1, k + 1);
    int southee = 0, easteast = 0, northee = 0, nnorthee = 0, nnwest = 0, nnorth = 0, nneast = 0;
    if (i < n) {
      southee = t.lm(i + 2, j - 1, k); easteast = t.lm(i + 2, j, k); northee = t.lm(i + 2, j + 1, k);
      if (j < m) nnorthee = t.lm(i + 2, j + 2, k);
    }
    if (j < m) { nnwest = t.lm(i, j + 2, k); nnorth = t.lm(i, j + 2, k); nneast = t.lm(i, j + 2, k); } // sic: all three read (i,j+2,k)
    if (center == OCEAN) {
      if (bottom == LAND) {
        if (westb == LAND && southwb == LAND && southb == LAND) { fold(1, 10, UU, i, j, k); fold(1, 10, VV, i, j, k); }
        zuv(10, i, j, k);
        if (westb == LAND && neastb == LAND && northb == LAND) { fold(2, 11, UU, i, j, k); fold(2, 11, VV, i, j, k); }
        zuv(11, i, j, k);
        if (eastb == LAND && southeb == LAND && southb == LAND) { fold(4, 13, UU, i, j, k); fold(4, 13, VV, i, j, k); }
        zuv(13, i, j, k);
        if (eastb == LAND && neastb == LAND && northb == LAND) { fold(5, 14, UU, i, j, k); fold(5, 14, VV, i, j, k); }
        fold(5, 14, TT, i, j, k); fold(5, 14, SS, i, j, k);
        zs(14, i, j, k);
      }
      if (southwb == LAND) zs(10, i, j, k);
      if (westb == LAND) zs(11, i, j, k);
      if (nwestb == LAND) zs(12, i, j, k);
      if (southb == LAND) zs(13, i, j, k);
      if (northb == LAND) zs(15, i, j, k);
      if (southeb == LAND) zs(16, i, j, k);
      if (eastb == LAND) zs(17, i, j, k);
      if (neastb == LAND) zs(18, i, j, k);
      if (top == LAND) {
        if (westt == LAND && southwt == LAND && southt == LAND) { fold(1, 19, UU, i, j, k); fold(1, 19, VV, i, j, k); }
        zuv(19, i, j, k);
        if (westt == LAND && nwestt == LAND && northt == LAND) { fold(2, 20, UU, i, j, k); fold(2, 20, VV, i, j, k); }
        zuv(20, i, j, k);
        if (eastt == LAND && southet == LAND && southt == LAND) { fold(4, 22, UU, i, j, k); fold(4, 22, VV, i, j, k); }
        zuv(22, i, j, k);
        if (eastt == LAND && neastt == LAND && northt == LAND) { fold(5, 23, UU, i, j, k); fold(5, 23, VV, i, j, k); }
        fold(5, 23, TT, i, j, k); fold(5, 23, SS, i, j, k);
        zs(23, i, j, k);
        t.Frc[t.find_row2(i, j, k, WW) - 1] = 0.0;
        zrow(WW, i, j, k);
        for (int A = 1; A <= 6; A++) {
          t.AN(5, A, WW, i, j, k) = 1.0e-10; t.AN(6, A, WW, i, j, k) = 1.0e-10;
          t.AN(8, A, WW, i, j, k) = 1.0e-10; t.AN(9, A, WW, i, j, k) = 1.0e-10;
        }
        t.AN(5, WW, WW, i, j, k) = 1.0;
      }
      if (southwt == LAND) zs(19, i, j, k);
      if (westt == LAND) zs(20, i, j, k);
      if (nwestt == LAND) zs(21, i, j, k);
      if (southt == LAND) zs(22, i, j, k);
      if (northt == LAND) zs(24, i, j, k);
      if (southet == LAND) zs(25, i, j, k);
      if (eastt == LAND) zs(26, i, j, k);
      if (neastt == LAND) zs(27, i, j, k);
      if (southw == LAND) zuv(1, i, j, k);
      if (west == LAND) {
        fold(5, 2, TT, i, j, k); fold(5, 2, SS, i, j, k);
        zs(2, i, j, k); zuv(1, i, j, k);
      }
      if (nwest == LAND) { zuv(2, i, j, k); zuv(3, i, j, k); }
      else if (j < m) { if (nnwest == LAND) zuv(3, i, j, k); }
      if (south == LAND) {
        fold(5, 4, SS, i, j, k); fold(5, 4, TT, i, j, k);
        zs(4, i, j, k); zuv(1, i, j, k);
      }
      if (north == LAND) {
        zuv(2, i, j, k);
        t.AN(2, PP, UU, i, j, k) = 0.0; t.AN(2, PP, VV, i, j, k) = 0.0; t.AN(5, PP, UU, i, j, k) = 0.0; t.AN(5, PP, VV, i, j, k) = 0.0;
        t.Frc[t.find_row2(i, j, k, VV) - 1] = 0.0;
        zrow(VV, i, j, k); zc(5, VV, i, j, k); t.AN(5, VV, VV, i, j, k) = 1.0;
        t.Frc[t.find_row2(i, j, k, UU) - 1] = 0.0;
        zrow(UU, i, j, k); zc(5, UU, i, j, k); t.AN(5, UU, UU, i, j, k) = 1.0;
        fold(5, 6, SS, i, j, k); fold(5, 6, TT, i, j, k);
        zs(6, i, j, k);
      } else if (j < m) {
        if (nnorth == LAND) { zuv(3, i, j, k); zuv(6, i, j, k); }
      }
      if (southe == LAND) { zuv(4, i, j, k); zuv(7, i, j, k); }
      else if (i < n) { if (southee == LAND) zuv(7, i, j, k); }
      if (east == LAND) {
        zuv(4, i, j, k);
        t.AN(4, PP, UU, i, j, k) = 0.0; t.AN(4, PP, VV, i, j, k) = 0.0; t.AN(5, PP, UU, i, j, k) = 0.0; t.AN(5, PP, VV, i, j, k) = 0.0;
        t.Frc[t.find_row2(i, j, k, UU) - 1] = 0.0;
        zrow(UU, i, j, k); zc(5, UU, i, j, k); t.AN(5, UU, UU, i, j, k) = 1.0;
        t.Frc[t.find_row2(i, j, k, VV) - 1] = 0.0;
        zrow(VV, i, j, k); zc(5, VV, i, j, k); t.AN(5, VV, VV, i, j, k) = 1.0;
        fold(5, 8, SS, i, j, k); fold(5, 8, TT, i, j, k);
        zs(8, i, j, k); zuv(7, i, j, k);
      } else if (i < n) {
        if (easteast == LAND) { zuv(7, i, j, k); zuv(8, i, j, k); }
      }
      if (neast == LAND) {
        t.Frc[t.find_row2(i, j, k, UU) - 1] = 0.0;
        zrow(UU, i, j, k); zc(5, UU, i, j, k); t.AN(5, UU, UU, i, j, k) = 1.0;
        t.Frc[t.find_row2(i, j, k, VV) - 1] = 0.0;
        zrow(VV, i, j, k); zc(5, VV, i, j, k); t.AN(5, VV, VV, i, j, k) = 1.0;
        zuv(7, i, j, k);
      } else if (i < n || j < m) {
        if (i < n) {
          if (northee == LAND) { zuv(8, i, j, k); zuv(9, i, j, k); }
          else if (j < m) { if (nnorthee == LAND) zuv(9, i, j, k); }
        }
        if (j < m) { if (nneast == LAND) { zuv(6, i, j, k); zuv(9, i, j, k); } }
      }
    } else {
      for (int B = 1; B <= 6; B++) for (int A = 1; A <= 6; A++) for (int s = 1; s <= 27; s++) t.AN(s, A, B, i, j, k) = 0.0;
      for (int ii = 1; ii <= NUN; ii++) { t.Frc[t.find_row2(i, j, k, ii) - 1] = 0.0; t.AN(5, ii, ii, i, j, k) = 1.0; }
    }
  }
}

// assemble.F90:142-179
void shift(T& t, int i, int j, int k, int& i2, int& j2, int& k2, int kk) {
  if (kk < 10) { k2 = k; j2 = j - 1 + (kk + 2) % 3; i2 = i - 1 + (kk - 1) / 3; }
  else if (kk < 19) { k2 = k - 1; j2 = j - 1 + (kk + 2) % 3; i2 = i - 1 + (kk - 10) / 3; }
  else { k2 = k + 1; j2 = j - 1 + (kk + 2) % 3; i2 = i - 1 + (kk - 19) / 3; }
  if (t.periodic) { if (i2 == 0) i2 = t.n; else if (i2 == t.n + 1) i2 = 1; }
}
// matetc.F90:51-88
void findex(T& t, int row, int& i, int& j, int& k, int& XX) {
  XX = (row - 1) % NUN + 1;
  int r = (row - XX) / NUN;
  int tt = r % (t.n * t.m);
  k = (r - tt) / (t.n * t.m) + 1;
  i = tt % t.n + 1;
  j = (tt - i + 1) / t.n + 1;
}
// assemble.F90:57-139
void fillcolA(T& t) {
  int n = t.n, m = t.m, l = t.l;
  t.begA.assign(t.ndim + 1, 0);
  t.jcoA.clear(); t.coA.clear();
  int v = 1, row = 1;
  for (int k = 1; k <= l; k++) for (int j = 1; j <= m; j++) for (int i = 1; i <= n; i++)
    for (int ii = 1; ii <= NUN; ii++) {
      t.begA[row - 1] = v;
      for (int kk = 1; kk <= NP; kk++) for (int jj = 1; jj <= NUN; jj++) {
        double a = t.AN(kk, ii, jj, i, j, k);
        if (std::fabs(a) > 1.0e-10) {
          int i2, j2, k2;
          shift(t, i, j, k, i2, j2, k2, kk);
          t.coA.push_back(a);
          t.jcoA.push_back(t.find_row2(i2, j2, k2, jj));
          v = v + 1;
        }
      }
      row = row + 1;
    }
  t.begA[t.ndim] = v;
}
// assemble.F90:18-54 (rowintcon is -1 inside the Fortran, usr.F90:36 / topo.F90:113)
void fillcolB(T& t) {
  std::fill(t.coB.begin(), t.coB.end(), 0.0);
  for (int k = 1; k <= t.l; k++) for (int j = 1; j <= t.m; j++) for (int i = 1; i <= t.n; i++)
    if (t.lm(i, j, k) == OCEAN) {
      if (t.lm(i + 1, j, k) != LAND) t.coB[t.find_row2(i, j, k, UU) - 1] = -t.par[ROSB];
      if (t.lm(i, j + 1, k) != LAND) t.coB[t.find_row2(i, j, k, VV) - 1] = -t.par[ROSB];
      t.coB[t.find_row2(i, j, k, TT) - 1] = -1.0;
      t.coB[t.find_row2(i, j, k, SS) - 1] = -1.0;
    }
}
// matetc.F90:147-166
void matAvec(T& t, const double* v1, double* v2) {
  for (int i = 1; i <= t.ndim; i++) {
    double s = 0.0;
    for (int v = t.begA[i - 1]; v <= t.begA[i] - 1; v++) s = t.coA[v - 1] * v1[t.jcoA[v - 1] - 1] + s;
    v2[i - 1] = s;
  }
}

// mix_imp.f:817-832
inline double isoc(T& t, int i, int j, int k) { int v = t.lm(i, j, k); return (v == OCEAN || v == PERIO) ? 1.0 : 0.0; }
// mix_imp.f:837-856
inline double tprstb(T& t, double grad, double spl) {
  double fac = t.cfg.alphaT * spl;
  double q = -grad * fac;
  return std::max(std::tanh(q * q * q), 0.0);
}
// mix_imp.f:666-727
void tprslp(T& t, double drdh, double drdz, double spl, double& slp, double& tpr) {
  const double width = 1.0, epsln = 1.0e-20;
  if (drdz == 0.0) drdz = epsln;
  slp = -drdh / drdz;
  double absslp = std::sqrt(slp * slp);
  double delta = (r0dim / t.cfg.hdim) * spl;
  double sd = width * delta;
  int tap = t.cfg.tap;
  if (tap == 1) { if (absslp > delta) { double q = delta / absslp; tpr = q * q; } else tpr = 1.0; }
  else if (tap == 2) tpr = 0.5 * (1.0 - std::tanh((absslp - delta) / sd));
  else if (tap == 3) {
    if (absslp < delta - sd && drdz < 0.0) tpr = 1.0;
    else if (absslp >= delta - sd && absslp < delta && drdz < 0.0) { double dum = (absslp - (delta - sd)) / sd; tpr = 1.0 - 3.0 * (dum * dum) + 2.0 * (dum * dum * dum); }
    else tpr = 0.0;
  } else tpr = 1.0;
}

struct MixWork {
  Fields f;
  F3 dtdxe, dsdxe, dtdyn, dsdyn, dtdzt, dsdzt, rho, drhods, drhodt, drhodzt;
  F3 Ftxe, Fsxe, Ftyn, Fsyn, Ftzt, Fszt, Ftimp, Fsimp;
  bool ready = false;
  void init(T& t) {
    if (ready) return;
    int n = t.n, m = t.m, l = t.l;
    alloc_fields(t, f);
    dtdxe.alloc(0, n, 0, m + 1, 0, l + 1); dsdxe.alloc(0, n, 0, m + 1, 0, l + 1);
    dtdyn.alloc(0, n + 1, 0, m, 0, l + 1); dsdyn.alloc(0, n + 1, 0, m, 0, l + 1);
    dtdzt.alloc(0, n + 1, 0, m + 1, 0, l); dsdzt.alloc(0, n + 1, 0, m + 1, 0, l);
    rho.alloc(0, n + 1, 0, m + 1, 0, l + 1); drhods.alloc(0, n + 1, 0, m + 1, 0, l + 1); drhodt.alloc(0, n + 1, 0, m + 1, 0, l + 1);
    drhodzt.alloc(0, n + 1, 0, m + 1, 0, l);
    Ftxe.alloc(0, n, 1, m, 1, l); Fsxe.alloc(0, n, 1, m, 1, l);
    Ftyn.alloc(1, n, 0, m, 1, l); Fsyn.alloc(1, n, 0, m, 1, l);
    Ftzt.alloc(1, n, 1, m, 0, l); Fszt.alloc(1, n, 1, m, 0, l);
    Ftimp.alloc(1, n, 1, m, 0, l); Fsimp.alloc(1, n, 1, m, 0, l);
    ready = true;
  }
};
// mix_imp.f:564-633
void dCdxt(T& t, F3& C, F3& d) {
  for (int k = 0; k <= t.l + 1; k++) for (int j = 0; j <= t.m + 1; j++) for (int i = 0; i <= t.n; i++)
    d(i, j, k) = isoc(t, i + 1, j, k) * isoc(t, i, j, k) * (C(i + 1, j, k) - C(i, j, k)) / (t.dx * std::cos(t.y[j]));
}
void dCdyt(T& t, F3& C, F3& d) {
  for (int k = 0; k <= t.l + 1; k++) for (int j = 0; j <= t.m; j++) for (int i = 0; i <= t.n + 1; i++)
    d(i, j, k) = isoc(t, i, j + 1, k) * isoc(t, i, j, k) * (C(i, j + 1, k) - C(i, j, k)) / t.dy;
}
void dCdzt(T& t, F3& C, F3& d) {
  for (int k = 0; k <= t.l; k++) for (int j = 0; j <= t.m + 1; j++) for (int i = 0; i <= t.n + 1; i++)
    d(i, j, k) = isoc(t, i, j, k + 1) * isoc(t, i, j, k) * (C(i, j, k + 1) - C(i, j, k)) / (t.dz * t.dfzW[k]);
}

// mix_imp.f:231-562
void vmix_fun(T& t, const double* un, double* mix, MixWork& W) {
  W.init(t);
  int n = t.n, m = t.m, l = t.l;
  double* par = t.par;
  Fields& f = W.f;
  usol(t, un, f);
  F3 &tt = f.t, &ss = f.s;
  double xes = par[NLES], lambda = par[LAMB], piso = par[MIXP] * par[PE_H], pgm = par[MKAP] * par[PE_H];
  double eps = (1.0 - par[ALPC]) * par[ENER] * par[PE_V], kvc = par[P_VC], sp1 = par[SPL1], sp2 = par[SPL2];
  const double epsln = 1.0e-20;
  dCdxt(t, tt, W.dtdxe); dCdxt(t, ss, W.dsdxe);
  dCdyt(t, tt, W.dtdyn); dCdyt(t, ss, W.dsdyn);
  dCdzt(t, tt, W.dtdzt); dCdzt(t, ss, W.dsdzt);
  for (int k = 0; k <= l + 1; k++) for (int j = 0; j <= m + 1; j++) for (int i = 0; i <= n + 1; i++) {
    double tv = tt(i, j, k);
    W.rho(i, j, k) = lambda * ss(i, j, k) - tv - xes * (alpt1 * tv + alpt2 * tv * tv - alpt3 * tv * tv * tv);
    W.drhodt(i, j, k) = -1.0 - xes * (alpt1 + 2.0 * alpt2 * tv - 3.0 * alpt3 * (tv * tv)); // drhodC, mix_imp.f:635-664
    W.drhods(i, j, k) = lambda;
  }
  dCdzt(t, W.rho, W.drhodzt);
  F3 &dtdxe = W.dtdxe, &dsdxe = W.dsdxe, &dtdyn = W.dtdyn, &dsdyn = W.dsdyn, &dtdzt = W.dtdzt, &dsdzt = W.dsdzt;
  F3 &drhodt = W.drhodt, &drhods = W.drhods, &drhodzt = W.drhodzt;
  W.Ftxe.zero(); W.Fsxe.zero(); W.Ftyn.zero(); W.Fsyn.zero(); W.Ftzt.zero(); W.Fszt.zero(); W.Ftimp.zero(); W.Fsimp.zero();
  std::vector<double>& dfzw = t.dfzW; std::vector<double>& dfzT = t.dfzT;
  auto eastface = [&](int i, int j, int k) {
    double dumt = 0.0, dums = 0.0, slp, tpr;
    for (int kr = 0; kr <= 1; kr++) for (int ip = 0; ip <= 1; ip++) {
      double drdh = (drhodt(i + ip, j, k) * dtdxe(i, j, k) + drhods(i + ip, j, k) * dsdxe(i, j, k));
      double drdz = (drhodt(i + ip, j, k) * dtdzt(i + ip, j, k - 1 + kr) + drhods(i + ip, j, k) * dsdzt(i + ip, j, k - 1 + kr));
      tprslp(t, drdh, drdz, sp2, slp, tpr);
      dumt = dumt + dfzw[k - 1 + kr] * (tpr * (piso)*dtdxe(i, j, k) + tpr * (piso - pgm) * slp * dtdzt(i + ip, j, k - 1 + kr));
      dums = dums + dfzw[k - 1 + kr] * (tpr * (piso)*dsdxe(i, j, k) + tpr * (piso - pgm) * slp * dsdzt(i + ip, j, k - 1 + kr));
    }
    W.Ftxe(i, j, k) = W.Ftxe(i, j, k) - dumt / (4 * dfzT[k]);
    W.Fsxe(i, j, k) = W.Fsxe(i, j, k) - dums / (4 * dfzT[k]);
  };
  for (int k = 1; k <= l; k++) for (int j = 1; j <= m; j++) {
    if (piso != 0.0 || pgm != 0.0) eastface(0, j, k);
    for (int i = 1; i <= n; i++) {
      if (piso != 0.0 || pgm != 0.0) {
        eastface(i, j, k);
        double dumt = 0.0, dums = 0.0, slp, tpr;
        for (int kr = 0; kr <= 1; kr++) for (int jq = 0; jq <= 1; jq++) {
          double drdh = (drhodt(i, j + jq, k) * dtdyn(i, j, k) + drhods(i, j + jq, k) * dsdyn(i, j, k));
          double drdz = (drhodt(i, j + jq, k) * dtdzt(i, j + jq, k - 1 + kr) + drhods(i, j + jq, k) * dsdzt(i, j + jq, k - 1 + kr));
          tprslp(t, drdh, drdz, sp2, slp, tpr);
          dumt = dumt + dfzw[k - 1 + kr] * std::cos(t.y[j + jq]) * (tpr * (piso)*dtdyn(i, j, k) + tpr * (piso - pgm) * slp * dtdzt(i, j + jq, k - 1 + kr));
          dums = dums + dfzw[k - 1 + kr] * std::cos(t.y[j + jq]) * (tpr * (piso)*dsdyn(i, j, k) + tpr * (piso - pgm) * slp * dsdzt(i, j + jq, k - 1 + kr));
        }
        W.Ftyn(i, j, k) = W.Ftyn(i, j, k) - dumt / (4 * dfzT[k] * std::cos(t.yv[j]));
        W.Fsyn(i, j, k) = W.Fsyn(i, j, k) - dums / (4 * dfzT[k] * std::cos(t.yv[j]));
        dumt = 0.0; dums = 0.0;
        for (int ip = 0; ip <= 1; ip++) for (int kr = 0; kr <= 1; kr++) {
          double drdh = (drhodt(i, j, k + kr) * dtdxe(i - 1 + ip, j, k + kr) + drhods(i, j, k + kr) * dsdxe(i - 1 + ip, j, k + kr));
          double drdz = (drhodt(i, j, k + kr) * dtdzt(i, j, k) + drhods(i, j, k + kr) * dsdzt(i, j, k));
          tprslp(t, drdh, drdz, sp2, slp, tpr);
          dumt = dumt + (tpr * (piso)*slp * slp * dtdzt(i, j, k) + tpr * (piso + pgm) * slp * dtdxe(i - 1 + ip, j, k + kr));
          dums = dums + (tpr * (piso)*slp * slp * dsdzt(i, j, k) + tpr * (piso + pgm) * slp * dsdxe(i - 1 + ip, j, k + kr));
        }
        W.Ftzt(i, j, k) = W.Ftzt(i, j, k) - dumt / 4;
        W.Fszt(i, j, k) = W.Fszt(i, j, k) - dums / 4;
        dumt = 0.0; dums = 0.0;
        for (int jq = 0; jq <= 1; jq++) for (int kr = 0; kr <= 1; kr++) {
          double drdh = (drhodt(i, j, k + kr) * dtdyn(i, j - 1 + jq, k + kr) + drhods(i, j, k + kr) * dsdyn(i, j - 1 + jq, k + kr));
          double drdz = (drhodt(i, j, k + kr) * dtdzt(i, j, k) + drhods(i, j, k + kr) * dsdzt(i, j, k));
          tprslp(t, drdh, drdz, sp2, slp, tpr);
          dumt = dumt + (tpr * (piso)*slp * slp * dtdzt(i, j, k) + tpr * (piso + pgm) * slp * dtdyn(i, j - 1 + jq, k + kr));
          dums = dums + (tpr * (piso)*slp * slp * dsdzt(i, j, k) + tpr * (piso + pgm) * slp * dsdyn(i, j - 1 + jq, k + kr));
        }
        W.Ftzt(i, j, k) = W.Ftzt(i, j, k) - dumt / 4;
        W.Fszt(i, j, k) = W.Fszt(i, j, k) - dums / 4;
      }
      if (eps != 0.0) {
        W.Ftzt(i, j, k) = W.Ftzt(i, j, k) + tprstb(t, drhodzt(i, j, k), sp1) * eps * dtdzt(i, j, k) / (drhodzt(i, j, k) - epsln);
        W.Fszt(i, j, k) = W.Fszt(i, j, k) + tprstb(t, drhodzt(i, j, k), sp1) * eps * dsdzt(i, j, k) / (drhodzt(i, j, k) - epsln);
      }
      if (kvc != 0.0) {
        W.Ftimp(i, j, k) = -tprstb(t, -drhodzt(i, j, k), sp1) * kvc * dtdzt(i, j, k);
        W.Fsimp(i, j, k) = -tprstb(t, -drhodzt(i, j, k), sp1) * kvc * dsdzt(i, j, k);
      }
    }
  }
  bool rhomix = t.cfg.rho_mixing && xes == 0.0;
  for (int k = 1; k <= l; k++) for (int j = 1; j <= m; j++) for (int i = 1; i <= n; i++) {
    double cy = std::cos(t.y[j]), cv = std::cos(t.yv[j]), cvm = std::cos(t.yv[j - 1]);
    int row = t.find_row2(i, j, k, TT) - 1;
    double mx = 0.0;
    if (t.vmix_temp == 1) {
      mx = (W.Ftxe(i, j, k) - W.Ftxe(i - 1, j, k)) / (t.dx * cy) + mx;
      mx = (W.Ftyn(i, j, k) * cv - W.Ftyn(i, j - 1, k) * cvm) / (t.dy * cy) + mx;
      mx = (W.Ftzt(i, j, k) - W.Ftzt(i, j, k - 1)) / (t.dz * dfzT[k]) + mx;
      if (rhomix) mx = ((W.Ftimp(i, j, k) - W.Ftimp(i, j, k - 1)) - (W.Fsimp(i, j, k) - W.Fsimp(i, j, k - 1)) * lambda) / (2.0 * t.dz * dfzT[k]) + mx;
      else mx = (W.Ftimp(i, j, k) - W.Ftimp(i, j, k - 1)) / (t.dz * dfzT[k]) + mx;
    }
    mix[row] = mx;
    row = t.find_row2(i, j, k, SS) - 1;
    mx = 0.0;
    if (t.vmix_salt == 1) {
      mx = (W.Fsxe(i, j, k) - W.Fsxe(i - 1, j, k)) / (t.dx * cy) + mx;
      mx = (W.Fsyn(i, j, k) * cv - W.Fsyn(i, j - 1, k) * cvm) / (t.dy * cy) + mx;
      mx = (W.Fszt(i, j, k) - W.Fszt(i, j, k - 1)) / (t.dz * dfzT[k]) + mx;
      if (rhomix) mx = ((W.Fsimp(i, j, k) - W.Fsimp(i, j, k - 1)) - (W.Ftimp(i, j, k) - W.Ftimp(i, j, k - 1)) / lambda) / (2.0 * t.dz * dfzT[k]) + mx;
      else mx = (W.Fsimp(i, j, k) - W.Fsimp(i, j, k - 1)) / (t.dz * dfzT[k]) + mx;
    }
    mix[row] = mx;
  }
}

// mix_imp.f:859-932 (vmix_el_1): row-oriented pattern, rows TT and SS of every OCEAN cell
struct MixPattern { std::vector<int> ipntr, col, ngrp; int maxgrp = 0; bool ready = false; };
void vmix_el_1(T& t, MixPattern& P) {
  int n = t.n, m = t.m, l = t.l;
  P.ipntr.assign(t.ndim + 2, 0);
  std::vector<std::vector<int>> rows(t.ndim + 1);
  int st[27][3]; int h = 0;
  for (int dk = -1; dk <= 1; dk++) for (int di = -1; di <= 1; di++) for (int dj = -1; dj <= 1; dj++) { st[h][0] = di; st[h][1] = dj; st[h][2] = dk; h++; }
  for (int k = 1; k <= l; k++) for (int j = 1; j <= m; j++) for (int i = 1; i <= n; i++)
    if (t.lm(i, j, k) == OCEAN) {
      std::vector<int> cols;
      for (h = 0; h < 27; h++) {
        int kind = t.lm(i + st[h][0], j + st[h][1], k + st[h][2]);
        if (kind == OCEAN) {
          cols.push_back(t.find_row2(i + st[h][0], j + st[h][1], k + st[h][2], TT));
          cols.push_back(t.find_row2(i + st[h][0], j + st[h][1], k + st[h][2], SS));
        } else if (kind == PERIO) {
          cols.push_back(t.find_row2(i - st[h][0] * (n - 1), j + st[h][1], k + st[h][2], TT));
          cols.push_back(t.find_row2(i - st[h][0] * (n - 1), j + st[h][1], k + st[h][2], SS));
        }
      }
      rows[t.find_row2(i, j, k, TT)] = cols;
      rows[t.find_row2(i, j, k, SS)] = cols;
    }
  P.col.clear();
  for (int r = 1; r <= t.ndim; r++) { P.ipntr[r] = (int)P.col.size() + 1; for (int c : rows[r]) P.col.push_back(c); }
  P.ipntr[t.ndim + 1] = (int)P.col.size() + 1;
  // Column partition.  The reference calls MINPACK DSM (mix_imp.f:187-192); any consistent partition gives
  // the same FDJS result (mix_sup.F90:1256), so a sequential greedy colouring is used.
  std::vector<std::vector<int>> colrows(t.ndim + 1);
  for (int r = 1; r <= t.ndim; r++) for (int c : rows[r]) colrows[c].push_back(r);
  P.ngrp.assign(t.ndim + 1, 0);
  P.maxgrp = 0;
  std::vector<int> mark(256, 0); int stamp = 0;
  for (int c = 1; c <= t.ndim; c++) {
    if (colrows[c].empty()) continue;
    stamp++;
    for (int r : colrows[c]) for (int c2 : rows[r]) { int g = P.ngrp[c2]; if (g > 0) { if ((int)mark.size() <= g) mark.resize(2 * g, 0); mark[g] = stamp; } }
    int g = 1; while ((int)mark.size() > g && mark[g] == stamp) g++;
    if ((int)mark.size() <= g) mark.resize(2 * g, 0);
    P.ngrp[c] = g; P.maxgrp = std::max(P.maxgrp, g);
  }
  P.ready = true;
}

// mix_imp.f:729-815 (vmix_diff = 1, forward differences) + FDJS row orientation mix_sup.F90:1380-1395
void vmix_jac(T& t, const double* un, MixPattern& P, MixWork& W) {
  int ndim = t.ndim;
  const double eps = 1.0e-08;
  std::vector<double> und(ndim), d(ndim), mix(ndim, 0.0), mixd(ndim, 0.0), fjac(P.col.size(), 0.0);
  vmix_fun(t, un, mix.data(), W);
  for (int numgrp = 1; numgrp <= P.maxgrp; numgrp++) {
    for (int j = 1; j <= ndim; j++) { d[j - 1] = 0.0; if (P.ngrp[j] == numgrp) d[j - 1] = eps; }
    for (int j = 0; j < ndim; j++) und[j] = un[j] + d[j];
    vmix_fun(t, und.data(), mixd.data(), W);
    for (int j = 0; j < ndim; j++) mixd[j] = mixd[j] - mix[j];
    for (int irow = 1; irow <= ndim; irow++)
      for (int ip = P.ipntr[irow]; ip <= P.ipntr[irow + 1] - 1; ip++) {
        int jcol = P.col[ip - 1];
        if (P.ngrp[jcol] == numgrp) { fjac[ip - 1] = mixd[irow - 1] / d[jcol - 1]; break; }
      }
  }
  int n = t.n;
  for (int i = 1; i <= ndim; i++) {
    int ix, iy, iz, ie; findex(t, i, ix, iy, iz, ie);
    for (int j = P.ipntr[i]; j <= P.ipntr[i + 1] - 1; j++) {
      int jx, jy, jz, je; findex(t, P.col[j - 1], jx, jy, jz, je);
      int s = 0;
      if (jz - iz == -1) s = s + 9;
      if (jz - iz == 1) s = s + 18;
      if (jx - ix == 0) s = s + 3;
      if (jx - ix == 1) s = s + 6;
      if (jx - ix == 1 - n) s = s + 6;
      if (jy - iy == -1) s = s + 1;
      if (jy - iy == 0) s = s + 2;
      if (jy - iy == 1) s = s + 3;
      t.AN(s, ie, je, ix, iy, iz) = t.AN(s, ie, je, ix, iy, iz) + fjac[j - 1];
    }
  }
}
} // namespace

struct thcm_oracle_priv { MixPattern P; MixWork W; };
static thcm_oracle_priv* priv(thcm_oracle* t);

namespace {
// usrc.F90:432-504
void matrix(T& t, const double* un) {
  t.An = t.Al;
  fillcolB(t);
  nlin_jac(t, un);
  if (t.vmix_flag >= 1) {
    thcm_oracle_priv* pv = priv(&t);
    if (!pv->P.ready) vmix_el_1(t, pv->P);
    bool vmix_dim = !pv->P.col.empty();
    if ((t.vmix_temp == 1 || t.vmix_salt == 1) && vmix_dim) vmix_jac(t, un, pv->P, pv->W);
  }
  boundaries(t);
  fillcolA(t);
}
// usrc.F90:506-586
void rhs(T& t, const double* un, double* B) {
  int ndim = t.ndim;
  std::vector<double> mix(ndim, 0.0), Au(ndim, 0.0);
  t.An = t.Al;
  nlin_rhs(t, un);
  boundaries(t);
  fillcolA(t);
  matAvec(t, un, Au.data());
  if (t.vmix_flag >= 1) {
    thcm_oracle_priv* pv = priv(&t);
    if (!pv->P.ready) vmix_el_1(t, pv->P);
    bool vmix_dim = !pv->P.col.empty();
    if ((t.vmix_temp == 1 || t.vmix_salt == 1) && vmix_dim) vmix_fun(t, un, mix.data(), pv->W);
  }
  // p0 = 0 (res.F90:22): the ures term vanishes; kept as  - 0*(1-par(RESC))*0
  for (int i = 0; i < ndim; i++) B[i] = -Au[i] - mix[i] + t.Frc[i] - 0.0 * (1 - t.par[RESC]) * 0.0;
  for (int i = 1; i <= t.n; i++) for (int j = 1; j <= t.m; j++) for (int k = 1; k <= t.l; k++) for (int k1 = 1; k1 <= NUN; k1++) {
    int row = t.find_row2(i, j, k, k1) - 1;
    B[row] = B[row] * (1 - t.lm(i, j, k));
  }
}

// THCM.C:2241-2539 CreateMaximalGraph + insert_graph_entry (0-based i,j,k); Epetra FillComplete sorts
// and de-duplicates the column indices of every row.
void create_graph(T& t) {
  int N = t.n, M = t.m, L = t.l;
  std::vector<std::vector<int>> rows(t.ndim);
  auto ins = [&](std::vector<int>& r, int i, int j, int k, int xx) {
    int ii = i;
    if (t.cfg.periodic) { ii = i % N; if (ii < 0) ii += N; } // MOD(double,double) of THCM.C:2528
    if (ii >= 0 && j >= 0 && k >= 0 && ii < N && j < M && k < L) r.push_back(NUN * (k * N * M + N * j + ii) + xx - 1);
  };
  for (int k = 0; k < L; k++) for (int j = 0; j < M; j++) for (int i = 0; i < N; i++) {
    int gid0 = NUN * (k * N * M + N * j + i) - 1 + 0; // gidU-1
    std::vector<int>* r;
    // U (THCM.C:2294-2323)
    r = &rows[gid0 + UU];
    for (int X : {UU}) { ins(*r, i, j, k, X); ins(*r, i - 1, j, k, X); ins(*r, i + 1, j, k, X); ins(*r, i, j - 1, k, X); ins(*r, i, j + 1, k, X); ins(*r, i, j, k - 1, X); ins(*r, i, j, k + 1, X); }
    ins(*r, i, j, k, VV); ins(*r, i - 1, j, k, VV); ins(*r, i + 1, j, k, VV); ins(*r, i, j - 1, k, VV); ins(*r, i, j + 1, k, VV);
    for (int kk = k; kk >= k - 1; kk--) { ins(*r, i, j, kk, WW); ins(*r, i + 1, j, kk, WW); ins(*r, i + 1, j + 1, kk, WW); ins(*r, i, j + 1, kk, WW); }
    ins(*r, i, j, k, PP); ins(*r, i + 1, j, k, PP); ins(*r, i, j + 1, k, PP); ins(*r, i + 1, j + 1, k, PP);
    // V (THCM.C:2342-2372)
    r = &rows[gid0 + VV];
    for (int X : {VV}) { ins(*r, i, j, k, X); ins(*r, i - 1, j, k, X); ins(*r, i + 1, j, k, X); ins(*r, i, j - 1, k, X); ins(*r, i, j + 1, k, X); ins(*r, i, j, k - 1, X); ins(*r, i, j, k + 1, X); }
    ins(*r, i, j, k, UU); ins(*r, i - 1, j, k, UU); ins(*r, i + 1, j, k, UU);
    for (int kk = k; kk >= k - 1; kk--) { ins(*r, i, j, kk, WW); ins(*r, i + 1, j, kk, WW); ins(*r, i + 1, j + 1, kk, WW); ins(*r, i, j + 1, kk, WW); }
    ins(*r, i, j, k, PP); ins(*r, i + 1, j, k, PP); ins(*r, i, j + 1, k, PP); ins(*r, i + 1, j + 1, k, PP);
    // W (THCM.C:2378-2386)
    r = &rows[gid0 + WW];
    ins(*r, i, j, k, WW); ins(*r, i, j, k, PP); ins(*r, i, j, k + 1, PP); ins(*r, i, j, k, TT); ins(*r, i, j, k + 1, TT); ins(*r, i, j, k, SS); ins(*r, i, j, k + 1, SS);
    // P (THCM.C:2392-2408)
    r = &rows[gid0 + PP];
    ins(*r, i, j, k, PP);
    for (int X : {UU, VV}) { ins(*r, i, j, k, X); ins(*r, i - 1, j, k, X); ins(*r, i, j - 1, k, X); ins(*r, i - 1, j - 1, k, X); }
    ins(*r, i, j, k, WW); ins(*r, i, j, k - 1, WW);
    // T, S (THCM.C:2414-2480)
    for (int X : {TT, SS}) {
      int Y = (X == TT) ? SS : TT;
      r = &rows[gid0 + X];
      ins(*r, i, j, k, X); ins(*r, i - 1, j, k, X); ins(*r, i + 1, j, k, X); ins(*r, i, j - 1, k, X); ins(*r, i, j + 1, k, X); ins(*r, i, j, k - 1, X); ins(*r, i, j, k + 1, X);
      for (int Z : {UU, VV}) { ins(*r, i, j, k, Z); ins(*r, i - 1, j, k, Z); ins(*r, i - 1, j - 1, k, Z); ins(*r, i, j - 1, k, Z); }
      ins(*r, i, j, k, WW); ins(*r, i, j, k - 1, WW);
      ins(*r, i, j, k, Y); ins(*r, i, j, k - 1, Y); ins(*r, i, j, k + 1, Y);
    }
  }
  if (t.rowintcon >= 0) { // THCM.C:2481-2513: the S row at rowintcon is replaced by the dense integral row
    std::vector<int>& r = rows[t.rowintcon];
    r.clear();
    for (int i = 0; i < N; i++) for (int j = 0; j < M; j++) for (int k = 0; k < L; k++) r.push_back(NUN * (k * N * M + N * j + i) + SS - 1);
  }
  t.g_rowptr.assign(t.ndim + 1, 0);
  t.g_col.clear();
  for (int r = 0; r < t.ndim; r++) {
    std::sort(rows[r].begin(), rows[r].end());
    rows[r].erase(std::unique(rows[r].begin(), rows[r].end()), rows[r].end());
    t.g_rowptr[r] = (int)t.g_col.size();
    t.g_col.insert(t.g_col.end(), rows[r].begin(), rows[r].end());
  }
  t.g_rowptr[t.ndim] = (int)t.g_col.size();
}

// THCM.C:2549-2577 getIntCondCoeff + thcm_utils.F90:285-323
void intcond_coeff(T& t) {
  t.intcondCoeff.assign(t.ndim, 0.0);
  for (int k = 1; k <= t.l; k++) for (int j = 1; j <= t.m; j++) for (int i = 1; i <= t.n; i++)
    if (t.lm(i, j, k) == OCEAN) t.intcondCoeff[t.find_row2(i, j, k, SS) - 1] = std::cos(t.y[j]) * t.dfzT[k];
}
} // namespace

// ------------------------------------------------------------------ C ABI
#include <map>
static std::map<thcm_oracle*, thcm_oracle_priv*> g_priv;
static thcm_oracle_priv* priv(thcm_oracle* t) { return g_priv[t]; }

extern "C" {

// usrc.F90:6-139 (init) preceded by the THCM.C:176-800 constructor logic that concerns this path
thcm_oracle* thcm_oracle_create(const thcm_oracle_config* cfg, const int* landm_in) {
  T* tp = new T();
  T& t = *tp;
  g_priv[tp] = new thcm_oracle_priv();
  t.cfg = *cfg;
  int n = t.n = cfg->n, m = t.m = cfg->m, l = t.l = cfg->l;
  t.ndim = n * m * l * NUN;
  t.periodic = cfg->periodic != 0;
  t.qdim = 0.01; t.nuq = 0.0; t.eta = 0.0; t.dqso = 0.0; t.eo0 = 0.0; t.albe0 = 0.0; t.albed = 0.0;
  t.zeta = 0.0; t.a0 = -0.0575; t.Lf = 3.347e+05; t.Qvar = 0.0; t.Q0 = 0.0;
  t.landm.assign((size_t)(n + 2) * (m + 2) * (l + 2), OCEAN);
  size_t pos = 0;
  for (int k = 0; k <= l + 1; k++) for (int j = 0; j <= m + 1; j++) for (int i = 0; i <= n + 1; i++) {
    int v = landm_in[pos++];
    if (!t.periodic && v == PERIO) v = OCEAN;
    t.lm(i, j, k) = v;
  }
  for (int k = 0; k <= l + 1; k++) for (int j = 0; j <= m + 1; j++) {
    if (!t.periodic) { t.lm(0, j, k) = LAND; t.lm(n + 1, j, k) = LAND; }
  }
  for (int k = 0; k <= l + 1; k++) for (int i = 0; i <= n + 1; i++) { t.lm(i, 0, k) = LAND; t.lm(i, m + 1, k) = LAND; }
  for (int j = 0; j <= m + 1; j++) for (int i = 0; i <= n + 1; i++) { t.lm(i, j, 0) = LAND; t.lm(i, j, l + 1) = LAND; }
  size_t nm = (size_t)n * m;
  t.Frc.assign(t.ndim, 0.0);
  for (auto* f : {&t.taux, &t.tauy, &t.tatm, &t.emip, &t.spert, &t.adapted_emip, &t.qatm, &t.albe, &t.patm, &t.msi, &t.gsi, &t.qsa}) f->assign(nm, 0.0);
  t.internal_temp.assign(nm * l, 0.0); t.internal_salt.assign(nm * l, 0.0);
  t.Al.assign((size_t)972 * n * m * l, 0.0);
  t.An.assign((size_t)972 * n * m * l, 0.0);
  t.coB.assign(t.ndim, 0.0);
  grid(t);
  double dzne = t.dz * t.dfzT[l];
  t.QTnd = r0dim / (udim * cp0 * rhodim * cfg->hdim * dzne);
  t.QSnd = s0 * r0dim / (deltas * udim * cfg->hdim * dzne);
  stpnt(t);
  vmix_init(t);
  atmos_coef(t);
  forcing(t);
  lin(t);
  // THCM.C:647-704: integral condition row
  t.rowintcon = -1; t.intCorrection = 0.0; t.dropped = 0;
  if (cfg->sres == 0) {
    int Nic = cfg->intcond_i == -1 ? n - 1 : cfg->intcond_i;
    int Mic = cfg->intcond_j == -1 ? m - 1 : cfg->intcond_j;
    if (landm_in[(size_t)(Nic + 1) + (size_t)(n + 2) * ((Mic + 1) + (size_t)(m + 2) * l)] != 0) {
      fprintf(stderr, "thcm_oracle: integral row coordinates (%d,%d) give a land point\n", Nic, Mic);
      delete tp; return nullptr;
    }
    t.rowintcon = NUN * ((l - 1) * n * m + n * Mic + Nic) + SS - 1;
  }
  if (cfg->fix_pressure_points) {
    t.rowPfix1 = NUN * ((l - 1) * n * m + n * (m - 1) + (n - 1)) + PP - 1;
    t.rowPfix2 = NUN * ((l - 1) * n * m + n * (m - 1) + (n - 2)) + PP - 1;
  } else { t.rowPfix1 = -1; t.rowPfix2 = -1; }
  intcond_coeff(t);
  create_graph(t);
  return tp;
}
void thcm_oracle_destroy(thcm_oracle* t) { if (!t) return; delete g_priv[t]; g_priv.erase(t); delete t; }

void thcm_oracle_set_par(thcm_oracle* t, int idx1, double value) { // usrc.F90:163-181
  if (idx1 >= 1 && idx1 <= NPAR) t->par[idx1] = value;
  forcing(*t);
  lin(*t);
}
// ---- coupling to an atmosphere / sea-ice model ----
// usrc.F90:237-293 (set_atmos_parameters): pars in the order of the bind(C) struct atmos_pars / Atmosphere::CommPars
void thcm_oracle_set_atmos_parameters(thcm_oracle* tp, const double* pars /* 18 */) {
  T& t = *tp;
  t.qdim = pars[1]; t.nuq = pars[2]; t.eta = pars[3]; t.dqso = pars[4]; t.eo0 = pars[7]; t.albe0 = pars[12]; t.albed = pars[13];
  t.nus = t.par[COMB] * t.par[SALT] * t.eta * t.qdim * t.QSnd;
  t.lvsc = t.par[COMB] * t.par[TEMP] * rhodim * lv * t.QTnd;
  forcing(t);
  lin(t);
}
// usrc.F90:296-333 (set_seaice_parameters): zeta, a0, Lf, s0, rhoo, qvar, q0
void thcm_oracle_set_seaice_parameters(thcm_oracle* tp, const double* pars /* 7 */) {
  T& t = *tp;
  t.zeta = pars[0]; t.a0 = pars[1]; t.Lf = pars[2]; t.Qvar = pars[5]; t.Q0 = pars[6];
  forcing(t);
  lin(t);
}
// inserts.F90:12-166; which: 0 tatm, 1 qatm, 2 albe, 3 patm, 4 qsa, 5 msi, 6 gsi.  Like the Fortran, an insert does not
// recompute forcing / lin: the next set_*_parameters or set_par does (Ocean::synchronize, Ocean.C:1443-1493).
void thcm_oracle_insert(thcm_oracle* tp, int which, const double* field /* n*m, i fastest */) {
  T& t = *tp;
  const bool cT = t.cfg.coupled_T == 1, cS = t.cfg.coupled_S == 1;
  std::vector<double>* dst = nullptr;
  switch (which) {
    case 0: if (cT) dst = &t.tatm; break;
    case 1: if (cT || cS) dst = &t.qatm; break;
    case 2: if (cT) dst = &t.albe; break;
    case 3: if (cS) dst = &t.patm; break;
    case 4: dst = &t.qsa; break;
    case 5: dst = &t.msi; break;
    case 6: if (cS) dst = &t.gsi; break;
  }
  if (dst) std::copy(field, field + (size_t)t.n * t.m, dst->begin());
}
// usrc.F90:201-219 (getdeps)
void thcm_oracle_getdeps(thcm_oracle* tp, double* out /* Ooa, Os, nus, eta, lvsc, qdim, pQSnd */) {
  T& t = *tp;
  out[0] = t.Ooa; out[1] = t.Os; out[2] = t.nus; out[3] = t.eta; out[4] = t.lvsc; out[5] = t.qdim;
  out[6] = t.par[COMB] * t.par[SALT] * t.QSnd;
}
void thcm_oracle_suno(thcm_oracle* tp, double* suno /* m, index j-1 */) { for (int j = 1; j <= tp->m; j++) suno[j - 1] = tp->suno[j]; }
// probe.F90:381-444 (get_derivatives), the coupled_T part: dftdm(pos) = QTos - QToa on OCEAN surface points, else 0
void thcm_oracle_dftdm(thcm_oracle* tp, const double* un, double* dftdm /* n*m */) {
  T& t = *tp;
  const int n = t.n, m = t.m, l = t.l;
  size_t pos = 0;
  t.nus = t.par[COMB] * t.par[SALT] * t.eta * t.qdim * t.QSnd; // probe.F90:414 assigns the module variable
  for (int j = 1; j <= m; j++)
    for (int i = 1; i <= n; i++, pos++) {
      dftdm[pos] = 0.0;
      if (t.lm(i, j, l) != OCEAN || t.cfg.coupled_T != 1) continue;
      // usol (usrc.F90:997-1104) leaves T, S of an OCEAN cell as they are in un
      const double To = un[t.find_row2(i, j, l, TT) - 1], So = un[t.find_row2(i, j, l, SS) - 1];
      const double Ta = t.s2(t.tatm, i, j), Ab = t.s2(t.albe, i, j), qa = t.s2(t.qatm, i, j);
      const double QTos = t.QTnd * t.zeta * (t.a0 * (So + s0) - (To + t0));
      const double QToa = t.par[COMB] * t.par[SUNP] * t.suno[j] * (1 - t.albe0 - t.albed * Ab) - t.Ooa * (To - Ta) -
                          t.lvsc * t.eta * t.qdim * (deltat / t.qdim * t.dqso * To - qa) - t.lvsc * t.eo0;
      dftdm[pos] = QTos - QToa;
    }
}

double thcm_oracle_get_par(thcm_oracle* t, int idx1) { return (idx1 >= 1 && idx1 <= NPAR) ? t->par[idx1] : 0.0; }
int thcm_oracle_ndim(thcm_oracle* t) { return t->ndim; }

void thcm_oracle_fortran_rhs(thcm_oracle* t, const double* un, double* B) { rhs(*t, un, B); }
void thcm_oracle_fortran_matrix(thcm_oracle* t, const double* un) { matrix(*t, un); }
int64_t thcm_oracle_fortran_nnz(thcm_oracle* t) { return (int64_t)t->coA.size(); }
void thcm_oracle_fortran_csr(thcm_oracle* t, int* begA, int* jcoA, double* coA, double* coB) {
  if (begA) memcpy(begA, t->begA.data(), sizeof(int) * t->begA.size());
  if (jcoA) memcpy(jcoA, t->jcoA.data(), sizeof(int) * t->jcoA.size());
  if (coA) memcpy(coA, t->coA.data(), sizeof(double) * t->coA.size());
  if (coB) memcpy(coB, t->coB.data(), sizeof(double) * t->coB.size());
}
void thcm_oracle_frc(thcm_oracle* t, double* frc) { memcpy(frc, t->Frc.data(), sizeof(double) * t->ndim); }

int64_t thcm_oracle_graph_nnz(thcm_oracle* t) { return (int64_t)t->g_col.size(); }
void thcm_oracle_graph(thcm_oracle* t, int* rowptr, int* col) {
  memcpy(rowptr, t->g_rowptr.data(), sizeof(int) * t->g_rowptr.size());
  memcpy(col, t->g_col.data(), sizeof(int) * t->g_col.size());
}
int thcm_oracle_rowintcon(thcm_oracle* t) { return t->rowintcon; }
void thcm_oracle_intcond_coeff(thcm_oracle* t, double* c) { memcpy(c, t->intcondCoeff.data(), sizeof(double) * t->ndim); }
void thcm_oracle_set_intcond_correction(thcm_oracle* t, double corr) { t->intCorrection = corr; }
int64_t thcm_oracle_dropped_entries(thcm_oracle* t) { return t->dropped; }

// THCM.C:949-1033 (rhs branch of THCM::evaluate), single process: maps are identities
void thcm_oracle_eval_rhs(thcm_oracle* t, const double* x, double* F) {
  rhs(*t, x, F);
  for (int i = 0; i < t->ndim; i++) F[i] = F[i] * -1.0; // tmp_rhs->Scale(-1.0), THCM.C:1003
  if (t->cfg.sres == 0 && t->rowintcon >= 0) {
    double intcond = 0.0;
    for (int i = 0; i < t->ndim; i++) intcond += t->intcondCoeff[i] * x[i];
    F[t->rowintcon] = t->cfg.int_sign * (intcond - t->intCorrection);
  }
  if (t->rowPfix1 >= 0) F[t->rowPfix1] = 0.0;
  if (t->rowPfix2 >= 0) F[t->rowPfix2] = 0.0;
}
// THCM.C:1035-1173 (Jacobian branch)
void thcm_oracle_eval_jac(thcm_oracle* t, const double* x, double* val, double* diagB) {
  matrix(*t, x);
  int64_t nnz = (int64_t)t->g_col.size();
  for (int64_t i = 0; i < nnz; i++) val[i] = 0.0;
  for (int i = 0; i < t->ndim; i++) diagB[i] = 0.0;
  t->dropped = 0;
  for (int i = 0; i < t->ndim; i++) {
    if (i == t->rowintcon) continue;
    const int* cb = &t->g_col[t->g_rowptr[i]];
    const int* ce = &t->g_col[t->g_rowptr[i + 1]];
    for (int v = t->begA[i]; v <= t->begA[i + 1] - 1; v++) {
      int c = t->jcoA[v - 1] - 1;
      const int* p = std::lower_bound(cb, ce, c);
      if (p != ce && *p == c) val[p - &t->g_col[0]] = t->coA[v - 1];
      else t->dropped++; // Epetra ReplaceGlobalValues ierr=2 "value excluded", THCM.C:1087-1092
    }
    diagB[i] = t->coB[i] * 1.0; // "Mass" parameter defaults to 1, THCM.C:1150-1153
  }
  if (t->cfg.sres == 0 && t->rowintcon >= 0) { // intcond_S, THCM.C:2121-2198
    int r = t->rowintcon;
    diagB[r] = 0.0;
    for (int p = t->g_rowptr[r]; p < t->g_rowptr[r + 1]; p++) val[p] = t->cfg.int_sign * t->intcondCoeff[t->g_col[p]];
  }
  for (int q = 0; q < 2; q++) { // fixPressurePoints, THCM.C:2201-2238
    int r = q == 0 ? t->rowPfix1 : t->rowPfix2;
    if (r < 0) continue;
    diagB[r] = 0.0;
    for (int p = t->g_rowptr[r]; p < t->g_rowptr[r + 1]; p++) val[p] = (t->g_col[p] == r) ? 1.0 : 0.0;
  }
}

// ---- m_scaling (src/ocean/scaling.F90) and THCM::RecomputeScaling (src/ocean/THCM.C:1694-1750), single process ----
// scaling.F90:29-62: average diagonal block over OCEAN cells, read from the Fortran CSR of the last matrix() call.
// db is the Fortran array db(nun,nun), column-major: db[(ii-1) + 6*(jj-1)].
void thcm_oracle_average_block(thcm_oracle* t, double* db) {
  for (int q = 0; q < NUN * NUN; q++) db[q] = 0.0;
  int nl = 0;
  for (int i = 1; i <= t->ndim; i++) {
    int ix, iy, iz, ii; findex(*t, i, ix, iy, iz, ii);
    if (t->lm(ix, iy, iz) == OCEAN) {
      if (ii == PP) nl = nl + 1;
      for (int j = t->begA[i - 1]; j <= t->begA[i] - 1; j++) {
        int jx, jy, jz, jj; findex(*t, t->jcoA[j - 1], jx, jy, jz, jj);
        if (ix == jx && iy == jy && iz == jz) db[(ii - 1) + NUN * (jj - 1)] = db[(ii - 1) + NUN * (jj - 1)] + t->coA[j - 1];
      }
    }
  }
  if (nl > 0) for (int q = 0; q < NUN * NUN; q++) db[q] = db[q] / nl;
}
// scaling.F90:198-284 (the LAPACK variant that is compiled, "#if 0 ... #else").  dgetrf/dgetri are restated as LU with
// partial pivoting (dgetf2's rule: first row of maximal modulus) and inv(A) from inv(U) and the unit lower factor, the
// way dgetri forms it; dgecon's estimate only feeds the test 1.0+rcond==1.0 and is replaced by the exact reciprocal
// condition number in the infinity norm.  tests/test_scaling_cpu.py pins this against LAPACK itself (scipy).
// Returns 0 if the block is singular to working precision (the Fortran returns with dr, dc unset).
int thcm_oracle_scal(const double* mat_in, double* dr, double* dc) {
  const int N = NUN;
  double a[NUN][NUN]; // a[i][j] = mat(i+1,j+1)
  for (int i = 0; i < N; i++) for (int j = 0; j < N; j++) a[i][j] = mat_in[i + N * j];
  double Anorm = 0.0;
  for (int i = 0; i < N; i++) { double s = 0.0; for (int j = 0; j < N; j++) s += std::fabs(a[i][j]); Anorm = std::max(Anorm, s); }
  int ipvt[NUN];
  for (int k = 0; k < N; k++) { // dgetf2
    int p = k;
    for (int i = k + 1; i < N; i++) if (std::fabs(a[i][k]) > std::fabs(a[p][k])) p = i;
    ipvt[k] = p;
    if (a[p][k] == 0.0) return 0;
    if (p != k) for (int j = 0; j < N; j++) std::swap(a[k][j], a[p][j]);
    const double r = 1.0 / a[k][k];
    for (int i = k + 1; i < N; i++) a[i][k] *= r;
    for (int j = k + 1; j < N; j++) for (int i = k + 1; i < N; i++) a[i][j] -= a[i][k] * a[k][j];
  }
  // dgetri: inv(U) in place (dtrtri, upper, non-unit) ...
  for (int j = 0; j < N; j++) {
    a[j][j] = 1.0 / a[j][j];
    const double ajj = -a[j][j];
    for (int i = 0; i < j; i++) { // column j above the diagonal := inv(U)(0:j,0:j) * U(0:j,j), then scaled by ajj
      double s = 0.0;
      for (int k2 = i; k2 < j; k2++) s += a[i][k2] * a[k2][j];
      a[i][j] = s;
    }
    for (int i = 0; i < j; i++) a[i][j] *= ajj;
  }
  // ... then solve inv(A)*L = inv(U) column by column from the right, and undo the row interchanges as column swaps
  for (int j = N - 2; j >= 0; j--) {
    double work[NUN];
    for (int i = j + 1; i < N; i++) { work[i] = a[i][j]; a[i][j] = 0.0; }
    for (int k2 = j + 1; k2 < N; k2++) for (int i = 0; i < N; i++) a[i][j] -= a[i][k2] * work[k2];
  }
  for (int j = N - 2; j >= 0; j--) {
    const int jp = ipvt[j];
    if (jp != j) for (int i = 0; i < N; i++) std::swap(a[i][j], a[i][jp]);
  }
  double Inorm = 0.0;
  for (int i = 0; i < N; i++) { double s = 0.0; for (int j = 0; j < N; j++) s += std::fabs(a[i][j]); Inorm = std::max(Inorm, s); }
  const double rcond = (Anorm > 0.0 && Inorm > 0.0) ? (1.0 / Anorm) / Inorm : 0.0;
  if (1.0 + rcond == 1.0) return 0;
#define MAT(i, j) a[(i) - 1][(j) - 1]
  double idc, idr;
  dr[0] = 1.0; dc[0] = 1.0;
  idc = std::sqrt(MAT(1, 1) / MAT(2, 2));
  dr[1] = 1 / idc; dc[1] = dr[1];
  idr = std::sqrt(std::fabs(MAT(1, 1) / MAT(4, 4)));
  dr[3] = 1 / idr; dc[3] = dr[3];
  if (std::fabs(MAT(4, 3)) > std::fabs(MAT(3, 3))) idr = MAT(1, 1) / (idr * MAT(4, 3));
  else idr = std::sqrt(std::fabs(MAT(1, 1) / MAT(3, 3)));
  dr[2] = 2 / idr; dc[2] = dr[2];
  if (std::fabs(MAT(4, 5) * MAT(5, 4)) < .01 * std::fabs(MAT(4, 4) * MAT(5, 5))) { MAT(4, 5) = 1; MAT(5, 4) = 1; }
  idc = std::sqrt(std::fabs(MAT(1, 1) * MAT(4, 5) / (MAT(5, 4) * MAT(5, 5))));
  idr = MAT(1, 1) / (idc * MAT(5, 5));
  dr[4] = 1 / idr; dc[4] = 1 / idc;
  if (std::fabs(MAT(4, 6) * MAT(6, 4)) < .01 * std::fabs(MAT(4, 4) * MAT(6, 6))) { MAT(4, 6) = 1; MAT(6, 4) = 1; }
  idc = std::sqrt(std::fabs(MAT(1, 1) * MAT(4, 6) / (MAT(6, 4) * MAT(6, 6))));
  idr = MAT(1, 1) / (idc * MAT(6, 6));
  dr[5] = 1 / idr; dc[5] = 1 / idc;
#undef MAT
  return 1;
}
// THCM.C:1694-1750 on one process (SumAll / NumProc are identities): average_block, scal + compute (scaling.F90:69-99),
// reciprocals, then T and S rows share the mean of their two factors.  Returns 0 if the block is singular.
int thcm_oracle_recompute_scaling(thcm_oracle* t, double* rowScaling, double* colScaling) {
  double gdb[NUN * NUN], rs[NUN], cs[NUN];
  thcm_oracle_average_block(t, gdb);
  if (!thcm_oracle_scal(gdb, rs, cs)) return 0;
  for (int i = 1; i <= t->ndim; i++) {
    int ix, iy, iz, ii; findex(*t, i, ix, iy, iz, ii);
    double r = 1.0, c = 1.0;
    if (t->lm(ix, iy, iz) == OCEAN) { r = rs[ii - 1]; c = cs[ii - 1]; }
    rowScaling[i - 1] = 1.0 / r;
    colScaling[i - 1] = 1.0 / c;
  }
  for (int i = TT - 1; i < t->ndim; i += NUN) {
    double mean = 0.5 * (rowScaling[i] + rowScaling[i + 1]);
    rowScaling[i] = mean; rowScaling[i + 1] = mean;
    mean = 0.5 * (colScaling[i] + colScaling[i + 1]);
    colScaling[i] = mean; colScaling[i + 1] = mean;
  }
  return 1;
}

void thcm_oracle_spmv(int nrows, const int* rowptr, const int* col, const double* val, const double* x, double* y) {
  for (int i = 0; i < nrows; i++) {
    double s = 0.0;
    for (int p = rowptr[i]; p < rowptr[i + 1]; p++) s = val[p] * x[col[p]] + s;
    y[i] = s;
  }
}

} // extern "C"
