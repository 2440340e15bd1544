// thcm_rows.cuh -- fused, matrix-free evaluation of ONE row of the THCM operator.
//
// This is the arithmetic core of the B200 path: instead of materialising the dense stencil tensor
// An(27,6,6,n,m,l) like the reference (src/ocean/usrc.F90:432-586), every (cell, unknown) row computes
// only its <=24 structurally possible entries (the CreateMaximalGraph pattern, src/ocean/THCM.C:2241-2521)
// in registers, applies the land-mask logic of src/ocean/boundary.F90:81-387 to them, and either stores
// them into the static CSR (Jacobian) or contracts them with the state (residual).
//
// The code is __host__ __device__ so that tests/ can compile it with g++ and compare it against the
// structure-faithful oracle without a GPU; the product only ever instantiates it inside CUDA kernels.
#pragma once
#include <stdint.h>
#include <math.h>

#ifdef __CUDACC__
#define THCM_HD __host__ __device__ __forceinline__
#define THCM_HD_NI static __host__ __device__ __noinline__   // kept out of line on the device: instruction-cache footprint
#else
#define THCM_HD inline __attribute__((always_inline))
#define THCM_HD_NI static inline
#endif

namespace thcm {

enum { UU = 1, VV = 2, WW = 3, PP = 4, TT = 5, SS = 6 };

// ---- per-cell static signature word (built once per land mask, thcm_host.cpp: build_signatures) ----
// bits 0..26 : neighbour in stencil slot s=bit+1 is LAND (slot numbering of src/ocean/par.F90:24-31)
enum : int {
  SIG_SOUTHEE = 27,  // i<n  and landm(i+2,j-1,k)==LAND   (boundary.F90:68-75)
  SIG_EASTEAST = 28, // i<n  and landm(i+2,j  ,k)==LAND
  SIG_NORTHEE = 29,  // i<n  and landm(i+2,j+1,k)==LAND
  SIG_NNORTHEE = 30, // i<n, j<m and landm(i+2,j+2,k)==LAND
  SIG_NN = 31,       // j<m  and landm(i,j+2,k)==LAND  (nnwest, nnorth, nneast all read (i,j+2,k), boundary.F90:77-81)
  SIG_NOTOCEAN = 33, // landm(i,j,k) /= OCEAN
  SIG_LIVE = 35,     // u,v at this cell's NE corner survive usol (usrc.F90:1043-1102)
  SIG_LIVEWRAP = 36, // same, for the periodic copy u(0,j,k)=u(N,j,k) (usrc.F90:1031-1032): only cell i=N
  SIG_FACE_UP = 37,  // isoc(i,j,k)*isoc(i,j,k+1) (mix_imp.f:617-627)
  SIG_FACE_DN = 38,  // isoc(i,j,k)*isoc(i,j,k-1)
  SIG_BU = 39,       // landm(i+1,j,k) /= LAND (assemble.F90:38)
  SIG_BV = 40,       // landm(i,j+1,k) /= LAND (assemble.F90:39)
  SIG_LANDVAL = 41   // landm(i,j,k)==1: the (1-landm) factor of usrc.F90:563-574
};
THCM_HD bool sbit(uint64_t sig, int b) { return (sig >> b) & 1ull; }
THCM_HD bool sland(uint64_t sig, int slot) { return (sig >> (slot - 1)) & 1ull; }

// ---- parameters and coefficient tables (filled on the host with the reference's expression order) ----
struct Scalars {
  int n, m, l, periodic;
  double epsr;                 // par(ROSB)
  double w_tt, w_ss;           // Al(5|23,WW,TT) , Al(5|23,WW,SS)            (usrc.F90:676-678)
  double ra_a2, ra_a3;         // Ra*xes*alpt2, Ra*xes*alpt3                   (usrc.F90:841-842)
  double tt_surf, ss_surf;     // TRES*bi, SRES*bi                              (usrc.F90:727,766)
  // coupled to an atmosphere / sea ice (coupled_T = 1, usrc.F90:725-741): tt_surf = Ooa*tc and, added one after the other,
  // dedt*sc and mc*(QTnd*zeta*tc - Ooa*tc - dedt*sc) with the sea-ice mask mc = msi(i,j); Al(5,TT,SS) = -QTnd*zeta*a0*mc
  int coupled_T;
  double tt_dedt, tt_mc, ts_mc;
  double dyi;                  // 1./(2*dy)                                     (spf.F90:337)
  double dzi;                  // 1./dz
  double tdzi2;                // 1.0/(2*dz)                                    (spf.F90:446)
  double kvc, lambda, xes;     // par(P_VC), par(LAMB), par(NLES)
  double fac;                  // alphaT*par(SPL1)                              (mix_imp.f:850)
  int vmix_on;                 // vmix_flag>=1 and kvc path supported
  int rho_mixing;              // rho_mixing .and. xes==0                       (mix_imp.f:516)
  double dz;
  // storage of this rank's sub-domain (TRIOS::Domain block, TRIOS_Domain.C:81-195): owned cells
  // io0..io0+nloc-1 x jo0..jo0+mloc-1 (global Fortran indices) x all levels, followed by one ring of
  // halo cells per level.  Single GPU: io0=jo0=1, nloc=n, mloc=m and the ring is never addressed.
  int io0, jo0, nloc, mloc;
  long long nown;              // nloc*mloc*l
};
// Tables indexed by the Fortran index (j in 0..m+1, k in 0..l); all double, one contiguous block.
struct Tables {
  // per latitude j
  const double *cdxu, *cdyu, *cyv, *tanv, *cdxt, *cdyt;
  const double *uu_we, *uu_s, *uu_n, *uu_ch, *uv_c, *uv_w, *uv_e;   // Al(.,UU,UU), Al(.,UU,VV)
  const double *vv_we, *vv_s, *vv_n, *vv_ch, *vu_c, *vu_w, *vu_e;   // Al(.,VV,VV), Al(.,VV,UU)
  const double *pu, *pv_s, *pv_n;                                   // Al(.,PP,UU), Al(.,PP,VV)
  const double *tt_we, *tt_s, *tt_n, *tt_ch;                        // Al(.,TT,TT) = Al(.,SS,SS) horizontal part
  // per level k
  const double *uu_b, *uu_t, *uu_cv;     // vertical friction  -EV*uzz(14|23), EV*uzz(5)
  const double *tt_b, *tt_t, *tt_cv;     // vertical diffusion -pv*tzz(14|23), pv*tzz(5)
  const double *wp;                      // dzi/dfzW(k)
  const double *pw;                      // dzi/dfzT(k)
  const double *tdzi8;                   // 1.0/(8*dfzT(k)*dz)
  const double *dfzT;                    // dfzT(k)
  const double *dzdfzT;                  // dz*dfzT(k)
  const double *dzdfzW;                  // dz*dfzW(k), k=0..l
  const double *r_dfzT, *r_dzdfzT, *r_dzdfzW; // correctly rounded reciprocals of the three divisors above (see thcm_div)
  // sea-ice mask msi(i,j) of the whole surface, index (j-1)*n + (i-1) with global Fortran i, j; null while no sea ice is attached
  const double* msi;
};
static const int NJTAB = 29, NKTAB = 15; // number of per-j / per-k tables above

// ---------------------------------------------------------------------------------------------
// Row patterns: entries sorted by (dk, dj, di, var) == ascending column index away from the periodic
// seam.  slot = 1 + (dj+1) + 3*(di+1) + {0: k, 9: k-1, 18: k+1}   (assemble.F90:142-169)
struct Ent { signed char slot, var; };

template <int A> struct RowDef;
template <> struct RowDef<UU> {
  static constexpr int NE = 24;
  static constexpr Ent e[NE] = {{14, UU}, {14, WW}, {17, WW}, {15, WW}, {18, WW},
                                {4, UU}, {4, VV},
                                {2, UU}, {2, VV}, {5, UU}, {5, VV}, {5, WW}, {5, PP}, {8, UU}, {8, VV}, {8, WW}, {8, PP},
                                {6, UU}, {6, VV}, {6, WW}, {6, PP}, {9, WW}, {9, PP},
                                {23, UU}};
};
template <> struct RowDef<VV> {
  static constexpr int NE = 22;
  static constexpr Ent e[NE] = {{14, VV}, {14, WW}, {17, WW}, {15, WW}, {18, WW},
                                {4, VV},
                                {2, UU}, {2, VV}, {5, UU}, {5, VV}, {5, WW}, {5, PP}, {8, UU}, {8, VV}, {8, WW}, {8, PP},
                                {6, VV}, {6, WW}, {6, PP}, {9, WW}, {9, PP},
                                {23, VV}};
};
template <> struct RowDef<WW> {
  static constexpr int NE = 7;
  static constexpr Ent e[NE] = {{5, WW}, {5, PP}, {5, TT}, {5, SS}, {23, PP}, {23, TT}, {23, SS}};
};
template <> struct RowDef<PP> {
  static constexpr int NE = 11;
  static constexpr Ent e[NE] = {{14, WW}, {1, UU}, {1, VV}, {4, UU}, {4, VV}, {2, UU}, {2, VV}, {5, UU}, {5, VV}, {5, WW}, {5, PP}};
};
template <> struct RowDef<TT> {
  static constexpr int NE = 20;
  static constexpr Ent e[NE] = {{14, WW}, {14, TT}, {14, SS},
                                {1, UU}, {1, VV}, {4, UU}, {4, VV}, {4, TT},
                                {2, UU}, {2, VV}, {2, TT}, {5, UU}, {5, VV}, {5, WW}, {5, TT}, {5, SS}, {8, TT},
                                {6, TT},
                                {23, TT}, {23, SS}};
};
template <> struct RowDef<SS> {
  static constexpr int NE = 20;
  static constexpr Ent e[NE] = {{14, WW}, {14, TT}, {14, SS},
                                {1, UU}, {1, VV}, {4, UU}, {4, VV}, {4, SS},
                                {2, UU}, {2, VV}, {2, SS}, {5, UU}, {5, VV}, {5, WW}, {5, TT}, {5, SS}, {8, SS},
                                {6, SS},
                                {23, TT}, {23, SS}};
};

template <int A> constexpr int find_ent(int slot, int var) {
  for (int q = 0; q < RowDef<A>::NE; q++)
    if (RowDef<A>::e[q].slot == slot && RowDef<A>::e[q].var == var) return q;
  return -1;
}
constexpr int slot_di(int s) { return ((s - 1) % 9) / 3 - 1; }
constexpr int slot_dj(int s) { return ((s - 1) % 3) - 1; }
constexpr int slot_dk(int s) { return s <= 9 ? 0 : (s <= 18 ? -1 : 1); }

template <int A> struct Row {
  static constexpr int NE = RowDef<A>::NE;
  double a[NE];
  THCM_HD void clear() {
#pragma unroll
    for (int q = 0; q < NE; q++) a[q] = 0.0;
  }
  template <int S, int B> static constexpr bool has() { return find_ent<A>(S, B) >= 0; }
  template <int S, int B> THCM_HD double& at() {
    static_assert(find_ent<A>(S, B) >= 0, "entry not in the static pattern");
    return a[find_ent<A>(S, B)];
  }
  // An(D,A,B) = An(D,A,B) + An(S,A,B)
  template <int D, int S, int B> THCM_HD void fold() {
    if constexpr (has<S, B>()) {
      static_assert(has<D, B>(), "fold target outside the static pattern");
      at<D, B>() = at<D, B>() + at<S, B>();
    }
  }
  template <int S, int B> THCM_HD void zc() {
    if constexpr (has<S, B>()) at<S, B>() = 0.0;
  }
  template <int S> THCM_HD void zs() { zc<S, 1>(); zc<S, 2>(); zc<S, 3>(); zc<S, 4>(); zc<S, 5>(); zc<S, 6>(); }
  template <int S> THCM_HD void zuv() { zc<S, UU>(); zc<S, VV>(); }
};


// ---------------------------------------------------------------------------------------------
// Cell coordinates use the Fortran convention (1-based) so that every formula reads like the reference.
struct Cell { int i, j, k; uint64_t sig; };

// Field accessor concept F (provided by the kernel from its shared-memory tile, or by the host test):
//   double F::u(di,dj,dk), v, w, t, s  -> the ghosted/land-zeroed fields that usol produces
//                                          (usrc.F90:997-1104) at cell (i+di, j+dj, k+dk).

// x / d with the divisor's reciprocal rd = RN(1/d) tabulated on the host: q = RN(x*rd), r = x - q*d (exact, one FMA),
// q' = RN(q + r*rd).  With a correctly rounded reciprocal this Markstein sequence returns the correctly rounded
// quotient -- the same bits as the Fortran's division -- in 3 instructions instead of the ~25 of a generic fp64
// divide (no overflow/underflow here: the operands are O(1e-6..1e6)).  Exact zeros stay exact zeros.
THCM_HD double thcm_div(double x, double d, double rd) {
  const double q = x * rd;
  const double r = fma(-q, d, x);
  return fma(r, rd, q);
}

// ---- vertical (convective-adjustment) mixing, the only non-zero branch with the stpnt defaults ----
// mix_imp.f:231-562 with piso=pgm=eps=0: mix = (Fimp(k)-Fimp(k-1))/(dz*dfzT(k)),
// Fimp(k) = -tprstb(-drhodzt(k),sp1)*kvc*dCdzt(k)   (mix_imp.f:486-492, :837-856)
THCM_HD double thcm_rho(const Scalars& S, double t, double s) {
  return S.lambda * s - t - S.xes * (2.93 * t + 8.3e-02 * t * t - 6.6e-04 * t * t * t); // alpt1..3, usr.F90:143-145
}
#ifdef __CUDACC__
// tanh(c) for 0 < c < 22 on one straight path: 2c = t ln2 + r with t = rint(2c log2 e),
// exp(r) = 1 + r + r^2 p(r) (degree-9 fit on |r| <= 0.354, 2.4e-17 relative), 2^t built in the exponent field,
// tanh = (E-1)/(E+1) with E-1 assembled without cancellation, reciprocal from MUFU.RCP64H plus one cubic Newton step
// and a final residual correction: relative error of a few 1e-16 over the whole range (tools/tanh_check.cu).  The
// convective-adjustment terms carry tolerances, not bit parity, already with libm's tanh (the forward-difference
// Jacobian entries amplify any tanh rounding by 1/eps = 1e8), but they do need RELATIVE accuracy.  Constants sit in the constant
// bank so that they are instruction operands; about 30 instructions instead of the 60-100 of the generic tanh.
static __constant__ double THCM_TANH_C[12] = {
    0.5000000000000001, 0.16666666666666669, 0.041666666666616864, 0.008333333333329503, 0.0013888888920770123,
    0.00019841269865788537, 2.4801515879478315e-05, 2.7557264294679126e-06, 2.762261407861969e-07, 2.5102327723609146e-08,
    -0.6931471805598903 /* -ln2, high part (low 11 bits clear) */, -5.497923018708371e-14 /* -ln2, low part */};
__device__ __forceinline__ double thcm_tanh_pos(double c) {
  const double* C = THCM_TANH_C;
  const double x2 = c + c;
  const float tf = rintf(__double2float_rn(x2) * 1.4426950408889634f);
  const double td = (double)tf;
  double r = fma(td, C[10], x2);
  r = fma(td, C[11], r);
  double p = C[9];
  p = fma(p, r, C[8]); p = fma(p, r, C[7]); p = fma(p, r, C[6]); p = fma(p, r, C[5]); p = fma(p, r, C[4]);
  p = fma(p, r, C[3]); p = fma(p, r, C[2]); p = fma(p, r, C[1]); p = fma(p, r, C[0]);
  const double em1 = fma(r * r, p, r);                                   // exp(r) - 1, full relative accuracy
  const double sc = __hiloint2double((__float2int_rn(tf) + 1023) << 20, 0); // 2^t, 0 <= t <= 64
  const double em = fma(sc, em1, sc - 1.0);                              // exp(2c) - 1 (sc - 1 is exact)
  const double d = em + 2.0;                                             // exp(2c) + 1
  double y;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(d));
  double e = fma(-d, y, 1.0);
  e = fma(e, e, e);
  y = fma(y, e, y);
  const double q = em * y;                                               // tanh = (E - 1)/(E + 1): relative accuracy for small c too
  return fma(fma(-q, d, em), y, q);
}
#endif
THCM_HD double thcm_tprstb(const Scalars& S, double grad) {
  double q = -grad * S.fac;
  double c = q * q * q;
  if (!(c > 0.0)) return 0.0;   // max(tanh(c),0) == 0 for c <= 0
  if (c >= 22.0) return 1.0;    // tanh rounds to exactly 1.0 in double precision beyond ~19.1
#ifdef __CUDA_ARCH__
  return thcm_tanh_pos(c);
#else
  return tanh(c);
#endif
}
// Implicit flux through one cell face between cell a (below) and cell b (above); flag = isoc(a)*isoc(b); ra, rb are
// thcm_rho of the two cells (passed in so that callers can reuse them).  Shortcuts that do not change any value:
// a closed face (flag = 0) and a stably stratified one (tprstb = 0) carry a zero flux whatever dT/dz, dS/dz are, so
// their two extra divisions are skipped (the sign of a zero flux cannot reach the result: thcm_mix_combine adds +0.0).
struct Face { double ft, fs; };
struct Div { double d, rd; }; // a divisor and its reciprocal
THCM_HD Face thcm_face(const Scalars& S, double flag, Div dzw, double ta, double sa, double ra, double tb, double sb, double rb) {
  Face f; f.ft = 0.0; f.fs = 0.0;
  if (flag == 0.0) return f;
  double dr = thcm_div(flag * (rb - ra), dzw.d, dzw.rd);
  double tp = thcm_tprstb(S, -dr);
  if (tp == 0.0) return f;
  double dt = thcm_div(flag * (tb - ta), dzw.d, dzw.rd), ds = thcm_div(flag * (sb - sa), dzw.d, dzw.rd);
  f.ft = -tp * S.kvc * dt; f.fs = -tp * S.kvc * ds;
  return f;
}
THCM_HD void thcm_mix_combine(const Scalars& S, Div dzt, const Face& u, const Face& d, double& mt, double& ms) {
  if (S.rho_mixing) {
    mt = ((u.ft - d.ft) - (u.fs - d.fs) * S.lambda) / (2.0 * dzt.d) + (0.0 + 0.0 + 0.0);
    ms = ((u.fs - d.fs) - (u.ft - d.ft) / S.lambda) / (2.0 * dzt.d) + (0.0 + 0.0 + 0.0);
  } else {
    mt = thcm_div(u.ft - d.ft, dzt.d, dzt.rd) + 0.0;
    ms = thcm_div(u.fs - d.fs, dzt.d, dzt.rd) + 0.0;
  }
}
// t0,s0: cell k-1; t1,s1: cell k; t2,s2: cell k+1.  Returns mix for the T row in mt and S row in ms.
THCM_HD void thcm_mix(const Scalars& S, const Tables& T, const Cell& c, double t0, double s0, double t1, double s1,
                      double t2, double s2, double& mt, double& ms) {
  const double up = sbit(c.sig, SIG_FACE_UP) ? 1.0 : 0.0, dn = sbit(c.sig, SIG_FACE_DN) ? 1.0 : 0.0;
  Face fu, fd; fu.ft = 0.0; fu.fs = 0.0; fd.ft = 0.0; fd.fs = 0.0;
  if (S.kvc != 0.0) {
    const double r1 = thcm_rho(S, t1, s1);
    fu = thcm_face(S, up, Div{T.dzdfzW[c.k], T.r_dzdfzW[c.k]}, t1, s1, r1, t2, s2, thcm_rho(S, t2, s2));
    if (c.k > 1) fd = thcm_face(S, dn, Div{T.dzdfzW[c.k - 1], T.r_dzdfzW[c.k - 1]}, t0, s0, thcm_rho(S, t0, s0), t1, s1, r1);
  }
  thcm_mix_combine(S, Div{T.dzdfzT[c.k], T.r_dzdfzT[c.k]}, fu, fd, mt, ms);
}
// vmix_jac (mix_imp.f:729-815): forward differences of mix with eps = 1e-8 with respect to t0,s0,t1,s1,t2,s2 (p = 0..5).
// A perturbation of the cell below only changes the lower face and one of the cell above only the upper face, so the
// unchanged face (and the unchanged densities) of the base evaluation are reused: 10 face evaluations and 9 densities
// instead of 14 full evaluations, same arithmetic.
struct MixFD { double bt, bs, mt[6], ms[6]; };
THCM_HD void thcm_mix_fd(const Scalars& S, const Tables& T, const Cell& c, double t0, double s0, double t1, double s1,
                         double t2, double s2, MixFD& o) {
  const double eps = 1.0e-08;
  const double up = sbit(c.sig, SIG_FACE_UP) ? 1.0 : 0.0, dn = sbit(c.sig, SIG_FACE_DN) ? 1.0 : 0.0;
  const int kd = c.k > 1 ? c.k - 1 : 0;
  const Div dzu{T.dzdfzW[c.k], T.r_dzdfzW[c.k]}, dzd{T.dzdfzW[kd], T.r_dzdfzW[kd]}, dzt{T.dzdfzT[c.k], T.r_dzdfzT[c.k]};
  const bool on = S.kvc != 0.0, lower = on && c.k > 1;
  Face zero; zero.ft = 0.0; zero.fs = 0.0;
  const double r0 = thcm_rho(S, t0, s0), r1 = thcm_rho(S, t1, s1), r2 = thcm_rho(S, t2, s2);
  const Face u0 = on ? thcm_face(S, up, dzu, t1, s1, r1, t2, s2, r2) : zero;
  const Face d0 = lower ? thcm_face(S, dn, dzd, t0, s0, r0, t1, s1, r1) : zero;
  thcm_mix_combine(S, dzt, u0, d0, o.bt, o.bs);
  Face u, d;
  // cell below: lower face only
  d = lower ? thcm_face(S, dn, dzd, t0 + eps, s0, thcm_rho(S, t0 + eps, s0), t1, s1, r1) : zero;
  thcm_mix_combine(S, dzt, u0, d, o.mt[0], o.ms[0]);
  d = lower ? thcm_face(S, dn, dzd, t0, s0 + eps, thcm_rho(S, t0, s0 + eps), t1, s1, r1) : zero;
  thcm_mix_combine(S, dzt, u0, d, o.mt[1], o.ms[1]);
  // the cell itself: both faces
  {
    const double rp = thcm_rho(S, t1 + eps, s1);
    u = on ? thcm_face(S, up, dzu, t1 + eps, s1, rp, t2, s2, r2) : zero;
    d = lower ? thcm_face(S, dn, dzd, t0, s0, r0, t1 + eps, s1, rp) : zero;
    thcm_mix_combine(S, dzt, u, d, o.mt[2], o.ms[2]);
  }
  {
    const double rp = thcm_rho(S, t1, s1 + eps);
    u = on ? thcm_face(S, up, dzu, t1, s1 + eps, rp, t2, s2, r2) : zero;
    d = lower ? thcm_face(S, dn, dzd, t0, s0, r0, t1, s1 + eps, rp) : zero;
    thcm_mix_combine(S, dzt, u, d, o.mt[3], o.ms[3]);
  }
  // cell above: upper face only
  u = on ? thcm_face(S, up, dzu, t1, s1, r1, t2 + eps, s2, thcm_rho(S, t2 + eps, s2)) : zero;
  thcm_mix_combine(S, dzt, u, d0, o.mt[4], o.ms[4]);
  u = on ? thcm_face(S, up, dzu, t1, s1, r1, t2, s2 + eps, thcm_rho(S, t2, s2 + eps)) : zero;
  thcm_mix_combine(S, dzt, u, d0, o.mt[5], o.ms[5]);
}

// ---- Jacobian rows: An = Al + nlin_jac (usrc.F90:873-995) + vmix_jac (mix_imp.f:729-815) ----
template <int A, class F> struct JacRow;

template <class F> struct JacRow<UU, F> {
  static THCM_HD void build(Row<UU>& r, const Scalars& S, const Tables& T, const Cell& c, const F& f, const MixFD* = nullptr) {
    const int j = c.j, k = c.k;
    const bool iL = c.i >= 2, iR = c.i <= S.n - 1, jS = j >= 2, jN = j <= S.m - 1;
    const double e = S.epsr, cx = T.cdxu[j], cy = T.cdyu[j], t8 = T.tdzi8[k];
    // UU-UU: Urux + uvy1 + uwz + uvy2  (spf.F90:561-640)
    double a23 = (f.w(0, 0, 0) + f.w(0, 1, 0) + f.w(1, 0, 0) + f.w(1, 1, 0)) * t8;
    double a14 = -(f.w(0, 0, -1) + f.w(0, 1, -1) + f.w(1, 0, -1) + f.w(1, 1, -1)) * t8;
    r.template at<2, UU>() = T.uu_we[j] + e * (iL ? -(2 * f.u(-1, 0, 0)) * cx : 0.0);
    r.template at<8, UU>() = T.uu_we[j] + e * (iR ? (2 * f.u(1, 0, 0)) * cx : 0.0);
    r.template at<4, UU>() = T.uu_s[j] + e * (jS ? -f.v(0, -1, 0) * T.cyv[j - 1] * cy : 0.0);
    r.template at<6, UU>() = T.uu_n[j] + e * (jN ? f.v(0, 1, 0) * T.cyv[j + 1] * cy : 0.0);
    r.template at<14, UU>() = T.uu_b[k] + e * a14;
    r.template at<23, UU>() = T.uu_t[k] + e * a23;
    r.template at<5, UU>() = (T.uu_ch[j] - T.uu_cv[k]) + e * ((a14 + a23) + f.v(0, 0, 0) * T.tanv[j]);
    // UU-VV: Urvy1 + Urvy2
    r.template at<2, VV>() = T.uv_w[j] + e * 0.0;
    r.template at<8, VV>() = T.uv_e[j] + e * 0.0;
    r.template at<4, VV>() = -0.0 + e * (jS ? -f.u(0, -1, 0) * T.cyv[j - 1] * cy : 0.0);
    r.template at<6, VV>() = -0.0 + e * (jN ? f.u(0, 1, 0) * T.cyv[j + 1] * cy : 0.0);
    r.template at<5, VV>() = T.uv_c[j] + e * (f.u(0, 0, 0) * T.tanv[j]);
    // UU-WW: Urwz
    double up = (f.u(0, 0, 0) + f.u(0, 0, 1)) * t8, dn = -(f.u(0, 0, 0) + f.u(0, 0, -1)) * t8;
    r.template at<5, WW>() = e * up; r.template at<6, WW>() = e * up; r.template at<8, WW>() = e * up; r.template at<9, WW>() = e * up;
    r.template at<14, WW>() = e * dn; r.template at<15, WW>() = e * dn; r.template at<17, WW>() = e * dn; r.template at<18, WW>() = e * dn;
    // UU-PP: gradp(1) (spf.F90:331-341)
    double px = jN ? cx : 0.0;
    r.template at<5, PP>() = -px; r.template at<6, PP>() = -px; r.template at<8, PP>() = px; r.template at<9, PP>() = px;
  }
};

template <class F> struct JacRow<VV, F> {
  static THCM_HD void build(Row<VV>& r, const Scalars& S, const Tables& T, const Cell& c, const F& f, const MixFD* = nullptr) {
    const int j = c.j, k = c.k;
    const bool iL = c.i >= 2, iR = c.i <= S.n - 1, jS = j >= 2, jN = j <= S.m - 1;
    const double e = S.epsr, cx = T.cdxu[j], cy = T.cdyu[j], t8 = T.tdzi8[k];
    double a23 = (f.w(0, 0, 0) + f.w(0, 1, 0) + f.w(1, 0, 0) + f.w(1, 1, 0)) * t8;
    double a14 = -(f.w(0, 0, -1) + f.w(0, 1, -1) + f.w(1, 0, -1) + f.w(1, 1, -1)) * t8;
    // VV-UU: Urt2 + uVrx (spf.F90:697-707, :781-789)
    r.template at<5, UU>() = T.vu_c[j] + e * ((2 * f.u(0, 0, 0)) * T.tanv[j]);
    r.template at<2, UU>() = T.vu_w[j] + e * (iL ? -f.v(-1, 0, 0) * cx : 0.0);
    r.template at<8, UU>() = T.vu_e[j] + e * (iR ? f.v(1, 0, 0) * cx : 0.0);
    // VV-VV: uvx + Vrvy + vwz
    r.template at<2, VV>() = T.vv_we[j] + e * (iL ? -f.u(-1, 0, 0) * cx : 0.0);
    r.template at<8, VV>() = T.vv_we[j] + e * (iR ? f.u(1, 0, 0) * cx : 0.0);
    r.template at<4, VV>() = T.vv_s[j] + e * (jS ? (-2 * f.v(0, -1, 0)) * T.cyv[j - 1] * cy : 0.0);
    r.template at<6, VV>() = T.vv_n[j] + e * (jN ? (2 * f.v(0, 1, 0)) * T.cyv[j + 1] * cy : 0.0);
    r.template at<14, VV>() = T.uu_b[k] + e * a14;
    r.template at<23, VV>() = T.uu_t[k] + e * a23;
    r.template at<5, VV>() = (T.vv_ch[j] - T.uu_cv[k]) + e * (a14 + a23);
    // VV-WW: Vrwz
    double up = (f.v(0, 0, 0) + f.v(0, 0, 1)) * t8, dn = -(f.v(0, 0, 0) + f.v(0, 0, -1)) * t8;
    r.template at<5, WW>() = e * up; r.template at<6, WW>() = e * up; r.template at<8, WW>() = e * up; r.template at<9, WW>() = e * up;
    r.template at<14, WW>() = e * dn; r.template at<15, WW>() = e * dn; r.template at<17, WW>() = e * dn; r.template at<18, WW>() = e * dn;
    // VV-PP: gradp(2)
    double py = jN ? S.dyi : 0.0;
    r.template at<5, PP>() = -py; r.template at<8, PP>() = -py; r.template at<6, PP>() = py; r.template at<9, PP>() = py;
  }
};

template <class F> struct JacRow<WW, F> {
  static THCM_HD void build(Row<WW>& r, const Scalars& S, const Tables& T, const Cell& c, const F& f, const MixFD* = nullptr) {
    const int k = c.k;
    r.template at<5, WW>() = 0.0;
    r.template at<5, PP>() = -T.wp[k];
    r.template at<23, PP>() = T.wp[k];
    double t2r = 0.0, t3r = 0.0;
    if (k <= S.l - 1) { // wnlin(1), wnlin(3) (spf.F90:496-527)
      double q = f.t(0, 0, 0) + f.t(0, 0, 1);
      t2r = q / 2.;
      t3r = 0.375 * (q * q);
    }
    double v = S.w_tt - S.ra_a2 * t2r + S.ra_a3 * t3r;
    r.template at<5, TT>() = v; r.template at<23, TT>() = v;
    r.template at<5, SS>() = S.w_ss; r.template at<23, SS>() = S.w_ss;
  }
};

template <class F> struct JacRow<PP, F> {
  static THCM_HD void build(Row<PP>& r, const Scalars& S, const Tables& T, const Cell& c, const F& f, const MixFD* = nullptr) {
    const int j = c.j, k = c.k;
    r.template at<5, PP>() = 0.0;
    r.template at<2, UU>() = -T.pu[j]; r.template at<4, UU>() = T.pu[j]; r.template at<1, UU>() = -T.pu[j]; r.template at<5, UU>() = T.pu[j];
    r.template at<4, VV>() = T.pv_s[j]; r.template at<2, VV>() = T.pv_n[j]; r.template at<1, VV>() = T.pv_s[j]; r.template at<5, VV>() = T.pv_n[j];
    r.template at<5, WW>() = T.pw[k]; r.template at<14, WW>() = -T.pw[k];
  }
};

// T and S rows share one implementation (X = own tracer, Y = the other one)
// Surface T row of an ocean coupled to an atmosphere / sea-ice model (lin, usrc.F90:725-741), the terms that follow Ooa*tc in the
// reference's expression, added in its order: + dedt*sc + mc*(QTnd*zeta*tc - Ooa*tc - dedt*sc); cross = Al(5,TT,SS).
// tc = sc = 1 on the surface level (tderiv 1, 2: spf.F90:206-209), mc = msi(i,j) there (masksi, spf.F90:347-359).
THCM_HD void thcm_coupled_surface(const Scalars& S, const Tables& T, const Cell& c, double& cen, double& cross) {
  cen = cen + S.tt_dedt;
  if (T.msi) {
    const double mc = T.msi[(size_t)(c.j - 1) * S.n + (c.i - 1)];
    cen = cen + mc * S.tt_mc;
    cross = S.ts_mc * mc;
  }
}

template <int X, class F> struct TracerJac {
  static THCM_HD double c_(const F& f, int di, int dj, int dk) { return X == TT ? f.t(di, dj, dk) : f.s(di, dj, dk); }
  static THCM_HD void build(Row<X>& r, const Scalars& S, const Tables& T, const Cell& c, const F& f, const MixFD* fd = nullptr) {
    constexpr int Y = (X == TT) ? SS : TT;
    const int j = c.j, k = c.k;
    const double cx = T.cdxt[j], cy = T.cdyt[j], td = S.tdzi2, dT = T.dfzT[k];
    const double c0 = c_(f, 0, 0, 0);
    // X-UU: urTx (tnlin 2), X-VV: vrTy (tnlin 4), X-WW: wrTz (tnlin 6)   (spf.F90:401-460)
    double uw = -(c0 + c_(f, -1, 0, 0)) * cx, ue = (c_(f, 1, 0, 0) + c0) * cx;
    r.template at<2, UU>() = uw; r.template at<4, UU>() = ue; r.template at<1, UU>() = uw; r.template at<5, UU>() = ue;
    double vs = -cy * (c0 + c_(f, 0, -1, 0)) * T.cyv[j - 1], vn = cy * (c_(f, 0, 1, 0) + c0) * T.cyv[j];
    r.template at<4, VV>() = vs; r.template at<1, VV>() = vs; r.template at<5, VV>() = vn; r.template at<2, VV>() = vn;
    const double rdT = T.r_dfzT[k];
    r.template at<14, WW>() = thcm_div(-td * (c0 + c_(f, 0, 0, -1)), dT, rdT);
    r.template at<5, WW>() = (k <= S.l - 1) ? thcm_div(td * (c_(f, 0, 0, 1) + c0), dT, rdT) : 0.0;
    // X-X: Utrx + Vtry + Wtrz (tnlin 3,5,7)
    double x2 = -(f.u(-1, 0, 0) + f.u(-1, -1, 0)) * cx, x8 = (f.u(0, 0, 0) + f.u(0, -1, 0)) * cx;
    double y4 = -(f.v(0, -1, 0) + f.v(-1, -1, 0)) * cy * T.cyv[j - 1], y6 = (f.v(0, 0, 0) + f.v(-1, 0, 0)) * cy * T.cyv[j];
    double z14 = thcm_div(-f.w(0, 0, -1) * td, dT, rdT), z23 = thcm_div(f.w(0, 0, 0) * td, dT, rdT);
    double surf = (k == S.l) ? (X == TT ? S.tt_surf : S.ss_surf) : 0.0;
    r.template at<2, X>() = T.tt_we[j] + x2;
    r.template at<8, X>() = T.tt_we[j] + x8;
    r.template at<4, X>() = T.tt_s[j] + y4;
    r.template at<6, X>() = T.tt_n[j] + y6;
    r.template at<14, X>() = T.tt_b[k] + z14;
    r.template at<23, X>() = T.tt_t[k] + z23;
    double cen = (T.tt_ch[j] - T.tt_cv[k]) + surf, cross = 0.0;
    if (X == TT && S.coupled_T && k == S.l) thcm_coupled_surface(S, T, c, cen, cross);
    r.template at<5, X>() = ((cen + (x2 + x8)) + (y4 + y6)) + (z14 + z23);
    r.template at<14, Y>() = 0.0; r.template at<5, Y>() = cross; r.template at<23, Y>() = 0.0;
    // vmix_jac: forward differences, eps = 1e-8 (mix_imp.f:745-769), only OCEAN rows carry a pattern (:891)
    if (S.vmix_on && !sbit(c.sig, SIG_NOTOCEAN)) {
      const double eps = 1.0e-08, reps = 1.0 / 1.0e-08;
      MixFD local;
      if (!fd) {
        thcm_mix_fd(S, T, c, f.t(0, 0, -1), f.s(0, 0, -1), f.t(0, 0, 0), f.s(0, 0, 0), f.t(0, 0, 1), f.s(0, 0, 1), local);
        fd = &local;
      }
      const double base = (X == TT) ? fd->bt : fd->bs;
      const double* mp = (X == TT) ? fd->mt : fd->ms;
      if (sbit(c.sig, SIG_FACE_DN)) { // neighbour k-1 is OCEAN (pattern of vmix_el_1, mix_imp.f:895-907)
        r.template at<14, TT>() = r.template at<14, TT>() + thcm_div(mp[0] - base, eps, reps);
        r.template at<14, SS>() = r.template at<14, SS>() + thcm_div(mp[1] - base, eps, reps);
      }
      r.template at<5, TT>() = r.template at<5, TT>() + thcm_div(mp[2] - base, eps, reps);
      r.template at<5, SS>() = r.template at<5, SS>() + thcm_div(mp[3] - base, eps, reps);
      if (sbit(c.sig, SIG_FACE_UP)) {
        r.template at<23, TT>() = r.template at<23, TT>() + thcm_div(mp[4] - base, eps, reps);
        r.template at<23, SS>() = r.template at<23, SS>() + thcm_div(mp[5] - base, eps, reps);
      }
    }
  }
};
template <class F> struct JacRow<TT, F> : TracerJac<TT, F> {};
template <class F> struct JacRow<SS, F> : TracerJac<SS, F> {};


// ---- residual rows: A_rhs(x) of nlin_rhs (usrc.F90:775-870); F = A_rhs(x) x + mix - Frc ----
template <int A, class F> struct RhsRow;

template <class F> struct RhsRow<UU, F> {
  static THCM_HD void build(Row<UU>& r, const Scalars& S, const Tables& T, const Cell& c, const F& f) {
    const int j = c.j, k = c.k;
    const bool iL = c.i >= 2, iR = c.i <= S.n - 1, jS = j >= 2, jN = j <= S.m - 1;
    const double e = S.epsr, cx = T.cdxu[j], cy = T.cdyu[j], t8 = T.tdzi8[k];
    double a23 = (f.w(0, 0, 0) + f.w(0, 1, 0) + f.w(1, 0, 0) + f.w(1, 1, 0)) * t8;
    double a14 = -(f.w(0, 0, -1) + f.w(0, 1, -1) + f.w(1, 0, -1) + f.w(1, 1, -1)) * t8;
    r.clear();
    r.template at<2, UU>() = T.uu_we[j] + e * (iL ? -f.u(-1, 0, 0) * cx : 0.0);   // uux (unlin 1)
    r.template at<8, UU>() = T.uu_we[j] + e * (iR ? f.u(1, 0, 0) * cx : 0.0);
    r.template at<4, UU>() = T.uu_s[j] + e * (jS ? -f.v(0, -1, 0) * T.cyv[j - 1] * cy : 0.0);
    r.template at<6, UU>() = T.uu_n[j] + e * (jN ? f.v(0, 1, 0) * T.cyv[j + 1] * cy : 0.0);
    r.template at<14, UU>() = T.uu_b[k] + e * a14;
    r.template at<23, UU>() = T.uu_t[k] + e * a23;
    r.template at<5, UU>() = (T.uu_ch[j] - T.uu_cv[k]) + e * ((a14 + a23) + f.v(0, 0, 0) * T.tanv[j]);
    r.template at<2, VV>() = T.uv_w[j]; r.template at<8, VV>() = T.uv_e[j]; r.template at<5, VV>() = T.uv_c[j];
    double px = jN ? cx : 0.0;
    r.template at<5, PP>() = -px; r.template at<6, PP>() = -px; r.template at<8, PP>() = px; r.template at<9, PP>() = px;
  }
};
template <class F> struct RhsRow<VV, F> {
  static THCM_HD void build(Row<VV>& r, const Scalars& S, const Tables& T, const Cell& c, const F& f) {
    const int j = c.j, k = c.k;
    const bool iL = c.i >= 2, iR = c.i <= S.n - 1, jS = j >= 2, jN = j <= S.m - 1;
    const double e = S.epsr, cx = T.cdxu[j], cy = T.cdyu[j], t8 = T.tdzi8[k];
    double a23 = (f.w(0, 0, 0) + f.w(0, 1, 0) + f.w(1, 0, 0) + f.w(1, 1, 0)) * t8;
    double a14 = -(f.w(0, 0, -1) + f.w(0, 1, -1) + f.w(1, 0, -1) + f.w(1, 1, -1)) * t8;
    r.clear();
    r.template at<5, UU>() = T.vu_c[j] + e * (f.u(0, 0, 0) * T.tanv[j]);           // ut2 (vnlin 7)
    r.template at<2, UU>() = T.vu_w[j]; r.template at<8, UU>() = T.vu_e[j];
    r.template at<2, VV>() = T.vv_we[j] + e * (iL ? -f.u(-1, 0, 0) * cx : 0.0);   // uvx (vnlin 1)
    r.template at<8, VV>() = T.vv_we[j] + e * (iR ? f.u(1, 0, 0) * cx : 0.0);
    r.template at<4, VV>() = T.vv_s[j] + e * (jS ? -f.v(0, -1, 0) * T.cyv[j - 1] * cy : 0.0); // vvy (vnlin 3)
    r.template at<6, VV>() = T.vv_n[j] + e * (jN ? f.v(0, 1, 0) * T.cyv[j + 1] * cy : 0.0);
    r.template at<14, VV>() = T.uu_b[k] + e * a14;
    r.template at<23, VV>() = T.uu_t[k] + e * a23;
    r.template at<5, VV>() = (T.vv_ch[j] - T.uu_cv[k]) + e * (a14 + a23);
    double py = jN ? S.dyi : 0.0;
    r.template at<5, PP>() = -py; r.template at<8, PP>() = -py; r.template at<6, PP>() = py; r.template at<9, PP>() = py;
  }
};
template <class F> struct RhsRow<WW, F> {
  static THCM_HD void build(Row<WW>& r, const Scalars& S, const Tables& T, const Cell& c, const F& f) {
    const int k = c.k;
    r.clear();
    r.template at<5, PP>() = -T.wp[k];
    r.template at<23, PP>() = T.wp[k];
    double q5 = 0.0, q23 = 0.0, c5 = 0.0, c23 = 0.0;
    if (k <= S.l - 1) { // wnlin(2), wnlin(4) (spf.F90:505-538)
      double t0 = f.t(0, 0, 0), t1 = f.t(0, 0, 1);
      q23 = t1 / 4.; q5 = (t0 + 2 * t1) / 4.;
      c5 = 0.125 * (t0 * t0 + 3 * t1 * t0 + 3 * t1 * t1);
      c23 = 0.125 * t1 * t1;
    }
    r.template at<5, TT>() = S.w_tt - S.ra_a2 * q5 + S.ra_a3 * c5;
    r.template at<23, TT>() = S.w_tt - S.ra_a2 * q23 + S.ra_a3 * c23;
    r.template at<5, SS>() = S.w_ss; r.template at<23, SS>() = S.w_ss;
  }
};
template <class F> struct RhsRow<PP, F> {
  static THCM_HD void build(Row<PP>& r, const Scalars& S, const Tables& T, const Cell& c, const F& f) {
    JacRow<PP, F>::build(r, S, T, c, f, nullptr); // linear row
  }
};
template <int X, class F> struct TracerRhs {
  static THCM_HD void build(Row<X>& r, const Scalars& S, const Tables& T, const Cell& c, const F& f) {
    const int j = c.j, k = c.k;
    const double cx = T.cdxt[j], cy = T.cdyt[j], td = S.tdzi2, dT = T.dfzT[k];
    double x2 = -(f.u(-1, 0, 0) + f.u(-1, -1, 0)) * cx, x8 = (f.u(0, 0, 0) + f.u(0, -1, 0)) * cx;
    double y4 = -(f.v(0, -1, 0) + f.v(-1, -1, 0)) * cy * T.cyv[j - 1], y6 = (f.v(0, 0, 0) + f.v(-1, 0, 0)) * cy * T.cyv[j];
    const double rdT = T.r_dfzT[k];
    double z14 = thcm_div(-f.w(0, 0, -1) * td, dT, rdT), z23 = thcm_div(f.w(0, 0, 0) * td, dT, rdT);
    double surf = (k == S.l) ? (X == TT ? S.tt_surf : S.ss_surf) : 0.0;
    r.clear();
    r.template at<2, X>() = T.tt_we[j] + x2;
    r.template at<8, X>() = T.tt_we[j] + x8;
    r.template at<4, X>() = T.tt_s[j] + y4;
    r.template at<6, X>() = T.tt_n[j] + y6;
    r.template at<14, X>() = T.tt_b[k] + z14;
    r.template at<23, X>() = T.tt_t[k] + z23;
    double cen = (T.tt_ch[j] - T.tt_cv[k]) + surf;
    if (X == TT && S.coupled_T && k == S.l) {
      double cross = 0.0;
      thcm_coupled_surface(S, T, c, cen, cross);
      r.template at<5, SS>() = cross;
    }
    r.template at<5, X>() = ((cen + (x2 + x8)) + (y4 + y6)) + (z14 + z23);
  }
};
template <class F> struct RhsRow<TT, F> : TracerRhs<TT, F> {};
template <class F> struct RhsRow<SS, F> : TracerRhs<SS, F> {};

// ---- boundary.F90:81-387, statement for statement, on the row's own entries ----
// Every statement of `boundaries` touches An(slot, A, B) for one A or for all A, never mixing rows, so it can be
// applied to each row independently.  The 1.0e-10 "inactive connection" values (boundary.F90:201-204) are
// dropped by the |a|>1e-10 filter of fillcolA (assemble.F90:115), hence written as 0 here.
template <int A> THCM_HD void apply_boundaries(Row<A>& r, uint64_t g) {
  auto L = [g](int slot) { return sland(g, slot); };
  if (!sbit(g, SIG_NOTOCEAN)) {
    if (L(14)) { // bottom
      if (L(11) && L(10) && L(13)) { r.template fold<1, 10, UU>(); r.template fold<1, 10, VV>(); }
      r.template zuv<10>();
      if (L(11) && L(18) && L(15)) { r.template fold<2, 11, UU>(); r.template fold<2, 11, VV>(); }
      r.template zuv<11>();
      if (L(17) && L(16) && L(13)) { r.template fold<4, 13, UU>(); r.template fold<4, 13, VV>(); }
      r.template zuv<13>();
      if (L(17) && L(18) && L(15)) { r.template fold<5, 14, UU>(); r.template fold<5, 14, VV>(); }
      r.template fold<5, 14, TT>(); r.template fold<5, 14, SS>();
      r.template zs<14>();
    }
    if (L(10)) r.template zs<10>();
    if (L(11)) r.template zs<11>();
    if (L(12)) r.template zs<12>();
    if (L(13)) r.template zs<13>();
    if (L(15)) r.template zs<15>();
    if (L(16)) r.template zs<16>();
    if (L(17)) r.template zs<17>();
    if (L(18)) r.template zs<18>();
    if (L(23)) { // top
      if (L(20) && L(19) && L(22)) { r.template fold<1, 19, UU>(); r.template fold<1, 19, VV>(); }
      r.template zuv<19>();
      if (L(20) && L(21) && L(24)) { r.template fold<2, 20, UU>(); r.template fold<2, 20, VV>(); }
      r.template zuv<20>();
      if (L(26) && L(25) && L(22)) { r.template fold<4, 22, UU>(); r.template fold<4, 22, VV>(); }
      r.template zuv<22>();
      if (L(26) && L(27) && L(24)) { r.template fold<5, 23, UU>(); r.template fold<5, 23, VV>(); }
      r.template fold<5, 23, TT>(); r.template fold<5, 23, SS>();
      r.template zs<23>();
      if (A == WW) r.clear();
      r.template zc<5, WW>(); r.template zc<6, WW>(); r.template zc<8, WW>(); r.template zc<9, WW>(); // = 1.0e-10, filtered
      if constexpr (A == WW) r.template at<5, WW>() = 1.0;
    }
    if (L(19)) r.template zs<19>();
    if (L(20)) r.template zs<20>();
    if (L(21)) r.template zs<21>();
    if (L(22)) r.template zs<22>();
    if (L(24)) r.template zs<24>();
    if (L(25)) r.template zs<25>();
    if (L(26)) r.template zs<26>();
    if (L(27)) r.template zs<27>();
    if (L(1)) r.template zuv<1>();
    if (L(2)) { // west
      r.template fold<5, 2, TT>(); r.template fold<5, 2, SS>();
      r.template zs<2>(); r.template zuv<1>();
    }
    if (L(3)) { r.template zuv<2>(); r.template zuv<3>(); }
    else if (sbit(g, SIG_NN)) r.template zuv<3>();
    if (L(4)) { // south
      r.template fold<5, 4, SS>(); r.template fold<5, 4, TT>();
      r.template zs<4>(); r.template zuv<1>();
    }
    if (L(6)) { // north
      r.template zuv<2>();
      if constexpr (A == PP) { r.template at<2, UU>() = 0.0; r.template at<2, VV>() = 0.0; r.template at<5, UU>() = 0.0; r.template at<5, VV>() = 0.0; }
      if constexpr (A == VV) { r.clear(); r.template at<5, VV>() = 1.0; } else r.template zc<5, VV>();
      if constexpr (A == UU) { r.clear(); r.template at<5, UU>() = 1.0; } else r.template zc<5, UU>();
      r.template fold<5, 6, SS>(); r.template fold<5, 6, TT>();
      r.template zs<6>();
    } else if (sbit(g, SIG_NN)) { r.template zuv<3>(); r.template zuv<6>(); }
    if (L(7)) { r.template zuv<4>(); r.template zuv<7>(); }
    else if (sbit(g, SIG_SOUTHEE)) r.template zuv<7>();
    if (L(8)) { // east
      r.template zuv<4>();
      if constexpr (A == PP) { r.template at<4, UU>() = 0.0; r.template at<4, VV>() = 0.0; r.template at<5, UU>() = 0.0; r.template at<5, VV>() = 0.0; }
      if constexpr (A == UU) { r.clear(); r.template at<5, UU>() = 1.0; } else r.template zc<5, UU>();
      if constexpr (A == VV) { r.clear(); r.template at<5, VV>() = 1.0; } else r.template zc<5, VV>();
      r.template fold<5, 8, SS>(); r.template fold<5, 8, TT>();
      r.template zs<8>(); r.template zuv<7>();
    } else if (sbit(g, SIG_EASTEAST)) { r.template zuv<7>(); r.template zuv<8>(); }
    if (L(9)) { // north-east
      if constexpr (A == UU) { r.clear(); r.template at<5, UU>() = 1.0; } else r.template zc<5, UU>();
      if constexpr (A == VV) { r.clear(); r.template at<5, VV>() = 1.0; } else r.template zc<5, VV>();
      r.template zuv<7>();
    } else {
      if (sbit(g, SIG_NORTHEE)) { r.template zuv<8>(); r.template zuv<9>(); }
      else if (sbit(g, SIG_NNORTHEE)) r.template zuv<9>();
      if (sbit(g, SIG_NN)) { r.template zuv<6>(); r.template zuv<9>(); }
    }
  } else {
    r.clear();
    r.template at<5, A>() = 1.0;
  }
}

// fillcolA filter (assemble.F90:115)
template <int A> THCM_HD void filter_row(Row<A>& r) {
#pragma unroll
  for (int q = 0; q < Row<A>::NE; q++) r.a[q] = (fabs(r.a[q]) > 1.0e-10) ? r.a[q] : 0.0;
}

// Frc rows zeroed by `boundaries` (boundary.F90:197,264,271,309,316,341,348,391)
template <int A> THCM_HD bool frc_masked(uint64_t g) {
  if (sbit(g, SIG_NOTOCEAN)) return true;
  if (A == WW) return sland(g, 23);
  if (A == UU || A == VV) return sland(g, 6) || sland(g, 8) || sland(g, 9);
  return false;
}


// ---- compile-time iteration helpers ----
template <int Q> struct IC { static constexpr int value = Q; };
template <int Q, int N, class Fn> THCM_HD void static_for(Fn&& fn) {
  if constexpr (Q < N) { fn(IC<Q>{}); static_for<Q + 1, N>(fn); }
}
// q-th entry of row A in Fortran CSR order (slot-major, then variable: assemble.F90:110-127)
template <int A> constexpr int slot_order(int q) {
  int cnt = 0;
  for (int s = 1; s <= 27; s++)
    for (int b = 1; b <= 6; b++) {
      int e = find_ent<A>(s, b);
      if (e >= 0) { if (cnt == q) return e; cnt++; }
    }
  return -1;
}

// Is the neighbour of stencil slot `slot` inside the global domain?  (insert_graph_entry, THCM.C:2522-2539)
THCM_HD bool slot_present(int slot, const Scalars& S, const Cell& c) {
  int di = slot_di(slot), dj = slot_dj(slot), dk = slot_dk(slot);
  if (!S.periodic && ((di < 0 && c.i == 1) || (di > 0 && c.i == S.n))) return false;
  if ((dj < 0 && c.j == 1) || (dj > 0 && c.j == S.m)) return false;
  if ((dk < 0 && c.k == 1) || (dk > 0 && c.k == S.l)) return false;
  return true;
}
// 0-based global column of entry (slot,var) seen from cell c (shift + find_row2, assemble.F90:142-179)
THCM_HD int slot_col(int slot, int var, const Scalars& S, const Cell& c) {
  int i2 = c.i + slot_di(slot), j2 = c.j + slot_dj(slot), k2 = c.k + slot_dk(slot);
  if (S.periodic) { if (i2 == 0) i2 = S.n; else if (i2 == S.n + 1) i2 = 1; }
  return 6 * (((k2 - 1) * S.m + (j2 - 1)) * S.n + (i2 - 1)) + var - 1;
}

// Rows whose stencil crosses the periodic seam inside ONE rank: their stored (= global) column order differs from the
// static (dk,dj,di,var) order and is sorted explicitly.  With several ranks in x the seam lies between ranks and the
// static order is kept everywhere.
THCM_HD bool row_has_seam(const Scalars& S, const Cell& c) { return S.periodic && S.nloc == S.n && (c.i == 1 || c.i == S.n); }

// Au(row) = sum coA*un(jcoA) in the reference's storage order (matAvec, matetc.F90:147-166).
// X::raw(di,dj,dk,var) returns the raw state of the neighbour cell.
template <int A, class X> THCM_HD double row_dot(const Row<A>& r, const X& x) {
  double s = 0.0;
  static_for<0, Row<A>::NE>([&](auto q) {
    constexpr int e = slot_order<A>(decltype(q)::value);
    constexpr int slot = RowDef<A>::e[e].slot, var = RowDef<A>::e[e].var;
    double a = r.a[e];
    if (a != 0.0) s = a * x.raw(slot_di(slot), slot_dj(slot), slot_dk(slot), var) + s;
  });
  return s;
}

// ---- storing a row into the static CSR ----
// Entries sort by (dk, dj) group first (9 groups), then by column inside the group.  At a periodic seam inside one
// rank the wrapped neighbour changes only the order INSIDE a group: at i=1 the west neighbours (di=-1) wrap to i=n and
// sort last, at i=n the east neighbours (di=+1) wrap to i=1 and sort first.  A group is present or absent as a whole
// there (di is never clipped on a periodic domain), so positions are group prefix sums plus compile-time ranks.
constexpr int slot_group(int s) { return (slot_dk(s) + 1) * 3 + (slot_dj(s) + 1); }
constexpr int seam_key(int di, int side) { // side 0: i==1, side 1: i==n
  return side == 0 ? (di == -1 ? 2 : (di == 0 ? 0 : 1)) : (di == 1 ? 0 : (di == -1 ? 1 : 2));
}
template <int A> constexpr int group_size(int g) {
  int cnt = 0;
  for (int q = 0; q < RowDef<A>::NE; q++) cnt += slot_group(RowDef<A>::e[q].slot) == g ? 1 : 0;
  return cnt;
}
template <int A> constexpr int seam_rank(int e, int side) {
  const int g = slot_group(RowDef<A>::e[e].slot);
  const int ke = seam_key(slot_di(RowDef<A>::e[e].slot), side) * 8 + RowDef<A>::e[e].var;
  int cnt = 0;
  for (int q = 0; q < RowDef<A>::NE; q++)
    if (slot_group(RowDef<A>::e[q].slot) == g) {
      const int kq = seam_key(slot_di(RowDef<A>::e[q].slot), side) * 8 + RowDef<A>::e[q].var;
      cnt += kq < ke ? 1 : 0;
    }
  return cnt;
}
// Seam cells whose nine (dk,dj) groups are all present (1 < j < m, 1 < k < l): positions are compile-time constants.
template <int A> constexpr int group_prefix(int g) {
  int acc = 0;
  for (int q = 0; q < g; q++) acc += group_size<A>(q);
  return acc;
}
template <int A> constexpr int seam_pos(int e, int side) { return group_prefix<A>(slot_group(RowDef<A>::e[e].slot)) + seam_rank<A>(e, side); }
template <int A, class Out> THCM_HD void emit_row_seam_inner(const Row<A>& r, bool west, Out out) {
  if (west) static_for<0, Row<A>::NE>([&](auto q) { constexpr int e = decltype(q)::value; out[seam_pos<A>(e, 0)] = r.a[e]; });
  else static_for<0, Row<A>::NE>([&](auto q) { constexpr int e = decltype(q)::value; out[seam_pos<A>(e, 1)] = r.a[e]; });
}
// `out` points at the first value of this row (rowptr[row]).
template <int A, class Out> THCM_HD void emit_row(const Row<A>& r, const Scalars& S, const Cell& c, Out out) {
  const bool seam = row_has_seam(S, c);
  const bool inner_jk = c.j > 1 && c.j < S.m && c.k > 1 && c.k < S.l;
  const bool inner_i = S.periodic || (c.i > 1 && c.i < S.n);
  if (!seam && inner_jk && inner_i) { // every slot present, static order == column order
    static_for<0, Row<A>::NE>([&](auto q) { constexpr int e = decltype(q)::value; out[e] = r.a[e]; });
  } else if (!seam) {
    int pos = 0;
    static_for<0, Row<A>::NE>([&](auto q) {
      constexpr int e = decltype(q)::value;
      if (slot_present(RowDef<A>::e[e].slot, S, c)) { out[pos] = r.a[e]; pos++; }
    });
  } else {
    int pre[9];
    int acc = 0;
    static_for<0, 9>([&](auto gq) {
      constexpr int g = decltype(gq)::value;
      constexpr int dk = g / 3 - 1, dj = g % 3 - 1;
      const bool present = !((dj < 0 && c.j == 1) || (dj > 0 && c.j == S.m) || (dk < 0 && c.k == 1) || (dk > 0 && c.k == S.l));
      pre[g] = present ? acc : -1;
      acc += present ? group_size<A>(g) : 0;
    });
    const bool west = c.i == 1;
    static_for<0, Row<A>::NE>([&](auto q) {
      constexpr int e = decltype(q)::value;
      constexpr int g = slot_group(RowDef<A>::e[e].slot);
      constexpr int rw = seam_rank<A>(e, 0), re = seam_rank<A>(e, 1);
      if (pre[g] >= 0) out[pre[g] + (west ? rw : re)] = r.a[e];
    });
  }
}
// number of stored entries of row A at cell c
template <int A> THCM_HD int row_len(const Scalars& S, const Cell& c) {
  int cnt = 0;
  static_for<0, Row<A>::NE>([&](auto q) {
    constexpr int e = decltype(q)::value;
    cnt += slot_present(RowDef<A>::e[e].slot, S, c) ? 1 : 0;
  });
  return cnt;
}

} // namespace thcm
