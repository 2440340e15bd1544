// thcm_cell.cuh -- one (cell, unknown) row from start to finish: build, boundaries, filter, store / contract.
// Shared by the CUDA kernels (thcm_kernels.cu) and by the host harness in tests/ (host_rowcheck.cpp).
#pragma once
#include "thcm_rows.cuh"

namespace thcm {

// Rows the C++ glue of the reference overwrites after the Fortran call (THCM.C:1006-1033, :1160-1165).
// `side` (optional, 18 doubles): the centre-slot entries An(5,A,1:6) of those three rows as the Fortran built them,
// which m_scaling::average_block (scaling.F90:29-62) reads from the Fortran CSR and the overwritten rows no longer hold.
struct SpecialRows { int rowintcon, rowPfix1, rowPfix2; double* side; };
template <int A> THCM_HD void save_centre(Row<A>& r, double* dst) {
  dst[0] = 0.0; dst[1] = 0.0; dst[2] = 0.0; dst[3] = 0.0; dst[4] = 0.0; dst[5] = 0.0;
  if constexpr (Row<A>::template has<5, UU>()) dst[0] = r.template at<5, UU>();
  if constexpr (Row<A>::template has<5, VV>()) dst[1] = r.template at<5, VV>();
  if constexpr (Row<A>::template has<5, WW>()) dst[2] = r.template at<5, WW>();
  if constexpr (Row<A>::template has<5, PP>()) dst[3] = r.template at<5, PP>();
  if constexpr (Row<A>::template has<5, TT>()) dst[4] = r.template at<5, TT>();
  if constexpr (Row<A>::template has<5, SS>()) dst[5] = r.template at<5, SS>();
}

// diag(B) entry of row A: fillcolB (assemble.F90:18-54) times "Mass" = 1 (THCM.C:1150-1153); the rows the C++ glue overwrites
// (integral condition THCM.C:2137-2139, fixed pressure points THCM.C:2201-2238) carry 0.
template <int A> THCM_HD double diagb_entry(const Scalars& S, uint64_t sig, const SpecialRows& sp, int row) {
  double b = 0.0;
  if (!sbit(sig, SIG_NOTOCEAN)) {
    if (A == UU) b = sbit(sig, SIG_BU) ? -S.epsr : 0.0;
    if (A == VV) b = sbit(sig, SIG_BV) ? -S.epsr : 0.0;
    if (A == TT || A == SS) b = -1.0;
  }
  if (row == sp.rowintcon) b = 0.0;
  if (A == PP && (row == sp.rowPfix1 || row == sp.rowPfix2)) b = 0.0;
  return b;
}

// Jacobian row A of cell c, built in registers (static pattern order).  Returns false when the row must not be stored:
// the integral-condition row is left untouched (its dense values are static, written once by the host).
template <int A, class F>
THCM_HD bool jac_row_build(Row<A>& r, const Scalars& S, const Tables& T, const Cell& c, const F& f, const SpecialRows& sp, int row,
                           const MixFD* fd = nullptr) {
  if (sbit(c.sig, SIG_NOTOCEAN)) { // `boundaries` turns every row of a non-ocean cell into an identity row (boundary.F90:383-391)
    r.clear();
    r.template at<5, A>() = 1.0;
  } else {
    JacRow<A, F>::build(r, S, T, c, f, fd);
    if ((uint32_t)c.sig != 0u) apply_boundaries<A>(r, c.sig); // bits 0..31: a land neighbour somewhere in the stencil
    filter_row<A>(r);
  }
  if (row == sp.rowintcon) {                              // intcond_S, THCM.C:2137-2139
    if constexpr (A == SS) { if (sp.side) save_centre<A>(r, sp.side); } // the integral row replaces an S row
    return false;
  }
  if constexpr (A == PP) {
    if (row == sp.rowPfix1 || row == sp.rowPfix2) { // fixPressurePoints, THCM.C:2201-2238
      if (sp.side) save_centre<A>(r, sp.side + (row == sp.rowPfix1 ? 6 : 12));
      r.clear();
      r.template at<5, PP>() = 1.0;
    }
  }
  return true;
}

// Jacobian row A of cell c -> `out` (first stored value of the row in the static CSR), returns diag(B) entry.
template <int A, class F, class Out>
THCM_HD double jac_row(const Scalars& S, const Tables& T, const Cell& c, const F& f, const SpecialRows& sp, int row, Out out,
                       const MixFD* fd = nullptr) {
  Row<A> r;
  if (jac_row_build<A, F>(r, S, T, c, f, sp, row, fd)) emit_row<A>(r, S, c, out);
  return diagb_entry<A>(S, c.sig, sp, row);
}

// Residual row A of cell c: F = -( (-Au - mix + Frc) * (1-landm) )   (usrc.F90:559-574, THCM.C:1003)
// X::raw(di,dj,dk,var) gives raw neighbour state; mixv is the mixing term of this row (0 unless T/S).
template <int A, class F, class X>
THCM_HD double rhs_row(const Scalars& S, const Tables& T, const Cell& c, const F& f, const X& x, double frc, double mixv) {
  Row<A> r;
  double Au;
  if (sbit(c.sig, SIG_NOTOCEAN)) {
    Au = 1.0 * x.raw(0, 0, 0, A) + 0.0; // identity row: what row_dot returns for it, without building the row
  } else {
    RhsRow<A, F>::build(r, S, T, c, f);
    if ((uint32_t)c.sig != 0u) apply_boundaries<A>(r, c.sig);
    filter_row<A>(r);
    Au = row_dot<A>(r, x);
  }
  double B = -Au - mixv + frc - 0.0;
  B = B * (sbit(c.sig, SIG_LANDVAL) ? 0.0 : 1.0);
  return B * -1.0;
}

} // namespace thcm
