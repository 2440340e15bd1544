// thcm_coupled.cuh -- arithmetic of the coupling row shared by the CUDA kernel (coupled.cu) and the host test harness.
#pragma once
#include "thcm_rows.cuh"

namespace thcm {

// scalars of m_probe::get_derivatives (probe.F90:429-444); the products are formed on the host in the Fortran's order
struct DftdmScal {
  double cs;    // par(COMB)*par(SUNP)
  double albe0, albed, Ooa;
  double leq;   // lvsc*eta*qdim
  double dq;    // deltat/qdim*dqso
  double lveo;  // lvsc*eo0
  double qtz;   // QTnd*zeta
  double a0, s0, t0;
};
// dF_T/dM_seaice = QTos - QToa at an OCEAN surface point (probe.F90:433-442)
THCM_HD double thcm_dftdm(const DftdmScal& a, double To, double So, double Ta, double Ab, double qa, double sunoj) {
  const double QTos = a.qtz * (a.a0 * (So + a.s0) - (To + a.t0));
  const double QToa = a.cs * sunoj * (1 - a.albe0 - a.albed * Ab) - a.Ooa * (To - Ta) - a.leq * (a.dq * To - qa) - a.lveo;
  return QTos - QToa;
}

} // namespace thcm
