// coupled.cu -- the ocean side of the coupling to an atmosphere / sea-ice model (Coupled Temperature = 1).
//
//   forcing_cells_kernel   surface T rows of Frc from the per-cell values the host evaluated with the reference's expression
//                          (forcing.F90:73-96); the kernel only copies and applies the Frc zeroing, so host and device agree
//   dftdm_kernel           m_probe::get_derivatives, coupled_T part (probe.F90:429-444): d F_T / d M_seaice per surface point
//   coupling_apply_kernel  CouplingBlock::applyMatrix (CouplingBlock.H:206 = Epetra_CrsMatrix::Apply) for the few rows a
//                          coupling block has, optionally accumulated into the result as CoupledModel::applyMatrix does
//                          (CoupledModel.C:436-470: out += z)
// Compiled with -fmad=false: the expressions follow the Fortran's order and must round like it.
#include "kernels.h"

namespace thcm {
namespace {

__global__ void __launch_bounds__(256) forcing_cells_kernel(const Scalars S, const uint64_t* __restrict__ sig, const double* __restrict__ ftcell,
                                                            double* __restrict__ frc) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= S.nloc * S.mloc) return;
  const int li = t % S.nloc, lj = t / S.nloc;
  const size_t c = ((size_t)(S.l - 1) * S.mloc + lj) * S.nloc + li;
  frc[6 * c + 4] = frc_masked<TT>(__ldg(sig + c)) ? 0.0 : __ldg(ftcell + t);
}

__global__ void __launch_bounds__(256) dftdm_kernel(const DftdmArgs a) {
  const Scalars& S = a.S;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= S.nloc * S.mloc) return;
  const int li = t % S.nloc, lj = t / S.nloc, i = S.io0 + li, j = S.jo0 + lj;
  const size_t c = ((size_t)(S.l - 1) * S.mloc + lj) * S.nloc + li;
  double out = 0.0;
  if (!sbit(__ldg(a.sig + c), SIG_NOTOCEAN)) { // landm(i,j,l) == OCEAN
    const size_t g = (size_t)(j - 1) * S.n + (i - 1);
    const double To = a.x[6 * c + 4], So = a.x[6 * c + 5];
    const double Ta = a.tatm ? a.tatm[g] : 0.0, Ab = a.albe ? a.albe[g] : 0.0, qa = a.qatm ? a.qatm[g] : 0.0;
    out = thcm_dftdm(a.c, To, So, Ta, Ab, qa, a.suno[j]);
  }
  a.out[t] = out;
}

// one thread per non-empty row, entries in stored (column-sorted) order
__global__ void __launch_bounds__(256) coupling_apply_kernel(int nrows, const int* __restrict__ rows, const int* __restrict__ beg, const int* __restrict__ jco,
                                                             const double* __restrict__ co, const double* __restrict__ x, double* __restrict__ y,
                                                             int accumulate) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nrows) return;
  double s = 0.0;
  for (int p = beg[t]; p < beg[t + 1]; p++) s = s + co[p] * x[jco[p]];
  const int r = rows[t];
  y[r] = accumulate ? y[r] + s : s;
}
} // namespace

void launch_forcing_cells(const Scalars& S, const uint64_t* sig, const double* ftcell, double* frc, cudaStream_t st) {
  const int cells = S.nloc * S.mloc;
  forcing_cells_kernel<<<(cells + 255) / 256, 256, 0, st>>>(S, sig, ftcell, frc);
}
void launch_dftdm(const DftdmArgs& a, cudaStream_t st) {
  const int cells = a.S.nloc * a.S.mloc;
  dftdm_kernel<<<(cells + 255) / 256, 256, 0, st>>>(a);
}
void launch_coupling_apply(int nrows, const int* rows, const int* beg, const int* jco, const double* co, const double* x, double* y, int accumulate,
                           cudaStream_t st) {
  if (nrows > 0) coupling_apply_kernel<<<(nrows + 255) / 256, 256, 0, st>>>(nrows, rows, beg, jco, co, x, y, accumulate);
}

} // namespace thcm
