// thcm_host.h -- host-side setup of the B200 THCM path: grid, parameters, coefficient tables, per-cell
// land signatures, forcing vector and the static CSR graph.  Everything here runs once per model /
// per parameter change; the per-evaluation arithmetic lives in the CUDA kernels (thcm_kernels.cu).
//
// Mirrors, for the hot path only, what the reference does in
//   src/ocean/THCM.C:176-800 (constructor), src/ocean/usrc.F90:6-139 (init), src/ocean/grid.F90,
//   src/ocean/usrc.F90:1136-1172 (stpnt), src/ocean/forcing.F90:4-218, src/ocean/THCM.C:2241-2539 (graph).
#pragma once
#include <stdint.h>
#include <string>
#include <vector>
#include "thcm_rows.cuh"
#include "thcm_coupled.cuh"

namespace thcm {

// Same fields as thcm_b200_config in include/iemic_b200.h (kept layout-compatible).
struct Config {
  int n, m, l;
  double xmin, xmax, ymin, ymax; // radians
  int periodic;
  double hdim, qz;
  double alphaT, alphaS;
  int ih, vmix, tap, rho_mixing, coriolis_on, forcing_type;
  int tres, sres, iza, ite, its;
  int coupled_T, coupled_S;
  int int_sign;
  int intcond_i, intcond_j;
  int fix_pressure_points;
};

enum { P_AL_T = 1, P_RAYL, P_EK_V, P_EK_H, P_ROSB, P_MIXP, P_RESC, P_SPL1, P_HMTP, P_SUNP, P_PE_H, P_PE_V, P_P_VC, P_LAMB,
       P_SALT, P_WIND, P_TEMP, P_BIOT, P_COMB, P_ARCL, P_NLES, P_IFRICB, P_CONT, P_ENER, P_ALPC, P_CMPR, P_FPER, P_SPER,
       P_MKAP, P_SPL2, NPAR = 30 };

class Host {
 public:
  // rank/nranks select one block of the 2-D decomposition (TRIOS::Domain::Decomp2D); throws std::runtime_error on
  // unsupported configurations
  Host(const Config& cfg, const int* landm_in, int rank = 0, int nranks = 1);
  void set_par(int idx1, double value);          // setparcs (usrc.F90:163-181): recomputes forcing + tables
  double get_par(int idx1) const { return (idx1 >= 1 && idx1 <= NPAR) ? par[idx1] : 0.0; }
  void set_landmask(const int* landm_in);        // set_landmask (usrc.F90:336-395)

  Config cfg;
  int n, m, l;
  size_t ncell, ndim;
  double par[NPAR + 1];
  double dx, dy, dz;
  std::vector<double> x, y, z, xu, yv, zw, dfzT, dfzW; // Fortran index ranges
  std::vector<int> landm;       // (0:n+1,0:m+1,0:l+1), after the init() border treatment
  std::vector<uint64_t> sig;    // per cell
  Scalars S;
  std::vector<double> tab;      // NJTAB*(m+2) + NKTAB*(l+1) doubles, order of thcm::Tables
  std::vector<double> frc;      // forcing vector with the rows `boundaries` zeroes already masked
  std::vector<int> rowptr, col; // static graph, 0-based, sorted columns
  int rowintcon, rowPfix1, rowPfix2;
  std::vector<double> intcond;  // integral condition coefficients (ndim, zero except ocean S dofs)
  double intCorrection;
  uint64_t version;             // bumped whenever tables/forcing change (device copies refresh lazily)
  uint64_t mask_version;

  // ---- decomposition (single block until dist_setup() is called) ----
  int rank = 0, nranks = 1, npN = 1, npM = 1, pidN = 0, pidM = 0;
  int io0 = 1, jo0 = 1, nloc = 0, mloc = 0;   // owned block: global Fortran index of its first cell, extent
  bool stored_cell(size_t c, int& fi, int& fj, int& fk) const;
  // halo exchange plan (the Epetra_Import of TRIOS::Domain, TRIOS_Domain.C:220-225, 517-664, as index lists):
  // for every peer rank the owned cells to pack for it and the halo-ring cells its data lands in, both in the
  // receiver's ring order so that the two sides agree without any handshake.
  struct HaloPeer { int rank; std::vector<int> send_cells, recv_cells; };
  std::vector<HaloPeer> halo_plan() const;
  bool owns_rowintcon = false;
  int rowintcon_local = -1, rowPfix1_local = -1, rowPfix2_local = -1;
  std::vector<long long> gid;   // global unknown index of every stored vector entry (owned rows, then halo ring; -1 unused)
  size_t nrows_local() const { return 6 * (size_t)S.nown; }
  size_t nvec_local() const { return gid.size(); }
  double intcond_col(int col) const { return intcond[(size_t)col]; } // coefficient of stored column `col`

  std::vector<unsigned> diag_info() const; // per row: where its own cell's 6x6 block entries sit (scaling.cu)

  // ---- coupling to an atmosphere / sea-ice model (Coupled Temperature = 1 only) ----
  // module variables of atm.F90 / ice.F90 with the reference's defaults (usrc.F90:62-65, ice.F90:6-10, atmos_coef usrc.F90:1175-1213)
  struct Coupling {
    double qdim = 0.01, nuq = 0.0, eta = 0.0, dqso = 0.0, eo0 = 0.0, albe0 = 0.0, albed = 0.0, nus = 0.0, lvsc = 0.0;
    double zeta = 0.0, a0 = -0.0575, Lf = 3.347e+05, Qvar = 0.0, Q0 = 0.0;
    double Ooa = 0.0, Os = 0.0, QTnd = 0.0, QSnd = 0.0;
    std::vector<double> suno;                    // (m + 2), Fortran j
    std::vector<double> tatm, qatm, albe, msi;   // whole surface, (j-1)*n + (i-1); empty = the zero field
    std::vector<double> msi_lin;                 // the mask the coefficients were last built with (`lin` reads msi when it runs, not at the insert)
    bool have_msi = false;                       // msi_lin has a non-zero entry
    uint64_t msi_version = 0;                    // bumped when msi_lin changes (device copy)
  } cpl;
  enum Field { F_TATM = 0, F_QATM, F_ALBE, F_PATM, F_QSA, F_MSI, F_GSI };
  void insert_field(int which, const double* field);       // inserts.F90:12-166 (takes effect at the next set_* call, as there)
  void set_atmos_parameters(const double* pars18);         // usrc.F90:237-293
  void set_seaice_parameters(const double* pars7);         // usrc.F90:296-333
  void getdeps(double* out7) const;                        // usrc.F90:201-219: Ooa, Os, nus, eta, lvsc, qdim, pQSnd
  DftdmScal dftdm_scalars();                               // probe.F90:414-440 (assigns nus like the Fortran does there)
  // Ocean::getBlock(Atmosphere) / getBlock(SeaIce) (Ocean.C:1538-1746), T rows, compact (see include/iemic_b200.h); return the row count
  int atmos_block(int* rows, int* beg, int* jco, double* co) const;
  int seaice_block(const double* dftdm /* n*m */, int* rows, int* beg, int* jco, double* co) const;
  // probe.F90:381-444, coupled_T part: the fields the formula needs per surface point; the arithmetic runs in a kernel (coupled.cu)
  std::vector<double> ftcell;   // coupled_T = 1: surface T forcing of the owned cells, (j-jo0)*nloc + (i-io0), before the Frc zeroing

  std::vector<double> fprof;    // surface forcing of an ocean cell per latitude: fu | fv | ft | fs, (m + 2) doubles each (forcing_profiles)
  bool forcing_on_device = false, frc_host_valid = true; // set_par leaves `frc` stale when a kernel fills the device copy
  void refresh_host_forcing() { if (!frc_host_valid) build_forcing(true); }
  size_t frc_dirty_from = 0; // first row of `frc` a device copy made before the last change does not have yet
  Tables tables(const double* base) const; // pointers into a (host or device) copy of `tab`
  Tables host_tables() const { Tables T = tables(tab.data()); T.msi = (cfg.coupled_T == 1 && cpl.have_msi) ? cpl.msi_lin.data() : nullptr; return T; }
  inline int lm(int i, int j, int k) const { return landm[(size_t)i + (size_t)(n + 2) * (j + (size_t)(m + 2) * k)]; }

 private:
  void decompose();
  void make_grid();
  void stpnt();
  void build_signatures();
  void build_tables();
  // surface_only: a parameter change touches the top level only (wind stress, surface T/S forcing); the body rows stay zero
  void build_forcing(bool surface_only = false);
  void forcing_profiles();
  void atmos_coef();
  void coupled_changed();
  void build_graph();
  void check_supported() const;
};

// m_scaling::scal (scaling.F90:198-284): db[ii+6*jj] averaged diagonal block -> six row and six column factors (THCM
// convention; Trilinos uses the reciprocals).  false: block singular to working precision (dr, dc untouched).
bool scal(const double* db, double* dr, double* dc);

} // namespace thcm
