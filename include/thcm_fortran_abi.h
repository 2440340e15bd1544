/* thcm_fortran_abi.h -- seam S1: the Fortran symbol surface of the reference's THCM kernel, served by the B200 path.
 *
 * The reference's C++ layer (src/ocean/THCM.C:47-174) binds the Fortran routines below as extern "C" symbols with GNU
 * name mangling (src/utils/my_f2c.H:15-42: `name_`, `__module_MOD_name`), every argument by pointer.
 * libthcm_b200_fortran.so exports ALL 64 of them, so that THCM.C links against it instead of the Fortran objects
 * (tests/test_fortran_link.py compiles that extern block and links it with -Wl,--no-undefined); the routines on the hot path
 * are served by the device path, the others are listed at the end of this header with what they do: same argument lists, same array conventions (interleaved unknowns 6*cell+var, cells i fastest,
 * land mask (n+2)(m+2)(l+2) with its border ring, CSR arrays 1-based and owned by the caller).  It is a host-buffer
 * adaptor over the native device API (iemic_b200.h): one process-global model, like the Fortran module state.
 *
 * Differences, all loud: the forcing is the idealised one (iza=2, ite=1, its=1: the wind / temperature / salinity
 * arrays passed to init_ are ignored and anything else is refused), mask files are not read (the mask comes through
 * init_ / set_landmask_), and instead of STOP an error prints `iemic_b200_last_error()` to stderr and calls abort()
 * after the optional callback thcm_throw_error_ (THCM.C:2741) -- resolved lazily, the library has no link-time
 * dependency on it.
 *
 *   symbol here                         reference definition
 *   __m_global_MOD_initialize           src/ocean/global.F90:66-170      (domain constants, forcing switches)
 *   __m_global_MOD_finalize             src/ocean/global.F90
 *   init_                               src/ocean/usrc.F90:6-139         (grid, stpnt, forcing, lin of this subdomain)
 *   finalize_                           src/ocean/usrc.F90:143-160
 *   setparcs_ / getparcs_               src/ocean/usrc.F90:163-196
 *   setsres_                            src/ocean/usrc.F90:417-429
 *   set_landmask_                       src/ocean/usrc.F90:336-401
 *   __m_mix_MOD_set_vmix_fix            src/ocean/mix_imp.f (flag only; the fixed-mixing mode vmix=2 is out of scope)
 *   __m_mat_MOD_get_array_sizes         src/ocean/mat.F90:56-70          (nrows = 6nml, nnz = nrows*(6*27+1))
 *   __m_mat_MOD_set_pointers            src/ocean/mat.F90:73-100         (caller-owned begA jcoA coA coB begF jcoF coF)
 *   rhs_                                src/ocean/usrc.F90:506-586       (B; THCM::evaluate flips the sign afterwards)
 *   matrix_                             src/ocean/usrc.F90:432-504       (coA/jcoA/begA by fillcolA's rule, coB)
 *   fillcolb_                           src/ocean/assemble.F90:18-54
 *   __m_thcm_utils_MOD_intcond_scaling  src/ocean/thcm_utils.F90:285-323
 *   __m_scaling_MOD_average_block       src/ocean/scaling.F90:29-62      (db(6,6) column-major; called at THCM.C:1709)
 *   __m_scaling_MOD_compute             src/ocean/scaling.F90:69-99      (THCM-convention factors; called at THCM.C:1721)
 */
#ifndef THCM_FORTRAN_ABI_H
#define THCM_FORTRAN_ABI_H

#ifdef __cplusplus
extern "C" {
#endif

void __m_global_MOD_initialize(int* N, int* M, int* L, double* Xmin, double* Xmax, double* Ymin, double* Ymax,
                               double* hdim, double* qz, int* periodic, int* itopo, int* flat, int* rd_mask,
                               int* TRES, int* SRES, int* iza, int* ite, int* its, int* rd_spertm,
                               int* coupled_T, int* coupled_S, int* forcing_type,
                               const char* maskfile, const char* spertmaskfile,
                               const char* windfile, const char* sstfile, const char* sssfile);
void __m_global_MOD_finalize(void);

void init_(int* n, int* m, int* l, int* nmlglob, double* xmin, double* xmax, double* ymin, double* ymax,
           double* alphaT, double* alphaS, int* ih, int* vmix, int* tap, int* rho_mixing, int* coriolis_on,
           int* periodic, int* landm, double* taux, double* tauy, double* tatm, double* emip, double* spert);
void finalize_(void);

void setparcs_(int* param, double* value);
void getparcs_(int* param, double* value);
void setsres_(int* sres);
void set_landmask_(int* landm, int* periodic, int* reinit);
void __m_mix_MOD_set_vmix_fix(int* vmix_fix);

void __m_mat_MOD_get_array_sizes(int* nrows, int* nnz);
void __m_mat_MOD_set_pointers(int* nrows, int* nnz, int* begA, int* jcoA, double* coA, double* coB,
                              int* begF, int* jcoF, double* coF);

void rhs_(double* un, double* b);
void matrix_(double* un);
void fillcolb_(void);

void __m_thcm_utils_MOD_intcond_scaling(double* values, int* indices, int* len);
void __m_scaling_MOD_average_block(double* db);
void __m_scaling_MOD_compute(double* db, double* row_scaling, double* col_scaling);

/* ---- the rest of the surface (src/ocean/THCM.C:47-174) ----
 * served: */
void writeparams_(void);                                  /* usrc.F90:199-234: prints par(1..30) */
void __m_global_MOD_set_landm(int* landm);               /* global.F90: keeps the GLOBAL (N+2)(M+2)(L+2) mask the host hands over */
void __m_global_MOD_get_landm(int* landm);               /* ... and returns it (mask FILES are not read: loud error if none was set) */
void __m_global_MOD_get_current_landm(int* landm);
void __m_thcm_utils_MOD_get_landm(int* landm);           /* the sub-domain mask init_ / set_landmask_ received */
void __m_global_MOD_get_windfield(double* taux, double* tauy); /* idealised forcing has no data fields: N x M zeros */
void __m_global_MOD_get_temforcing(double* tatm);
void __m_global_MOD_get_salforcing(double* emip);
void __m_global_MOD_get_spert(double* spert);
/* refused loudly (message, thcm_throw_error_ when the host program has it, abort): data-file / time-dependent forcing and the
 * coupled model's inserts and probes (SURVEY.md rows out of scope / N2) */
void stochastic_forcing_(void);
void write_levitus_(const char* file);
void __m_global_MOD_set_maskfile(const char* maskfile);
void __m_global_MOD_get_internal_temforcing(double* temp);
void __m_global_MOD_get_internal_salforcing(double* salt);
void __m_global_MOD_get_monthly_forcing(double* tatm, double* emip, double* taux, double* tauy, int* month);
void __m_global_MOD_get_monthly_internal_forcing(double* temp, double* salt, int* month);
void __m_global_MOD_get_land_temp(double* land);
void __m_monthly_MOD_set_forcing(double* tatm, double* emip, double* taux, double* tauy, int* month);
void __m_monthly_MOD_set_internal_forcing(double* temp, double* salt, int* month);
void __m_monthly_MOD_update_forcing(double* t, double* gammaw, double* gammat, double* gammas);
void __m_monthly_MOD_update_internal_forcing(double* t, double* gammat, double* gammas);
void __m_usr_MOD_set_internal_forcing(double* temp, double* salt);
void __m_thcm_utils_MOD_loadbal_weights(double* weights, double* a, double* b, double* c);
/* Ocean.C's own Fortran bindings for the coupled model (Ocean.C:1469, 1490, 1545; usrc.F90:201-333), served for Coupled Temperature = 1 */
void set_atmos_parameters_(void* pars /* Atmosphere::CommPars */); void set_seaice_parameters_(void* pars /* SeaIce::CommPars */);
void getdeps_(double* Ooa, double* Os, double* nus, double* eta, double* lvsc, double* qdim, double* pQSnd);
void __m_inserts_MOD_insert_atmosphere_t(double* f); void __m_inserts_MOD_insert_atmosphere_q(double* f);
void __m_inserts_MOD_insert_atmosphere_a(double* f); void __m_inserts_MOD_insert_atmosphere_p(double* f);
void __m_inserts_MOD_insert_seaice_q(double* f); void __m_inserts_MOD_insert_seaice_m(double* f); void __m_inserts_MOD_insert_seaice_g(double* f);
void __m_inserts_MOD_insert_emip(double* f); void __m_inserts_MOD_insert_adapted_emip(double* f);
void __m_inserts_MOD_insert_emip_pert(double* f); void __m_inserts_MOD_insert_tatm(double* f);
void __m_integrals_MOD_salt_advection(double* un, double* check); void __m_integrals_MOD_salt_diffusion(double* un, double* check);
void __m_probe_MOD_get_emip(double* f); void __m_probe_MOD_get_suno(double* f); void __m_probe_MOD_get_adapted_emip(double* f);
void __m_probe_MOD_get_emip_pert(double* f); void __m_probe_MOD_get_atmosphere_t(double* f); void __m_probe_MOD_get_atmosphere_q(double* f);
void __m_probe_MOD_get_atmosphere_p(double* f); void __m_probe_MOD_compute_evap(double* oceanE, double* x);
void __m_probe_MOD_get_derivatives(double* sol, double* dftdm, double* dfsdq, double* dfsdm, double* dfsdg);
void __m_probe_MOD_get_salflux(double* sol, double* salflux, double* scorr, double* qsoaflux, double* qsosflux);
void __m_probe_MOD_get_temflux(double* sol, double* totflux, double* swflux, double* shflux, double* lhflux, double* siflux, double* simask);

#ifdef __cplusplus
}
#endif
#endif
