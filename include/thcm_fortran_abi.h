/* thcm_fortran_abi.h -- seam S1: the Fortran symbol surface of the reference's THCM kernel, served by the B200 path.
 *
 * The reference's C++ layer (src/ocean/THCM.C:47-174) binds the Fortran routines below as extern "C" symbols with GNU
 * name mangling (src/utils/my_f2c.H:15-42: `name_`, `__module_MOD_name`), every argument by pointer.
 * libthcm_b200_fortran.so exports the subset that lies on the hot path, so that THCM.C can link against it instead of the
 * Fortran objects: same argument lists, same array conventions (interleaved unknowns 6*cell+var, cells i fastest,
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

#ifdef __cplusplus
}
#endif
#endif
