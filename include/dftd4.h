/*
 * dftd4.h -- C-API of libdftd4 as served by the B200 (sm_100a) engine.
 *
 * Binary-compatible with the reference library's public header (reference: include/dftd4.h:37-230, implemented
 * there by src/dftd4/api.f90:98-891): same 22 symbols, same argument order and meaning, same opaque handles and
 * error-handle convention, so programs and the CFFI Python package written against the reference link against this
 * library unchanged.  Each prototype below names the reference routine it replaces.
 *
 * Conventions (reference dftd4.h:94, api.f90:200-246):
 *   - lengths in Bohr, energies in Hartree;
 *   - positions[natoms][3], lattice[3][3] with lattice[k][:] the k-th lattice vector, gradient[natoms][3],
 *     sigma[3][3], pair matrices [natoms][natoms];
 *   - the first argument of every call is the error handle; a NULL error handle makes the call a silent no-op;
 *   - objects are created by the library and released with dftd4_delete_*(&handle), which resets the handle to NULL.
 *
 * Define DFTD4_CFFI before inclusion to skip <stdbool.h> (used when the header is fed to cffi's cdef parser).
 */
#ifndef DFTD4_B200_DFTD4_H
#define DFTD4_B200_DFTD4_H

#ifdef __cplusplus
#define DFTD4_API_ENTRY extern "C"
#else
#define DFTD4_API_ENTRY extern
#ifndef DFTD4_CFFI
#include <stdbool.h>
#endif
#endif
#define DFTD4_API_CALL

typedef struct _dftd4_error* dftd4_error;         /* error handle                     */
typedef struct _dftd4_structure* dftd4_structure; /* molecular / periodic structure   */
typedef struct _dftd4_model* dftd4_model;         /* dispersion model (+ its cutoffs) */
typedef struct _dftd4_param* dftd4_param;         /* damping parameters               */

#ifndef DFTD4_CFFI
/* type-generic delete, as in the reference header (dftd4.h:53-58) */
#define dftd4_delete(ptr)                     \
  _Generic((ptr), dftd4_error                 \
           : dftd4_delete_error, dftd4_structure \
           : dftd4_delete_structure, dftd4_model \
           : dftd4_delete_model, dftd4_param  \
           : dftd4_delete_param)(&ptr)
#endif

/* ---- library ------------------------------------------------------------------------------------------------ */
/* api.f90:98-107: 10000*major + 100*minor + patch */
DFTD4_API_ENTRY int DFTD4_API_CALL dftd4_get_version(void);

/* ---- error handle (api.f90:111-196) --------------------------------------------------------------------------- */
DFTD4_API_ENTRY dftd4_error DFTD4_API_CALL dftd4_new_error(void);
/* 0: no error, 1: error set, 2: handle is NULL */
DFTD4_API_ENTRY int DFTD4_API_CALL dftd4_check_error(dftd4_error error);
/* copies the NUL-terminated message into buffer; *buffersize limits the copy when given */
DFTD4_API_ENTRY void DFTD4_API_CALL dftd4_get_error(dftd4_error error, char* buffer, const int* buffersize);
DFTD4_API_ENTRY void DFTD4_API_CALL dftd4_delete_error(dftd4_error* error);

/* ---- structure (api.f90:200-307) ------------------------------------------------------------------------------ */
/* charge, lattice and periodic may be NULL; positions are wrapped into the central cell and checked for
 * coinciding atoms ("Too close interatomic distances found") */
DFTD4_API_ENTRY dftd4_structure DFTD4_API_CALL dftd4_new_structure(dftd4_error error, const int natoms,
                                                                   const int* numbers, const double* positions,
                                                                   const double* charge, const double* lattice,
                                                                   const bool* periodic);
DFTD4_API_ENTRY void DFTD4_API_CALL dftd4_delete_structure(dftd4_structure* mol);
/* lattice may be NULL */
DFTD4_API_ENTRY void DFTD4_API_CALL dftd4_update_structure(dftd4_error error, dftd4_structure mol,
                                                           const double* positions, const double* lattice);

/* ---- dispersion model (api.f90:311-569) ----------------------------------------------------------------------- */
DFTD4_API_ENTRY dftd4_model DFTD4_API_CALL dftd4_new_d4_model(dftd4_error error, dftd4_structure mol);
DFTD4_API_ENTRY dftd4_model DFTD4_API_CALL dftd4_new_d4s_model(dftd4_error error, dftd4_structure mol);
DFTD4_API_ENTRY dftd4_model DFTD4_API_CALL dftd4_custom_d4_model(dftd4_error error, dftd4_structure mol, double ga,
                                                                 double gc, double wf);
DFTD4_API_ENTRY dftd4_model DFTD4_API_CALL dftd4_custom_d4s_model(dftd4_error error, dftd4_structure mol, double ga,
                                                                  double gc);
DFTD4_API_ENTRY void DFTD4_API_CALL dftd4_delete_model(dftd4_model* disp);
DFTD4_API_ENTRY void DFTD4_API_CALL dftd4_set_model_realspace_cutoff(dftd4_error error, dftd4_model disp, double disp2,
                                                                     double disp3, double cn);
DFTD4_API_ENTRY void DFTD4_API_CALL dftd4_set_model_realspace_cutoff_smooth(dftd4_error error, dftd4_model disp,
                                                                            double disp2, double disp3, double cn,
                                                                            double width2, double width3);

/* ---- damping parameters (api.f90:573-640) --------------------------------------------------------------------- */
DFTD4_API_ENTRY dftd4_param DFTD4_API_CALL dftd4_new_rational_damping(dftd4_error error, double s6, double s8,
                                                                      double s9, double a1, double a2, double alp);
/* method: functional name (case-insensitive, text after '/' ignored); mbd=false switches the ATM term off */
DFTD4_API_ENTRY dftd4_param DFTD4_API_CALL dftd4_load_rational_damping(dftd4_error error, char* method, bool mbd);
DFTD4_API_ENTRY void DFTD4_API_CALL dftd4_delete_param(dftd4_param* param);

/* ---- evaluations (api.f90:644-891) ---------------------------------------------------------------------------- */
/* cn[n], charges[n], c6[n*n], alpha[n] */
DFTD4_API_ENTRY void DFTD4_API_CALL dftd4_get_properties(dftd4_error error, dftd4_structure mol, dftd4_model disp,
                                                         double* cn, double* charges, double* c6, double* alpha);
/* THE HOT PATH (src/dftd4/disp.f90:40-134).  gradient[n][3] and sigma[3][3] may be NULL. */
DFTD4_API_ENTRY void DFTD4_API_CALL dftd4_get_dispersion(dftd4_error error, dftd4_structure mol, dftd4_model disp,
                                                         dftd4_param param, double* energy, double* gradient,
                                                         double* sigma);
/* hess[n][3][n][3] by central differences of the analytic gradient (src/dftd4/numdiff.f90:35-79) */
DFTD4_API_ENTRY void DFTD4_API_CALL dftd4_get_numerical_hessian(dftd4_error error, dftd4_structure mol, dftd4_model disp,
                                                                dftd4_param param, double* hess);
/* pair_energy2[n][n], pair_energy3[n][n] */
DFTD4_API_ENTRY void DFTD4_API_CALL dftd4_get_pairwise_dispersion(dftd4_error error, dftd4_structure mol,
                                                                  dftd4_model disp, dftd4_param param,
                                                                  double* pair_energy2, double* pair_energy3);

#endif
