/* slate_b200_scalapack.h -- ScaLAPACK-calling-convention entry points of the B200 tile-update
 * path (N ranks x 1 GPU, user-owned 2D block-cyclic host arrays).
 *
 * Each function replaces the one of the same name in the reference's scalapack_api/
 * directory; arguments, 1-based (ia, ja) offsets and descriptor layout are ScaLAPACK's:
 *
 *   slate_pdgemm / slate_pzgemm   scalapack_api/scalapack_gemm.cc:108-125 (generic body :132-183)
 *   slate_pdpotrf / slate_pzpotrf scalapack_api/scalapack_potrf.cc:94-124
 *   slate_pdpotrs / slate_pzpotrs scalapack_api/scalapack_potrs.cc
 *   slate_pdposv / slate_pzposv   scalapack_api/scalapack_posv.cc:94-130
 *   slate_pdtrsm / slate_pztrsm   scalapack_api/scalapack_trsm.cc
 *   slate_pdsyrk / slate_pzsyrk   scalapack_api/scalapack_syrk.cc
 *   slate_pzherk                  scalapack_api/scalapack_herk.cc
 *   slate_pdsymm / slate_pzsymm   scalapack_api/scalapack_symm.cc
 *   slate_pzhemm                  scalapack_api/scalapack_hemm.cc
 *   slate_pdsyr2k / slate_pzsyr2k scalapack_api/scalapack_syr2k.cc
 *   slate_pzher2k                 scalapack_api/scalapack_her2k.cc
 *
 * Behaviour kept from the reference: arrays are wrapped in place with fromScaLAPACK (no
 * redistribution), sub-matrix offsets must fall on tile boundaries, SLATE_SCALAPACK_LOOKAHEAD
 * and SLATE_SCALAPACK_VERBOSE are read from the environment.  Differences:
 *   - the target is always Devices (this library has no CPU path); SLATE_SCALAPACK_TARGET set
 *     to a host target raises slate::NotImplemented instead of silently computing on the CPU;
 *   - *info of potrf / posv is the real LAPACK info (the reference returns 0, "todo");
 *   - RSRC / CSRC of the descriptors must be 0 (the reference ignores them);
 *   - the process grid comes from BLACS when the process has one linked
 *     (Cblacs_gridinfo, Cblacs_pinfo, Cblacs_get, Cblacs_pcoord are weak references), else
 *     from slate_scalapack_grid_set() below.
 * slate_b200 must have been initialised for this rank (sb_initialize / slate::initialize:
 * device, rank, size, NCCL id) before the first call.
 *
 * libslate_b200_scalapack.so additionally exports the Fortran names (pdgemm_, pdgemm, PDGEMM,
 * ...) forwarding to these, for link-order / LD_PRELOAD interposition in front of a real
 * ScaLAPACK, as the reference's libslate_scalapack_api does.
 */
#ifndef SLATE_B200_SCALAPACK_H
#define SLATE_B200_SCALAPACK_H

#ifdef __cplusplus
extern "C" {
#endif

/* Without BLACS: declare the grid that descriptor context `ctxt` refers to.
 * order 'C' (ranks column-major over the grid, BLACS "Col") or 'R'.  Returns 0, or -1 on a bad
 * argument.  No reference counterpart (there BLACS answers). */
int slate_scalapack_grid_set(int ctxt, char order, int nprow, int npcol);

/* 0 if the last slate_p* call finished, else non-zero with the text in
 * slate_scalapack_last_error().  The Fortran convention has no error return, so a failed call
 * also prints the text to stderr. */
int slate_scalapack_status(void);
const char* slate_scalapack_last_error(void);

void slate_pdgemm(const char* transa, const char* transb, int* m, int* n, int* k, double* alpha,
                  double* a, int* ia, int* ja, int* desca, double* b, int* ib, int* jb, int* descb,
                  double* beta, double* c, int* ic, int* jc, int* descc);
void slate_pzgemm(const char* transa, const char* transb, int* m, int* n, int* k, void* alpha,
                  void* a, int* ia, int* ja, int* desca, void* b, int* ib, int* jb, int* descb,
                  void* beta, void* c, int* ic, int* jc, int* descc);

void slate_pdpotrf(const char* uplo, int* n, double* a, int* ia, int* ja, int* desca, int* info);
void slate_pzpotrf(const char* uplo, int* n, void* a, int* ia, int* ja, int* desca, int* info);

void slate_pdpotrs(const char* uplo, int* n, int* nrhs, double* a, int* ia, int* ja, int* desca,
                   double* b, int* ib, int* jb, int* descb, int* info);
void slate_pzpotrs(const char* uplo, int* n, int* nrhs, void* a, int* ia, int* ja, int* desca,
                   void* b, int* ib, int* jb, int* descb, int* info);

void slate_pdposv(const char* uplo, int* n, int* nrhs, double* a, int* ia, int* ja, int* desca,
                  double* b, int* ib, int* jb, int* descb, int* info);
void slate_pzposv(const char* uplo, int* n, int* nrhs, void* a, int* ia, int* ja, int* desca,
                  void* b, int* ib, int* jb, int* descb, int* info);

void slate_pdtrsm(const char* side, const char* uplo, const char* transa, const char* diag, int* m,
                  int* n, double* alpha, double* a, int* ia, int* ja, int* desca, double* b,
                  int* ib, int* jb, int* descb);
void slate_pztrsm(const char* side, const char* uplo, const char* transa, const char* diag, int* m,
                  int* n, void* alpha, void* a, int* ia, int* ja, int* desca, void* b, int* ib,
                  int* jb, int* descb);

void slate_pdsyrk(const char* uplo, const char* trans, int* n, int* k, double* alpha, double* a,
                  int* ia, int* ja, int* desca, double* beta, double* c, int* ic, int* jc,
                  int* descc);
void slate_pzsyrk(const char* uplo, const char* trans, int* n, int* k, void* alpha, void* a,
                  int* ia, int* ja, int* desca, void* beta, void* c, int* ic, int* jc, int* descc);
void slate_pzherk(const char* uplo, const char* trans, int* n, int* k, double* alpha, void* a,
                  int* ia, int* ja, int* desca, double* beta, void* c, int* ic, int* jc,
                  int* descc);

void slate_pdsymm(const char* side, const char* uplo, int* m, int* n, double* alpha, double* a,
                  int* ia, int* ja, int* desca, double* b, int* ib, int* jb, int* descb, double* beta,
                  double* c, int* ic, int* jc, int* descc);
void slate_pzsymm(const char* side, const char* uplo, int* m, int* n, void* alpha, void* a, int* ia,
                  int* ja, int* desca, void* b, int* ib, int* jb, int* descb, void* beta, void* c,
                  int* ic, int* jc, int* descc);
void slate_pzhemm(const char* side, const char* uplo, int* m, int* n, void* alpha, void* a, int* ia,
                  int* ja, int* desca, void* b, int* ib, int* jb, int* descb, void* beta, void* c,
                  int* ic, int* jc, int* descc);

void slate_pdsyr2k(const char* uplo, const char* trans, int* n, int* k, double* alpha, double* a,
                   int* ia, int* ja, int* desca, double* b, int* ib, int* jb, int* descb,
                   double* beta, double* c, int* ic, int* jc, int* descc);
void slate_pzsyr2k(const char* uplo, const char* trans, int* n, int* k, void* alpha, void* a, int* ia,
                   int* ja, int* desca, void* b, int* ib, int* jb, int* descb, void* beta, void* c,
                   int* ic, int* jc, int* descc);
void slate_pzher2k(const char* uplo, const char* trans, int* n, int* k, void* alpha, void* a, int* ia,
                   int* ja, int* desca, void* b, int* ib, int* jb, int* descb, double* beta, void* c,
                   int* ic, int* jc, int* descc);

#ifdef __cplusplus
}
#endif
#endif
