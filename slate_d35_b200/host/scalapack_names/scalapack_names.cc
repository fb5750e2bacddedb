// scalapack_names.cc -- libslate_b200_scalapack.so: the Fortran-callable ScaLAPACK names of
// the routines the B200 path implements, forwarding to slate_p* (slate_b200_scalapack.h).
//
// Linked in front of (or LD_PRELOADed over) a real ScaLAPACK, an unmodified caller's
// pdgemm_/pdpotrf_/... resolve here and everything else falls through to ScaLAPACK -- the
// interposition scheme of the reference's libslate_scalapack_api
// (scalapack_api/scalapack_gemm.cc:23-38: upper-case, lower-case and underscore spellings).
// Only the FP64 real and complex routines exist on this path; ps*/pc* are left to ScaLAPACK.
#include "slate_b200_scalapack.h"

#define THREE_NAMES(UPPER, lower, PARAMS, ARGS)                      \
    extern "C" void UPPER PARAMS { slate_##lower ARGS; }             \
    extern "C" void lower PARAMS { slate_##lower ARGS; }             \
    extern "C" void lower##_ PARAMS { slate_##lower ARGS; }

#define GEMM_PARAMS(T)                                                                          \
    (const char* ta, const char* tb, int* m, int* n, int* k, T* alpha, T* a, int* ia, int* ja,   \
     int* da, T* b, int* ib, int* jb, int* db, T* beta, T* c, int* ic, int* jc, int* dc)
#define GEMM_ARGS (ta, tb, m, n, k, alpha, a, ia, ja, da, b, ib, jb, db, beta, c, ic, jc, dc)
THREE_NAMES(PDGEMM, pdgemm, GEMM_PARAMS(double), GEMM_ARGS)
THREE_NAMES(PZGEMM, pzgemm, GEMM_PARAMS(void), GEMM_ARGS)

#define POTRF_PARAMS(T) (const char* uplo, int* n, T* a, int* ia, int* ja, int* da, int* info)
#define POTRF_ARGS (uplo, n, a, ia, ja, da, info)
THREE_NAMES(PDPOTRF, pdpotrf, POTRF_PARAMS(double), POTRF_ARGS)
THREE_NAMES(PZPOTRF, pzpotrf, POTRF_PARAMS(void), POTRF_ARGS)

#define SOLVE_PARAMS(T)                                                                          \
    (const char* uplo, int* n, int* nrhs, T* a, int* ia, int* ja, int* da, T* b, int* ib, int* jb, \
     int* db, int* info)
#define SOLVE_ARGS (uplo, n, nrhs, a, ia, ja, da, b, ib, jb, db, info)
THREE_NAMES(PDPOTRS, pdpotrs, SOLVE_PARAMS(double), SOLVE_ARGS)
THREE_NAMES(PZPOTRS, pzpotrs, SOLVE_PARAMS(void), SOLVE_ARGS)
THREE_NAMES(PDPOSV, pdposv, SOLVE_PARAMS(double), SOLVE_ARGS)
THREE_NAMES(PZPOSV, pzposv, SOLVE_PARAMS(void), SOLVE_ARGS)

#define TRSM_PARAMS(T)                                                                           \
    (const char* side, const char* uplo, const char* ta, const char* diag, int* m, int* n,        \
     T* alpha, T* a, int* ia, int* ja, int* da, T* b, int* ib, int* jb, int* db)
#define TRSM_ARGS (side, uplo, ta, diag, m, n, alpha, a, ia, ja, da, b, ib, jb, db)
THREE_NAMES(PDTRSM, pdtrsm, TRSM_PARAMS(double), TRSM_ARGS)
THREE_NAMES(PZTRSM, pztrsm, TRSM_PARAMS(void), TRSM_ARGS)

#define RANKK_PARAMS(S, T)                                                                       \
    (const char* uplo, const char* trans, int* n, int* k, S* alpha, T* a, int* ia, int* ja,       \
     int* da, S* beta, T* c, int* ic, int* jc, int* dc)
#define RANKK_ARGS (uplo, trans, n, k, alpha, a, ia, ja, da, beta, c, ic, jc, dc)
THREE_NAMES(PDSYRK, pdsyrk, RANKK_PARAMS(double, double), RANKK_ARGS)
THREE_NAMES(PZSYRK, pzsyrk, RANKK_PARAMS(void, void), RANKK_ARGS)
THREE_NAMES(PZHERK, pzherk, RANKK_PARAMS(double, void), RANKK_ARGS)

#define HEMM_PARAMS(T)                                                                           \
    (const char* side, const char* uplo, int* m, int* n, T* alpha, T* a, int* ia, int* ja,        \
     int* da, T* b, int* ib, int* jb, int* db, T* beta, T* c, int* ic, int* jc, int* dc)
#define HEMM_ARGS (side, uplo, m, n, alpha, a, ia, ja, da, b, ib, jb, db, beta, c, ic, jc, dc)
THREE_NAMES(PDSYMM, pdsymm, HEMM_PARAMS(double), HEMM_ARGS)
THREE_NAMES(PZSYMM, pzsymm, HEMM_PARAMS(void), HEMM_ARGS)
THREE_NAMES(PZHEMM, pzhemm, HEMM_PARAMS(void), HEMM_ARGS)

#define RANK2K_PARAMS(S, T)                                                                      \
    (const char* uplo, const char* trans, int* n, int* k, T* alpha, T* a, int* ia, int* ja,       \
     int* da, T* b, int* ib, int* jb, int* db, S* beta, T* c, int* ic, int* jc, int* dc)
#define RANK2K_ARGS (uplo, trans, n, k, alpha, a, ia, ja, da, b, ib, jb, db, beta, c, ic, jc, dc)
THREE_NAMES(PDSYR2K, pdsyr2k, RANK2K_PARAMS(double, double), RANK2K_ARGS)
THREE_NAMES(PZSYR2K, pzsyr2k, RANK2K_PARAMS(void, void), RANK2K_ARGS)
THREE_NAMES(PZHER2K, pzher2k, RANK2K_PARAMS(double, void), RANK2K_ARGS)
