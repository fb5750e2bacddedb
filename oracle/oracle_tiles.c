/* oracle_tiles.c -- TEST INFRASTRUCTURE.  See oracle.h. */
#include "oracle.h"
#include <complex.h>
#include <math.h>
#include <omp.h>

typedef double _Complex zc;

#define T double
#define R double
#define CONJ(x) (x)
#define REAL(x) (x)
#define SQRT(x) sqrt(x)
#define FN(x) d_##x
#include "oracle_tiles_body.inc"
#undef T
#undef R
#undef CONJ
#undef REAL
#undef SQRT
#undef FN

#define T zc
#define R double
#define CONJ(x) conj(x)
#define REAL(x) creal(x)
#define SQRT(x) sqrt(x)
#define FN(x) z_##x
#include "oracle_tiles_body.inc"
#undef T
#undef R
#undef CONJ
#undef REAL
#undef SQRT
#undef FN

static zc mkz(const double* p) { return p[0] + p[1] * I; }

int oracle_num_threads(void) { return omp_get_max_threads(); }
void oracle_set_num_threads(int n) { if (n > 0) omp_set_num_threads(n); }

void oracle_tile_dgemm(int opa, int opb, int64_t m, int64_t n, int64_t k, double alpha,
                       const double* A, int64_t lda, const double* B, int64_t ldb,
                       double beta, double* C, int64_t ldc)
{ d_tile_gemm(opa, opb, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc); }

void oracle_tile_zgemm(int opa, int opb, int64_t m, int64_t n, int64_t k, const double* alpha,
                       const double* A, int64_t lda, const double* B, int64_t ldb,
                       const double* beta, double* C, int64_t ldc)
{ z_tile_gemm(opa, opb, m, n, k, mkz(alpha), (const zc*)A, lda, (const zc*)B, ldb, mkz(beta),
              (zc*)C, ldc); }

void oracle_tile_dsyrk(int uplo, int opa, int64_t n, int64_t k, double alpha,
                       const double* A, int64_t lda, double beta, double* C, int64_t ldc)
{ d_tile_herk(uplo, opa, n, k, alpha, A, lda, beta, C, ldc); }

void oracle_tile_zherk(int uplo, int opa, int64_t n, int64_t k, double alpha,
                       const double* A, int64_t lda, double beta, double* C, int64_t ldc)
{ z_tile_herk(uplo, opa, n, k, alpha, (const zc*)A, lda, beta, (zc*)C, ldc); }

void oracle_tile_dtrsm(int side, int uplo, int opa, int diag, int64_t m, int64_t n, double alpha,
                       const double* A, int64_t lda, double* B, int64_t ldb)
{ d_tile_trsm(side, uplo, opa, diag, m, n, alpha, A, lda, B, ldb); }

void oracle_tile_ztrsm(int side, int uplo, int opa, int diag, int64_t m, int64_t n,
                       const double* alpha, const double* A, int64_t lda, double* B, int64_t ldb)
{ z_tile_trsm(side, uplo, opa, diag, m, n, mkz(alpha), (const zc*)A, lda, (zc*)B, ldb); }

int64_t oracle_tile_dpotrf(int uplo, int64_t n, double* A, int64_t lda)
{ return d_tile_potrf(uplo, n, A, lda); }

int64_t oracle_tile_zpotrf(int uplo, int64_t n, double* A, int64_t lda)
{ return z_tile_potrf(uplo, n, (zc*)A, lda); }

void oracle_dgemm(int opa, int opb, int64_t m, int64_t n, int64_t k, int64_t nb, double alpha,
                  const double* A, int64_t lda, const double* B, int64_t ldb, double beta,
                  double* C, int64_t ldc)
{ d_gemm_summa(opa, opb, m, n, k, nb, alpha, A, lda, B, ldb, beta, C, ldc); }

void oracle_zgemm(int opa, int opb, int64_t m, int64_t n, int64_t k, int64_t nb,
                  const double* alpha, const double* A, int64_t lda, const double* B,
                  int64_t ldb, const double* beta, double* C, int64_t ldc)
{ z_gemm_summa(opa, opb, m, n, k, nb, mkz(alpha), (const zc*)A, lda, (const zc*)B, ldb,
               mkz(beta), (zc*)C, ldc); }

int64_t oracle_dpotrf(int uplo, int64_t n, int64_t nb, double* A, int64_t lda)
{ return d_potrf_tiled(uplo, n, nb, A, lda); }

int64_t oracle_zpotrf(int uplo, int64_t n, int64_t nb, double* A, int64_t lda)
{ return z_potrf_tiled(uplo, n, nb, (zc*)A, lda); }

/* ---- SURVEY.md section 8f ---------------------------------------------------------------- */
void oracle_drank2k(int conjugate, int uplo, int opa, int64_t n, int64_t k, int64_t nb, double alpha,
                    const double* A, int64_t lda, const double* B, int64_t ldb, double beta, double* C,
                    int64_t ldc)
{ d_rank2k_tiled(conjugate, uplo, opa, n, k, nb, alpha, A, lda, B, ldb, beta, C, ldc); }

void oracle_zrank2k(int conjugate, int uplo, int opa, int64_t n, int64_t k, int64_t nb,
                    const double* alpha, const double* A, int64_t lda, const double* B, int64_t ldb,
                    const double* beta, double* C, int64_t ldc)
{ z_rank2k_tiled(conjugate, uplo, opa, n, k, nb, mkz(alpha), (const zc*)A, lda, (const zc*)B, ldb,
                 mkz(beta), (zc*)C, ldc); }

void oracle_dhemm(int herm, int side, int uplo, int64_t m, int64_t n, int64_t nb, double alpha,
                  const double* A, int64_t lda, const double* B, int64_t ldb, double beta, double* C,
                  int64_t ldc)
{ d_hemm_tiled(herm, side, uplo, m, n, nb, alpha, A, lda, B, ldb, beta, C, ldc); }

void oracle_zhemm(int herm, int side, int uplo, int64_t m, int64_t n, int64_t nb, const double* alpha,
                  const double* A, int64_t lda, const double* B, int64_t ldb, const double* beta,
                  double* C, int64_t ldc)
{ z_hemm_tiled(herm, side, uplo, m, n, nb, mkz(alpha), (const zc*)A, lda, (const zc*)B, ldb,
               mkz(beta), (zc*)C, ldc); }

int64_t oracle_dgetrf_nopiv(int64_t m, int64_t n, int64_t nb, double* A, int64_t lda)
{ return d_getrf_nopiv_tiled(m, n, nb, A, lda); }

int64_t oracle_zgetrf_nopiv(int64_t m, int64_t n, int64_t nb, double* A, int64_t lda)
{ return z_getrf_nopiv_tiled(m, n, nb, (zc*)A, lda); }
