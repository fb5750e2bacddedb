/*
 * oracle.h -- TEST INFRASTRUCTURE, not product code.
 *
 * Plain-C CPU restatement of the reference algorithms on the hot path
 * (SURVEY.md section 8a): matgen's counter-based generator, the per-tile
 * Level-3 operations of include/slate/Tile_blas.hh, and the two tiled drivers
 * (gemmC SUMMA, right-looking potrf) of src/gemmC.cc and src/potrf.cc.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may call this; the product path never does.
 *
 * PARITY PINNING: the reference holds no golden vectors for this path
 * (every reference test is tolerance based).  The oracle
 * is pinned instead against outputs of the reference itself, run in the
 * authoring container through oracle/_ref/libslate_ref.so (oracle/Makefile
 * target `ref`): tests/test_oracle.py compares the two directly when _ref is
 * present, and tests/golden/*.npz holds outputs generated from _ref by
 * tests/golden/make_golden.py for boxes where _ref cannot be rebuilt.  The
 * generator (matgen) is integer arithmetic and is pinned bit-exactly; the
 * floating-point routines are pinned to the tolerances the reference's own
 * tests use (3*sqrt(k)*eps elementwise for one block update,
 * unit_test/test_internal_blas.cc:446-447).
 *
 * All matrices are column-major doubles; complex matrices are interleaved
 * (re, im) pairs and leading dimensions count complex elements.
 */
#ifndef ORACLE_H
#define ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* matgen/random.cc:53-75, 93-159: Philox-2x64-10 keyed by seed, counter (i, j).
 * dist 1 = uniform [0,1), 2 = uniform (-1,1). */
void oracle_matgen_d(int dist, int64_t seed, int64_t m, int64_t n, int64_t ioff, int64_t joff,
                     double* A, int64_t lda);
void oracle_matgen_z(int dist, int64_t seed, int64_t m, int64_t n, int64_t ioff, int64_t joff,
                     double* A, int64_t lda);

/* include/slate/Tile_blas.hh:30-96 (tile::gemm):  C = alpha op(A) op(B) + beta C. */
void oracle_tile_dgemm(int opa, int opb, int64_t m, int64_t n, int64_t k, double alpha,
                       const double* A, int64_t lda, const double* B, int64_t ldb,
                       double beta, double* C, int64_t ldc);
void oracle_tile_zgemm(int opa, int opb, int64_t m, int64_t n, int64_t k, const double* alpha,
                       const double* A, int64_t lda, const double* B, int64_t ldb,
                       const double* beta, double* C, int64_t ldc);

/* include/slate/Tile_blas.hh:273-290 (tile::herk): C = alpha op(A) op(A)^H + beta C,
 * `uplo` triangle only; complex: imaginary part of the diagonal set to zero. */
void oracle_tile_dsyrk(int uplo, int opa, int64_t n, int64_t k, double alpha,
                       const double* A, int64_t lda, double beta, double* C, int64_t ldc);
void oracle_tile_zherk(int uplo, int opa, int64_t n, int64_t k, double alpha,
                       const double* A, int64_t lda, double beta, double* C, int64_t ldc);

/* include/slate/Tile_blas.hh:682-728 (tile::trsm): B = alpha op(A)^-1 B (side L) or
 * alpha B op(A)^-1 (side R); A triangular `uplo`, diag 'N' or 'U'. */
void oracle_tile_dtrsm(int side, int uplo, int opa, int diag, int64_t m, int64_t n, double alpha,
                       const double* A, int64_t lda, double* B, int64_t ldb);
void oracle_tile_ztrsm(int side, int uplo, int opa, int diag, int64_t m, int64_t n,
                       const double* alpha, const double* A, int64_t lda, double* B, int64_t ldb);

/* src/internal/Tile_lapack.hh:266-272 (tile::potrf -> lapack::potrf): Cholesky of one tile;
 * returns 0 or the 1-based index of the first non-positive pivot. */
int64_t oracle_tile_dpotrf(int uplo, int64_t n, double* A, int64_t lda);
int64_t oracle_tile_zpotrf(int uplo, int64_t n, double* A, int64_t lda);

/* src/gemmC.cc:113-171: SUMMA over block columns k of A / block rows of B with tile
 * size nb (ragged last tile); beta applied at k == 0 only. */
void oracle_dgemm(int opa, int opb, int64_t m, int64_t n, int64_t k, int64_t nb, double alpha,
                  const double* A, int64_t lda, const double* B, int64_t ldb, double beta,
                  double* C, int64_t ldc);
void oracle_zgemm(int opa, int opb, int64_t m, int64_t n, int64_t k, int64_t nb,
                  const double* alpha, const double* A, int64_t lda, const double* B,
                  int64_t ldb, const double* beta, double* C, int64_t ldc);

/* src/potrf.cc:88-194: right-looking tiled Cholesky (lookahead only reorders independent
 * updates, so the result does not depend on it).  Returns info as src/potrf.cc:104-105. */
int64_t oracle_dpotrf(int uplo, int64_t n, int64_t nb, double* A, int64_t lda);
int64_t oracle_zpotrf(int uplo, int64_t n, int64_t nb, double* A, int64_t lda);

/* SURVEY.md section 8f.  src/her2k.cc:25-190 / src/syr2k.cc (conjugate = 1 / 0): rank-2k
 * update of the uplo triangle by block columns of nb.  For her2k beta must be real. */
void oracle_drank2k(int conjugate, int uplo, int opa, int64_t n, int64_t k, int64_t nb, double alpha,
                    const double* A, int64_t lda, const double* B, int64_t ldb, double beta, double* C,
                    int64_t ldc);
void oracle_zrank2k(int conjugate, int uplo, int opa, int64_t n, int64_t k, int64_t nb,
                    const double* alpha, const double* A, int64_t lda, const double* B, int64_t ldb,
                    const double* beta, double* C, int64_t ldc);

/* src/hemmC.cc:28-420 / src/symm.cc (herm = 1 / 0): C = alpha A B + beta C (side L) or
 * alpha B A + beta C (side R); only the uplo triangle of A is read. */
void oracle_dhemm(int herm, int side, int uplo, int64_t m, int64_t n, int64_t nb, double alpha,
                  const double* A, int64_t lda, const double* B, int64_t ldb, double beta, double* C,
                  int64_t ldc);
void oracle_zhemm(int herm, int side, int uplo, int64_t m, int64_t n, int64_t nb, const double* alpha,
                  const double* A, int64_t lda, const double* B, int64_t ldb, const double* beta,
                  double* C, int64_t ldc);

/* src/getrf_nopiv.cc:26-260 with src/internal/Tile_getrf_nopiv.hh on the diagonal tiles:
 * A = L U in place, no pivoting; returns info (first exactly zero pivot, 1-based) */
int64_t oracle_dgetrf_nopiv(int64_t m, int64_t n, int64_t nb, double* A, int64_t lda);
int64_t oracle_zgetrf_nopiv(int64_t m, int64_t n, int64_t nb, double* A, int64_t lda);

int oracle_num_threads(void);
void oracle_set_num_threads(int n);

#ifdef __cplusplus
}
#endif
#endif
