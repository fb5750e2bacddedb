// slate/slate.hh -- public drivers of the B200 engine: slate::gemm and slate::potrf with
// the reference's signatures (include/slate/slate.hh:189-194, 764-767) and the
// simplified-API aliases (include/slate/simplified_api.hh:30-37, 560-570).
#ifndef SLATE_B200_SLATE_HH
#define SLATE_B200_SLATE_HH

#include "slate/internal.hh"

namespace slate {

/// C = alpha A B + beta C (src/gemm.cc:66-88): gemmA when B has a single block column,
/// gemmC otherwise, or as Option::MethodGemm says.
/// Options: Target (only Target::Devices is built), Lookahead (default 1), MethodGemm.
template <typename scalar_t>
void gemm(scalar_t alpha, Matrix<scalar_t>& A, Matrix<scalar_t>& B, scalar_t beta,
          Matrix<scalar_t>& C, Options const& opts = Options());

/// SUMMA with C stationary: panels of A and B are broadcast (src/gemmC.cc:29-191).
template <typename scalar_t>
void gemmC(scalar_t alpha, Matrix<scalar_t>& A, Matrix<scalar_t>& B, scalar_t beta,
           Matrix<scalar_t>& C, Options const& opts = Options());

/// A stationary: B is broadcast to where A is, partial C tiles are reduced
/// (src/gemmA.cc:32-210).  For A much larger than B and C.
template <typename scalar_t>
void gemmA(scalar_t alpha, Matrix<scalar_t>& A, Matrix<scalar_t>& B, scalar_t beta,
           Matrix<scalar_t>& C, Options const& opts = Options());

/// Cholesky factorisation A = L L^H (lower) or U^H U (upper), in place
/// (src/potrf.cc:23-210, 262-281).  Returns info: 0, or i if the leading minor of order i
/// is not positive definite.  Options: Target, Lookahead (default 1), HoldLocalWorkspace.
template <typename scalar_t>
int64_t potrf(HermitianMatrix<scalar_t>& A, Options const& opts = Options());

/// C = alpha A A^H + beta C, C Hermitian (src/herk.cc:15-160; include/slate/slate.hh:340-345).
template <typename scalar_t>
void herk(real_type<scalar_t> alpha, Matrix<scalar_t>& A, real_type<scalar_t> beta,
          HermitianMatrix<scalar_t>& C, Options const& opts = Options());

/// C = alpha A A^T + beta C, C symmetric (src/syrk.cc:15-163).
template <typename scalar_t>
void syrk(scalar_t alpha, Matrix<scalar_t>& A, scalar_t beta, SymmetricMatrix<scalar_t>& C,
          Options const& opts = Options());

/// C = alpha A B^H + conj(alpha) B A^H + beta C, C Hermitian (src/her2k.cc:25-190).
template <typename scalar_t>
void her2k(scalar_t alpha, Matrix<scalar_t>& A, Matrix<scalar_t>& B, real_type<scalar_t> beta,
           HermitianMatrix<scalar_t>& C, Options const& opts = Options());

/// C = alpha A B^T + alpha B A^T + beta C, C symmetric (src/syr2k.cc).
template <typename scalar_t>
void syr2k(scalar_t alpha, Matrix<scalar_t>& A, Matrix<scalar_t>& B, scalar_t beta,
           SymmetricMatrix<scalar_t>& C, Options const& opts = Options());

/// C = alpha A B + beta C (side L) or alpha B A + beta C (side R), A Hermitian / symmetric with
/// one triangle stored (src/hemm.cc, src/hemmC.cc:28-420, src/symm.cc).
template <typename scalar_t>
void hemm(Side side, scalar_t alpha, HermitianMatrix<scalar_t>& A, Matrix<scalar_t>& B,
          scalar_t beta, Matrix<scalar_t>& C, Options const& opts = Options());
template <typename scalar_t>
void symm(Side side, scalar_t alpha, SymmetricMatrix<scalar_t>& A, Matrix<scalar_t>& B,
          scalar_t beta, Matrix<scalar_t>& C, Options const& opts = Options());

/// LU factorisation without pivoting, A = L U in place, L unit lower (src/getrf_nopiv.cc:26-260).
/// Returns info: 0, or j if U(j, j) is exactly zero.  Options: Target, Lookahead.
template <typename scalar_t>
int64_t getrf_nopiv(Matrix<scalar_t>& A, Options const& opts = Options());

/// Triangular solve with a distributed triangular matrix: op(A) X = alpha B (side L) or
/// X op(A) = alpha B (side R); src/trsm.cc, src/trsmB.cc, src/work/work_trsm.cc.
template <typename scalar_t>
void trsm(Side side, scalar_t alpha, TriangularMatrix<scalar_t>& A, Matrix<scalar_t>& B,
          Options const& opts = Options());

/// Solve with the Cholesky factor (src/potrs.cc:41-62) / factor and solve (src/posv.cc).
template <typename scalar_t>
void potrs(HermitianMatrix<scalar_t>& A, Matrix<scalar_t>& B, Options const& opts = Options());
template <typename scalar_t>
int64_t posv(HermitianMatrix<scalar_t>& A, Matrix<scalar_t>& B, Options const& opts = Options());

// ---- simplified API -----------------------------------------------------------------------
template <typename scalar_t>
void multiply(scalar_t alpha, Matrix<scalar_t>& A, Matrix<scalar_t>& B, scalar_t beta,
              Matrix<scalar_t>& C, Options const& opts = Options())
{
    gemm(alpha, A, B, beta, C, opts);
}

template <typename scalar_t>
int64_t chol_factor(HermitianMatrix<scalar_t>& A, Options const& opts = Options())
{
    return potrf(A, opts);
}

// ---- the broadcast lists of one driver step (shared by the drivers and the planners) -----
/// src/hemmC.cc:98-137: the broadcasts of step k of hemm / symm, side Left (which = 0: tiles
/// of A, 1: block row k of B)
template <typename scalar_t, typename AMatrix>
typename Matrix<scalar_t>::BcastListTag hemm_bcast_list(int which, AMatrix& A, Matrix<scalar_t>& B,
                                                        Matrix<scalar_t>& C, int64_t k);
/// src/getrf_nopiv.cc:96-100 (which = 0), 113-120 (1), 178-188 (2): broadcasts of LU step k
template <typename scalar_t>
typename Matrix<scalar_t>::BcastListTag getrf_nopiv_bcast_list(int which, Matrix<scalar_t>& A,
                                                               int64_t k);
/// src/gemmA.cc:72-78 and :110-118: the broadcast of block column k of B and the reduction of
/// block column k of C in slate::gemmA
template <typename scalar_t>
typename Matrix<scalar_t>::BcastListTag gemmA_bcast_list(Matrix<scalar_t>& A, Matrix<scalar_t>& B,
                                                         int64_t k);
template <typename scalar_t>
typename Matrix<scalar_t>::ReduceList gemmA_reduce_list(Matrix<scalar_t>& A, Matrix<scalar_t>& C,
                                                        int64_t k);

/// src/gemmC.cc:143-157: block column k of A to the owners of each row of C (which = 0) or
/// block row k of B to the owners of each column of C (which = 1).
template <typename scalar_t>
typename Matrix<scalar_t>::BcastListTag gemm_bcast_list(int which, Matrix<scalar_t>& A,
                                                        Matrix<scalar_t>& B, Matrix<scalar_t>& C,
                                                        int64_t k);
/// src/potrf.cc:109 (which = 0: A(k,k) down its column) and :122-132 (which = 1: A(i,k) across
/// row i and down column i); A is the lower view.
template <typename scalar_t>
typename Matrix<scalar_t>::BcastListTag potrf_bcast_list(int which, HermitianMatrix<scalar_t>& A,
                                                         int64_t k);

// ---- harness helpers (not part of the reference surface) --------------------------------
/// Fills the local tiles with matgen's "rand" family (matgen/generate_type_rand.hh:40-135):
/// entry (i, j) is a pure function of (seed, i, j); `dominant` adds n to the diagonal.
/// Device-origin tiles are generated on the device (te_matgen_tile), host-origin tiles
/// are generated on the device and copied back, so both see identical bits.
template <typename scalar_t>
void generate_rand(BaseMatrix<scalar_t>& A, int64_t seed, bool dominant, int dist = 1);

/// sum over local stored tiles of |a_ij|^2 (lower/upper triangle only for trapezoid
/// matrices, diagonal tiles masked accordingly); not reduced over ranks.
template <typename scalar_t>
double local_sumsq(BaseMatrix<scalar_t>& A);

/// || A0 - L L^H ||_F^2 over this rank's tiles of the lower triangle, where L is the
/// factor held in A (lower) and A0 is regenerated from (seed, dominant).  Uses n^3/3
/// flops of the engine's own GEMM; the north-star backward-error metric
/// ||A - L L^H|| / (n ||A||) is assembled from it by the caller.
template <typename scalar_t>
double potrf_residual_sumsq(HermitianMatrix<scalar_t>& A, int64_t seed, bool dominant,
                            double* a0_sumsq);

}  // namespace slate

#endif
