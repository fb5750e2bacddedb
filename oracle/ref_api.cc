// ref_api.cc -- TEST INFRASTRUCTURE.  C entry points into the UNMODIFIED reference
// (SLATE 2023.11.05 under /root/reference), compiled single-rank, CPU only, by
// oracle/Makefile into oracle/_ref/libslate_ref.so.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
// reference legs may load this library; it is the checker and the timed CPU
// baseline, never part of the product path.
//
// Every function wraps user column-major arrays with Matrix::fromLAPACK
// (include/slate/Matrix.hh:355-384) and calls the reference's own driver or
// internal op with Target::HostTask, the path SURVEY.md section 8(c) names as
// the oracle.
#include "slate/slate.hh"
#include "internal/internal.hh"
#include "../matgen/random.hh"

#include <complex>
#include <cstdint>
#include <cstring>
#include <omp.h>

using slate::Op;
using slate::Uplo;
using slate::Target;
using slate::Option;
using slate::Layout;
typedef std::complex<double> zdouble;

namespace {

template <typename M>
M apply_op(M A, int op)
{
    if (op == 'T') return slate::transpose(A);
    if (op == 'C') return slate::conj_transpose(A);
    return A;
}

slate::Options host_opts(int lookahead)
{
    return {{Option::Target, Target::HostTask}, {Option::Lookahead, (int64_t)lookahead}};
}

// C = alpha op(A) op(B) + beta C through slate::gemm (src/gemm.cc:66-88 -> gemmC).
// A is stored (opa=='N' ? m x k : k x m), B likewise.
template <typename T>
int ref_gemm(int opa, int opb, int64_t m, int64_t n, int64_t k, int64_t nb,
             T alpha, T* A, int64_t lda, T* B, int64_t ldb,
             T beta, T* C, int64_t ldc, int lookahead)
{
    try {
        int64_t am = (opa == 'N') ? m : k, an = (opa == 'N') ? k : m;
        int64_t bm = (opb == 'N') ? k : n, bn = (opb == 'N') ? n : k;
        auto Am = slate::Matrix<T>::fromLAPACK(am, an, A, lda, nb, 1, 1, MPI_COMM_WORLD);
        auto Bm = slate::Matrix<T>::fromLAPACK(bm, bn, B, ldb, nb, 1, 1, MPI_COMM_WORLD);
        auto Cm = slate::Matrix<T>::fromLAPACK(m, n, C, ldc, nb, 1, 1, MPI_COMM_WORLD);
        auto opA = apply_op(Am, opa);
        auto opB = apply_op(Bm, opb);
        slate::gemm(alpha, opA, opB, beta, Cm, host_opts(lookahead));
        return 0;
    }
    catch (std::exception const& e) {
        fprintf(stderr, "ref_gemm: %s\n", e.what());
        return -1;
    }
}

// Cholesky through slate::potrf (src/potrf.cc:262-281); returns info.
template <typename T>
int64_t ref_potrf(int uplo, int64_t n, int64_t nb, T* A, int64_t lda, int lookahead)
{
    try {
        auto Am = slate::HermitianMatrix<T>::fromLAPACK(
            uplo == 'U' ? Uplo::Upper : Uplo::Lower, n, A, lda, nb, 1, 1, MPI_COMM_WORLD);
        return slate::potrf(Am, host_opts(lookahead));
    }
    catch (std::exception const& e) {
        fprintf(stderr, "ref_potrf: %s\n", e.what());
        return -1000000;
    }
}

// One block outer product, internal::gemm<HostTask> (src/internal/internal_gemm.cc:59-115):
// A must be one block column (k <= nb) and B one block row after the ops.
template <typename T>
int ref_internal_gemm(int opa, int opb, int64_t m, int64_t n, int64_t k, int64_t nb,
                      T alpha, T* A, int64_t lda, T* B, int64_t ldb,
                      T beta, T* C, int64_t ldc)
{
    try {
        int64_t am = (opa == 'N') ? m : k, an = (opa == 'N') ? k : m;
        int64_t bm = (opb == 'N') ? k : n, bn = (opb == 'N') ? n : k;
        auto Am = slate::Matrix<T>::fromLAPACK(am, an, A, lda, nb, 1, 1, MPI_COMM_WORLD);
        auto Bm = slate::Matrix<T>::fromLAPACK(bm, bn, B, ldb, nb, 1, 1, MPI_COMM_WORLD);
        auto Cm = slate::Matrix<T>::fromLAPACK(m, n, C, ldc, nb, 1, 1, MPI_COMM_WORLD);
        auto opA = apply_op(Am, opa);
        auto opB = apply_op(Bm, opb);
        #pragma omp parallel
        #pragma omp master
        slate::internal::gemm<Target::HostTask>(
            alpha, std::move(opA), std::move(opB), beta, std::move(Cm), Layout::ColMajor);
        return 0;
    }
    catch (std::exception const& e) {
        fprintf(stderr, "ref_internal_gemm: %s\n", e.what());
        return -1;
    }
}

// C = alpha op(A) op(A)^H + beta C on the `uplo` triangle, internal::herk<HostTask>
// (src/internal/internal_herk.cc:47-110).  op(A) is n x k, k <= nb.
template <typename T>
int ref_internal_herk(int uplo, int opa, int64_t n, int64_t k, int64_t nb,
                      double alpha, T* A, int64_t lda, double beta, T* C, int64_t ldc)
{
    try {
        int64_t am = (opa == 'N') ? n : k, an = (opa == 'N') ? k : n;
        auto Am = slate::Matrix<T>::fromLAPACK(am, an, A, lda, nb, 1, 1, MPI_COMM_WORLD);
        auto Cm = slate::HermitianMatrix<T>::fromLAPACK(
            uplo == 'U' ? Uplo::Upper : Uplo::Lower, n, C, ldc, nb, 1, 1, MPI_COMM_WORLD);
        auto opA = apply_op(Am, opa);
        // as the driver does (src/herk.cc: "if upper, change to lower")
        if (Cm.uplo() == Uplo::Upper)
            Cm = slate::conj_transpose(Cm);
        int err = 0;
        #pragma omp parallel
        #pragma omp master
        {
            try {
                slate::internal::herk<Target::HostTask>(
                    alpha, std::move(opA), beta, std::move(Cm), 0, 0, Layout::ColMajor);
            }
            catch (std::exception const& e) { err = 1; }
        }
        return err ? -1 : 0;
    }
    catch (std::exception const& e) {
        fprintf(stderr, "ref_internal_herk: %s\n", e.what());
        return -1;
    }
}

// B = alpha B op(A)^{-1} (side R) or alpha op(A)^{-1} B (side L), A one nb x nb
// triangular tile; internal::trsm<HostTask> (src/internal/internal_trsm.cc:40-92).
template <typename T>
int ref_internal_trsm(int side, int uplo, int opa, int diag, int64_t m, int64_t n, int64_t nb,
                      T alpha, T* A, int64_t lda, T* B, int64_t ldb)
{
    try {
        int64_t an = (side == 'R') ? n : m;
        auto Am = slate::TriangularMatrix<T>::fromLAPACK(
            uplo == 'U' ? Uplo::Upper : Uplo::Lower,
            diag == 'U' ? slate::Diag::Unit : slate::Diag::NonUnit,
            an, A, lda, nb, 1, 1, MPI_COMM_WORLD);
        auto Bm = slate::Matrix<T>::fromLAPACK(m, n, B, ldb, nb, 1, 1, MPI_COMM_WORLD);
        auto opA = apply_op(Am, opa);
        #pragma omp parallel
        #pragma omp master
        slate::internal::trsm<Target::HostTask>(
            side == 'R' ? slate::Side::Right : slate::Side::Left,
            alpha, std::move(opA), std::move(Bm), 0, Layout::ColMajor, 0);
        return 0;
    }
    catch (std::exception const& e) {
        fprintf(stderr, "ref_internal_trsm: %s\n", e.what());
        return -1;
    }
}

// Cholesky of one tile, internal::potrf<HostTask> (src/internal/internal_potrf.cc:35-50).
template <typename T>
int64_t ref_internal_potrf(int uplo, int64_t n, T* A, int64_t lda)
{
    try {
        auto Am = slate::HermitianMatrix<T>::fromLAPACK(
            uplo == 'U' ? Uplo::Upper : Uplo::Lower, n, A, lda, n, 1, 1, MPI_COMM_WORLD);
        int64_t info = 0;
        #pragma omp parallel
        #pragma omp master
        info = slate::internal::potrf<Target::HostTask>(std::move(Am), 0, 0);
        return info;
    }
    catch (std::exception const& e) {
        fprintf(stderr, "ref_internal_potrf: %s\n", e.what());
        return -1000000;
    }
}

// ---- the drivers of SURVEY.md section 8f, through the reference's own public functions ----
template <typename T> T scalar_of(const double* v);
template <> double scalar_of<double>(const double* v) { return v[0]; }
template <> zdouble scalar_of<zdouble>(const double* v) { return zdouble(v[0], v[1]); }

inline Uplo uplo_of(int c) { return c == 'U' ? Uplo::Upper : Uplo::Lower; }
inline slate::Side side_of(int c) { return c == 'R' ? slate::Side::Right : slate::Side::Left; }

template <typename F>
int64_t guarded(const char* name, F&& f)
{
    try { return f(); }
    catch (std::exception const& e) {
        fprintf(stderr, "%s: %s\n", name, e.what());
        return -1000000;
    }
}

// kind 0: slate::herk (src/herk.cc), 1: slate::syrk (src/syrk.cc); op(A) is n x k
template <typename T>
int64_t ref_rank_k(int kind, int uplo, int opa, int64_t n, int64_t k, int64_t nb, const double* alpha,
                   T* A, int64_t lda, const double* beta, T* C, int64_t ldc, int la)
{
    return guarded("ref_rank_k", [&]() -> int64_t {
        int64_t am = (opa == 'N') ? n : k, an = (opa == 'N') ? k : n;
        auto Am = slate::Matrix<T>::fromLAPACK(am, an, A, lda, nb, 1, 1, MPI_COMM_WORLD);
        auto opA = apply_op(Am, opa);
        if (kind == 0) {
            auto Cm = slate::HermitianMatrix<T>::fromLAPACK(uplo_of(uplo), n, C, ldc, nb, 1, 1, MPI_COMM_WORLD);
            slate::herk(alpha[0], opA, beta[0], Cm, host_opts(la));
        }
        else {
            auto Cm = slate::SymmetricMatrix<T>::fromLAPACK(uplo_of(uplo), n, C, ldc, nb, 1, 1, MPI_COMM_WORLD);
            slate::syrk(scalar_of<T>(alpha), opA, scalar_of<T>(beta), Cm, host_opts(la));
        }
        return 0;
    });
}

// kind 0: slate::her2k (src/her2k.cc), 1: slate::syr2k (src/syr2k.cc)
template <typename T>
int64_t ref_rank_2k(int kind, int uplo, int opa, int64_t n, int64_t k, int64_t nb, const double* alpha,
                    T* A, int64_t lda, T* B, int64_t ldb, const double* beta, T* C, int64_t ldc, int la)
{
    return guarded("ref_rank_2k", [&]() -> int64_t {
        int64_t am = (opa == 'N') ? n : k, an = (opa == 'N') ? k : n;
        auto Am = slate::Matrix<T>::fromLAPACK(am, an, A, lda, nb, 1, 1, MPI_COMM_WORLD);
        auto Bm = slate::Matrix<T>::fromLAPACK(am, an, B, ldb, nb, 1, 1, MPI_COMM_WORLD);
        auto opA = apply_op(Am, opa);
        auto opB = apply_op(Bm, opa);
        if (kind == 0) {
            auto Cm = slate::HermitianMatrix<T>::fromLAPACK(uplo_of(uplo), n, C, ldc, nb, 1, 1, MPI_COMM_WORLD);
            slate::her2k(scalar_of<T>(alpha), opA, opB, beta[0], Cm, host_opts(la));
        }
        else {
            auto Cm = slate::SymmetricMatrix<T>::fromLAPACK(uplo_of(uplo), n, C, ldc, nb, 1, 1, MPI_COMM_WORLD);
            slate::syr2k(scalar_of<T>(alpha), opA, opB, scalar_of<T>(beta), Cm, host_opts(la));
        }
        return 0;
    });
}

// kind 0: slate::hemm (src/hemm.cc -> hemmC / hemmA), 1: slate::symm (src/symm.cc)
template <typename T>
int64_t ref_hemm(int kind, int side, int uplo, int64_t m, int64_t n, int64_t nb, const double* alpha,
                 T* A, int64_t lda, T* B, int64_t ldb, const double* beta, T* C, int64_t ldc, int la)
{
    return guarded("ref_hemm", [&]() -> int64_t {
        int64_t an = (side == 'R') ? n : m;
        auto Bm = slate::Matrix<T>::fromLAPACK(m, n, B, ldb, nb, 1, 1, MPI_COMM_WORLD);
        auto Cm = slate::Matrix<T>::fromLAPACK(m, n, C, ldc, nb, 1, 1, MPI_COMM_WORLD);
        if (kind == 0) {
            auto Am = slate::HermitianMatrix<T>::fromLAPACK(uplo_of(uplo), an, A, lda, nb, 1, 1, MPI_COMM_WORLD);
            slate::hemm(side_of(side), scalar_of<T>(alpha), Am, Bm, scalar_of<T>(beta), Cm, host_opts(la));
        }
        else {
            auto Am = slate::SymmetricMatrix<T>::fromLAPACK(uplo_of(uplo), an, A, lda, nb, 1, 1, MPI_COMM_WORLD);
            slate::symm(side_of(side), scalar_of<T>(alpha), Am, Bm, scalar_of<T>(beta), Cm, host_opts(la));
        }
        return 0;
    });
}

// slate::trsm (src/trsm.cc -> trsmB / trsmA, src/work/work_trsm.cc)
template <typename T>
int64_t ref_trsm(int side, int uplo, int opa, int diag, int64_t m, int64_t n, int64_t nb,
                 const double* alpha, T* A, int64_t lda, T* B, int64_t ldb, int la)
{
    return guarded("ref_trsm", [&]() -> int64_t {
        int64_t an = (side == 'R') ? n : m;
        auto Am = slate::TriangularMatrix<T>::fromLAPACK(
            uplo_of(uplo), diag == 'U' ? slate::Diag::Unit : slate::Diag::NonUnit, an, A, lda, nb, 1, 1,
            MPI_COMM_WORLD);
        auto Bm = slate::Matrix<T>::fromLAPACK(m, n, B, ldb, nb, 1, 1, MPI_COMM_WORLD);
        auto opA = apply_op(Am, opa);
        slate::trsm(side_of(side), scalar_of<T>(alpha), opA, Bm, host_opts(la));
        return 0;
    });
}

// slate::posv (src/posv.cc): factor and solve; returns info
template <typename T>
int64_t ref_posv(int uplo, int64_t n, int64_t nrhs, int64_t nb, T* A, int64_t lda, T* B, int64_t ldb,
                 int la)
{
    return guarded("ref_posv", [&]() -> int64_t {
        auto Am = slate::HermitianMatrix<T>::fromLAPACK(uplo_of(uplo), n, A, lda, nb, 1, 1, MPI_COMM_WORLD);
        auto Bm = slate::Matrix<T>::fromLAPACK(n, nrhs, B, ldb, nb, 1, 1, MPI_COMM_WORLD);
        return slate::posv(Am, Bm, host_opts(la));
    });
}

// slate::gemm forced onto gemmA (src/gemmA.cc) by Option::MethodGemm
template <typename T>
int64_t ref_gemmA(int opa, int opb, int64_t m, int64_t n, int64_t k, int64_t nb, const double* alpha,
                  T* A, int64_t lda, T* B, int64_t ldb, const double* beta, T* C, int64_t ldc, int la)
{
    return guarded("ref_gemmA", [&]() -> int64_t {
        int64_t am = (opa == 'N') ? m : k, an = (opa == 'N') ? k : m;
        int64_t bm = (opb == 'N') ? k : n, bn = (opb == 'N') ? n : k;
        auto Am = slate::Matrix<T>::fromLAPACK(am, an, A, lda, nb, 1, 1, MPI_COMM_WORLD);
        auto Bm = slate::Matrix<T>::fromLAPACK(bm, bn, B, ldb, nb, 1, 1, MPI_COMM_WORLD);
        auto Cm = slate::Matrix<T>::fromLAPACK(m, n, C, ldc, nb, 1, 1, MPI_COMM_WORLD);
        auto opA = apply_op(Am, opa);
        auto opB = apply_op(Bm, opb);
        slate::Options opts = host_opts(la);
        opts[Option::MethodGemm] = slate::MethodGemm::GemmA;
        slate::gemm(scalar_of<T>(alpha), opA, opB, scalar_of<T>(beta), Cm, opts);
        return 0;
    });
}

// slate::getrf_nopiv (src/getrf_nopiv.cc); returns info
template <typename T>
int64_t ref_getrf_nopiv(int64_t m, int64_t n, int64_t nb, T* A, int64_t lda, int la)
{
    return guarded("ref_getrf_nopiv", [&]() -> int64_t {
        auto Am = slate::Matrix<T>::fromLAPACK(m, n, A, lda, nb, 1, 1, MPI_COMM_WORLD);
        return slate::getrf_nopiv(Am, host_opts(la));
    });
}

}  // namespace

extern "C" {

const char* ref_version(void) { return "slate-ref 2023.11.05 HostTask single-rank"; }
int ref_num_threads(void) { return omp_get_max_threads(); }
void ref_set_num_threads(int n) { if (n > 0) omp_set_num_threads(n); }

int ref_dgemm(int opa, int opb, int64_t m, int64_t n, int64_t k, int64_t nb, double alpha,
              double* A, int64_t lda, double* B, int64_t ldb, double beta, double* C, int64_t ldc,
              int lookahead)
{ return ref_gemm<double>(opa, opb, m, n, k, nb, alpha, A, lda, B, ldb, beta, C, ldc, lookahead); }

int ref_zgemm(int opa, int opb, int64_t m, int64_t n, int64_t k, int64_t nb, const double* alpha,
              double* A, int64_t lda, double* B, int64_t ldb, const double* beta, double* C,
              int64_t ldc, int lookahead)
{
    return ref_gemm<zdouble>(opa, opb, m, n, k, nb, zdouble(alpha[0], alpha[1]),
                             (zdouble*)A, lda, (zdouble*)B, ldb, zdouble(beta[0], beta[1]),
                             (zdouble*)C, ldc, lookahead);
}

int64_t ref_dpotrf(int uplo, int64_t n, int64_t nb, double* A, int64_t lda, int lookahead)
{ return ref_potrf<double>(uplo, n, nb, A, lda, lookahead); }

int64_t ref_zpotrf(int uplo, int64_t n, int64_t nb, double* A, int64_t lda, int lookahead)
{ return ref_potrf<zdouble>(uplo, n, nb, (zdouble*)A, lda, lookahead); }

int ref_internal_dgemm(int opa, int opb, int64_t m, int64_t n, int64_t k, int64_t nb, double alpha,
                       double* A, int64_t lda, double* B, int64_t ldb, double beta, double* C,
                       int64_t ldc)
{ return ref_internal_gemm<double>(opa, opb, m, n, k, nb, alpha, A, lda, B, ldb, beta, C, ldc); }

int ref_internal_zgemm(int opa, int opb, int64_t m, int64_t n, int64_t k, int64_t nb,
                       const double* alpha, double* A, int64_t lda, double* B, int64_t ldb,
                       const double* beta, double* C, int64_t ldc)
{
    return ref_internal_gemm<zdouble>(opa, opb, m, n, k, nb, zdouble(alpha[0], alpha[1]),
                                      (zdouble*)A, lda, (zdouble*)B, ldb,
                                      zdouble(beta[0], beta[1]), (zdouble*)C, ldc);
}

int ref_internal_dherk(int uplo, int opa, int64_t n, int64_t k, int64_t nb, double alpha,
                       double* A, int64_t lda, double beta, double* C, int64_t ldc)
{ return ref_internal_herk<double>(uplo, opa, n, k, nb, alpha, A, lda, beta, C, ldc); }

int ref_internal_zherk(int uplo, int opa, int64_t n, int64_t k, int64_t nb, double alpha,
                       double* A, int64_t lda, double beta, double* C, int64_t ldc)
{ return ref_internal_herk<zdouble>(uplo, opa, n, k, nb, alpha, (zdouble*)A, lda, beta,
                                    (zdouble*)C, ldc); }

int ref_internal_dtrsm(int side, int uplo, int opa, int diag, int64_t m, int64_t n, int64_t nb,
                       double alpha, double* A, int64_t lda, double* B, int64_t ldb)
{ return ref_internal_trsm<double>(side, uplo, opa, diag, m, n, nb, alpha, A, lda, B, ldb); }

int ref_internal_ztrsm(int side, int uplo, int opa, int diag, int64_t m, int64_t n, int64_t nb,
                       const double* alpha, double* A, int64_t lda, double* B, int64_t ldb)
{
    return ref_internal_trsm<zdouble>(side, uplo, opa, diag, m, n, nb,
                                      zdouble(alpha[0], alpha[1]), (zdouble*)A, lda,
                                      (zdouble*)B, ldb);
}

int64_t ref_internal_dpotrf(int uplo, int64_t n, double* A, int64_t lda)
{ return ref_internal_potrf<double>(uplo, n, A, lda); }

int64_t ref_internal_zpotrf(int uplo, int64_t n, double* A, int64_t lda)
{ return ref_internal_potrf<zdouble>(uplo, n, (zdouble*)A, lda); }

// matgen's counter-based generator (matgen/random.cc:93-159, 201-232): entry (i, j)
// is a pure function of (seed, i + ioffset, j + joffset).  dist 1 = uniform [0,1).
void ref_matgen_d(int dist, int64_t seed, int64_t m, int64_t n, int64_t ioff, int64_t joff,
                  double* A, int64_t lda)
{ slate::random::generate<double>(slate::random::Dist(dist), seed, m, n, ioff, joff, A, lda); }

void ref_matgen_z(int dist, int64_t seed, int64_t m, int64_t n, int64_t ioff, int64_t joff,
                  double* A, int64_t lda)
{
    slate::random::generate<zdouble>(slate::random::Dist(dist), seed, m, n, ioff, joff,
                                     (zdouble*)A, lda);
}

// ---- widened drivers: dtype 0 = double, 1 = complex<double>; scalars as 2 doubles ----
#define BY_TYPE(call_d, call_z) (dtype == 1 ? (call_z) : (call_d))

int64_t ref_rank_k_any(int dtype, int kind, int uplo, int opa, int64_t n, int64_t k, int64_t nb,
                       const double* alpha, void* A, int64_t lda, const double* beta, void* C,
                       int64_t ldc, int la)
{
    return BY_TYPE(ref_rank_k<double>(kind, uplo, opa, n, k, nb, alpha, (double*)A, lda, beta, (double*)C, ldc, la),
                   ref_rank_k<zdouble>(kind, uplo, opa, n, k, nb, alpha, (zdouble*)A, lda, beta, (zdouble*)C, ldc, la));
}

int64_t ref_rank_2k_any(int dtype, int kind, int uplo, int opa, int64_t n, int64_t k, int64_t nb,
                        const double* alpha, void* A, int64_t lda, void* B, int64_t ldb,
                        const double* beta, void* C, int64_t ldc, int la)
{
    return BY_TYPE(ref_rank_2k<double>(kind, uplo, opa, n, k, nb, alpha, (double*)A, lda, (double*)B, ldb, beta, (double*)C, ldc, la),
                   ref_rank_2k<zdouble>(kind, uplo, opa, n, k, nb, alpha, (zdouble*)A, lda, (zdouble*)B, ldb, beta, (zdouble*)C, ldc, la));
}

int64_t ref_hemm_any(int dtype, int kind, int side, int uplo, int64_t m, int64_t n, int64_t nb,
                     const double* alpha, void* A, int64_t lda, void* B, int64_t ldb,
                     const double* beta, void* C, int64_t ldc, int la)
{
    return BY_TYPE(ref_hemm<double>(kind, side, uplo, m, n, nb, alpha, (double*)A, lda, (double*)B, ldb, beta, (double*)C, ldc, la),
                   ref_hemm<zdouble>(kind, side, uplo, m, n, nb, alpha, (zdouble*)A, lda, (zdouble*)B, ldb, beta, (zdouble*)C, ldc, la));
}

int64_t ref_trsm_any(int dtype, int side, int uplo, int opa, int diag, int64_t m, int64_t n, int64_t nb,
                     const double* alpha, void* A, int64_t lda, void* B, int64_t ldb, int la)
{
    return BY_TYPE(ref_trsm<double>(side, uplo, opa, diag, m, n, nb, alpha, (double*)A, lda, (double*)B, ldb, la),
                   ref_trsm<zdouble>(side, uplo, opa, diag, m, n, nb, alpha, (zdouble*)A, lda, (zdouble*)B, ldb, la));
}

int64_t ref_posv_any(int dtype, int uplo, int64_t n, int64_t nrhs, int64_t nb, void* A, int64_t lda,
                     void* B, int64_t ldb, int la)
{
    return BY_TYPE(ref_posv<double>(uplo, n, nrhs, nb, (double*)A, lda, (double*)B, ldb, la),
                   ref_posv<zdouble>(uplo, n, nrhs, nb, (zdouble*)A, lda, (zdouble*)B, ldb, la));
}

int64_t ref_gemmA_any(int dtype, int opa, int opb, int64_t m, int64_t n, int64_t k, int64_t nb,
                      const double* alpha, void* A, int64_t lda, void* B, int64_t ldb,
                      const double* beta, void* C, int64_t ldc, int la)
{
    return BY_TYPE(ref_gemmA<double>(opa, opb, m, n, k, nb, alpha, (double*)A, lda, (double*)B, ldb, beta, (double*)C, ldc, la),
                   ref_gemmA<zdouble>(opa, opb, m, n, k, nb, alpha, (zdouble*)A, lda, (zdouble*)B, ldb, beta, (zdouble*)C, ldc, la));
}

int64_t ref_getrf_nopiv_any(int dtype, int64_t m, int64_t n, int64_t nb, void* A, int64_t lda, int la)
{
    return BY_TYPE(ref_getrf_nopiv<double>(m, n, nb, (double*)A, lda, la),
                   ref_getrf_nopiv<zdouble>(m, n, nb, (zdouble*)A, lda, la));
}

}  // extern "C"
