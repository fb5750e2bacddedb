// binding_api.cc -- C entry points of the drop-in build (see Makefile): the reference's OWN
// slate::gemm and slate::potrf (src/gemm.cc, src/gemmC.cc, src/potrf.cc, compiled unmodified)
// called with Target::Devices or Target::HostTask on the same library, single rank.
//
// With Target::Devices the reference's internal::gemm / herk / trsm / potrf Devices overloads
// run as written and end in blas::batch::* / lapack::potrf(queue) -- which in this build are
// te_blas_binding.cc / te_lapack_binding.cc over include/tile_engine.h (libslate_ref_te.so),
// or, in the sibling build, BLAS++'s own cuBLAS / cuSOLVER wrappers (libslate_ref_cublas.so:
// the reference's stock GPU path, the yardstick).
//
// origin 'H': matrices wrap the caller's column-major arrays (Matrix::fromLAPACK,
// include/slate/Matrix.hh:355-384); SLATE moves tiles to the device and back itself.
// origin 'D': SLATE-owned tiles born on the device (insertLocalTiles(Target::Devices),
// include/slate/Matrix.hh:832-847, the tester's --origin d); filled from / read back to the
// caller's arrays outside the timed call.  *seconds = wall time of the driver call alone.
#include "slate/slate.hh"

#include <chrono>
#include <complex>
#include <cstdint>
#include <cstdio>

using slate::LayoutConvert;
using slate::Option;
using slate::Target;
using slate::Uplo;
typedef std::complex<double> zdouble;

namespace {

slate::Options opts(int target, int lookahead)
{
    return {{Option::Target, target == 'D' ? Target::Devices : Target::HostTask},
            {Option::Lookahead, (int64_t)lookahead}};
}

double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

/// copies between the caller's array and the (device-origin) matrix through host tile copies
template <typename MatrixType, typename T>
void fill(MatrixType& M, T const* src, int64_t ld, bool to_matrix)
{
    int64_t joff = 0;
    for (int64_t j = 0; j < M.nt(); ++j) {
        int64_t ioff = 0;
        for (int64_t i = 0; i < M.mt(); ++i) {
            const bool stored = M.uplo() == Uplo::General || (M.uplo() == Uplo::Lower ? i >= j : i <= j);
            if (stored && M.tileIsLocal(i, j)) {
                if (to_matrix) M.tileGetForWriting(i, j, slate::HostNum, LayoutConvert::ColMajor);
                else M.tileGetForReading(i, j, slate::HostNum, LayoutConvert::ColMajor);
                auto tile = M(i, j);
                for (int64_t jj = 0; jj < tile.nb(); ++jj)
                    for (int64_t ii = 0; ii < tile.mb(); ++ii) {
                        T* user = const_cast<T*>(src) + (ioff + ii) + (joff + jj) * ld;
                        if (to_matrix) tile.at(ii, jj) = *user;
                        else *user = tile(ii, jj);
                    }
            }
            ioff += M.tileMb(i);
        }
        joff += M.tileNb(j);
    }
    if (to_matrix) {
        M.tileUpdateAllOrigin();
        M.releaseWorkspace();
    }
}

template <typename T>
int gemm_any(int target, int origin, int64_t m, int64_t n, int64_t k, int64_t nb, T alpha, T* A, int64_t lda,
             T* B, int64_t ldb, T beta, T* C, int64_t ldc, int lookahead, double* seconds)
{
    try {
        double t = 0;
        if (origin == 'D') {
            slate::Matrix<T> Am(m, k, nb, 1, 1, MPI_COMM_WORLD), Bm(k, n, nb, 1, 1, MPI_COMM_WORLD),
                Cm(m, n, nb, 1, 1, MPI_COMM_WORLD);
            Am.insertLocalTiles(Target::Devices);
            Bm.insertLocalTiles(Target::Devices);
            Cm.insertLocalTiles(Target::Devices);
            fill(Am, A, lda, true);
            fill(Bm, B, ldb, true);
            fill(Cm, C, ldc, true);
            t = now();
            slate::gemm(alpha, Am, Bm, beta, Cm, opts(target, lookahead));
            t = now() - t;
            fill(Cm, C, ldc, false);
        }
        else {
            auto Am = slate::Matrix<T>::fromLAPACK(m, k, A, lda, nb, 1, 1, MPI_COMM_WORLD);
            auto Bm = slate::Matrix<T>::fromLAPACK(k, n, B, ldb, nb, 1, 1, MPI_COMM_WORLD);
            auto Cm = slate::Matrix<T>::fromLAPACK(m, n, C, ldc, nb, 1, 1, MPI_COMM_WORLD);
            t = now();
            slate::gemm(alpha, Am, Bm, beta, Cm, opts(target, lookahead));
            t = now() - t;
        }
        if (seconds) *seconds = t;
        return 0;
    }
    catch (std::exception const& e) {
        fprintf(stderr, "rb gemm: %s\n", e.what());
        return -1;
    }
}

template <typename T>
int64_t potrf_any(int target, int origin, int uplo, int64_t n, int64_t nb, T* A, int64_t lda, int lookahead,
                  double* seconds)
{
    try {
        const Uplo ul = uplo == 'U' ? Uplo::Upper : Uplo::Lower;
        double t = 0;
        int64_t info;
        if (origin == 'D') {
            slate::HermitianMatrix<T> Am(ul, n, nb, 1, 1, MPI_COMM_WORLD);
            Am.insertLocalTiles(Target::Devices);
            fill(Am, A, lda, true);
            t = now();
            info = slate::potrf(Am, opts(target, lookahead));
            t = now() - t;
            fill(Am, A, lda, false);
        }
        else {
            auto Am = slate::HermitianMatrix<T>::fromLAPACK(ul, n, A, lda, nb, 1, 1, MPI_COMM_WORLD);
            t = now();
            info = slate::potrf(Am, opts(target, lookahead));
            t = now() - t;
        }
        if (seconds) *seconds = t;
        return info;
    }
    catch (std::exception const& e) {
        fprintf(stderr, "rb potrf: %s\n", e.what());
        return -1000000;
    }
}

}  // namespace

extern "C" {

#ifndef RB_FLAVOUR
#define RB_FLAVOUR "?"
#endif
/// "tile_engine" or "cublas": which library sits under blas::batch::* in this build
const char* rb_flavour(void) { return RB_FLAVOUR; }
int rb_num_devices(void) { return slate::Memory::num_devices_; }

int rb_dgemm(int target, int origin, int64_t m, int64_t n, int64_t k, int64_t nb, double alpha, double* A,
             int64_t lda, double* B, int64_t ldb, double beta, double* C, int64_t ldc, int lookahead,
             double* seconds)
{ return gemm_any<double>(target, origin, m, n, k, nb, alpha, A, lda, B, ldb, beta, C, ldc, lookahead, seconds); }

int rb_zgemm(int target, int origin, int64_t m, int64_t n, int64_t k, int64_t nb, const double* alpha, double* A,
             int64_t lda, double* B, int64_t ldb, const double* beta, double* C, int64_t ldc, int lookahead,
             double* seconds)
{
    return gemm_any<zdouble>(target, origin, m, n, k, nb, zdouble(alpha[0], alpha[1]), (zdouble*)A, lda,
                             (zdouble*)B, ldb, zdouble(beta[0], beta[1]), (zdouble*)C, ldc, lookahead, seconds);
}

int64_t rb_dpotrf(int target, int origin, int uplo, int64_t n, int64_t nb, double* A, int64_t lda, int lookahead,
                  double* seconds)
{ return potrf_any<double>(target, origin, uplo, n, nb, A, lda, lookahead, seconds); }

int64_t rb_zpotrf(int target, int origin, int uplo, int64_t n, int64_t nb, double* A, int64_t lda, int lookahead,
                  double* seconds)
{ return potrf_any<zdouble>(target, origin, uplo, n, nb, (zdouble*)A, lda, lookahead, seconds); }

}  // extern "C"
