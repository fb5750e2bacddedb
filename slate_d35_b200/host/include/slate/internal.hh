// slate/internal.hh -- the Level-3 internal operations on the Target::Devices path:
// one block outer product / panel op on this rank's local tiles, launched on the
// tile engine (tile_engine.h) instead of blas::batch / cuBLAS / cuSOLVER.
//
// Signatures follow src/internal/internal.hh:173-177 (gemm), :221-224 (herk),
// :286-289 (syrk), :336-341 (trsm), :646-650 (potrf).  Only Target::Devices is built:
// there is no CPU fallback on this path, other targets throw NotImplemented.
// The calls enqueue and return; ordering against later launches, broadcasts and copies
// is by events (the reference blocks in queue->sync() at the end of every call).
#ifndef SLATE_B200_INTERNAL_HH
#define SLATE_B200_INTERNAL_HH

#include "slate/Matrix.hh"

namespace slate {
namespace internal {

using device_info_int = int;      // lapack::device_info_int of the reference

template <typename scalar_t> struct te_dtype_of;
template <> struct te_dtype_of<double> { static const int value = TE_F64; };
template <> struct te_dtype_of<std::complex<double>> { static const int value = TE_C64; };

// ---- gemm: C = alpha A B + beta C, A one block column, B one block row ------------------
template <typename scalar_t>
void gemm(TargetType<Target::Devices>, scalar_t alpha, Matrix<scalar_t>& A, Matrix<scalar_t>& B,
          scalar_t beta, Matrix<scalar_t>& C, Layout layout, int priority, int64_t queue_index);

template <Target target = Target::HostTask, typename scalar_t>
void gemm(scalar_t alpha, Matrix<scalar_t>&& A, Matrix<scalar_t>&& B, scalar_t beta,
          Matrix<scalar_t>&& C, Layout layout, int priority = 0, int64_t queue_index = 0)
{
    if (target != Target::Devices)
        slate_not_implemented("internal::gemm: only Target::Devices is built (no CPU fallback)");
    gemm(TargetType<Target::Devices>(), alpha, A, B, beta, C, layout, priority, queue_index);
}

/// one block outer product of a multi-part launch (gemm_multi)
template <typename scalar_t>
struct GemmPart { Matrix<scalar_t> A, B, C; };

template <typename scalar_t>
void gemm_multi(TargetType<Target::Devices>, scalar_t alpha, std::vector<GemmPart<scalar_t>>& parts,
                scalar_t beta, Layout layout, int64_t queue_index);

// ---- gemmA: C(:,0) = alpha A B(:,0) + beta C(:,0), computed where the tiles of A are -------
// (src/internal/internal.hh gemmA, Devices body src/internal/internal_gemmA.cc:224-560)
template <typename scalar_t>
void gemmA(TargetType<Target::Devices>, scalar_t alpha, Matrix<scalar_t>& A, Matrix<scalar_t>& B,
           scalar_t beta, Matrix<scalar_t>& C, Layout layout, int priority, int64_t queue_index);

template <Target target = Target::HostTask, typename scalar_t>
void gemmA(scalar_t alpha, Matrix<scalar_t>&& A, Matrix<scalar_t>&& B, scalar_t beta,
           Matrix<scalar_t>&& C, Layout layout, int priority = 0, int64_t queue_index = 0)
{
    if (target != Target::Devices)
        slate_not_implemented("internal::gemmA: only Target::Devices is built (no CPU fallback)");
    gemmA(TargetType<Target::Devices>(), alpha, A, B, beta, C, layout, priority, queue_index);
}

// ---- hemm / symm: C(0,:) = alpha A B(0,:) + beta C(0,:), A one diagonal tile ----------------
// (src/internal/internal.hh hemm/symm; the reference has no Devices body for these and runs
// the tile product on the host, src/internal/internal_hemm.cc -- here it runs on the GPU)
template <typename scalar_t>
void hemm(TargetType<Target::Devices>, Side side, scalar_t alpha, HermitianMatrix<scalar_t>& A,
          Matrix<scalar_t>& B, scalar_t beta, Matrix<scalar_t>& C, int priority, int queue_index);
template <typename scalar_t>
void symm(TargetType<Target::Devices>, Side side, scalar_t alpha, SymmetricMatrix<scalar_t>& A,
          Matrix<scalar_t>& B, scalar_t beta, Matrix<scalar_t>& C, int priority, int queue_index);

template <Target target = Target::HostTask, typename scalar_t>
void hemm(Side side, scalar_t alpha, HermitianMatrix<scalar_t>&& A, Matrix<scalar_t>&& B,
          scalar_t beta, Matrix<scalar_t>&& C, int priority = 0, int queue_index = 0)
{
    if (target != Target::Devices)
        slate_not_implemented("internal::hemm: only Target::Devices is built (no CPU fallback)");
    hemm(TargetType<Target::Devices>(), side, alpha, A, B, beta, C, priority, queue_index);
}
template <Target target = Target::HostTask, typename scalar_t>
void symm(Side side, scalar_t alpha, SymmetricMatrix<scalar_t>&& A, Matrix<scalar_t>&& B,
          scalar_t beta, Matrix<scalar_t>&& C, int priority = 0, int queue_index = 0)
{
    if (target != Target::Devices)
        slate_not_implemented("internal::symm: only Target::Devices is built (no CPU fallback)");
    symm(TargetType<Target::Devices>(), side, alpha, A, B, beta, C, priority, queue_index);
}

// ---- herk: C = alpha A A^H + beta C on the lower triangle of C ---------------------------
template <typename scalar_t>
void herk(TargetType<Target::Devices>, real_type<scalar_t> alpha, Matrix<scalar_t>& A,
          real_type<scalar_t> beta, HermitianMatrix<scalar_t>& C, int priority, int queue_index,
          Layout layout);

template <Target target = Target::HostTask, typename scalar_t>
void herk(real_type<scalar_t> alpha, Matrix<scalar_t>&& A, real_type<scalar_t> beta,
          HermitianMatrix<scalar_t>&& C, int priority = 0, int queue_index = 0,
          Layout layout = Layout::ColMajor)
{
    if (target != Target::Devices)
        slate_not_implemented("internal::herk: only Target::Devices is built (no CPU fallback)");
    herk(TargetType<Target::Devices>(), alpha, A, beta, C, priority, queue_index, layout);
}

// ---- syrk: C = alpha A A^T + beta C on the lower triangle of C ---------------------------
template <typename scalar_t>
void syrk(TargetType<Target::Devices>, scalar_t alpha, Matrix<scalar_t>& A, scalar_t beta,
          SymmetricMatrix<scalar_t>& C, int priority, int queue_index, Layout layout);

template <Target target = Target::HostTask, typename scalar_t>
void syrk(scalar_t alpha, Matrix<scalar_t>&& A, scalar_t beta, SymmetricMatrix<scalar_t>&& C,
          int priority = 0, int queue_index = 0, Layout layout = Layout::ColMajor)
{
    if (target != Target::Devices)
        slate_not_implemented("internal::syrk: only Target::Devices is built (no CPU fallback)");
    syrk(TargetType<Target::Devices>(), alpha, A, beta, C, priority, queue_index, layout);
}

// ---- her2k: C = alpha A B^H + conj(alpha) B A^H + beta C on the lower triangle of C -------
// (src/internal/internal.hh:240-258, Devices body src/internal/internal_her2k.cc:408-660)
template <typename scalar_t>
void her2k(TargetType<Target::Devices>, scalar_t alpha, Matrix<scalar_t>& A, Matrix<scalar_t>& B,
           real_type<scalar_t> beta, HermitianMatrix<scalar_t>& C, int priority, int queue_index,
           Layout layout);

template <Target target = Target::HostTask, typename scalar_t>
void her2k(scalar_t alpha, Matrix<scalar_t>&& A, Matrix<scalar_t>&& B, real_type<scalar_t> beta,
           HermitianMatrix<scalar_t>&& C, int priority = 0, int queue_index = 0,
           Layout layout = Layout::ColMajor)
{
    if (target != Target::Devices)
        slate_not_implemented("internal::her2k: only Target::Devices is built (no CPU fallback)");
    her2k(TargetType<Target::Devices>(), alpha, A, B, beta, C, priority, queue_index, layout);
}

// ---- syr2k: C = alpha A B^T + alpha B A^T + beta C on the lower triangle of C -------------
// (src/internal/internal.hh:305-323, src/internal/internal_syr2k.cc:391-640)
template <typename scalar_t>
void syr2k(TargetType<Target::Devices>, scalar_t alpha, Matrix<scalar_t>& A, Matrix<scalar_t>& B,
           scalar_t beta, SymmetricMatrix<scalar_t>& C, int priority, int queue_index, Layout layout);

template <Target target = Target::HostTask, typename scalar_t>
void syr2k(scalar_t alpha, Matrix<scalar_t>&& A, Matrix<scalar_t>&& B, scalar_t beta,
           SymmetricMatrix<scalar_t>&& C, int priority = 0, int queue_index = 0,
           Layout layout = Layout::ColMajor)
{
    if (target != Target::Devices)
        slate_not_implemented("internal::syr2k: only Target::Devices is built (no CPU fallback)");
    syr2k(TargetType<Target::Devices>(), alpha, A, B, beta, C, priority, queue_index, layout);
}

// ---- trsm: B = alpha op(A)^-1 B or alpha B op(A)^-1, A one triangular tile ----------------
template <typename scalar_t>
void trsm(TargetType<Target::Devices>, Side side, scalar_t alpha, TriangularMatrix<scalar_t>& A,
          Matrix<scalar_t>& B, int priority, Layout layout, int64_t queue_index);

template <Target target = Target::HostTask, typename scalar_t>
void trsm(Side side, scalar_t alpha, TriangularMatrix<scalar_t>&& A, Matrix<scalar_t>&& B,
          int priority = 0, Layout layout = Layout::ColMajor, int64_t queue_index = 0)
{
    if (target != Target::Devices)
        slate_not_implemented("internal::trsm: only Target::Devices is built (no CPU fallback)");
    trsm(TargetType<Target::Devices>(), side, alpha, A, B, priority, layout, queue_index);
}

// ---- potrf of one diagonal tile ------------------------------------------------------------
/// device_info != nullptr: the factorisation is only enqueued, 0 is returned and the LAPACK
/// info lands in *device_info (device memory) on the queue.  device_info == nullptr: waits
/// and returns info, as the reference does (src/internal/internal_potrf.cc:72-80).
template <typename scalar_t>
int64_t potrf(TargetType<Target::Devices>, HermitianMatrix<scalar_t>& A, int priority,
              int64_t queue_index, device_info_int* device_info);

template <Target target = Target::HostTask, typename scalar_t>
int64_t potrf(HermitianMatrix<scalar_t>&& A, int priority = 0, int64_t queue_index = 0,
              device_info_int* device_info = nullptr)
{
    if (target != Target::Devices)
        slate_not_implemented("internal::potrf: only Target::Devices is built (no CPU fallback)");
    return potrf(TargetType<Target::Devices>(), A, priority, queue_index, device_info);
}

// ---- LU without pivoting of one diagonal tile ------------------------------------------------
/// src/internal/internal.hh getrf_nopiv(A, ib, priority, &info); device_info as for potrf.
template <typename scalar_t>
int64_t getrf_nopiv(TargetType<Target::Devices>, Matrix<scalar_t>& A, int64_t ib, int priority,
                    int64_t queue_index, device_info_int* device_info);

template <Target target = Target::HostTask, typename scalar_t>
int64_t getrf_nopiv(Matrix<scalar_t>&& A, int64_t ib, int priority = 0, int64_t queue_index = 0,
                    device_info_int* device_info = nullptr)
{
    if (target != Target::Devices)
        slate_not_implemented("internal::getrf_nopiv: only Target::Devices is built (no CPU fallback)");
    return getrf_nopiv(TargetType<Target::Devices>(), A, ib, priority, queue_index, device_info);
}

/// smallest non-zero info over the ranks (src/internal/internal_reduce_info.cc:23-39)
void reduce_info(int64_t* info, MPI_Comm comm);

}  // namespace internal
}  // namespace slate

#endif
