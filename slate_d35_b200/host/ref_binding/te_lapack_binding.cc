// te_lapack_binding.cc -- lapack::potrf on a device queue, bound to the tile engine.
//
// Compiled IN PLACE OF lapackpp/src/cuda/cuda_potrf.cc (cusolverDnXpotrf).  The caller is the
// reference's internal::potrf<Target::Devices> (src/internal/internal_potrf.cc:57-81), compiled
// unmodified: it passes the diagonal tile and a device info word, then copies that word back and
// synchronises.  te_potrf_tile writes the same LAPACK info (0, or the 1-based index of the first
// pivot that is not positive) into dev_info on the queue's stream; its scratch (the inverse of
// one diagonal block) comes from the queue's device workspace.  No cuSOLVER routine is called.
#include "lapack/defines.h"
#include "lapack/device.hh"

#include "tile_engine.h"

#include <complex>
#include <string>

namespace {
template <typename T> struct te_scalar { static constexpr int dtype = -1; };
template <> struct te_scalar<double> { static constexpr int dtype = TE_F64; };
template <> struct te_scalar<std::complex<double>> { static constexpr int dtype = TE_C64; };
}  // namespace

namespace lapack {

template <typename scalar_t>
void potrf(lapack::Uplo uplo, int64_t n, scalar_t* dA, int64_t ldda, device_info_int* dev_info, lapack::Queue& queue)
{
    const int dtype = te_scalar<scalar_t>::dtype;
    if (dtype < 0)
        throw lapack::Error("single precision is not built on the B200 tile engine", __func__);
    static_assert(sizeof(device_info_int) == sizeof(int), "te_potrf_tile writes an int");
    const size_t bytes = te_potrf_scratch_bytes(dtype);
    queue.work_ensure_size<char>(bytes);      // syncs only if it has to grow
    int st = te_potrf_tile(queue.device(), queue.stream(), dtype,
                           uplo == lapack::Uplo::Lower ? TE_UPLO_LOWER : TE_UPLO_UPPER, (int)n, dA, (int)ldda,
                           (int*)dev_info, queue.work(), bytes);
    if (st != TE_OK)
        throw lapack::Error((std::string("te_potrf_tile failed: ") + te_last_error()).c_str(), __func__);
}

#define TE_POTRF_INST(T) \
    template void potrf(lapack::Uplo, int64_t, T*, int64_t, device_info_int*, lapack::Queue&);

TE_POTRF_INST(float)
TE_POTRF_INST(double)
TE_POTRF_INST(std::complex<float>)
TE_POTRF_INST(std::complex<double>)

}  // namespace lapack
