// te_transpose.cu -- tile transposes and the symmetric / Hermitian fill of a diagonal tile.
//
// Reference counterparts: src/cuda/device_transpose.cu (device::transpose in place for square
// tiles, out of place for rectangular ones, and their batched forms, used by
// tileLayoutConvert between ColMajor and RowMajor).  The fill has no single counterpart: it is
// what lets a diagonal tile of a Hermitian / symmetric matrix, of which only one triangle is
// stored, be used as a plain operand of the batched GEMM kernel (slate::hemm / symm,
// src/internal/internal_hemm.cc calls a tile hemm there).
//
// HBM-bound, 2 x 8 (16) bytes per element; 32 x 32 blocks staged through padded shared memory
// so that both the global reads and the global writes are 256-byte row segments.
#include "te_common.cuh"

namespace te {

__device__ __forceinline__ double  conj_if(double x, bool) { return x; }
__device__ __forceinline__ double2 conj_if(double2 x, bool c) { return c ? make_double2(x.x, -x.y) : x; }

constexpr int TB = 32;        // block edge
constexpr int TROWS = 8;      // thread rows: 32 x 8 threads, 4 elements each

// AT (n x m) = op(A (m x n)); blockIdx.z walks the batch
template <typename T>
__global__ void transpose_out_kernel(int m, int n, bool cj, const T* A, const T* const* A_array,
                                     int64_t lda, T* AT, T* const* AT_array, int64_t ldat)
{
    __shared__ T blk[TB][TB + 1];
    const T* a = A_array ? A_array[blockIdx.z] : A;
    T* at = AT_array ? AT_array[blockIdx.z] : AT;
    const int i0 = blockIdx.x * TB, j0 = blockIdx.y * TB;
    for (int jj = threadIdx.y; jj < TB; jj += TROWS) {
        const int i = i0 + threadIdx.x, j = j0 + jj;
        if (i < m && j < n) blk[jj][threadIdx.x] = a[i + (int64_t)j * lda];
    }
    __syncthreads();
    for (int ii = threadIdx.y; ii < TB; ii += TROWS) {
        const int j = j0 + threadIdx.x, i = i0 + ii;
        if (i < m && j < n) at[j + (int64_t)i * ldat] = conj_if(blk[threadIdx.x][ii], cj);
    }
}

// A (n x n) = op(A) in place: block (bi, bj), bi >= bj, is exchanged with block (bj, bi)
template <typename T>
__global__ void transpose_in_kernel(int n, bool cj, T* A, T* const* A_array, int64_t lda)
{
    if (blockIdx.y > blockIdx.x) return;
    __shared__ T lo[TB][TB + 1];
    __shared__ T up[TB][TB + 1];
    T* a = A_array ? A_array[blockIdx.z] : A;
    const int i0 = blockIdx.x * TB, j0 = blockIdx.y * TB;   // lower block rows i0.., cols j0..
    const bool diag = (blockIdx.x == blockIdx.y);
    for (int jj = threadIdx.y; jj < TB; jj += TROWS) {
        const int i = i0 + threadIdx.x, j = j0 + jj;
        if (i < n && j < n) lo[jj][threadIdx.x] = a[i + (int64_t)j * lda];
        if (!diag) {
            const int iu = j0 + threadIdx.x, ju = i0 + jj;   // upper block rows j0.., cols i0..
            if (iu < n && ju < n) up[jj][threadIdx.x] = a[iu + (int64_t)ju * lda];
        }
    }
    __syncthreads();
    for (int jj = threadIdx.y; jj < TB; jj += TROWS) {
        // upper block (iu, ju) takes lower (ju, iu) = lo[iu - j0][ju - i0]
        const int iu = j0 + threadIdx.x, ju = i0 + jj;
        if (iu < n && ju < n) a[iu + (int64_t)ju * lda] = conj_if(lo[threadIdx.x][jj], cj);
        if (!diag) {
            // lower block (i, j) takes upper (j, i) = up[i - i0][j - j0]
            const int i = i0 + threadIdx.x, j = j0 + jj;
            if (i < n && j < n) a[i + (int64_t)j * lda] = conj_if(up[threadIdx.x][jj], cj);
        }
    }
}

__device__ __forceinline__ void drop_imag(double&) {}
__device__ __forceinline__ void drop_imag(double2& x) { x.y = 0.0; }

// Fill the other triangle of a square tile from the stored one: A(j,i) = op(A(i,j)).
// herm: op conjugates, and the imaginary part of the diagonal is dropped.
// Block (bx, by), bx >= by, names a block pair; the source block is the one in the stored
// triangle, the destination its mirror image (the same block for bx == by).
template <typename T>
__global__ void symmetrize_kernel(int n, bool from_lower, bool herm, T* A, int64_t lda)
{
    if (blockIdx.y > blockIdx.x) return;
    __shared__ T blk[TB][TB + 1];
    const int r0 = blockIdx.x * TB, c0 = blockIdx.y * TB;
    const int src_r0 = from_lower ? r0 : c0, src_c0 = from_lower ? c0 : r0;
    const bool diag = (blockIdx.x == blockIdx.y);
    for (int jj = threadIdx.y; jj < TB; jj += TROWS) {
        const int i = src_r0 + threadIdx.x, j = src_c0 + jj;
        if (i < n && j < n) blk[jj][threadIdx.x] = A[i + (int64_t)j * lda];
    }
    __syncthreads();
    for (int jj = threadIdx.y; jj < TB; jj += TROWS) {
        // destination (i, j) mirrors source (j, i), which sits in blk[i - src_c0][j - src_r0]
        const int i = src_c0 + threadIdx.x, j = src_r0 + jj;
        if (i >= n || j >= n) continue;
        if (diag) {
            const bool other = from_lower ? (i < j) : (i > j);
            if (!other) {
                if (herm && i == j) drop_imag(A[i + (int64_t)j * lda]);
                continue;
            }
        }
        A[i + (int64_t)j * lda] = conj_if(blk[threadIdx.x][jj], herm);
    }
}

}  // namespace te

// ---------------------------------------------------------------------------- C ABI
namespace {

template <typename T>
int transpose_any(cudaStream_t stream, bool cj, int m, int n, const void* A, const void* const* A_array,
                  int64_t lda, void* AT, void* const* AT_array, int64_t ldat, int batch, bool in_place)
{
    const dim3 threads(te::TB, te::TROWS);
    if (in_place) {
        const dim3 grid((n + te::TB - 1) / te::TB, (n + te::TB - 1) / te::TB, batch);
        te::transpose_in_kernel<T><<<grid, threads, 0, stream>>>(n, cj, (T*)AT, (T* const*)AT_array, lda);
    }
    else {
        const dim3 grid((m + te::TB - 1) / te::TB, (n + te::TB - 1) / te::TB, batch);
        te::transpose_out_kernel<T><<<grid, threads, 0, stream>>>(
            m, n, cj, (const T*)A, (const T* const*)A_array, lda, (T*)AT, (T* const*)AT_array, ldat);
    }
    te::count_launch();
    TE_CUDA_CHECK(cudaGetLastError());
    return TE_OK;
}

}  // namespace

extern "C" int te_tile_transpose(int device, void* stream, int dtype, int conj, int m, int n,
                                 const void* A, int64_t lda, void* AT, int64_t ldat)
{
    if (!A || !AT || m < 0 || n < 0 || lda < m || ldat < n) return TE_ERR_ARG;
    const bool in_place = (A == AT);
    if (in_place && (m != n || lda != ldat)) return TE_ERR_ARG;
    if (m == 0 || n == 0) return TE_OK;
    int st = te::set_device(device);
    if (st != TE_OK) return st;
    if (dtype == TE_C64)
        return transpose_any<double2>((cudaStream_t)stream, conj != 0, m, n, A, nullptr, lda, AT, nullptr,
                                      ldat, 1, in_place);
    return transpose_any<double>((cudaStream_t)stream, false, m, n, A, nullptr, lda, AT, nullptr, ldat, 1,
                                 in_place);
}

extern "C" int te_tile_transpose_batch(int device, void* stream, int dtype, int conj, int m, int n,
                                       const void* const* A_array_dev, int64_t lda,
                                       void* const* AT_array_dev, int64_t ldat, int batch)
{
    if (!A_array_dev || !AT_array_dev || m < 0 || n < 0 || lda < m || ldat < n || batch < 0 ||
        batch > 65535)
        return TE_ERR_ARG;
    const bool in_place = ((const void*)A_array_dev == (const void*)AT_array_dev);
    if (in_place && (m != n || lda != ldat)) return TE_ERR_ARG;
    if (m == 0 || n == 0 || batch == 0) return TE_OK;
    int st = te::set_device(device);
    if (st != TE_OK) return st;
    if (dtype == TE_C64)
        return transpose_any<double2>((cudaStream_t)stream, conj != 0, m, n, nullptr, A_array_dev, lda,
                                      nullptr, AT_array_dev, ldat, batch, in_place);
    return transpose_any<double>((cudaStream_t)stream, false, m, n, nullptr, A_array_dev, lda, nullptr,
                                 AT_array_dev, ldat, batch, in_place);
}

extern "C" int te_tile_symmetrize(int device, void* stream, int dtype, int uplo, int hermitian, int n,
                                  void* A, int64_t lda)
{
    if (!A || n < 0 || lda < n || (uplo != TE_UPLO_LOWER && uplo != TE_UPLO_UPPER)) return TE_ERR_ARG;
    if (n == 0) return TE_OK;
    int st = te::set_device(device);
    if (st != TE_OK) return st;
    const dim3 threads(te::TB, te::TROWS);
    const dim3 grid((n + te::TB - 1) / te::TB, (n + te::TB - 1) / te::TB);
    if (dtype == TE_C64)
        te::symmetrize_kernel<double2><<<grid, threads, 0, (cudaStream_t)stream>>>(
            n, uplo == TE_UPLO_LOWER, hermitian != 0, (double2*)A, lda);
    else
        te::symmetrize_kernel<double><<<grid, threads, 0, (cudaStream_t)stream>>>(
            n, uplo == TE_UPLO_LOWER, false, (double*)A, lda);
    te::count_launch();
    TE_CUDA_CHECK(cudaGetLastError());
    return TE_OK;
}
