// te_aux.cu -- small tile kernels used by the harness around the hot path:
// the matgen generator (so device-origin matrices need no 32 GiB host staging),
// tile add / sum of squares / triangle masking for the residual checks.
//
// Reference counterparts: matgen/random.cc:53-159 (generator),
// src/cuda/device_geadd.cu, device_genorm.cu, device_tzset.cu (aux tile kernels).
#include "te_common.cuh"

namespace te {

// matgen/random.cc:53-75: Philox-2x64, 10 rounds, counter (i, j), key = seed.
__device__ __forceinline__ void philox_2x64(uint64_t c0, uint64_t c1, uint64_t key,
                                            uint64_t& o0, uint64_t& o1)
{
    const uint64_t bump = 0xD2B74407B1CE6E93ull;
    const uint64_t mult = 0x9E3779B97F4A7C15ull;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        if (r != 0) key += bump;
        const uint64_t lo = c1 * mult;
        const uint64_t hi = __umul64hi(c1, mult);
        const uint64_t L = c0;
        c0 = lo;
        c1 = hi ^ key ^ L;
    }
    o0 = c0;
    o1 = c1;
}

// matgen/random.cc:80-89
__device__ __forceinline__ double bits_to_real(uint64_t b)
{
    return (double)(b >> 11) * (1.0 / 9007199254740992.0);
}

template <bool CPLX>
__global__ void matgen_kernel(double* A, int64_t lda, int m, int n, int64_t ioff, int64_t joff,
                              uint64_t seed, int dist, double diag_add)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y;
    if (i >= m || j >= n) return;
    uint64_t b0, b1;
    philox_2x64((uint64_t)(i + ioff), (uint64_t)(j + joff), seed, b0, b1);
    double re = bits_to_real(b0), im = bits_to_real(b1);
    if (dist == 2) { re = 2 * re - 1; im = 2 * im - 1; }
    if (i + ioff == j + joff) re += diag_add;        // generate_type_rand.hh:70-78 (dominant)
    if (CPLX) {
        A[2 * (i + j * lda)] = re;
        A[2 * (i + j * lda) + 1] = im;
    }
    else {
        A[i + j * lda] = re;
    }
}

// B = alpha A + beta B (real view: complex tiles pass 2m rows)
__global__ void tile_add_kernel(int m, int n, double alpha, const double* A, int64_t lda,
                                double beta, double* B, int64_t ldb)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y;
    if (i >= m || j >= n) return;
    B[i + j * ldb] = alpha * A[i + j * lda] + beta * B[i + j * ldb];
}

// A = beta A, beta = (br, bi); esz = doubles per element.  beta == 0 stores zeros whatever A
// held (NaN and Inf included), as a BLAS beta = 0 does.
__global__ void tile_scale_kernel(int m, int n, int esz, double br, double bi, double* A, int64_t lda)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y;
    if (i >= m || j >= n) return;
    double* a = A + (int64_t)esz * (i + j * lda);
    if (br == 0.0 && bi == 0.0) {
        a[0] = 0.0;
        if (esz == 2) a[1] = 0.0;
    }
    else if (esz == 1) {
        a[0] = br * a[0];
    }
    else {
        const double re = a[0], im = a[1];
        a[0] = br * re - bi * im;
        a[1] = br * im + bi * re;
    }
}

// *accum += sum |a_ij|^2 over the masked part (mask 0 full, 1 lower incl. diag, 2 upper incl.
// diag, 3 strictly lower).  esz = doubles per element.
__global__ void tile_sumsq_kernel(int m, int n, int esz, int mask, const double* A, int64_t lda,
                                  double* accum)
{
    double s = 0;
    for (int j = blockIdx.y; j < n; j += gridDim.y) {
        for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < m; i += gridDim.x * blockDim.x) {
            const bool in = (mask == 0) || (mask == 1 && i >= j) || (mask == 2 && i <= j) ||
                            (mask == 3 && i > j);
            if (in) {
                for (int e = 0; e < esz; ++e) {
                    const double v = A[esz * (i + j * lda) + e];
                    s += v * v;
                }
            }
        }
    }
    for (int off = 16; off > 0; off >>= 1)
        s += __shfl_xor_sync(0xffffffffu, s, off);
    if ((threadIdx.x & 31) == 0 && s != 0.0)
        atomicAdd(accum, s);
}

// zero the strict upper (keep 1) or strict lower (keep 2) part of a tile
__global__ void tile_keep_triangle_kernel(int m, int n, int esz, int keep, double* A, int64_t lda)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y;
    if (i >= m || j >= n) return;
    const bool zero = (keep == 1) ? (i < j) : (i > j);
    if (zero)
        for (int e = 0; e < esz; ++e) A[esz * (i + j * lda) + e] = 0.0;
}

}  // namespace te

extern "C" int te_matgen_tile(int device, void* stream, int dtype, int dist, int64_t seed,
                              int m, int n, int64_t ioff, int64_t joff, void* A, int64_t lda,
                              double diag_add)
{
    if (!A || m < 0 || n < 0 || (dist != 1 && dist != 2)) return TE_ERR_ARG;
    if (m == 0 || n == 0) return TE_OK;
    int st = te::set_device(device);
    if (st != TE_OK) return st;
    dim3 grid((m + 127) / 128, n);
    if (dtype == TE_C64)
        te::matgen_kernel<true><<<grid, 128, 0, (cudaStream_t)stream>>>(
            (double*)A, lda, m, n, ioff, joff, (uint64_t)seed, dist, diag_add);
    else
        te::matgen_kernel<false><<<grid, 128, 0, (cudaStream_t)stream>>>(
            (double*)A, lda, m, n, ioff, joff, (uint64_t)seed, dist, diag_add);
    te::count_launch();
    TE_CUDA_CHECK(cudaGetLastError());
    return TE_OK;
}

extern "C" int te_tile_add(int device, void* stream, int dtype, int m, int n, double alpha,
                           const void* A, int64_t lda, double beta, void* B, int64_t ldb)
{
    if (!A || !B || m < 0 || n < 0) return TE_ERR_ARG;
    if (m == 0 || n == 0) return TE_OK;
    int st = te::set_device(device);
    if (st != TE_OK) return st;
    const int esz = (dtype == TE_C64) ? 2 : 1;
    dim3 grid((m * esz + 127) / 128, n);
    te::tile_add_kernel<<<grid, 128, 0, (cudaStream_t)stream>>>(
        m * esz, n, alpha, (const double*)A, lda * esz, beta, (double*)B, ldb * esz);
    te::count_launch();
    TE_CUDA_CHECK(cudaGetLastError());
    return TE_OK;
}

extern "C" int te_tile_scale(int device, void* stream, int dtype, int m, int n, const void* beta,
                             void* A, int64_t lda)
{
    if (!A || !beta || m < 0 || n < 0 || lda < m) return TE_ERR_ARG;
    if (m == 0 || n == 0) return TE_OK;
    int st = te::set_device(device);
    if (st != TE_OK) return st;
    const double* b = (const double*)beta;
    dim3 grid((m + 127) / 128, n);
    te::tile_scale_kernel<<<grid, 128, 0, (cudaStream_t)stream>>>(
        m, n, (dtype == TE_C64) ? 2 : 1, b[0], (dtype == TE_C64) ? b[1] : 0.0, (double*)A, lda);
    te::count_launch();
    TE_CUDA_CHECK(cudaGetLastError());
    return TE_OK;
}

extern "C" int te_tile_sumsq(int device, void* stream, int dtype, int mask, int m, int n,
                             const void* A, int64_t lda, double* accum_dev)
{
    if (!A || !accum_dev || m < 0 || n < 0 || mask < 0 || mask > 3) return TE_ERR_ARG;
    if (m == 0 || n == 0) return TE_OK;
    int st = te::set_device(device);
    if (st != TE_OK) return st;
    dim3 grid((m + 127) / 128 > 8 ? 8 : (m + 127) / 128, n > 64 ? 64 : n);
    te::tile_sumsq_kernel<<<grid, 128, 0, (cudaStream_t)stream>>>(
        m, n, (dtype == TE_C64) ? 2 : 1, mask, (const double*)A, lda, accum_dev);
    te::count_launch();
    TE_CUDA_CHECK(cudaGetLastError());
    return TE_OK;
}

extern "C" int te_tile_keep_triangle(int device, void* stream, int dtype, int uplo, int m, int n,
                                     void* A, int64_t lda)
{
    if (!A || m < 0 || n < 0 || (uplo != TE_UPLO_LOWER && uplo != TE_UPLO_UPPER)) return TE_ERR_ARG;
    if (m == 0 || n == 0) return TE_OK;
    int st = te::set_device(device);
    if (st != TE_OK) return st;
    dim3 grid((m + 127) / 128, n);
    te::tile_keep_triangle_kernel<<<grid, 128, 0, (cudaStream_t)stream>>>(
        m, n, (dtype == TE_C64) ? 2 : 1, uplo == TE_UPLO_LOWER ? 1 : 2, (double*)A, lda);
    te::count_launch();
    TE_CUDA_CHECK(cudaGetLastError());
    return TE_OK;
}

// ---------------------------------------------------------------- panel <-> tiles
// A run of host-origin tiles that lie next to each other in the caller's array (a block column
// of a ScaLAPACK-layout local matrix: tiles stacked in rows; a block row: tiles side by side)
// crosses PCIe as ONE copy into / out of a device panel; these kernels move the panel's pieces
// to / from the tiles' own blocks.  Why: every cudaMemcpy(2D)Async between host and device costs
// the enqueueing host thread ~21 us whatever its size (scripts/copy_batch_probe.py), so
// fetching a slab of 2048 tiles one by one took 43 ms of host time per direction and the host,
// not PCIe, set the pace of the host-origin gemm.
namespace te {
struct PanelPieces {
    te_panel_piece piece[TE_MAX_PANEL_PIECES];
    int count;
};

// blockIdx.x = piece, blockIdx.y = slice of its columns.  d = doubles per element.
template <bool TO_TILES>
__global__ void panel_pieces_kernel(double* __restrict__ panel, int64_t ld_panel, int d, const __grid_constant__ PanelPieces pp)
{
    const te_panel_piece& pc = pp.piece[blockIdx.x];
    const int64_t rows = (int64_t)pc.m * d;                // doubles per column
    double* tile = (double*)pc.tile;
    const int64_t ldt = (int64_t)pc.ld * d;
    double* pan = panel + pc.panel_offset * d;
    const int64_t ldp = ld_panel * d;
    const bool vec = ((rows | ldt | ldp) & 1) == 0 && (((uintptr_t)tile | (uintptr_t)pan) & 15) == 0;
    for (int j = blockIdx.y; j < pc.n; j += gridDim.y) {
        double* t = tile + (int64_t)j * ldt;
        double* q = pan + (int64_t)j * ldp;
        if (vec) {
            for (int64_t i = threadIdx.x; i < rows / 2; i += blockDim.x) {
                if (TO_TILES) reinterpret_cast<double2*>(t)[i] = reinterpret_cast<const double2*>(q)[i];
                else reinterpret_cast<double2*>(q)[i] = reinterpret_cast<const double2*>(t)[i];
            }
        }
        else {
            for (int64_t i = threadIdx.x; i < rows; i += blockDim.x) {
                if (TO_TILES) t[i] = q[i];
                else q[i] = t[i];
            }
        }
    }
}
}  // namespace te

extern "C" int te_panel_pieces(int device, void* stream, int dtype, int to_tiles, void* panel,
                               int64_t ld_panel, const te_panel_piece* pieces_host, int count)
{
    if (!panel || count < 0 || (count > 0 && !pieces_host) || ld_panel < 1) return TE_ERR_ARG;
    int st = te::set_device(device);
    if (st != TE_OK) return st;
    const int d = (dtype == TE_C64) ? 2 : 1;
    for (int first = 0; first < count; first += TE_MAX_PANEL_PIECES) {
        te::PanelPieces pp;
        pp.count = count - first < TE_MAX_PANEL_PIECES ? count - first : TE_MAX_PANEL_PIECES;
        int max_n = 1;
        for (int i = 0; i < pp.count; ++i) {
            pp.piece[i] = pieces_host[first + i];
            if (pp.piece[i].m < 0 || pp.piece[i].n < 0 || !pp.piece[i].tile) return TE_ERR_ARG;
            if (pp.piece[i].n > max_n) max_n = pp.piece[i].n;
        }
        dim3 grid(pp.count, max_n < 64 ? max_n : 64);
        if (to_tiles)
            te::panel_pieces_kernel<true><<<grid, 128, 0, (cudaStream_t)stream>>>((double*)panel, ld_panel, d, pp);
        else
            te::panel_pieces_kernel<false><<<grid, 128, 0, (cudaStream_t)stream>>>((double*)panel, ld_panel, d, pp);
        te::count_launch();
        TE_CUDA_CHECK(cudaGetLastError());
    }
    return TE_OK;
}
