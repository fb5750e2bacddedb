// te_common.cuh -- shared device/host helpers of the tile engine (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include "tile_engine.h"

namespace te {

// ---- error plumbing ---------------------------------------------------------
void set_last_error(const char* what, const char* detail);

#define TE_CUDA_CHECK(expr)                                                  \
    do {                                                                     \
        cudaError_t e_ = (expr);                                             \
        if (e_ != cudaSuccess) {                                             \
            te::set_last_error(#expr, cudaGetErrorString(e_));               \
            return TE_ERR_CUDA;                                              \
        }                                                                    \
    } while (0)

int num_sms(int device);          // cached cudaDevAttrMultiProcessorCount
void count_launch(int n = 1);     // kernels launched by this library (te_launch_count)
int set_device(int device);       // cudaSetDevice with status code

// ---- PTX wrappers -----------------------------------------------------------
#ifdef __CUDACC__

__device__ __forceinline__ uint32_t smem_u32(const void* p)
{
    return (uint32_t)__cvta_generic_to_shared(p);
}

// 16-byte async copy global->shared, L2 only (.cg); bytes beyond src_bytes are
// zero-filled.  src must be 16-byte aligned.
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, int src_bytes)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n"
                 :: "r"(dst), "l"(src), "r"(src_bytes) : "memory");
}

// 8-byte async copy (.ca is the only variant for < 16 bytes).
__device__ __forceinline__ void cp_async8(uint32_t dst, const void* src, int src_bytes)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;\n"
                 :: "r"(dst), "l"(src), "r"(src_bytes) : "memory");
}

__device__ __forceinline__ void cp_async_commit()
{
    asm volatile("cp.async.commit_group;\n" ::: "memory");
}

template <int N>
__device__ __forceinline__ void cp_async_wait()
{
    asm volatile("cp.async.wait_group %0;\n" :: "n"(N) : "memory");
}

__device__ __forceinline__ void cp_async_wait_all()
{
    asm volatile("cp.async.wait_all;\n" ::: "memory");
}

// ---- mbarrier (shared::cta) ---------------------------------------------------
__device__ __forceinline__ void mbar_init(uint32_t bar, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" :: "r"(bar), "r"(count) : "memory");
}

__device__ __forceinline__ void mbar_arrive(uint32_t bar)
{
    asm volatile("{\n .reg .b64 st;\n mbarrier.arrive.shared::cta.b64 st, [%0];\n}\n"
                 :: "r"(bar) : "memory");
}

// Blocks until the phase with the given parity has completed.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        " .reg .pred p;\n"
        "WAIT_%=:\n"
        " mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        " @p bra DONE_%=;\n"
        " bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" :: "r"(bar), "r"(parity) : "memory");
}

// The barrier receives one arrival (counted in its expected count, .noinc) when
// all cp.async issued so far by this thread have landed in shared memory.
__device__ __forceinline__ void cp_async_mbar_arrive(uint32_t bar)
{
    asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];\n" :: "r"(bar) : "memory");
}

__device__ __forceinline__ void prefetch_l2(const void* p)
{
    asm volatile("prefetch.global.L2 [%0];\n" :: "l"(p));
}

// FP64 tensor-core MMA, D(8x8) += A(8x4, row) * B(4x8, col).
// Lane l = 4*g + t holds A[g][t], B[t][g], D[g][2t], D[g][2t+1].
// On sm_100a this is one DMMA.8x8x4 SASS instruction.
__device__ __forceinline__ void dmma884(double& d0, double& d1, double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 "
                 "{%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

#endif  // __CUDACC__

}  // namespace te

// ---- scalar helpers shared by the panel kernels (double / complex double) ----
#ifdef __CUDACC__
namespace te {

struct zd { double x, y; };      // interleaved complex double

__host__ __device__ __forceinline__ zd make_zd(double x, double y) { zd r; r.x = x; r.y = y; return r; }
__device__ __forceinline__ zd operator+(zd a, zd b) { return make_zd(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ zd operator-(zd a, zd b) { return make_zd(a.x - b.x, a.y - b.y); }
__device__ __forceinline__ zd operator-(zd a) { return make_zd(-a.x, -a.y); }
__device__ __forceinline__ zd operator*(zd a, zd b)
{
    return make_zd(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}
__device__ __forceinline__ zd operator*(zd a, double s) { return make_zd(a.x * s, a.y * s); }
__device__ __forceinline__ zd conj_(zd a) { return make_zd(a.x, -a.y); }
__device__ __forceinline__ double conj_(double a) { return a; }
__device__ __forceinline__ double real_(zd a) { return a.x; }
__device__ __forceinline__ double real_(double a) { return a; }
__device__ __forceinline__ void set_real(zd& a, double r) { a.x = r; a.y = 0.0; }
__device__ __forceinline__ void set_real(double& a, double r) { a = r; }
template <typename T> __device__ __forceinline__ T zero_();
template <> __device__ __forceinline__ double zero_<double>() { return 0.0; }
template <> __device__ __forceinline__ zd zero_<zd>() { return make_zd(0.0, 0.0); }
template <typename T> __device__ __forceinline__ T from_real(double r);
template <> __device__ __forceinline__ double from_real<double>(double r) { return r; }
template <> __device__ __forceinline__ zd from_real<zd>(double r) { return make_zd(r, 0.0); }
__device__ __forceinline__ double shfl_xor_(double v, int m) { return __shfl_xor_sync(0xffffffffu, v, m); }
__device__ __forceinline__ zd shfl_xor_(zd v, int m)
{
    return make_zd(__shfl_xor_sync(0xffffffffu, v.x, m), __shfl_xor_sync(0xffffffffu, v.y, m));
}

}  // namespace te
#endif
