// te_peak.cu -- register-resident FP64 issue-rate microbenchmarks.
//
// The driver-written MEASURED_PEAKS.json has no FP64 figure, so the roofline
// denominator of every FP64 kernel here is measured by this file: a loop of
// independent DMMA.8x8x4 (kind 0) or DFMA (kind 1) chains with no memory
// traffic, on every SM, timed with CUDA events.
#include "te_common.cuh"

namespace te {

template <int CHAINS>
__global__ void peak_dmma_kernel(int iters, double seed, double* sink)
{
    double c[CHAINS][2];
    double a = seed + threadIdx.x * 1e-9, b = 1.0 - seed;
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) { c[i][0] = i; c[i][1] = -i; }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < CHAINS; ++i)
            dmma884(c[i][0], c[i][1], a, b);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) s += c[i][0] + c[i][1];
    if (s == 123.456) sink[0] = s;   // never true; keeps the chains alive
}

template <int CHAINS>
__global__ void peak_dfma_kernel(int iters, double seed, double* sink)
{
    double c[CHAINS];
    double a = seed + threadIdx.x * 1e-9, b = 1.0 - seed;
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) c[i] = i;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < CHAINS; ++i)
            c[i] = fma(a, c[i], b);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) s += c[i];
    if (s == 123.456) sink[0] = s;
}

template <int CHAINS>
static void launch_peak(int kind, int grid, int threads, int iters, double* sink)
{
    if (kind == 0)
        peak_dmma_kernel<CHAINS><<<grid, threads>>>(iters, 0.25, sink);
    else
        peak_dfma_kernel<CHAINS><<<grid, threads>>>(iters, 0.25, sink);
    te::count_launch();
}

}  // namespace te

extern "C" int te_peak_fp64_ex(int device, int kind, int iters, int warps_per_cta, int ctas_per_sm,
                               int chains, double* tflops)
{
    using namespace te;
    if (!tflops || iters <= 0 || (kind != 0 && kind != 1) || warps_per_cta < 1 ||
        warps_per_cta > 32 || ctas_per_sm < 1 || ctas_per_sm > 32)
        return TE_ERR_ARG;
    int st = set_device(device);
    if (st != TE_OK) return st;
    double* sink = nullptr;
    TE_CUDA_CHECK(cudaMalloc(&sink, sizeof(double)));
    const int sms = num_sms(device);
    const int grid = sms * ctas_per_sm;
    const int threads = warps_per_cta * 32;
    cudaEvent_t e0, e1;
    TE_CUDA_CHECK(cudaEventCreate(&e0));
    TE_CUDA_CHECK(cudaEventCreate(&e1));
    float best_ms = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {       // rep 0-1 warm up clocks
        TE_CUDA_CHECK(cudaEventRecord(e0, 0));
        switch (chains) {
            case 1:  launch_peak<1>(kind, grid, threads, iters, sink); break;
            case 2:  launch_peak<2>(kind, grid, threads, iters, sink); break;
            case 4:  launch_peak<4>(kind, grid, threads, iters, sink); break;
            case 8:  launch_peak<8>(kind, grid, threads, iters, sink); break;
            case 16: launch_peak<16>(kind, grid, threads, iters, sink); break;
            case 32: launch_peak<32>(kind, grid, threads, iters, sink); break;
            default: cudaFree(sink); return TE_ERR_ARG;
        }
        TE_CUDA_CHECK(cudaEventRecord(e1, 0));
        TE_CUDA_CHECK(cudaEventSynchronize(e1));
        TE_CUDA_CHECK(cudaGetLastError());
        float ms = 0;
        TE_CUDA_CHECK(cudaEventElapsedTime(&ms, e0, e1));
        if (rep >= 2 && ms < best_ms) best_ms = ms;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(sink);
    const double warps = (double)grid * warps_per_cta;
    double flops;
    if (kind == 0)
        flops = warps * (double)iters * chains * (2.0 * 8 * 8 * 4);
    else
        flops = warps * 32.0 * (double)iters * chains * 2.0;
    *tflops = flops / (best_ms * 1e-3) * 1e-12;
    return TE_OK;
}

extern "C" int te_peak_fp64(int device, int kind, int iters, double* tflops)
{
    return te_peak_fp64_ex(device, kind, iters, 8, 4, 16, tflops);
}
