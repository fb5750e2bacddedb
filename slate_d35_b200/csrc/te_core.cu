// te_core.cu -- error text, device helpers and identification of the tile engine.
#include "te_common.cuh"
#include <string.h>
#include <atomic>

namespace te {

static thread_local char g_err[512] = "";

long long launches();

void set_last_error(const char* what, const char* detail)
{
    snprintf(g_err, sizeof(g_err), "%s: %s", what ? what : "?", detail ? detail : "?");
}

int num_sms(int device)
{
    static int cached[64] = {};
    if (device >= 0 && device < 64 && cached[device] > 0)
        return cached[device];
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || n <= 0)
        n = 148;
    if (device >= 0 && device < 64)
        cached[device] = n;
    return n;
}

static std::atomic<long long> g_launches{0};
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
long long launches() { return g_launches.load(std::memory_order_relaxed); }

int set_device(int device)
{
    TE_CUDA_CHECK(cudaSetDevice(device));
    return TE_OK;
}

}  // namespace te

extern "C" const char* te_last_error(void) { return te::g_err; }

extern "C" const char* te_version(void) { return "tile_engine 0.1 sm_100a"; }

extern "C" long long te_launch_count(void) { return te::launches(); }

extern "C" int te_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}
