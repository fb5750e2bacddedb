// te_runtime.cu -- streams, events, copies, pinned memory and the device-tile pool.
//
// This is the part of the C ABI that stands in for blaspp's blas::Queue /
// device_malloc / device_memcpy (reference: blaspp/include/blas/device.hh:445-726,
// blaspp/src/device_queue.cc) and for slate::Memory
// (reference: include/slate/internal/Memory.hh:29-101, src/core/Memory.cc:17-218),
// so that the host C++ above it needs no CUDA header.
#include "te_common.cuh"
#include <string.h>
#include <stdlib.h>
#include <mutex>
#include <vector>

// ---------------------------------------------------------------- streams / events
extern "C" int te_stream_create(int device, int high_priority, void** stream)
{
    if (!stream) return TE_ERR_ARG;
    int st = te::set_device(device);
    if (st != TE_OK) return st;
    int lo = 0, hi = 0;
    TE_CUDA_CHECK(cudaDeviceGetStreamPriorityRange(&lo, &hi));
    cudaStream_t s;
    TE_CUDA_CHECK(cudaStreamCreateWithPriority(&s, cudaStreamNonBlocking, high_priority ? hi : lo));
    *stream = (void*)s;
    return TE_OK;
}

extern "C" int te_stream_destroy(int device, void* stream)
{
    int st = te::set_device(device);
    if (st != TE_OK) return st;
    TE_CUDA_CHECK(cudaStreamDestroy((cudaStream_t)stream));
    return TE_OK;
}

extern "C" int te_stream_sync(int device, void* stream)
{
    int st = te::set_device(device);
    if (st != TE_OK) return st;
    TE_CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream));
    return TE_OK;
}

extern "C" int te_device_sync(int device)
{
    int st = te::set_device(device);
    if (st != TE_OK) return st;
    TE_CUDA_CHECK(cudaDeviceSynchronize());
    return TE_OK;
}

extern "C" int te_event_create(int device, int timing, void** event)
{
    if (!event) return TE_ERR_ARG;
    int st = te::set_device(device);
    if (st != TE_OK) return st;
    cudaEvent_t e;
    TE_CUDA_CHECK(cudaEventCreateWithFlags(&e, timing ? cudaEventDefault : cudaEventDisableTiming));
    *event = (void*)e;
    return TE_OK;
}

extern "C" int te_event_destroy(void* event)
{
    TE_CUDA_CHECK(cudaEventDestroy((cudaEvent_t)event));
    return TE_OK;
}

extern "C" int te_event_record(void* event, void* stream)
{
    TE_CUDA_CHECK(cudaEventRecord((cudaEvent_t)event, (cudaStream_t)stream));
    return TE_OK;
}

extern "C" int te_stream_wait_event(void* stream, void* event)
{
    TE_CUDA_CHECK(cudaStreamWaitEvent((cudaStream_t)stream, (cudaEvent_t)event, 0));
    return TE_OK;
}

extern "C" int te_event_sync(void* event)
{
    TE_CUDA_CHECK(cudaEventSynchronize((cudaEvent_t)event));
    return TE_OK;
}

extern "C" int te_event_elapsed_ms(void* start, void* stop, float* ms)
{
    if (!ms) return TE_ERR_ARG;
    TE_CUDA_CHECK(cudaEventElapsedTime(ms, (cudaEvent_t)start, (cudaEvent_t)stop));
    return TE_OK;
}

// ---------------------------------------------------------------- memory / copies
extern "C" int te_device_malloc(int device, size_t bytes, void** ptr)
{
    if (!ptr) return TE_ERR_ARG;
    int st = te::set_device(device);
    if (st != TE_OK) return st;
    cudaError_t e = cudaMalloc(ptr, bytes);
    if (e == cudaErrorMemoryAllocation) {
        cudaGetLastError();
        te::set_last_error("cudaMalloc", cudaGetErrorString(e));
        return TE_ERR_NOMEM;
    }
    TE_CUDA_CHECK(e);
    return TE_OK;
}

extern "C" int te_device_free(int device, void* ptr)
{
    int st = te::set_device(device);
    if (st != TE_OK) return st;
    TE_CUDA_CHECK(cudaFree(ptr));
    return TE_OK;
}

extern "C" int te_device_mem_info(int device, size_t* free_bytes, size_t* total_bytes)
{
    int st = te::set_device(device);
    if (st != TE_OK) return st;
    TE_CUDA_CHECK(cudaMemGetInfo(free_bytes, total_bytes));
    return TE_OK;
}

extern "C" int te_host_alloc_pinned(size_t bytes, void** ptr)
{
    if (!ptr) return TE_ERR_ARG;
    TE_CUDA_CHECK(cudaHostAlloc(ptr, bytes, cudaHostAllocPortable | cudaHostAllocMapped));
    return TE_OK;
}

extern "C" int te_host_free_pinned(void* ptr)
{
    TE_CUDA_CHECK(cudaFreeHost(ptr));
    return TE_OK;
}

extern "C" int te_host_register(void* ptr, size_t bytes)
{
    TE_CUDA_CHECK(cudaHostRegister(ptr, bytes, cudaHostRegisterPortable));
    return TE_OK;
}

extern "C" int te_host_unregister(void* ptr)
{
    TE_CUDA_CHECK(cudaHostUnregister(ptr));
    return TE_OK;
}

static cudaMemcpyKind to_kind(int kind)
{
    switch (kind) {
        case TE_COPY_H2D: return cudaMemcpyHostToDevice;
        case TE_COPY_D2H: return cudaMemcpyDeviceToHost;
        case TE_COPY_D2D: return cudaMemcpyDeviceToDevice;
        default: return cudaMemcpyDefault;
    }
}

extern "C" int te_memcpy_async(void* dst, const void* src, size_t bytes, int kind, void* stream)
{
    if (bytes == 0) return TE_OK;
    TE_CUDA_CHECK(cudaMemcpyAsync(dst, src, bytes, to_kind(kind), (cudaStream_t)stream));
    return TE_OK;
}

extern "C" int te_memcpy2d_async(void* dst, size_t dpitch, const void* src, size_t spitch,
                                 size_t width_bytes, size_t height, int kind, void* stream)
{
    if (width_bytes == 0 || height == 0) return TE_OK;
    TE_CUDA_CHECK(cudaMemcpy2DAsync(dst, dpitch, src, spitch, width_bytes, height, to_kind(kind),
                                    (cudaStream_t)stream));
    return TE_OK;
}

extern "C" int te_memset_async(void* dst, int value, size_t bytes, void* stream)
{
    TE_CUDA_CHECK(cudaMemsetAsync(dst, value, bytes, (cudaStream_t)stream));
    return TE_OK;
}

// ---------------------------------------------------------------- device-tile pool
// Fixed-size blocks carved from slabs.  reserve() adds ONE slab (a single cudaMalloc)
// holding the requested blocks; alloc() pops a free block and, like the reference
// (Memory.cc:118-141), grows the pool by one block when the free list is empty -- as far
// as available / capacity show.  Underneath, growth does NOT cost a cudaMalloc per block
// (a synchronous call in the middle of a factorisation, the weakness of the reference's
// pool): the pool mallocs max(16, capacity / 8) blocks at once and keeps the rest as
// spares, which later growth steps take one at a time.  Blocks are 256-byte aligned
// (block_bytes is rounded up), so every tile keeps the 16-byte cp.async fast path.
struct te_pool {
    int device;
    size_t block_bytes;
    std::mutex mu;
    std::vector<void*> free_blocks;      // LIFO
    std::vector<void*> spare;            // malloc'ed ahead by growth, not yet in circulation
    std::vector<void*> slabs;
    size_t capacity;
};

extern "C" int te_pool_create(int device, size_t block_bytes, te_pool** pool)
{
    if (!pool || block_bytes == 0) return TE_ERR_ARG;
    te_pool* p = new te_pool;
    p->device = device;
    p->block_bytes = (block_bytes + 255) & ~(size_t)255;
    p->capacity = 0;
    *pool = p;
    return TE_OK;
}

extern "C" int te_pool_reserve(te_pool* p, int64_t nblocks)
{
    if (!p || nblocks < 0) return TE_ERR_ARG;
    if (nblocks == 0) return TE_OK;
    void* slab = nullptr;
    int st = te_device_malloc(p->device, p->block_bytes * (size_t)nblocks, &slab);
    if (st != TE_OK) return st;
    std::lock_guard<std::mutex> lock(p->mu);
    p->slabs.push_back(slab);
    // push in reverse so that alloc() hands blocks out in ascending address order
    for (int64_t i = nblocks - 1; i >= 0; --i)
        p->free_blocks.push_back((char*)slab + (size_t)i * p->block_bytes);
    p->capacity += (size_t)nblocks;
    return TE_OK;
}

extern "C" int te_pool_alloc(te_pool* p, void** block)
{
    if (!p || !block) return TE_ERR_ARG;
    std::lock_guard<std::mutex> lock(p->mu);
    if (!p->free_blocks.empty()) {
        *block = p->free_blocks.back();
        p->free_blocks.pop_back();
        return TE_OK;
    }
    if (p->spare.empty()) {
        // grow: one cudaMalloc for a chunk of blocks (fall back to a single block if the
        // device cannot hold the chunk)
        size_t chunk = p->capacity / 8 > 16 ? p->capacity / 8 : 16;
        void* slab = nullptr;
        int st = te_device_malloc(p->device, p->block_bytes * chunk, &slab);
        if (st != TE_OK) {
            chunk = 1;
            st = te_device_malloc(p->device, p->block_bytes, &slab);
            if (st != TE_OK) return st;
        }
        p->slabs.push_back(slab);
        for (size_t i = chunk; i-- > 0;)
            p->spare.push_back((char*)slab + i * p->block_bytes);
    }
    *block = p->spare.back();
    p->spare.pop_back();
    p->capacity += 1;            // Memory.cc:196-202: capacity counts the blocks in circulation
    return TE_OK;
}

extern "C" int te_pool_free(te_pool* p, void* block)
{
    if (!p || !block) return TE_ERR_ARG;
    std::lock_guard<std::mutex> lock(p->mu);
    p->free_blocks.push_back(block);
    return TE_OK;
}

extern "C" int te_pool_stats(te_pool* p, size_t* available, size_t* capacity, size_t* block_bytes)
{
    if (!p) return TE_ERR_ARG;
    std::lock_guard<std::mutex> lock(p->mu);
    if (available) *available = p->free_blocks.size();
    if (capacity) *capacity = p->capacity;
    if (block_bytes) *block_bytes = p->block_bytes;
    return TE_OK;
}

extern "C" int te_pool_clear(te_pool* p)
{
    if (!p) return TE_ERR_ARG;
    std::lock_guard<std::mutex> lock(p->mu);
    int st = te::set_device(p->device);
    if (st != TE_OK) return st;
    for (void* s : p->slabs)
        TE_CUDA_CHECK(cudaFree(s));
    p->slabs.clear();
    p->free_blocks.clear();
    p->spare.clear();
    p->capacity = 0;
    return TE_OK;
}

extern "C" int te_pool_destroy(te_pool* p)
{
    if (!p) return TE_OK;
    int st = te_pool_clear(p);
    delete p;
    return st;
}

// 1 = everything recorded in the event has finished, 0 = not yet, -1 = error
extern "C" int te_event_query_done(void* event)
{
    cudaError_t e = cudaEventQuery((cudaEvent_t)event);
    if (e == cudaSuccess) return 1;
    if (e == cudaErrorNotReady) return 0;
    te::set_last_error("cudaEventQuery", cudaGetErrorString(e));
    return -1;
}

// ---------------------------------------------------------------- small uploads by kernel
// A descriptor array (a few hundred KB) copied with cudaMemcpyAsync queues on the host-to-
// device copy engine BEHIND whatever bulk tile traffic another stream has already submitted
// (measured: the launch that needed it sat idle for ~65 ms behind a 4 GiB slab of C).  This
// copy is done by a kernel instead: it reads the pinned host buffer directly over PCIe
// (pinned allocations are device-accessible under unified addressing) and is ordered only by
// its own stream.
namespace te {
__global__ void upload_small_kernel(uint4* __restrict__ dst, const uint4* __restrict__ src, size_t n16)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16;
         i += (size_t)gridDim.x * blockDim.x)
        dst[i] = src[i];
}
}  // namespace te

extern "C" int te_upload_small(int device, void* stream, void* dst_dev, const void* src_pinned,
                               size_t bytes)
{
    if (bytes == 0) return TE_OK;
    if (!dst_dev || !src_pinned || (((uintptr_t)dst_dev | (uintptr_t)src_pinned) & 15))
        return TE_ERR_ARG;
    int st = te::set_device(device);
    if (st != TE_OK) return st;
    const size_t n16 = (bytes + 15) / 16;      // buffers are allocated with slack
    int blocks = (int)((n16 + 255) / 256);
    if (blocks > 64) blocks = 64;
    te::upload_small_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>((uint4*)dst_dev,
                                                                     (const uint4*)src_pinned, n16);
    te::count_launch();
    TE_CUDA_CHECK(cudaGetLastError());
    return TE_OK;
}
