// te_memory_binding.cc -- slate::Memory, the device-tile pool, bound to the tile engine.
//
// Compiled IN PLACE OF the reference's src/core/Memory.cc.  The class (and its inline
// available() / capacity() / allocated(), which read free_blocks_ and capacity_) is the
// reference's own include/slate/internal/Memory.hh, unmodified, so the free list stays where the
// header puts it; what changes is where device memory comes from and how the pool grows:
//   * slabs come from te_device_malloc (tile_engine.h) instead of blas::device_malloc;
//   * when the free list is empty the reference cudaMallocs ONE block (Memory.cc:118-141,
//     196-202), a synchronous call in the middle of the task DAG; here growth mallocs
//     max(16, capacity / 8) blocks at once and keeps the rest as spares, which later growth
//     steps take one at a time.  capacity() still counts blocks in circulation, so
//     unit_test/test_Memory.cc's expectations (capacity grows by one per extra alloc) hold.
#include "auxiliary/Debug.hh"
#include "slate/internal/Memory.hh"
#include "slate/Exception.hh"

#include "tile_engine.h"

#include <map>
#include <mutex>
#include <utility>
#include <vector>

namespace slate {

namespace {
std::mutex g_spare_mu;
std::map<std::pair<const Memory*, int>, std::vector<void*>> g_spares;

void* te_slab(int device, size_t bytes)
{
    void* p = nullptr;
    if (te_device_malloc(device, bytes, &p) != TE_OK)
        throw Exception(std::string("te_device_malloc failed: ") + te_last_error(), __func__, __FILE__, __LINE__);
    return p;
}
}  // namespace

int Memory::num_devices_;
Memory::StaticConstructor Memory::static_constructor_;

Memory::Memory(size_t block_size)
    : block_size_((block_size + 255) & ~(size_t)255),     // tiles stay on the 16-byte cp.async path
      free_blocks_(num_devices_), allocated_mem_(num_devices_), capacity_(num_devices_)
{}

Memory::~Memory()
{
    for (int device = 0; device < num_devices_; ++device)
        assert(capacity_[device] == 0);
    std::lock_guard<std::mutex> lock(g_spare_mu);
    for (int device = 0; device < num_devices_; ++device)
        g_spares.erase({this, device});
}

void Memory::addDeviceBlocks(int device, int64_t num_blocks, blas::Queue* queue)
{
    if (num_blocks <= 0) return;
    uint8_t* dev_mem = (uint8_t*)allocDeviceMemory(device, block_size_ * num_blocks, queue);
    capacity_[device] += num_blocks;
    // pushed in reverse so that alloc() hands blocks out in ascending address order
    for (int64_t i = num_blocks - 1; i >= 0; --i)
        free_blocks_[device].push(dev_mem + i * block_size_);
}

void Memory::clearDeviceBlocks(int device, blas::Queue* queue)
{
    while (!free_blocks_[device].empty())
        free_blocks_[device].pop();
    {
        std::lock_guard<std::mutex> lock(g_spare_mu);
        g_spares.erase({this, device});
    }
    while (!allocated_mem_[device].empty()) {
        freeDeviceMemory(device, allocated_mem_[device].top(), queue);
        allocated_mem_[device].pop();
    }
    capacity_[device] = 0;
    Debug::checkDeviceMemoryLeaks(*this, device);
}

void* Memory::alloc(int device, size_t size, blas::Queue* queue)
{
    void* block;
    if (device == HostNum) {
        block = new char[size];
    }
    else {
        slate_assert(size <= block_size_);
        #pragma omp critical(slate_memory)
        {
            if (free_blocks_[device].size() > 0) {
                block = free_blocks_[device].top();
                free_blocks_[device].pop();
            }
            else {
                block = allocBlock(device, queue);
            }
        }
    }
    return block;
}

void Memory::free(void* block, int device)
{
    if (device == HostNum) {
        delete[] (char*)block;
    }
    else {
        #pragma omp critical(slate_memory)
        {
            free_blocks_[device].push(block);
        }
    }
}

/// One more block in circulation; the cudaMalloc behind it covers a whole chunk.
void* Memory::allocBlock(int device, blas::Queue* queue)
{
    if (device == HostNum)
        return allocHostMemory(block_size_);
    std::lock_guard<std::mutex> lock(g_spare_mu);
    std::vector<void*>& spare = g_spares[{this, device}];
    if (spare.empty()) {
        const size_t chunk = capacity_[device] / 8 > 16 ? capacity_[device] / 8 : 16;
        uint8_t* dev_mem = (uint8_t*)allocDeviceMemory(device, block_size_ * chunk, queue);
        for (size_t i = chunk; i-- > 0;)
            spare.push_back(dev_mem + i * block_size_);
    }
    void* block = spare.back();
    spare.pop_back();
    capacity_[device] += 1;
    return block;
}

void* Memory::allocHostMemory(size_t size)
{
    slate_not_implemented("Memory pool currently doesn't handle host memory");
    return nullptr;
}

void* Memory::allocDeviceMemory(int device, size_t size, blas::Queue* queue)
{
    void* dev_mem = te_slab(device, size);
    allocated_mem_[device].push(dev_mem);
    return dev_mem;
}

void Memory::freeHostMemory(void* host_mem) { std::free(host_mem); }

void Memory::freeDeviceMemory(int device, void* dev_mem, blas::Queue* queue)
{
    te_device_free(device, dev_mem);
}

}  // namespace slate
