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

#include <algorithm>
#include <map>
#include <mutex>
#include <utility>
#include <vector>

namespace slate {

namespace {

// One lock for every pool of the process: pool operations are a handful of pointer moves, and
// the reference serialises them too (omp critical(slate_memory)).
std::mutex g_pool_mu;
std::mutex g_slab_mu;        // the list of slabs (reserve and growth can run on different threads)

// Blocks that a growth step malloc'ed ahead and that are not in circulation yet, per pool and
// device.  Kept outside the class because its data members are fixed by the reference's header.
std::map<std::pair<const Memory*, int>, std::vector<void*>> g_ahead;

void* engine_malloc(int device, size_t bytes)
{
    void* p = nullptr;
    const int st = te_device_malloc(device, bytes, &p);
    if (st != TE_OK)
        throw Exception(std::string("tile engine: te_device_malloc of ") + std::to_string(bytes)
                        + " bytes failed: " + te_last_error(), __func__, __FILE__, __LINE__);
    return p;
}

}  // namespace

int Memory::num_devices_;
Memory::StaticConstructor Memory::static_constructor_;

// Block size rounded up to 256 bytes: every tile keeps the 16-byte cp.async fast path.
Memory::Memory(size_t block_size)
    : block_size_((block_size + 255) / 256 * 256)
{
    free_blocks_.resize(num_devices_);
    allocated_mem_.resize(num_devices_);
    capacity_.assign(num_devices_, 0);
}

// As in the reference, device slabs must have been given back (clearDeviceBlocks) by now.
Memory::~Memory()
{
    std::lock_guard<std::mutex> lock(g_pool_mu);
    for (int dev = 0; dev < num_devices_; ++dev) {
        assert(capacity_[dev] == 0);
        g_ahead.erase(std::make_pair((const Memory*)this, dev));
    }
}

// Reserve: ONE slab of num_blocks blocks; the free list hands them out in ascending address order,
// so the tiles of a device-origin matrix lie next to each other.
void Memory::addDeviceBlocks(int device, int64_t num_blocks, blas::Queue* queue)
{
    if (num_blocks < 1) return;
    char* slab = static_cast<char*>(allocDeviceMemory(device, block_size_ * (size_t)num_blocks, queue));
    std::lock_guard<std::mutex> lock(g_pool_mu);
    for (int64_t b = num_blocks; b-- > 0;)
        free_blocks_[device].push(slab + (size_t)b * block_size_);
    capacity_[device] += (size_t)num_blocks;
}

void Memory::clearDeviceBlocks(int device, blas::Queue* queue)
{
    {
        std::lock_guard<std::mutex> lock(g_pool_mu);
        free_blocks_[device] = std::stack<void*>();
        g_ahead.erase(std::make_pair((const Memory*)this, device));
        capacity_[device] = 0;
    }
    for (; !allocated_mem_[device].empty(); allocated_mem_[device].pop())
        freeDeviceMemory(device, allocated_mem_[device].top(), queue);
    Debug::checkDeviceMemoryLeaks(*this, device);
}

void* Memory::alloc(int device, size_t size, blas::Queue* queue)
{
    if (device == HostNum)
        return new char[size];                       // the reference keeps no host pool either
    slate_assert(size <= block_size_);
    {
        std::lock_guard<std::mutex> lock(g_pool_mu);
        std::stack<void*>& idle = free_blocks_[device];
        if (!idle.empty()) {
            void* block = idle.top();
            idle.pop();
            return block;
        }
    }
    return allocBlock(device, queue);
}

void Memory::free(void* block, int device)
{
    if (device == HostNum) {
        delete[] static_cast<char*>(block);
        return;
    }
    std::lock_guard<std::mutex> lock(g_pool_mu);
    free_blocks_[device].push(block);
}

// Growth.  capacity() rises by ONE per call, as unit_test/test_Memory.cc expects of the reference,
// but the device malloc behind it covers max(16, capacity / 8) blocks; the others wait in g_ahead.
void* Memory::allocBlock(int device, blas::Queue* queue)
{
    if (device == HostNum)
        return allocHostMemory(block_size_);
    std::lock_guard<std::mutex> lock(g_pool_mu);
    std::vector<void*>& ahead = g_ahead[std::make_pair((const Memory*)this, device)];
    if (ahead.empty()) {
        const size_t chunk = std::max<size_t>(16, capacity_[device] / 8);
        char* slab = static_cast<char*>(allocDeviceMemory(device, block_size_ * chunk, queue));
        for (size_t b = chunk; b-- > 0;)
            ahead.push_back(slab + b * block_size_);
    }
    void* block = ahead.back();
    ahead.pop_back();
    capacity_[device] += 1;
    return block;
}

void* Memory::allocHostMemory(size_t)
{
    slate_not_implemented("host blocks come from new[] (Memory::alloc), there is no host pool");
    return nullptr;
}

void Memory::freeHostMemory(void* host_mem)
{
    std::free(host_mem);
}

void* Memory::allocDeviceMemory(int device, size_t size, blas::Queue*)
{
    void* slab = engine_malloc(device, size);
    std::lock_guard<std::mutex> lock(g_slab_mu);
    allocated_mem_[device].push(slab);
    return slab;
}

void Memory::freeDeviceMemory(int device, void* dev_mem, blas::Queue*)
{
    te_device_free(device, dev_mem);
}

}  // namespace slate
