// te_peer.cu -- tile transport between the box's GPUs on the COPY ENGINES (no SM is used).
//
// Second data plane for the panel broadcasts of BaseMatrix::listBcastMT
//   (reference: tileBcastToSet / tileIbcastToSet, include/slate/BaseMatrix.hh:2283-2387;
//    Tile::isend / irecv, include/slate/Tile.hh:1084-1200),
// next to the NCCL send/recv plane of te_comm.cu.  Why: NCCL's point-to-point kernels need
// SMs, and during a factorisation every SM is held by a CTA of the trailing update, so a
// 190-tile panel broadcast crawled at 11-17 GB/s and every message waited for a CTA to
// retire (profiles/r01_trace_potrf_n4_rank0.txt).  Here a tile travels as one
// cudaMemcpyAsync from the root's tile into a landing zone in the destination GPU's memory
// (mapped into the root's process with CUDA IPC), i.e. on a copy engine over NVLink, and
// "it has arrived" is a sequence number written behind it and waited for with a stream
// memory operation (cuStreamWriteValue64 / cuStreamWaitValue64): no kernel on either side.
//
// One rank per GPU, all ranks on one box.  Per rank:
//   arena     one cudaMalloc'd block: a header of per-root arrival flags, then the landing
//             zones (a ring, reclaimed in order when the destination has copied a message
//             out to its workspace tiles).  Exported once with cudaIpcGetMemHandle; every
//             peer opens it once at te_peer_open.
//   mailbox   POSIX shared memory, one ring per ordered (root, destination) pair, through
//             which the DESTINATION tells the root where in its arena the next message of
//             that pair goes (host to host; the hosts enqueue far ahead of the GPUs, so
//             this wait is off the critical path).
// Protocol for message number s of pair (root -> dest), both sides derive it from the same
// broadcast plan:
//   dest   te_peer_post(root, bytes)        reserve a landing zone, publish its offset
//   root   te_peer_fetch(dest, bytes)       wait (host) for that offset
//          te_peer_put(...) x tiles         copy-engine copies into the landing zone
//          te_peer_signal(dest, s)          flag[root] on dest = s, ordered behind the copies
//   dest   te_peer_wait(root, s)            its stream waits for flag[root] >= s
//          ... copies the tiles out ...     (device-to-device, its own copy engine)
//          te_peer_consumed(root, s)        the landing zone may be reused after this point
#include "te_comm_internal.cuh"

#include <atomic>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <mutex>
#include <thread>
#include <vector>

#include <errno.h>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

namespace te {

constexpr int kBoxRing = 256;            // announcements a destination may be ahead of a root
constexpr size_t kHeaderBytes = 65536;   // arrival flags (8 bytes per root) + scratch

// One direction of one pair, in shared memory.  `posted` is written by the destination only,
// `fetched` by the root only.
struct PairBox {
    std::atomic<uint64_t> posted;
    std::atomic<uint64_t> fetched;
    struct Slot { uint64_t offset, bytes; } slot[kBoxRing];
};
static_assert(std::atomic<uint64_t>::is_always_lock_free, "mailbox counters must be lock free");

struct Segment {
    size_t begin, end;       // [begin, end) of the data region
    int root;
    uint64_t seq;
    cudaEvent_t ev;
    bool recorded;           // te_peer_consumed was called
};

// CUresult (*)(CUstream, CUdeviceptr, cuuint64_t, unsigned int)
typedef int (*StreamValue64Fn)(cudaStream_t, unsigned long long, unsigned long long, unsigned int);

struct PeerState {
    size_t data_bytes = 0;            // landing zones
    size_t max_message = 0;
    char* arena = nullptr;            // header + data, this rank
    std::vector<char*> peer_base;     // the peers' arenas as mapped here (nullptr for self)
    PairBox* boxes = nullptr;         // [root * size + dest]
    size_t shm_bytes = 0;
    std::vector<uint64_t> posted, fetched;   // my side of the counters (per root / per dest)
    std::deque<Segment> live;
    size_t head = 0;
    std::vector<cudaEvent_t> ev_pool;
    StreamValue64Fn wait_value = nullptr, write_value = nullptr;
    bool memop_wait = false, memop_signal = false;
    uint64_t* host_vals = nullptr;    // pinned ring of sequence numbers (signal by copy)
    size_t host_vals_next = 0;
    double timeout_s = 120.0;
    std::mutex mu;
    // statistics
    uint64_t n_put = 0, n_put_bytes = 0, n_signal = 0, n_wait = 0;
};
constexpr size_t kHostVals = 1 << 16;

__global__ void peer_spin_kernel(const volatile unsigned long long* flag, unsigned long long seq)
{
    while (*flag < seq) __nanosleep(100);
    __threadfence_system();
}

static double now_s()
{
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

static inline size_t up256(size_t b) { return (b + 255) & ~(size_t)255; }

// Pops finished segments; with `wait`, blocks (bounded) until the oldest one has finished.
static int reclaim(PeerState* ps, bool wait)
{
    while (!ps->live.empty()) {
        Segment& s = ps->live.front();
        if (!s.recorded) {
            if (wait) {
                set_last_error("te_peer_post", "arena too small: the oldest landing zone belongs to "
                               "a message whose tiles have not been copied out yet");
                return TE_ERR_LIMIT;
            }
            return TE_OK;
        }
        cudaError_t e = cudaEventQuery(s.ev);
        if (e == cudaErrorNotReady) {
            if (!wait) return TE_OK;
            const double t0 = now_s();
            while ((e = cudaEventQuery(s.ev)) == cudaErrorNotReady) {
                if (now_s() - t0 > ps->timeout_s) {
                    set_last_error("te_peer_post", "timed out waiting for a landing zone to drain");
                    return TE_ERR_CUDA;
                }
                std::this_thread::sleep_for(std::chrono::microseconds(20));
            }
            wait = false;
        }
        if (e != cudaSuccess) {
            set_last_error("cudaEventQuery", cudaGetErrorString(e));
            return TE_ERR_CUDA;
        }
        ps->ev_pool.push_back(s.ev);
        ps->live.pop_front();
    }
    return TE_OK;
}

// Ring allocation of `bytes` (a multiple of 256) in the data region.  Zones never wrap around
// the end; `head` is the end of the newest zone, the oldest zone starts at `tail`.  While
// zones are live, head != tail is kept (strict comparisons), so "full" and "empty" differ.
static int ring_alloc(PeerState* ps, size_t bytes, size_t* offset)
{
    if (bytes > ps->data_bytes) {
        set_last_error("te_peer_post", "message larger than the arena");
        return TE_ERR_LIMIT;
    }
    int st = reclaim(ps, false);
    if (st != TE_OK) return st;
    for (;;) {
        if (ps->live.empty()) {
            *offset = 0;
            ps->head = bytes;
            return TE_OK;
        }
        const size_t tail = ps->live.front().begin;
        if (ps->head > tail) {
            // live zones lie in [tail, head): free are [head, end) and [0, tail)
            if (ps->data_bytes - ps->head >= bytes) {
                *offset = ps->head;
                ps->head += bytes;
                return TE_OK;
            }
            if (tail > bytes) {
                *offset = 0;
                ps->head = bytes;
                return TE_OK;
            }
        }
        else if (tail - ps->head > bytes) {
            // wrapped: live zones lie in [tail, end) and [0, head): free is [head, tail)
            *offset = ps->head;
            ps->head += bytes;
            return TE_OK;
        }
        st = reclaim(ps, true);
        if (st != TE_OK) return st;
    }
}

}  // namespace te

using te::PeerState;

extern "C" int te_peer_is_open(const te_comm* c) { return (c && c->peer) ? 1 : 0; }

extern "C" size_t te_peer_max_message(const te_comm* c) { return (c && c->peer) ? c->peer->max_message : 0; }

extern "C" int te_peer_open(te_comm* c, size_t arena_bytes)
{
    if (!c) return TE_ERR_ARG;
    if (c->size == 1 || c->peer) return TE_OK;
    if (arena_bytes < ((size_t)32 << 20)) return TE_ERR_ARG;
    int st = te::set_device(c->device);
    if (st != TE_OK) return st;

    PeerState* ps = new PeerState;
    ps->data_bytes = arena_bytes & ~(size_t)255;
    ps->max_message = (ps->data_bytes / (2 * (size_t)c->size)) & ~(size_t)255;
    ps->peer_base.assign(c->size, nullptr);
    ps->posted.assign(c->size, 0);
    ps->fetched.assign(c->size, 0);
    if (const char* e = std::getenv("TE_PEER_TIMEOUT_S")) ps->timeout_s = std::atof(e);

    // ---- arena, exported to and opened by every peer ----
    TE_CUDA_CHECK(cudaMalloc(&ps->arena, te::kHeaderBytes + ps->data_bytes));
    TE_CUDA_CHECK(cudaMemset(ps->arena, 0, te::kHeaderBytes));
    cudaIpcMemHandle_t mine;
    TE_CUDA_CHECK(cudaIpcGetMemHandle(&mine, ps->arena));
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    cudaStream_t s;
    TE_CUDA_CHECK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    char* dbuf = nullptr;
    TE_CUDA_CHECK(cudaMalloc(&dbuf, 64 * (size_t)c->size));
    TE_CUDA_CHECK(cudaMemcpyAsync(dbuf + 64 * c->rank, &mine, 64, cudaMemcpyHostToDevice, s));
    TE_NCCL_CHECK(ncclAllGather(dbuf + 64 * c->rank, dbuf, 64, ncclChar, c->nccl, s));
    std::vector<cudaIpcMemHandle_t> handles(c->size);
    TE_CUDA_CHECK(cudaMemcpyAsync(handles.data(), dbuf, 64 * (size_t)c->size, cudaMemcpyDeviceToHost, s));
    TE_CUDA_CHECK(cudaStreamSynchronize(s));
    for (int r = 0; r < c->size; ++r) {
        if (r == c->rank) continue;
        void* p = nullptr;
        TE_CUDA_CHECK(cudaIpcOpenMemHandle(&p, handles[r], cudaIpcMemLazyEnablePeerAccess));
        ps->peer_base[r] = (char*)p;
    }

    // ---- mailbox in POSIX shared memory, named after the job ----
    char name[64];
    snprintf(name, sizeof(name), "/te_peer_%016llx", (unsigned long long)c->job_tag);
    ps->shm_bytes = sizeof(te::PairBox) * (size_t)c->size * (size_t)c->size;
    int fd = shm_open(name, O_CREAT | O_RDWR, 0600);
    if (fd < 0) {
        te::set_last_error("shm_open", strerror(errno));
        return TE_ERR_CUDA;
    }
    if (ftruncate(fd, (off_t)ps->shm_bytes) != 0) {
        te::set_last_error("ftruncate(shm)", strerror(errno));
        close(fd);
        return TE_ERR_CUDA;
    }
    void* m = mmap(nullptr, ps->shm_bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    close(fd);
    if (m == MAP_FAILED) {
        te::set_last_error("mmap(shm)", strerror(errno));
        return TE_ERR_CUDA;
    }
    ps->boxes = (te::PairBox*)m;      // a fresh object is zero filled: all counters start at 0

    // everyone has mapped it (and opened every arena) before anyone goes on or unlinks the name
    TE_NCCL_CHECK(ncclAllGather(dbuf + 64 * c->rank, dbuf, 64, ncclChar, c->nccl, s));
    TE_CUDA_CHECK(cudaStreamSynchronize(s));
    shm_unlink(name);
    TE_CUDA_CHECK(cudaFree(dbuf));
    TE_CUDA_CHECK(cudaStreamDestroy(s));

    // ---- stream memory operations (driver API, looked up at run time) ----
    const char* mode = std::getenv("TE_PEER_MEMOPS");       // "0": kernels / copies instead
    if (!(mode && mode[0] == '0')) {
        // (64-bit stream memory operations are supported by every GPU this library targets;
        //  a failing first call falls back to the kernel / copy form, see te_peer_wait)
        const int can64 = 1;
        void* fw = nullptr;
        void* fs = nullptr;
        cudaDriverEntryPointQueryResult q1, q2;
        if (can64
            && cudaGetDriverEntryPoint("cuStreamWaitValue64", &fw, cudaEnableDefault, &q1) == cudaSuccess
            && cudaGetDriverEntryPoint("cuStreamWriteValue64", &fs, cudaEnableDefault, &q2) == cudaSuccess
            && q1 == cudaDriverEntryPointSuccess && q2 == cudaDriverEntryPointSuccess && fw && fs) {
            ps->wait_value = (te::StreamValue64Fn)fw;
            ps->write_value = (te::StreamValue64Fn)fs;
            ps->memop_wait = ps->memop_signal = true;
        }
        cudaGetLastError();
    }
    TE_CUDA_CHECK(cudaHostAlloc((void**)&ps->host_vals, sizeof(uint64_t) * te::kHostVals, cudaHostAllocPortable));
    c->peer = ps;
    return TE_OK;
}

extern "C" int te_peer_close(te_comm* c)
{
    if (!c || !c->peer) return TE_OK;
    PeerState* ps = c->peer;
    int st = te::set_device(c->device);
    if (st != TE_OK) return st;
    cudaDeviceSynchronize();
    for (int r = 0; r < c->size; ++r)
        if (ps->peer_base[r]) cudaIpcCloseMemHandle(ps->peer_base[r]);
    // nobody frees an arena that a peer still has mapped
    ncclAllReduce(ps->arena + 4096, ps->arena + 4096, 1, ncclInt64, ncclSum, c->nccl, 0);
    cudaDeviceSynchronize();
    cudaFree(ps->arena);
    if (ps->boxes) munmap(ps->boxes, ps->shm_bytes);
    for (auto& s : ps->live) cudaEventDestroy(s.ev);
    for (auto e : ps->ev_pool) cudaEventDestroy(e);
    if (ps->host_vals) cudaFreeHost(ps->host_vals);
    delete ps;
    c->peer = nullptr;
    cudaGetLastError();
    return TE_OK;
}

extern "C" int te_peer_post(te_comm* c, int root, size_t bytes, uint64_t* seq, void** landing)
{
    if (!c || !c->peer || root < 0 || root >= c->size || root == c->rank || !seq || !landing || bytes == 0)
        return TE_ERR_ARG;
    PeerState* ps = c->peer;
    std::lock_guard<std::mutex> lock(ps->mu);
    if (bytes > ps->max_message) {
        te::set_last_error("te_peer_post", "message larger than te_peer_max_message()");
        return TE_ERR_LIMIT;
    }
    size_t off = 0;
    int st = te::ring_alloc(ps, te::up256(bytes), &off);
    if (st != TE_OK) return st;
    te::Segment sg;
    sg.begin = off;
    sg.end = off + te::up256(bytes);
    sg.root = root;
    sg.seq = ps->posted[root] + 1;
    sg.recorded = false;
    if (!ps->ev_pool.empty()) {
        sg.ev = ps->ev_pool.back();
        ps->ev_pool.pop_back();
    }
    else
        TE_CUDA_CHECK(cudaEventCreateWithFlags(&sg.ev, cudaEventDisableTiming));
    ps->live.push_back(sg);

    te::PairBox& box = ps->boxes[(size_t)root * c->size + c->rank];
    const double t0 = te::now_s();
    while (sg.seq - box.fetched.load(std::memory_order_acquire) > (uint64_t)te::kBoxRing) {
        if (te::now_s() - t0 > ps->timeout_s) {
            te::set_last_error("te_peer_post", "timed out: the root has not fetched earlier messages");
            return TE_ERR_CUDA;
        }
        std::this_thread::yield();
    }
    box.slot[(sg.seq - 1) % te::kBoxRing].offset = off;
    box.slot[(sg.seq - 1) % te::kBoxRing].bytes = bytes;
    box.posted.store(sg.seq, std::memory_order_release);
    ps->posted[root] = sg.seq;
    *seq = sg.seq;
    *landing = ps->arena + te::kHeaderBytes + off;
    return TE_OK;
}

extern "C" int te_peer_fetch(te_comm* c, int dest, size_t bytes, uint64_t* seq, uint64_t* offset)
{
    if (!c || !c->peer || dest < 0 || dest >= c->size || dest == c->rank || !seq || !offset)
        return TE_ERR_ARG;
    PeerState* ps = c->peer;
    te::PairBox& box = ps->boxes[(size_t)c->rank * c->size + dest];
    const uint64_t want = ps->fetched[dest] + 1;
    const double t0 = te::now_s();
    int spins = 0;
    while (box.posted.load(std::memory_order_acquire) < want) {
        if (++spins > 2000) {
            if (te::now_s() - t0 > ps->timeout_s) {
                te::set_last_error("te_peer_fetch", "timed out waiting for the destination to announce "
                                   "its landing zone (ranks not walking the same broadcast lists?)");
                return TE_ERR_CUDA;
            }
            std::this_thread::yield();
        }
    }
    const uint64_t off = box.slot[(want - 1) % te::kBoxRing].offset;
    const uint64_t got = box.slot[(want - 1) % te::kBoxRing].bytes;
    box.fetched.store(want, std::memory_order_release);
    ps->fetched[dest] = want;
    if (got != bytes) {
        te::set_last_error("te_peer_fetch", "root and destination disagree on the size of a message");
        return TE_ERR_ARG;
    }
    *seq = want;
    *offset = off;
    return TE_OK;
}

extern "C" int te_peer_put(te_comm* c, int dest, uint64_t offset, const void* src, size_t bytes, void* stream)
{
    if (!c || !c->peer || dest < 0 || dest >= c->size || !c->peer->peer_base[dest] || !src)
        return TE_ERR_ARG;
    PeerState* ps = c->peer;
    if (offset + bytes > ps->data_bytes) return TE_ERR_ARG;
    TE_CUDA_CHECK(cudaMemcpyAsync(ps->peer_base[dest] + te::kHeaderBytes + offset, src, bytes,
                                  cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    ps->n_put += 1;
    ps->n_put_bytes += bytes;
    return TE_OK;
}

extern "C" int te_peer_signal(te_comm* c, int dest, uint64_t seq, void* stream)
{
    if (!c || !c->peer || dest < 0 || dest >= c->size || !c->peer->peer_base[dest]) return TE_ERR_ARG;
    PeerState* ps = c->peer;
    char* flag = ps->peer_base[dest] + sizeof(uint64_t) * (size_t)c->rank;
    ps->n_signal += 1;
    if (ps->memop_signal) {
        int r = ps->write_value((cudaStream_t)stream, (unsigned long long)(uintptr_t)flag, seq, 0);
        if (r == 0) return TE_OK;
        ps->memop_signal = false;         // not supported on peer memory here: copy instead
        fprintf(stderr, "[tile_engine] cuStreamWriteValue64 on peer memory failed (%d): signalling by copy\n", r);
    }
    std::lock_guard<std::mutex> lock(ps->mu);
    uint64_t* v = &ps->host_vals[ps->host_vals_next];
    ps->host_vals_next = (ps->host_vals_next + 1) % te::kHostVals;
    *v = seq;
    TE_CUDA_CHECK(cudaMemcpyAsync(flag, v, sizeof(uint64_t), cudaMemcpyHostToDevice, (cudaStream_t)stream));
    return TE_OK;
}

extern "C" int te_peer_wait(te_comm* c, int root, uint64_t seq, void* stream)
{
    if (!c || !c->peer || root < 0 || root >= c->size) return TE_ERR_ARG;
    PeerState* ps = c->peer;
    char* flag = ps->arena + sizeof(uint64_t) * (size_t)root;
    ps->n_wait += 1;
    if (ps->memop_wait) {
        int r = ps->wait_value((cudaStream_t)stream, (unsigned long long)(uintptr_t)flag, seq, 0 /* GEQ */);
        if (r == 0) return TE_OK;
        ps->memop_wait = false;
        fprintf(stderr, "[tile_engine] cuStreamWaitValue64 failed (%d): waiting by kernel\n", r);
    }
    te::peer_spin_kernel<<<1, 1, 0, (cudaStream_t)stream>>>((const volatile unsigned long long*)flag, seq);
    te::count_launch();
    TE_CUDA_CHECK(cudaGetLastError());
    return TE_OK;
}

extern "C" int te_peer_consumed(te_comm* c, int root, uint64_t seq, void* stream)
{
    if (!c || !c->peer) return TE_ERR_ARG;
    PeerState* ps = c->peer;
    std::lock_guard<std::mutex> lock(ps->mu);
    for (auto it = ps->live.rbegin(); it != ps->live.rend(); ++it) {
        if (it->root == root && it->seq == seq) {
            TE_CUDA_CHECK(cudaEventRecord(it->ev, (cudaStream_t)stream));
            it->recorded = true;
            return TE_OK;
        }
    }
    te::set_last_error("te_peer_consumed", "no such message");
    return TE_ERR_ARG;
}

extern "C" int te_peer_stats(const te_comm* c, uint64_t* puts, uint64_t* put_bytes, uint64_t* signals,
                             uint64_t* waits, int* memops)
{
    if (!c || !c->peer) return TE_ERR_ARG;
    if (puts) *puts = c->peer->n_put;
    if (put_bytes) *put_bytes = c->peer->n_put_bytes;
    if (signals) *signals = c->peer->n_signal;
    if (waits) *waits = c->peer->n_wait;
    if (memops) *memops = (c->peer->memop_wait ? 1 : 0) | (c->peer->memop_signal ? 2 : 0);
    return TE_OK;
}
