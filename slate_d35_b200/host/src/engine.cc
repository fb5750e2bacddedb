// engine.cc -- see slate/engine.hh.
#include "slate/engine.hh"

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

namespace slate {

static Comm g_world;
static Comm g_self;
Comm* comm_world() { return &g_world; }
Comm* comm_self() { return &g_self; }

//------------------------------------------------------------------------------
int Memory::num_devices_ = -1;

Memory::Memory(size_t block_size) : block_size_(block_size), pool_(nullptr)
{
    Engine& eng = Engine::instance();
    if (!eng.initialized()) eng.initialize();
    device_ = eng.device();
    if (num_devices_ < 0) num_devices_ = eng.num_devices();
    if (num_devices_ > 0)
        slate_te_call(te_pool_create(device_, block_size_ ? block_size_ : 256, &pool_));
}

Memory::~Memory()
{
    if (pool_) te_pool_destroy(pool_);
}

/// src/core/Memory.cc:63-72: one cudaMalloc of num_blocks blocks.
void Memory::addDeviceBlocks(int device, int64_t num_blocks)
{
    slate_assert(pool_ != nullptr);
    slate_te_call(te_pool_reserve(pool_, num_blocks));
}

void Memory::clearDeviceBlocks(int device)
{
    if (pool_) slate_te_call(te_pool_clear(pool_));
}

/// src/core/Memory.cc:118-141: host = new[], device = pop a block (grow by one if empty).
void* Memory::alloc(int device, size_t size)
{
    if (device == HostNum)
        return new char[size];
    slate_assert(size <= block_size_);
    slate_assert(pool_ != nullptr);
    // Blocks released by earlier steps may still be waiting for their kernels to finish
    // (Engine::defer).  Collect those first; growing the pool means a cudaMalloc in the
    // middle of the run.
    if (available(device) == 0) {
        Engine& eng = Engine::instance();
        eng.poll_deferred(false);
        while (available(device) == 0 && eng.has_deferred())
            eng.poll_deferred(true);
    }
    void* block = nullptr;
    slate_te_call(te_pool_alloc(pool_, &block));
    return block;
}

void Memory::free(void* block, int device)
{
    if (device == HostNum) {
        delete[] (char*)block;
        return;
    }
    slate_te_call(te_pool_free(pool_, block));
}

size_t Memory::available(int device) const
{
    if (device == HostNum || !pool_) return 0;
    size_t a = 0;
    te_pool_stats(pool_, &a, nullptr, nullptr);
    return a;
}

size_t Memory::capacity(int device) const
{
    if (device == HostNum || !pool_) return 0;
    size_t c = 0;
    te_pool_stats(pool_, nullptr, &c, nullptr);
    return c;
}

//------------------------------------------------------------------------------
Engine& Engine::instance()
{
    static Engine e;
    return e;
}

void Engine::initialize(int device)
{
    std::lock_guard<std::recursive_mutex> lock(mu_);
    if (device < 0) {
        const char* lr = std::getenv("LOCAL_RANK");
        device = lr ? std::atoi(lr) : 0;
        int n = te_device_count();
        if (n > 0) device %= n;
    }
    if (initialized_) {
        slate_assert(device == device_);
        return;
    }
    device_ = device;
    initialized_ = true;
    tracing_ = std::getenv("SLATE_B200_TIMING") != nullptr;
    if (te_device_count() == 0)
        return;     // host-only process: matrices can exist, device paths throw
    slate_te_call(te_stream_create(device_, 1, &comm_stream_));
    slate_te_call(te_stream_create(device_, 1, &home_stream_));
    slate_te_call(te_event_create(device_, 0, &comm_event_));
    void* p = nullptr;
    slate_te_call(te_device_malloc(device_, sizeof(int) * (size_t)(info_cap_ + 1), &p));
    info_base_ = (int*)p;
}

void* Engine::comm_stream()
{
    slate_assert(comm_stream_ != nullptr);
    return comm_stream_;
}

void* Engine::home_stream()
{
    slate_assert(home_stream_ != nullptr);
    home_used_ = true;
    return home_stream_;
}

void Engine::home_wait_queue_seq(int other, uint64_t seq)
{
    std::lock_guard<std::recursive_mutex> lock(mu_);
    if (other < 0 || !queue_[other] || seq == 0) return;
    slate_te_call(te_stream_wait_event(home_stream(), queue_mark_event(other, seq)));
}

void Engine::home_wait_comm_seq(uint64_t seq)
{
    std::lock_guard<std::recursive_mutex> lock(mu_);
    if (seq == 0) return;
    if (seq > comm_seq_) mark_comm();
    const int slot = (int)(seq % CommRing);
    slate_te_call(te_stream_wait_event(home_stream(),
                                       comm_ring_seq_[slot] == seq ? comm_ring_event_[slot] : comm_event_));
}

void Engine::sync_home()
{
    if (home_stream_ && home_used_) slate_te_call(te_stream_sync(device_, home_stream_));
}

void* Engine::compute_stream(int q)
{
    std::lock_guard<std::recursive_mutex> lock(mu_);
    slate_assert(q >= 0 && q < MaxQueues);
    if (!queue_[q]) {
        // queue 0 carries the bulk trailing update; every other queue (1 = panel of potrf,
        // 2.. = lookahead columns) is on the critical path and gets the high stream priority,
        // so that its CTAs are dispatched ahead of the thousands of pending trailing CTAs.
        // (With equal priorities an 85-tile lookahead update took as long as the 1806-tile
        //  trailing update it shared the GPU with: 15 ms instead of < 1 ms, measured on 8 GPUs.)
        slate_te_call(te_stream_create(device_, q >= 1 ? 1 : 0, &queue_[q]));
        slate_te_call(te_event_create(device_, 0, &queue_event_[q]));
    }
    return queue_[q];
}

void Engine::mark_queue(int q)
{
    std::lock_guard<std::recursive_mutex> lock(mu_);
    void* s = compute_stream(q);
    ++queue_seq_[q];
    slate_te_call(te_event_record(queue_event_[q], s));
    int slot = (int)(queue_seq_[q] % QueueRing);
    if (!queue_ring_event_[q][slot])
        slate_te_call(te_event_create(device_, 0, &queue_ring_event_[q][slot]));
    slate_te_call(te_event_record(queue_ring_event_[q][slot], s));
    queue_ring_seq_[q][slot] = queue_seq_[q];
}

/// Event that covers mark `seq` of queue `other`: that mark itself if it is still in the
/// ring, else the oldest mark kept (which is later, so it covers it too).
void* Engine::queue_mark_event(int other, uint64_t seq)
{
    if (seq > queue_seq_[other])
        mark_queue(other);                 // written but not marked yet: cover it now
    if (seq > queue_seq_[other]) seq = queue_seq_[other];
    int slot = (int)(seq % QueueRing);
    if (queue_ring_seq_[other][slot] == seq) return queue_ring_event_[other][slot];
    uint64_t oldest = queue_seq_[other] >= (uint64_t)QueueRing ? queue_seq_[other] - QueueRing + 1 : 1;
    slot = (int)(oldest % QueueRing);
    if (queue_ring_seq_[other][slot] == oldest) return queue_ring_event_[other][slot];
    return queue_event_[other];
}

bool Engine::queue_mark_done(int q, uint64_t seq)
{
    std::lock_guard<std::recursive_mutex> lock(mu_);
    if (q < 0 || !queue_[q] || seq == 0) return true;
    if (seq > queue_seq_[q]) return false;                   // not even marked yet
    const int slot = (int)(seq % QueueRing);
    if (queue_ring_seq_[q][slot] != seq)
        return true;    // left the ring: >= QueueRing marks ago, and the host is never that far ahead
    return te_event_query_done(queue_ring_event_[q][slot]) == 1;
}

void Engine::queue_wait_queue_seq(int q, int other, uint64_t seq)
{
    std::lock_guard<std::recursive_mutex> lock(mu_);
    if (q == other || other < 0 || !queue_[other] || seq == 0) return;
    slate_te_call(te_stream_wait_event(compute_stream(q), queue_mark_event(other, seq)));
}

void Engine::comm_wait_queue_seq(int other, uint64_t seq)
{
    std::lock_guard<std::recursive_mutex> lock(mu_);
    if (other < 0 || !queue_[other] || seq == 0) return;
    slate_te_call(te_stream_wait_event(comm_stream(), queue_mark_event(other, seq)));
}

uint64_t Engine::mark_comm()
{
    std::lock_guard<std::recursive_mutex> lock(mu_);
    ++comm_seq_;
    slate_te_call(te_event_record(comm_event_, comm_stream()));
    // a separate event per mark, kept for the last CommRing marks
    int slot = (int)(comm_seq_ % CommRing);
    if (!comm_ring_event_[slot])
        slate_te_call(te_event_create(device_, 0, &comm_ring_event_[slot]));
    slate_te_call(te_event_record(comm_ring_event_[slot], comm_stream()));
    comm_ring_seq_[slot] = comm_seq_;
    return comm_seq_;
}

void Engine::queue_wait_comm_seq(int q, uint64_t seq)
{
    std::lock_guard<std::recursive_mutex> lock(mu_);
    if (seq == 0 || queue_saw_comm_[q] >= seq) return;
    void* s = compute_stream(q);
    if (seq > comm_seq_) {
        // not marked yet: mark now so that the wait covers it
        mark_comm();
    }
    int slot = (int)(seq % CommRing);
    if (comm_ring_seq_[slot] == seq) {
        slate_te_call(te_stream_wait_event(s, comm_ring_event_[slot]));
        queue_saw_comm_[q] = seq;
    }
    else {
        // too old to be in the ring (or overwritten): wait for everything so far
        slate_te_call(te_stream_wait_event(s, comm_event_));
        queue_saw_comm_[q] = comm_seq_;
    }
}

void Engine::queue_wait_comm(int q)
{
    void* s = compute_stream(q);
    if (comm_seq_ == 0 || queue_saw_comm_[q] == comm_seq_) return;
    slate_te_call(te_stream_wait_event(s, comm_event_));
    queue_saw_comm_[q] = comm_seq_;
}

void Engine::queue_wait_queue(int q, int other)
{
    if (q == other || other < 0 || !queue_[other] || queue_seq_[other] == 0) return;
    slate_te_call(te_stream_wait_event(compute_stream(q), queue_event_[other]));
}

void Engine::comm_wait_queue(int q)
{
    if (q < 0 || !queue_[q] || queue_seq_[q] == 0) return;
    slate_te_call(te_stream_wait_event(comm_stream(), queue_event_[q]));
}

void Engine::comm_wait_all_queues()
{
    for (int q = 0; q < MaxQueues; ++q) comm_wait_queue(q);
}

void Engine::sync_comm() { slate_te_call(te_stream_sync(device_, comm_stream())); }
void Engine::sync_queue(int q) { if (queue_[q]) slate_te_call(te_stream_sync(device_, queue_[q])); }

void Engine::sync_all()
{
    if (te_device_count() == 0) return;
    slate_te_call(te_device_sync(device_));
    drain_deferred();
}

Engine::DescSlot& Engine::acquire_desc(int q, size_t bytes)
{
    std::lock_guard<std::recursive_mutex> lock(mu_);
    compute_stream(q);
    DescSlot& s = ring_[q][ring_next_[q]];
    ring_next_[q] = (ring_next_[q] + 1) % RingDepth;
    if (s.pending) {
        slate_te_call(te_event_sync(s.event));
        s.pending = false;
    }
    if (s.capacity < bytes) {
        if (s.host) slate_te_call(te_host_free_pinned(s.host));
        if (s.dev) slate_te_call(te_device_free(device_, s.dev));
        size_t cap = bytes + bytes / 4 + 4096;
        cap = (cap + 255) & ~(size_t)255;
        slate_te_call(te_host_alloc_pinned(cap, &s.host));
        slate_te_call(te_device_malloc(device_, cap, &s.dev));
        s.capacity = cap;
    }
    if (!s.event) slate_te_call(te_event_create(device_, 0, &s.event));
    return s;
}

Engine::CommSlab& Engine::acquire_slab(bool home, size_t bytes)
{
    std::lock_guard<std::recursive_mutex> lock(mu_);
    const int w = home ? 1 : 0;
    CommSlab& s = comm_slab_[w][comm_slab_next_[w]];
    comm_slab_next_[w] = (comm_slab_next_[w] + 1) % CommSlabs;
    if (s.pending) {
        slate_te_call(te_event_sync(s.event));
        s.pending = false;
    }
    if (s.capacity < bytes) {
        if (s.dev) slate_te_call(te_device_free(device_, s.dev));
        size_t cap = bytes + bytes / 4;
        cap = (cap + ((size_t)1 << 20) - 1) & ~(((size_t)1 << 20) - 1);
        slate_te_call(te_device_malloc(device_, cap, &s.dev));
        s.capacity = cap;
    }
    if (!s.event) slate_te_call(te_event_create(device_, 0, &s.event));
    return s;
}

void Engine::retire_slab(bool home, CommSlab& s)
{
    slate_te_call(te_event_record(s.event, home ? home_stream() : comm_stream()));
    s.pending = true;
}

void Engine::upload_desc(int q, DescSlot& s, size_t bytes)
{
    // By kernel while bulk host<->device tile traffic is in flight (a descriptor copy would
    // queue behind it on the copy engine, see te_upload_small); by the copy engine otherwise,
    // which needs no SM slot -- on potrf's panel chain every extra kernel is one more wait for
    // a slot among the trailing update's CTAs.
    static const char* mode = std::getenv("SLATE_B200_DESC_UPLOAD");
    bool by_kernel = mode ? (mode[0] == 'k') : bulk_host_traffic_;
    if (by_kernel)
        slate_te_call(te_upload_small(device_, compute_stream(q), s.dev, s.host, bytes));
    else
        slate_te_call(te_memcpy_async(s.dev, s.host, bytes, TE_COPY_H2D, compute_stream(q)));
}

void Engine::retire_desc(int q, DescSlot& s)
{
    slate_te_call(te_event_record(s.event, compute_stream(q)));
    s.pending = true;
}

void* Engine::scratch(int q, size_t bytes)
{
    std::lock_guard<std::recursive_mutex> lock(mu_);
    if (scratch_bytes_[q] < bytes) {
        if (scratch_[q]) {
            // the old buffer may still be in use by launches on this queue
            sync_queue(q);
            slate_te_call(te_device_free(device_, scratch_[q]));
        }
        size_t cap = bytes + bytes / 2;
        slate_te_call(te_device_malloc(device_, cap, &scratch_[q]));
        scratch_bytes_[q] = cap;
    }
    return scratch_[q];
}

int* Engine::info_slot(int index)
{
    slate_assert(index >= 0 && index < info_cap_ && info_base_);
    return info_base_ + index;
}

/// The slot internal::potrf / getrf_nopiv use when the caller passes no device_info.
int* Engine::info_scratch_slot()
{
    slate_assert(info_base_);
    return info_base_ + info_cap_;
}

/// Makes info_slot(0 .. n-1) valid (one slot per block column of a factorisation).  Growing
/// waits for the device: the old array may still be written by enqueued panel kernels.
void Engine::reserve_info_slots(int64_t n)
{
    std::lock_guard<std::recursive_mutex> lock(mu_);
    if (n <= info_cap_ || !info_base_) return;
    sync_all();
    slate_te_call(te_device_free(device_, info_base_));
    info_cap_ = (int)(n + n / 2);
    void* p = nullptr;
    slate_te_call(te_device_malloc(device_, sizeof(int) * (size_t)(info_cap_ + 1), &p));
    info_base_ = (int*)p;
}

void* Engine::new_event(bool timing)
{
    if (!timing && !event_pool_.empty()) {
        void* e = event_pool_.back();
        event_pool_.pop_back();
        return e;
    }
    void* e = nullptr;
    slate_te_call(te_event_create(device_, timing ? 1 : 0, &e));
    return e;
}

void Engine::recycle_event(void* ev) { event_pool_.push_back(ev); }

void Engine::defer(unsigned queue_mask, bool comm, std::function<void()> fn)
{
    std::lock_guard<std::recursive_mutex> lock(mu_);
    Deferred d;
    d.fn = std::move(fn);
    for (int q = 0; q < MaxQueues; ++q) {
        if ((queue_mask >> q) & 1u) {
            if (!queue_[q] || queue_seq_[q] == 0) continue;
            void* e = new_event();
            slate_te_call(te_event_record(e, queue_[q]));
            d.events.push_back(e);
        }
    }
    if (comm && comm_stream_) {
        void* e = new_event();
        slate_te_call(te_event_record(e, comm_stream_));
        d.events.push_back(e);
        if (home_used_) {            // a tile still being copied home must not be recycled either
            e = new_event();
            slate_te_call(te_event_record(e, home_stream_));
            d.events.push_back(e);
        }
    }
    if (d.events.empty()) {
        d.fn();
        return;
    }
    deferred_.push_back(std::move(d));
}

void Engine::poll_deferred(bool wait_oldest)
{
    std::lock_guard<std::recursive_mutex> lock(mu_);
    while (!deferred_.empty()) {
        Deferred& d = deferred_.front();
        bool done = true;
        for (void* e : d.events) {
            if (wait_oldest) {
                slate_te_call(te_event_sync(e));
            }
            else if (te_event_query_done(e) != 1) {
                done = false;
                break;
            }
        }
        if (!done) break;
        for (void* e : d.events) recycle_event(e);
        auto fn = std::move(d.fn);
        deferred_.pop_front();
        fn();
        wait_oldest = false;
    }
}

void Engine::drain_deferred()
{
    std::lock_guard<std::recursive_mutex> lock(mu_);
    while (!deferred_.empty())
        poll_deferred(true);
}

void* Engine::timer_begin()
{
    sync_all();
    void* e = new_event(true);
    slate_te_call(te_event_record(e, compute_stream(0)));
    // every other stream starts after the start event
    for (int q = 1; q < MaxQueues; ++q)
        if (queue_[q]) slate_te_call(te_stream_wait_event(queue_[q], e));
    slate_te_call(te_stream_wait_event(comm_stream(), e));
    slate_te_call(te_stream_wait_event(home_stream_, e));
    return e;
}

float Engine::timer_end(void* start)
{
    void* s0 = compute_stream(0);
    for (int q = 1; q < MaxQueues; ++q) {
        if (queue_[q]) {
            void* e = new_event();
            slate_te_call(te_event_record(e, queue_[q]));
            slate_te_call(te_stream_wait_event(s0, e));
            recycle_event(e);
        }
    }
    for (void* side : {comm_stream(), home_stream_}) {
        void* e = new_event();
        slate_te_call(te_event_record(e, side));
        slate_te_call(te_stream_wait_event(s0, e));
        recycle_event(e);
    }
    void* stop = new_event(true);
    slate_te_call(te_event_record(stop, s0));
    slate_te_call(te_event_sync(stop));
    float ms = 0;
    slate_te_call(te_event_elapsed_ms(start, stop, &ms));
    te_event_destroy(start);
    te_event_destroy(stop);
    drain_deferred();
    return ms;
}

void Engine::trace_begin(int q, const char* what, int64_t count)
{
    if (!tracing_) return;
    TraceRec r;
    r.what = what;  r.count = count;  r.q = q;
    r.e0 = new_event(true);
    r.e1 = new_event(true);
    r.host_s = std::chrono::duration<double>(
                   std::chrono::steady_clock::now().time_since_epoch()).count();
    void* stream = (q < 0) ? comm_stream() : compute_stream(q);
    if (!trace_epoch_) {
        trace_epoch_ = new_event(true);
        slate_te_call(te_event_record(trace_epoch_, stream));
    }
    slate_te_call(te_event_record(r.e0, stream));
    trace_.push_back(r);
}

void Engine::trace_end(int q)
{
    if (!tracing_ || trace_.empty()) return;
    void* stream = (q < 0) ? comm_stream() : compute_stream(q);
    // the matching begin is the last record of that queue
    for (size_t i = trace_.size(); i-- > 0;) {
        if (trace_[i].q == q) {
            slate_te_call(te_event_record(trace_[i].e1, stream));
            break;
        }
    }
}

void Engine::trace_dump(const char* title)
{
    if (!tracing_ || trace_.empty()) return;
    sync_all();
    std::fprintf(stderr, "[slate_b200 rank %d] %s: %zu launches (start ms, duration ms, host enqueue ms)\n",
                 comm_world()->rank, title, trace_.size());
    // SLATE_B200_TRACE_FILE=<prefix>: every record, one file per rank and call
    std::FILE* tf = nullptr;
    if (const char* prefix = std::getenv("SLATE_B200_TRACE_FILE")) {
        static int call = 0;
        std::string name = std::string(prefix) + "_r" + std::to_string(comm_world()->rank) + "_"
                           + std::to_string(call++) + ".txt";
        tf = std::fopen(name.c_str(), "w");
        if (tf) std::fprintf(tf, "# %s: queue what count start_ms dur_ms host_ms\n", title);
    }
    double h0 = trace_.front().host_s;
    size_t shown = 0;
    for (size_t i = 0; i < trace_.size(); ++i) {
        float t0 = 0, dt = 0;
        te_event_elapsed_ms(trace_epoch_, trace_[i].e0, &t0);
        te_event_elapsed_ms(trace_[i].e0, trace_[i].e1, &dt);
        if (tf)
            std::fprintf(tf, "%d %s %lld %.3f %.3f %.3f\n", trace_[i].q, trace_[i].what,
                         (long long)trace_[i].count, t0, dt, (trace_[i].host_s - h0) * 1e3);
        if (i < 40 || i + 4 >= trace_.size()) {
            std::fprintf(stderr, "   q%d %-10s n=%-6lld start %9.3f  dur %8.3f  host %9.3f\n", trace_[i].q,
                         trace_[i].what, (long long)trace_[i].count, t0, dt, (trace_[i].host_s - h0) * 1e3);
            ++shown;
        }
        te_event_destroy(trace_[i].e0);
        te_event_destroy(trace_[i].e1);
    }
    if (tf) std::fclose(tf);
    te_event_destroy(trace_epoch_);
    trace_epoch_ = nullptr;
    trace_.clear();
}

void Engine::finalize()
{
    if (!initialized_ || te_device_count() == 0) return;
    sync_all();
}

//------------------------------------------------------------------------------
void initialize(int device, int rank, int size, const void* nccl_id)
{
    Engine::instance().initialize(device);
    g_world.rank = rank;
    g_world.size = size;
    // a process without a GPU can still build matrices and plans (tests of the host logic)
    if (size > 1 && g_world.nccl == nullptr && te_device_count() > 0) {
        slate_assert(nccl_id != nullptr);
        slate_te_call(te_comm_init_rank(nccl_id, TE_COMM_ID_BYTES, size, rank,
                                        Engine::instance().device(), &g_world.nccl));
        slate_te_call(te_comm_connect_all(g_world.nccl, Engine::instance().comm_stream()));
        // Broadcast data plane.  Default: copy engines over CUDA IPC, no SM used (te_peer_*).
        const char* plane = std::getenv("SLATE_B200_BCAST");
        const char* old_switch = std::getenv("SLATE_B200_BCAST_COALESCE");
        std::string mode = plane ? plane : (old_switch && old_switch[0] == '1') ? "coalesce" : "peer";
        g_world.peer_plane = g_world.nccl_coalesce = false;
        if (mode == "peer") {
            size_t mb = 512;
            if (const char* e = std::getenv("SLATE_B200_ARENA_MB")) mb = (size_t)std::atoll(e);
            slate_te_call(te_peer_open(g_world.nccl, mb << 20));
            g_world.peer_plane = te_peer_is_open(g_world.nccl) != 0;
        }
        else if (mode == "coalesce")
            g_world.nccl_coalesce = true;
        else if (mode != "nccl")
            slate_error("SLATE_B200_BCAST must be peer, nccl or coalesce");
    }
}

void finalize()
{
    Engine::instance().finalize();
    if (g_world.nccl) {
        te_comm_destroy(g_world.nccl);
        g_world.nccl = nullptr;
    }
    g_world.rank = 0;
    g_world.size = 1;
    g_world.peer_plane = g_world.nccl_coalesce = false;
}

}  // namespace slate
