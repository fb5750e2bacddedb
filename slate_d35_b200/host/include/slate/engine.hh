// slate/engine.hh -- per-process runtime under the matrix classes: the device, its
// streams ("queues"), event-ordered hand-offs, descriptor staging, deferred release of
// workspace tiles, the communicator, and slate::Memory (the device-tile pool).
//
// Mirrors, B200-first, what the reference spreads over
//   MatrixStorage::initQueues / compute_queue / comm_queue / allocateBatchArrays
//       (include/slate/internal/MatrixStorage.hh:198-245, 574-696)
//   slate::Memory (include/slate/internal/Memory.hh:29-101, src/core/Memory.cc)
//   the MPI communicator argument of every matrix constructor.
// Differences by design: one process drives ONE GPU (one rank per GPU, NCCL between
// them); nothing synchronises the host per step -- ordering between the comm stream
// and the compute queues is by CUDA events, and a workspace tile goes back to the pool
// only once the kernels that read it have finished (deferred release).
//
// Everything device-side goes through the C ABI of include/tile_engine.h.
#ifndef SLATE_B200_ENGINE_HH
#define SLATE_B200_ENGINE_HH

#include "slate/enums.hh"
#include "tile_engine.h"

#include <deque>
#include <functional>
#include <mutex>
#include <vector>

namespace slate {

#define slate_te_call(expr)                                                             \
    do {                                                                                \
        int te_status_ = (expr);                                                        \
        if (te_status_ != TE_OK)                                                        \
            throw slate::EngineException(#expr, te_status_, te_last_error(), __func__,  \
                                         __FILE__, __LINE__);                           \
    } while (0)

//------------------------------------------------------------------------------
/// Stand-in for MPI_Comm: the set of ranks (one per GPU) and their NCCL communicator.
/// A single-rank Comm needs no NCCL.
struct Comm {
    te_comm* nccl = nullptr;
    int rank = 0;
    int size = 1;
    /// data plane of the tile broadcasts (slate::initialize; SLATE_B200_BCAST=peer|nccl|coalesce):
    /// copy engines over CUDA IPC (te_peer_*, the default), NCCL send/recv per tile, or NCCL
    /// send/recv per (root, destination) pair
    bool peer_plane = false;
    bool nccl_coalesce = false;
};

/// MPI_COMM_WORLD of this process (set by slate::initialize; single rank by default).
Comm* comm_world();
Comm* comm_self();

//------------------------------------------------------------------------------
/// include/slate/internal/Memory.hh:29-101.  Fixed-size blocks, block_size bytes each.
class Memory {
public:
    Memory(size_t block_size);
    ~Memory();
    Memory(Memory const&) = delete;

    void addHostBlocks(int64_t) {}
    void addDeviceBlocks(int device, int64_t num_blocks);
    void clearHostBlocks() {}
    void clearDeviceBlocks(int device);

    void* alloc(int device, size_t size);
    void free(void* block, int device);

    size_t available(int device) const;
    size_t capacity(int device) const;
    size_t allocated(int device) const { return capacity(device) - available(device); }
    size_t block_size() const { return block_size_; }

    static int num_devices_;

private:
    size_t block_size_;
    te_pool* pool_;       // the one device this process drives
    int device_;
};

//------------------------------------------------------------------------------
class Engine {
public:
    static Engine& instance();

    /// Binds this process to a CUDA device (default: $LOCAL_RANK, else 0) and creates
    /// the comm stream.  Idempotent for the same device.
    void initialize(int device = -1);
    bool initialized() const { return initialized_; }
    int device() const { return device_; }
    int num_devices() const { return te_device_count() > 0 ? 1 : 0; }

    void* comm_stream();
    /// Stream of the device-to-host copies that bring tiles home (tileGet on HostNum).  Separate
    /// from the comm stream: a write-back waits for the launch that produced its tile, and on the
    /// comm stream that wait would hold up every host-to-device prefetch and broadcast queued
    /// behind it (PCIe is full duplex; the two directions have nothing to wait for in each other).
    void* home_stream();
    void* compute_stream(int queue);
    static const int MaxQueues = 16;

    // ---- event-ordered hand-offs ----
    /// call after enqueueing work on queue q / on the comm stream
    void mark_queue(int q);
    /// sequence number the NEXT mark_queue(q) will carry: a tile written by the launch being
    /// prepared remembers (q, that number) and later consumers wait for exactly that mark
    uint64_t next_queue_seq(int q) const { return queue_seq_[q] + 1; }
    /// true once mark `seq` of queue q has been reached on the device (never blocks)
    bool queue_mark_done(int q, uint64_t seq);
    void queue_wait_queue_seq(int q, int other, uint64_t seq);
    void comm_wait_queue_seq(int other, uint64_t seq);
    /// the home stream behind mark `seq` of compute queue `other` / mark `seq` of the comm stream
    void home_wait_queue_seq(int other, uint64_t seq);
    void home_wait_comm_seq(uint64_t seq);
    void sync_home();
    /// Records a point on the comm stream and returns its sequence number; tiles filled by
    /// the comm work enqueued before it carry that number (TileNode::comm_seq).
    uint64_t mark_comm();
    /// sequence number the NEXT mark_comm() will return
    uint64_t next_comm_seq() const { return comm_seq_ + 1; }
    /// make queue q wait for everything enqueued so far on the comm stream / other queue
    void queue_wait_comm(int q);
    /// make queue q wait for the comm work up to sequence number `seq` only (so that a
    /// broadcast issued ahead for a later step does not hold this launch back)
    void queue_wait_comm_seq(int q, uint64_t seq);
    void queue_wait_queue(int q, int other);
    /// make the comm stream wait for everything enqueued so far on queue q
    void comm_wait_queue(int q);
    void comm_wait_all_queues();
    void sync_comm();
    void sync_queue(int q);
    void sync_all();

    // ---- descriptor staging: pinned host + device ring per queue ----
    struct DescSlot {
        void* host = nullptr;
        void* dev = nullptr;
        size_t capacity = 0;
        void* event = nullptr;
        bool pending = false;
    };
    /// A slot with room for `bytes`; blocks only if the ring is full of unfinished launches.
    DescSlot& acquire_desc(int q, size_t bytes);
    /// Uploads the first `bytes` of the slot on queue q.
    void upload_desc(int q, DescSlot& slot, size_t bytes);
    /// Call after the kernel that reads the slot was launched on queue q.
    void retire_desc(int q, DescSlot& slot);

    // ---- staging slabs of the comm stream (coalesced broadcasts, SLATE_B200_BCAST_COALESCE) ----
    struct CommSlab {
        void* dev = nullptr;
        size_t capacity = 0;
        void* event = nullptr;
        bool pending = false;
    };
    /// A device buffer of at least `bytes` for pack / unpack copies on the comm stream; blocks
    /// only if all slabs of the ring still belong to unfinished broadcasts.
    CommSlab& acquire_comm_slab(size_t bytes) { return acquire_slab(false, bytes); }
    /// Call after the last comm-stream operation that touches the slab was enqueued.
    void retire_comm_slab(CommSlab& slab) { retire_slab(false, slab); }
    /// The same for either side stream: home == false the comm stream, true the home stream
    /// (panels of host-origin tiles on their way to / from the host, BaseMatrix::tileGetRun).
    CommSlab& acquire_slab(bool home, size_t bytes);
    void retire_slab(bool home, CommSlab& slab);

    /// Device scratch private to queue q (grows; stream-ordered reuse on that queue).
    void* scratch(int q, size_t bytes);
    /// One device int per (queue, slot) for potrf's info.
    int* info_slot(int index);
    int* info_scratch_slot();
    void reserve_info_slots(int64_t n);

    // ---- deferred actions (release of workspace tiles) ----
    /// Runs fn once everything enqueued so far on the given queues (bit q) and, if
    /// `comm`, on the comm stream has finished.  Polled by poll_deferred().
    void defer(unsigned queue_mask, bool comm, std::function<void()> fn);
    void poll_deferred(bool wait_oldest = false);
    void drain_deferred();
    bool has_deferred() const { return !deferred_.empty(); }

    // ---- timing across all streams ----
    void* timer_begin();             // joins all streams, returns start event
    float timer_end(void* start);    // joins all streams, returns elapsed ms (synchronises)

    void finalize();

    // ---- optional per-launch device timing (SLATE_B200_TIMING) ----
    bool tracing() const { return tracing_; }

    /// Set by a driver while it streams origin tiles between host and device (host-origin
    /// gemm): descriptor arrays are then uploaded by kernel instead of by the copy engine.
    void set_bulk_host_traffic(bool on) { bulk_host_traffic_ = on; }
    /// q >= 0: compute queue q; q == -1: the comm stream
    void trace_begin(int q, const char* what, int64_t count);
    void trace_end(int q);
    void trace_dump(const char* title);

private:
    Engine() {}
    struct Deferred { std::vector<void*> events; std::function<void()> fn; };
    void* new_event(bool timing = false);
    void recycle_event(void* ev);

    bool initialized_ = false;
    int device_ = 0;
    void* comm_stream_ = nullptr;
    void* home_stream_ = nullptr;
    bool home_used_ = false;
    void* comm_event_ = nullptr;
    uint64_t comm_seq_ = 0;
    static const int QueueRing = 64;
    void* queue_ring_event_[MaxQueues][QueueRing] = {};
    uint64_t queue_ring_seq_[MaxQueues][QueueRing] = {};
    void* queue_mark_event(int other, uint64_t seq);
    static const int CommRing = 256;
    void* comm_ring_event_[CommRing] = {};
    uint64_t comm_ring_seq_[CommRing] = {};
    void* queue_[MaxQueues] = {};
    void* queue_event_[MaxQueues] = {};
    uint64_t queue_seq_[MaxQueues] = {};
    uint64_t queue_saw_comm_[MaxQueues] = {};
    // descriptor slots per queue = launches the host may run ahead of the device.  16: a burst of
    // host work (enqueueing the 2 x 2048 tile copies of a slab of host-origin C takes ~45 ms) must
    // not drain the device's queue (measured: with 4 slots the update idled 45 ms per slab)
    static const int RingDepth = 16;
    DescSlot ring_[MaxQueues][RingDepth];
    int ring_next_[MaxQueues] = {};
    void* scratch_[MaxQueues] = {};
    size_t scratch_bytes_[MaxQueues] = {};
    int* info_base_ = nullptr;
    int info_cap_ = 4096;
    static const int CommSlabs = 4;
    CommSlab comm_slab_[2][CommSlabs];       // [0] comm stream, [1] home stream
    int comm_slab_next_[2] = {0, 0};
    std::deque<Deferred> deferred_;
    std::vector<void*> event_pool_;
    std::recursive_mutex mu_;
    bool tracing_ = false;
    bool bulk_host_traffic_ = false;
    struct TraceRec { void* e0; void* e1; const char* what; int64_t count; int q; double host_s; };
    std::vector<TraceRec> trace_;
    void* trace_epoch_ = nullptr;
};

/// Process set-up for the harness: device binding and (size > 1) the NCCL communicator.
/// nccl_id: TE_COMM_ID_BYTES bytes from te_comm_unique_id on rank 0, or nullptr if size == 1.
void initialize(int device, int rank, int size, const void* nccl_id);
void finalize();

}  // namespace slate

// The reference's constructors take an MPI_Comm; keep the call text compiling.
#ifndef MPI_VERSION
typedef slate::Comm* MPI_Comm;
#define MPI_COMM_WORLD (slate::comm_world())
#define MPI_COMM_SELF  (slate::comm_self())
#endif

#endif
