// slate/BaseMatrix.hh -- distributed tiled matrix views over a 2D block-cyclic p x q
// grid: sub / transpose views, tile access, MOSI coherence between the host and the
// device instance of a tile, tile broadcasts, workspace release.
//
// Mirrors the parts of include/slate/BaseMatrix.hh that slate::gemm and slate::potrf
// touch (reference lines cited per method).  Deliberate differences:
//   * one rank drives one GPU; broadcasts go GPU to GPU over NCCL
//     (tile_engine.h te_comm_*), never through host staging or MPI;
//   * nothing here blocks the host except copies TO the host: device-side ordering is
//     by events (Engine), and erased workspace blocks return to the pool deferred.
// Methods that take a compute `queue` extend the reference signature with a trailing
// defaulted argument; -1 means "no device consumer yet".
#ifndef SLATE_B200_BASE_MATRIX_HH
#define SLATE_B200_BASE_MATRIX_HH

#include "slate/MatrixStorage.hh"

#include <algorithm>
#include <cstdlib>
#include <list>
#include <map>
#include <set>
#include <vector>

namespace slate {

template <typename scalar_t> class BaseMatrix;
template <typename scalar_t> class Matrix;

template <typename scalar_t>
class BaseMatrix {
public:
    using ij_tuple = slate::ij_tuple;
    /// include/slate/BaseMatrix.hh:42-50
    using BcastList =
        std::vector<std::tuple<int64_t, int64_t, std::list<BaseMatrix<scalar_t>>>>;
    using BcastListTag =
        std::vector<std::tuple<int64_t, int64_t, std::list<BaseMatrix<scalar_t>>, int64_t>>;
    /// include/slate/BaseMatrix.hh:52-56: tile (i, j), destination view, views whose owners
    /// hold partial sums
    using ReduceList =
        std::vector<std::tuple<int64_t, int64_t, BaseMatrix<scalar_t>, std::list<BaseMatrix<scalar_t>>>>;
    static constexpr bool is_complex = slate::is_complex<scalar_t>::value;
    static constexpr bool is_real = !is_complex;

    BaseMatrix()
        : ioffset_(0), joffset_(0), mt_(0), nt_(0), uplo_(Uplo::General), op_(Op::NoTrans),
          layout_(Layout::ColMajor), origin_(Target::Host)
    {}

protected:
    BaseMatrix(int64_t m, int64_t n, int64_t mb, int64_t nb, GridOrder order, int p, int q,
               MPI_Comm comm)
        : ioffset_(0), joffset_(0), uplo_(Uplo::General), op_(Op::NoTrans),
          layout_(Layout::ColMajor), origin_(Target::Host),
          storage_(std::make_shared<MatrixStorage<scalar_t>>(m, n, mb, nb, order, p, q, comm))
    {
        mt_ = storage_->mt_;
        nt_ = storage_->nt_;
    }

    /// include/slate/BaseMatrix.hh (lambda constructor): caller-supplied tile sizes and owners
    BaseMatrix(int64_t m, int64_t n, std::function<int64_t(int64_t)>& inTileMb,
               std::function<int64_t(int64_t)>& inTileNb, std::function<int(ij_tuple)>& inTileRank,
               std::function<int(ij_tuple)>& inTileDevice, MPI_Comm comm)
        : ioffset_(0), joffset_(0), uplo_(Uplo::General), op_(Op::NoTrans),
          layout_(Layout::ColMajor), origin_(Target::Host),
          storage_(std::make_shared<MatrixStorage<scalar_t>>(m, n, inTileMb, inTileNb, inTileRank,
                                                             inTileDevice, comm))
    {
        mt_ = storage_->mt_;
        nt_ = storage_->nt_;
    }

    /// sub-matrix view of tiles [i1..i2] x [j1..j2] (include/slate/BaseMatrix.hh:863-905)
    BaseMatrix(BaseMatrix const& orig, int64_t i1, int64_t i2, int64_t j1, int64_t j2)
        : BaseMatrix(orig)
    {
        if (i2 < i1 || j2 < j1) {       // empty view
            mt_ = nt_ = 0;
            return;
        }
        slate_assert(0 <= i1 && i2 < orig.mt() && 0 <= j1 && j2 < orig.nt());
        if (op_ == Op::NoTrans) {
            ioffset_ += i1;  joffset_ += j1;
            mt_ = i2 - i1 + 1;  nt_ = j2 - j1 + 1;
        }
        else {
            ioffset_ += j1;  joffset_ += i1;
            mt_ = j2 - j1 + 1;  nt_ = i2 - i1 + 1;
        }
    }

public:
    // ---- shape -------------------------------------------------------------------------
    int64_t mt() const { return op_ == Op::NoTrans ? mt_ : nt_; }
    int64_t nt() const { return op_ == Op::NoTrans ? nt_ : mt_; }
    Op op() const { return op_; }
    Uplo uplo() const { return uploLogical(); }
    Uplo uploLogical() const
    {
        if (uplo_ == Uplo::General) return Uplo::General;
        if ((uplo_ == Uplo::Lower) == (op_ == Op::NoTrans)) return Uplo::Lower;
        return Uplo::Upper;
    }
    Uplo uploPhysical() const { return uplo_; }
    Layout layout() const { return layout_; }
    Target origin() const { return origin_; }

    int64_t tileMb(int64_t i) const
    {
        return op_ == Op::NoTrans ? storage_->tileMb(ioffset_ + i) : storage_->tileNb(joffset_ + i);
    }
    int64_t tileNb(int64_t j) const
    {
        return op_ == Op::NoTrans ? storage_->tileNb(joffset_ + j) : storage_->tileMb(ioffset_ + j);
    }
    int64_t m() const { int64_t s = 0; for (int64_t i = 0; i < mt(); ++i) s += tileMb(i); return s; }
    int64_t n() const { int64_t s = 0; for (int64_t j = 0; j < nt(); ++j) s += tileNb(j); return s; }

    /// (i, j) of the view -> global storage index (BaseMatrix.hh:3975-3990)
    ij_tuple globalIndex(int64_t i, int64_t j) const
    {
        return op_ == Op::NoTrans ? ij_tuple{ioffset_ + i, joffset_ + j}
                                  : ij_tuple{ioffset_ + j, joffset_ + i};
    }

    int tileRank(int64_t i, int64_t j) const { return storage_->tileRank(globalIndex(i, j)); }
    int tileDevice(int64_t i, int64_t j) const { return storage_->tileDevice(globalIndex(i, j)); }
    bool tileIsLocal(int64_t i, int64_t j) const { return storage_->tileIsLocal(globalIndex(i, j)); }
    bool tileExists(int64_t i, int64_t j, int device = AnyDevice) const
    {
        auto* node = storage_->find(globalIndex(i, j));
        if (!node) return false;
        if (device == AnyDevice) return true;
        return device == HostNum ? node->host != nullptr : node->dev != nullptr;
    }
    int num_devices() const { return Engine::instance().num_devices(); }
    MPI_Comm mpiComm() const { return storage_->comm_; }
    int mpiRank() const { return storage_->comm_->rank; }
    void gridinfo(GridOrder* order, int* p, int* q, int* myrow, int* mycol) const
    {
        *order = storage_->order_;  *p = storage_->p_;  *q = storage_->q_;
        int r = storage_->comm_->rank;
        if (storage_->order_ == GridOrder::Col) { *myrow = r % *p;  *mycol = r / *p; }
        else { *myrow = r / *q;  *mycol = r % *q; }
    }
    void* compute_queue(int device, int queue_index) { return Engine::instance().compute_stream(queue_index); }
    void* comm_queue(int device) { return Engine::instance().comm_stream(); }

    /// Tile handle with this view's op / uplo applied (BaseMatrix.hh:1066-1090).
    Tile<scalar_t> operator()(int64_t i, int64_t j, int device = HostNum)
    {
        auto& node = storage_->at(globalIndex(i, j));
        Tile<scalar_t>* t = (device == HostNum) ? node.host : node.dev;
        if (!t)
            slate_error("tile (" + std::to_string(i) + ", " + std::to_string(j) + ") has no instance on "
                        + (device == HostNum ? std::string("the host") : std::string("the device")));
        Tile<scalar_t> tile = *t;
        tile.op(op_);
        if (i == j) tile.uplo(uplo_);
        return tile;
    }

    // ---- tile management -----------------------------------------------------------------
    /// workspace tile (BaseMatrix.hh:1253-1270)
    Tile<scalar_t>* tileInsertWorkspace(int64_t i, int64_t j, int device)
    {
        return storage_->tileInsert(globalIndex(i, j), device, TileKind::Workspace);
    }
    Tile<scalar_t>* tileInsert(int64_t i, int64_t j, int device = HostNum)
    {
        auto* t = storage_->tileInsert(globalIndex(i, j), device, TileKind::SlateOwned);
        auto& node = storage_->at(globalIndex(i, j));
        (device == HostNum ? node.host_state : node.dev_state) = MOSI::Shared;
        return t;
    }
    Tile<scalar_t>* tileInsert(int64_t i, int64_t j, int device, scalar_t* A, int64_t ld)
    {
        return storage_->tileInsert(globalIndex(i, j), device, A, ld);
    }
    void tileErase(int64_t i, int64_t j, int device)
    {
        storage_->tileErase(globalIndex(i, j), device);
    }

    /// MOSI state of one instance (BaseMatrix.hh:1719-1740)
    short tileState(int64_t i, int64_t j, int device = HostNum)
    {
        auto& node = storage_->at(globalIndex(i, j));
        return device == HostNum ? node.host_state : node.dev_state;
    }

    /// Marks the instance on `device` Modified and the other Invalid (BaseMatrix.hh:2811-2840).
    void tileModified(int64_t i, int64_t j, int device = HostNum, int queue = -1)
    {
        auto& node = storage_->at(globalIndex(i, j));
        if (device == HostNum) {
            slate_assert(node.host);
            node.host_state = MOSI::Modified;
            if (node.dev) node.dev_state = MOSI::Invalid;
        }
        else {
            slate_assert(node.dev);
            node.dev_state = MOSI::Modified;
            if (node.host) node.host_state = MOSI::Invalid;
            if (queue >= 0) {
                node.write_queue = queue;
                node.write_seq = Engine::instance().next_queue_seq(queue);
            }
        }
    }

    /// Makes the instance on `device` valid, copying from the other one if needed
    /// (tileGet, BaseMatrix.hh:2640-2720).  Host-to-device copies are enqueued on the comm
    /// stream (call Engine::mark_comm + queue_wait_comm before the consuming launch, or use
    /// the set overloads which do it); device-to-host copies wait for the producing queue
    /// and are complete on return.
    void tileGetForReading(int64_t i, int64_t j, int device = HostNum,
                           LayoutConvert lc = LayoutConvert::None, int queue = -1)
    {
        tileGet(i, j, device, false, queue);
        if (device == HostNum) finishHostCopies();
        else convertIfNeeded(i, j, device, lc, queue);
    }
    void tileGetForWriting(int64_t i, int64_t j, int device = HostNum,
                           LayoutConvert lc = LayoutConvert::None, int queue = -1)
    {
        // a conversion is itself a launch with its own mark: do it before the tile is stamped
        // with the mark of the launch that is being prepared
        if (device != HostNum && needsConvert(i, j, lc)) {
            tileGet(i, j, device, false, queue);
            finishSet(device, queue);
            convertIfNeeded(i, j, device, lc, queue);
        }
        tileGet(i, j, device, true, queue);
        if (device == HostNum) finishHostCopies();
    }
    void tileGetForReading(std::set<ij_tuple>& tiles, int device,
                           LayoutConvert lc = LayoutConvert::None, int queue = -1)
    {
        for (auto const& ij : tiles) tileGet(std::get<0>(ij), std::get<1>(ij), device, false, queue);
        finishSet(device, queue);
        if (device != HostNum && lc != LayoutConvert::None && storage_->any_rowmajor_)
            for (auto const& ij : tiles) convertIfNeeded(std::get<0>(ij), std::get<1>(ij), device, lc, queue);
    }
    void tileGetForWriting(std::set<ij_tuple>& tiles, int device,
                           LayoutConvert lc = LayoutConvert::None, int queue = -1)
    {
        if (device != HostNum && lc != LayoutConvert::None && storage_->any_rowmajor_) {
            // conversions first (see the single-tile overload)
            bool any = false;
            for (auto const& ij : tiles)
                if (needsConvert(std::get<0>(ij), std::get<1>(ij), lc)) {
                    tileGet(std::get<0>(ij), std::get<1>(ij), device, false, queue);
                    any = true;
                }
            if (any) {
                finishSet(device, queue);
                for (auto const& ij : tiles)
                    convertIfNeeded(std::get<0>(ij), std::get<1>(ij), device, lc, queue);
            }
        }
        for (auto const& ij : tiles) tileGet(std::get<0>(ij), std::get<1>(ij), device, true, queue);
        finishSet(device, queue);
    }

    /// Layout of the instance on `device` (BaseMatrix.hh tileLayout).  Host instances are
    /// always ColMajor in this library.
    Layout tileLayout(int64_t i, int64_t j, int device = HostNum)
    {
        auto& node = storage_->at(globalIndex(i, j));
        Tile<scalar_t>* t = (device == HostNum) ? node.host : node.dev;
        slate_assert(t != nullptr);
        return t->layout();
    }

    /// Transposes the stored array of the device instance so that it has `layout`
    /// (BaseMatrix.hh tileLayoutConvert; the kernels of src/cuda/device_transpose.cu become
    /// te_tile_transpose).  Square contiguous tiles are converted in place; rectangular ones
    /// into a fresh pool block that replaces the old one.  Enqueued on compute queue `queue`,
    /// or on the comm stream if queue < 0, behind the last writer of the tile.
    void tileLayoutConvert(int64_t i, int64_t j, int device, Layout layout, int queue = -1);
    /// BaseMatrix.hh:3081-3104 (fetch_C in gemmC.cc:70-74)
    void tileGetAllForWritingOnDevices(LayoutConvert = LayoutConvert::None, int queue = -1)
    {
        tileGetAllOnDevices(true, queue);
    }
    void tileGetAllForReadingOnDevices(LayoutConvert = LayoutConvert::None, int queue = -1)
    {
        tileGetAllOnDevices(false, queue);
    }
    /// block column by block column: local tiles that are neighbours in the caller's host array
    /// travel as one copy (tileGetRun)
    void tileGetAllOnDevices(bool modify, int queue)
    {
        std::vector<std::pair<int64_t, int64_t>> run;
        const bool by_column = (op_ == Op::NoTrans);
        for (int64_t a = 0; a < (by_column ? nt() : mt()); ++a) {
            run.clear();
            for (int64_t b = 0; b < (by_column ? mt() : nt()); ++b) {
                const int64_t i = by_column ? b : a, j = by_column ? a : b;
                if (tileIsLocal(i, j) && inStoredTriangle(i, j)) run.push_back({i, j});
            }
            tileGetRun(run, modify, queue);
        }
        finishSet(0, queue);
    }

    /// Refreshes the origin instance from the Modified replica (BaseMatrix.hh:3190-3243).
    void tileUpdateOrigin(int64_t i, int64_t j)
    {
        updateOrigin(i, j);
        finishHostCopies();
    }
    /// Same, but only enqueues the copies to the host (each behind the launch that produced
    /// its tile); a later tileUpdateAllOrigin() / Engine::sync_all() completes them.
    void tileUpdateAllOriginAsync()
    {
        updateAllOrigin();
        host_copies_pending_ = false;
    }
    void tileUpdateAllOrigin()
    {
        updateAllOrigin();
        finishHostCopies();
    }
    /// block column by block column (of the stored matrix): neighbours in the caller's host
    /// array go home as one copy (updateOriginRun)
    void updateAllOrigin()
    {
        std::vector<std::pair<int64_t, int64_t>> run;
        const bool by_column = (op_ == Op::NoTrans);
        for (int64_t a = 0; a < (by_column ? nt() : mt()); ++a) {
            run.clear();
            for (int64_t b = 0; b < (by_column ? mt() : nt()); ++b) {
                const int64_t i = by_column ? b : a, j = by_column ? a : b;
                if (tileIsLocal(i, j) && inStoredTriangle(i, j) && tileExists(i, j)) run.push_back({i, j});
            }
            updateOriginRun(run);
        }
    }

    // ---- workspace -----------------------------------------------------------------------
    /// gemmC.cc:57-60 / potrf.cc:69-71.  Queues and descriptor rings are created lazily by
    /// the engine, so this only keeps the call sites compiling.
    void allocateBatchArrays(int64_t batch_size = 0, int64_t num_arrays = 1)
    {
        for (int q = 0; q < num_arrays && q < Engine::MaxQueues; ++q)
            Engine::instance().compute_stream(q);
    }
    void clearBatchArrays() {}
    int64_t getMaxDeviceTiles(int device)
    {
        int64_t count = 0;
        for (int64_t j = 0; j < nt(); ++j)
            for (int64_t i = 0; i < mt(); ++i)
                if (tileIsLocal(i, j) && inStoredTriangle(i, j)) ++count;
        return count;
    }
    /// Matrix.hh:745-768 / MatrixStorage.hh:697-757: make sure the pool holds a block for
    /// every local tile not born on the device plus `extra` blocks for broadcast panels.
    void reserveDeviceWorkspace(int64_t extra_blocks = -1)
    {
        if (num_devices() == 0) return;
        int64_t need = 0;
        if (origin_ != Target::Devices) need += getMaxDeviceTiles(0);
        if (extra_blocks < 0) {
            // (lookahead + 2) panels of one block row + one block column of remote tiles
            extra_blocks = (storage_->comm_->size > 1) ? 4 * (storage_->mt_ + storage_->nt_) : 0;
        }
        need += extra_blocks;
        int64_t have = (int64_t)storage_->memory_.available(0);
        if (need > have) storage_->memory_.addDeviceBlocks(0, need - have);
    }

    /// MatrixStorage::ensureDeviceWorkspace (MatrixStorage.hh:759-790): at least `blocks`
    /// free blocks in the pool, for broadcast panels and host-origin staging.  A matrix whose
    /// tiles were born on the device of a single-rank run needs none.
    void ensureDeviceWorkspace(int64_t blocks)
    {
        if (num_devices() == 0) return;
        if (origin_ == Target::Devices && storage_->comm_->size == 1) return;
        int64_t have = (int64_t)storage_->memory_.available(0);
        if (blocks > have) storage_->memory_.addDeviceBlocks(0, blocks - have);
    }

    /// Erases tiles that are not local (received by broadcast) once their receive count
    /// drops to zero (BaseMatrix.hh:3875-3937).
    void releaseRemoteWorkspace(int64_t receive_count = 1)
    {
        for (int64_t j = 0; j < nt(); ++j)
            for (int64_t i = 0; i < mt(); ++i)
                if (!tileIsLocal(i, j)) releaseRemoteWorkspaceTile(i, j, receive_count);
        storage_->flushErased();
    }
    void releaseRemoteWorkspaceTile(int64_t i, int64_t j, int64_t receive_count = 1)
    {
        auto* node = storage_->find(globalIndex(i, j));
        if (!node) return;
        node->receive_count -= receive_count;
        if (node->receive_count <= 0) {
            storage_->tileErase(globalIndex(i, j), 0);
            storage_->tileErase(globalIndex(i, j), HostNum);
        }
    }
    /// Erases workspace replicas of local tiles unless they hold the only valid copy
    /// (MatrixStorage.hh:894-933).
    void releaseLocalWorkspace()
    {
        for (int64_t j = 0; j < nt(); ++j)
            for (int64_t i = 0; i < mt(); ++i)
                if (tileIsLocal(i, j)) releaseLocalWorkspaceTile(i, j);
        storage_->flushErased();
    }
    void releaseLocalWorkspaceTile(int64_t i, int64_t j)
    {
        auto* node = storage_->find(globalIndex(i, j));
        if (!node) return;
        if (node->dev && node->dev->workspace() && node->dev_state != MOSI::Modified
            && node->host && node->host_state != MOSI::Invalid)
            storage_->tileErase(globalIndex(i, j), 0);
        else if (node->host && node->host->workspace() && node->host_state != MOSI::Modified
                 && node->dev && node->dev_state != MOSI::Invalid)
            storage_->tileErase(globalIndex(i, j), HostNum);
    }
    void releaseWorkspace()
    {
        releaseRemoteWorkspace(1 << 30);
        releaseLocalWorkspace();
    }
    void clearWorkspace() { releaseWorkspace(); }

    // ---- broadcasts ------------------------------------------------------------------------
    /// Send tile (i, j) to every rank owning a tile of B (BaseMatrix.hh:1961-1990).
    template <Target target = Target::Host>
    void tileBcast(int64_t i, int64_t j, BaseMatrix<scalar_t> const& B,
                   Layout layout = Layout::ColMajor, int tag = 0, int64_t life_factor = 1)
    {
        BcastList list;
        list.push_back({i, j, {B}});
        listBcast<target>(list, layout, tag, life_factor);
    }

    /// BaseMatrix.hh:2028-2106.  One NCCL group for the whole list, on the comm stream.
    template <Target target = Target::Host>
    void listBcast(BcastList& bcast_list, Layout layout = Layout::ColMajor, int tag = 0,
                   int64_t life_factor = 1, bool is_shared = false)
    {
        BcastListTag tagged;
        tagged.reserve(bcast_list.size());
        for (auto& e : bcast_list)
            tagged.push_back({std::get<0>(e), std::get<1>(e), std::get<2>(e), (int64_t)tag});
        listBcastMT<target>(tagged, layout, life_factor, is_shared);
    }

    /// BaseMatrix.hh:2129-2213.  Tags are not needed: NCCL matches sends and receives in
    /// issue order per pair of ranks, and every rank walks the list in the same order.
    template <Target target = Target::Host>
    void listBcastMT(BcastListTag& bcast_list, Layout layout = Layout::ColMajor,
                     int64_t life_factor = 1, bool is_shared = false);

    /// One tile broadcast of a plan: tile (i, j) of this view goes from rank `root` to `dests`.
    struct BcastOp { int64_t i, j; int root; std::vector<int> dests; };

    /// The host-side half of listBcast (BaseMatrix.hh:2129-2213 + getRanks :3702-3756): which
    /// rank sends which tile to which ranks.  Pure index arithmetic on the 2D block-cyclic map,
    /// no device and no communicator needed; listBcastMT executes exactly this plan, and the
    /// CPU tests execute it over gloo.  Entries whose destination set is empty are dropped.
    std::vector<BcastOp> bcastPlan(BcastListTag const& bcast_list) const
    {
        std::vector<BcastOp> plan;
        const int nranks = storage_->comm_->size;
        std::vector<char> in_set(nranks);
        for (auto const& e : bcast_list) {
            BcastOp op;
            op.i = std::get<0>(e);
            op.j = std::get<1>(e);
            op.root = tileRank(op.i, op.j);
            std::fill(in_set.begin(), in_set.end(), 0);
            for (auto const& B : std::get<2>(e))
                for (int64_t jj = 0; jj < B.nt(); ++jj)
                    for (int64_t ii = 0; ii < B.mt(); ++ii)
                        if (B.inStoredTriangle(ii, jj)) in_set[B.tileRank(ii, jj)] = 1;
            in_set[op.root] = 0;
            for (int r = 0; r < nranks; ++r)
                if (in_set[r]) op.dests.push_back(r);
            if (!op.dests.empty()) plan.push_back(std::move(op));
        }
        return plan;
    }

    /// All tiles one rank sends to (receives from) one peer within a list, in plan order.
    struct PairMsg { int peer; std::vector<size_t> ops; size_t bytes; };

    /// Groups a broadcast plan by (root, destination) pair as seen from rank `me`: the messages
    /// `me` sends as a root and those it receives.  Sender and receiver derive the same message
    /// (same tiles, same order, same byte count) from the same plan, so a whole pair can travel
    /// as ONE NCCL message instead of one per tile (95 GB/s per pair at 2 MiB against 316 GB/s at
    /// 128 MiB on this pool's NVLink, scripts/mp_commbench.py).  Pure index arithmetic.
    void bcastPairs(std::vector<BcastOp> const& plan, int me, std::vector<PairMsg>* sends,
                    std::vector<PairMsg>* recvs) const
    {
        std::map<int, size_t> send_at, recv_at;
        for (size_t o = 0; o < plan.size(); ++o) {
            auto const& op = plan[o];
            const size_t bytes = sizeof(scalar_t) * (size_t)tileMb(op.i) * (size_t)tileNb(op.j);
            if (op.root == me) {
                for (int d : op.dests) {
                    auto it = send_at.find(d);
                    if (it == send_at.end()) {
                        it = send_at.emplace(d, sends->size()).first;
                        sends->push_back(PairMsg{d, {}, 0});
                    }
                    (*sends)[it->second].ops.push_back(o);
                    (*sends)[it->second].bytes += bytes;
                }
            }
            else if (std::find(op.dests.begin(), op.dests.end(), me) != op.dests.end()) {
                auto it = recv_at.find(op.root);
                if (it == recv_at.end()) {
                    it = recv_at.emplace(op.root, recvs->size()).first;
                    recvs->push_back(PairMsg{op.root, {}, 0});
                }
                (*recvs)[it->second].ops.push_back(o);
                (*recvs)[it->second].bytes += bytes;
            }
        }
    }

    /// One tile reduction of a plan: the partial sums of tile (i, j) held by `sources` are
    /// added into the instance on `root`.
    struct ReduceOp { int64_t i, j; int root; std::vector<int> sources; };

    /// The host-side half of listReduce (BaseMatrix.hh:2219-2255): root = owner of the
    /// destination view's first tile, sources = owners of the listed views, root excluded.
    /// Entries without a remote source are dropped.
    std::vector<ReduceOp> reducePlan(ReduceList const& reduce_list) const
    {
        std::vector<ReduceOp> plan;
        for (auto const& e : reduce_list) {
            ReduceOp op;
            op.i = std::get<0>(e);
            op.j = std::get<1>(e);
            op.root = std::get<2>(e).tileRank(0, 0);
            std::set<int> from;
            for (auto const& S : std::get<3>(e)) S.getRanks(&from);
            from.erase(op.root);
            op.sources.assign(from.begin(), from.end());
            if (!op.sources.empty()) plan.push_back(std::move(op));
        }
        return plan;
    }

    /// BaseMatrix.hh:2219-2255, device-resident: every source sends its partial tile to the
    /// root over NCCL (one group for the whole list) and drops it; the root receives into
    /// pool blocks and adds them to its tile on the comm stream.  The reference reduces on
    /// the host over an MPI hypercube; with NVSwitch every pair has full bandwidth and the
    /// direct pattern needs one hop.
    void listReduce(ReduceList& reduce_list, Layout layout = Layout::ColMajor, int tag = 0);

    /// ranks owning at least one tile of this view (BaseMatrix.hh:3702-3756)
    void getRanks(std::set<int>* bcast_set) const
    {
        for (int64_t j = 0; j < nt(); ++j)
            for (int64_t i = 0; i < mt(); ++i)
                if (inStoredTriangle(i, j)) bcast_set->insert(tileRank(i, j));
    }

    std::shared_ptr<MatrixStorage<scalar_t>> const& storage() const { return storage_; }

protected:
    /// general matrices store everything; trapezoid ones override through uplo_
    bool inStoredTriangle(int64_t i, int64_t j) const
    {
        if (uplo_ == Uplo::General) return true;
        Uplo u = uploLogical();
        return u == Uplo::Lower ? i >= j : i <= j;
    }

    void tileGet(int64_t i, int64_t j, int device, bool modify, int queue);
    void updateOrigin(int64_t i, int64_t j);
    /// tileGet(.., device 0, ..) / updateOrigin for a list of tiles in the order they lie in the
    /// caller's host array; neighbours cross PCIe as ONE copy through a device panel
    void tileGetRun(std::vector<std::pair<int64_t, int64_t>> const& tiles, bool modify, int queue);
    void updateOriginRun(std::vector<std::pair<int64_t, int64_t>> const& tiles);
    /// roots of a broadcast plan: their host-origin tiles are fetched run by run
    void fetchRootRuns(std::vector<BcastOp> const& plan, int me);
    /// 1: b lies right below a in the same host array (block column); 2: right of it (block
    /// row); 0: neither
    /// SLATE_B200_TILE_RUNS=1 turns the panel path on (default off: every tile its own copy).
    /// Measured with bench.py's e2e arm at 20 steps (round 2): no gain where it was expected --
    /// 8 GPUs 137.9 against 137.4 TFLOP/s, 1 GPU 29.3 against 28.3-30.4 across boxes -- and 3 %
    /// slower on 2 GPUs (56.0 against 57.7: the scatter / gather kernels wait for SMs the update
    /// holds, and the tiles cross HBM twice).  The host's copy-enqueue time is real (21 us per
    /// copy) but, once the write-backs are enqueued late (drivers.cc), it is hidden behind the
    /// 16-deep launch queue.
    static bool tile_runs_enabled()
    {
        static const bool on = [] {
            const char* e = std::getenv("SLATE_B200_TILE_RUNS");
            return e && e[0] == '1';
        }();
        return on;
    }
    static int hostNeighbours(Tile<scalar_t> const* a, Tile<scalar_t> const* b)
    {
        if (a->stride() != b->stride()) return 0;
        if (a->nb() == b->nb() && b->data() == a->data() + a->mb()) return 1;
        if (a->mb() == b->mb() && b->data() == a->data() + (int64_t)a->nb() * a->stride()) return 2;
        return 0;
    }
    void bcastCoalesced(std::vector<BcastOp> const& plan);
    void bcastPeer(std::vector<BcastOp> const& plan);
    Tile<scalar_t>* prepareToReceive(int64_t i, int64_t j);
    /// true if tile (i, j) has a valid device instance in the other layout
    bool needsConvert(int64_t i, int64_t j, LayoutConvert lc)
    {
        if (lc == LayoutConvert::None || !storage_->any_rowmajor_) return false;
        auto* node = storage_->find(globalIndex(i, j));
        return node && node->dev && node->dev_state != MOSI::Invalid
               && node->dev->layout() != Layout(char(lc));
    }
    void convertIfNeeded(int64_t i, int64_t j, int device, LayoutConvert lc, int queue)
    {
        if (lc == LayoutConvert::None || !storage_->any_rowmajor_) return;
        auto* node = storage_->find(globalIndex(i, j));
        if (node && node->dev && node->dev_state != MOSI::Invalid
            && node->dev->layout() != Layout(char(lc)))
            tileLayoutConvert(i, j, device, Layout(char(lc)), queue);
    }
    void finishHostCopies()
    {
        if (host_copies_pending_) {
            Engine::instance().sync_home();
            host_copies_pending_ = false;
        }
    }
    void finishSet(int device, int queue)
    {
        if (device == HostNum) { finishHostCopies(); return; }
        if (device_copies_pending_) {
            Engine::instance().mark_comm();
            device_copies_pending_ = false;
        }
        if (queue >= 0) Engine::instance().queue_wait_comm_seq(queue, needed_comm_seq_);
        needed_comm_seq_ = 0;
    }

    int64_t ioffset_, joffset_, mt_, nt_;
    Uplo uplo_;
    Op op_;
    Layout layout_;
    Target origin_;
    std::shared_ptr<MatrixStorage<scalar_t>> storage_;
    bool host_copies_pending_ = false;
    bool device_copies_pending_ = false;
    uint64_t needed_comm_seq_ = 0;

    template <typename T> friend class BaseMatrix;
};

//------------------------------------------------------------------------------
template <typename scalar_t>
void BaseMatrix<scalar_t>::tileGet(int64_t i, int64_t j, int device, bool modify, int queue)
{
    Engine& eng = Engine::instance();
    auto& node = storage_->at(globalIndex(i, j));
    if (device == HostNum) {
        if (!(node.host && node.host_state != MOSI::Invalid)) {
            slate_assert(node.dev && node.dev_state != MOSI::Invalid);
            if (!node.host) storage_->tileInsert(globalIndex(i, j), HostNum, TileKind::Workspace);
            // host instances are ColMajor: bring the device array back first (comm stream)
            if (node.dev->layout() != Layout::ColMajor)
                tileLayoutConvert(i, j, 0, Layout::ColMajor, -1);
            // device -> host, on the home stream: after the launch that produced the data and
            // after the comm-stream work (broadcast, layout conversion) that filled the tile
            eng.home_wait_queue_seq(node.write_queue, node.write_seq);
            eng.home_wait_comm_seq(node.comm_seq);
            Tile<scalar_t>* s = node.dev;
            Tile<scalar_t>* d = node.host;
            slate_te_call(te_memcpy2d_async(d->data(), sizeof(scalar_t) * d->stride(), s->data(),
                                            sizeof(scalar_t) * s->stride(),
                                            sizeof(scalar_t) * s->mb(), s->nb(), TE_COPY_D2H,
                                            eng.home_stream()));
            host_copies_pending_ = true;
            node.host_state = MOSI::Shared;
            if (node.dev_state == MOSI::Modified) node.dev_state = MOSI::Shared;
        }
        if (modify) {
            node.host_state = MOSI::Modified;
            if (node.dev) node.dev_state = MOSI::Invalid;
        }
        return;
    }
    // ---- device instance ----
    if (node.dev && node.dev_state != MOSI::Invalid) {
        // read-after-write / write-after-write across compute queues
        if (queue >= 0 && node.write_queue >= 0 && node.write_queue != queue)
            eng.queue_wait_queue_seq(queue, node.write_queue, node.write_seq);
        // filled by a broadcast or a host-to-device copy: the consumer waits for that mark
        if (node.comm_seq > needed_comm_seq_) needed_comm_seq_ = node.comm_seq;
    }
    else {
        slate_assert(node.host && node.host_state != MOSI::Invalid);
        if (!node.dev) storage_->tileInsert(globalIndex(i, j), 0, TileKind::Workspace);
        Tile<scalar_t>* s = node.host;
        Tile<scalar_t>* d = node.dev;
        // an invalid device array left in RowMajor is simply overwritten in ColMajor
        if (d->layout() != Layout::ColMajor && d->kind() != TileKind::UserOwned)
            d->setLayout(Layout::ColMajor, d->mb());
        slate_te_call(te_memcpy2d_async(d->data(), sizeof(scalar_t) * d->stride(), s->data(),
                                        sizeof(scalar_t) * s->stride(), sizeof(scalar_t) * s->mb(),
                                        s->nb(), TE_COPY_H2D, eng.comm_stream()));
        device_copies_pending_ = true;
        node.dev_state = MOSI::Shared;
        node.write_queue = -1;
        node.comm_seq = eng.next_comm_seq();
        if (node.comm_seq > needed_comm_seq_) needed_comm_seq_ = node.comm_seq;
        if (node.host_state == MOSI::Modified) node.host_state = MOSI::Shared;
    }
    if (modify) {
        node.dev_state = MOSI::Modified;
        if (node.host) node.host_state = MOSI::Invalid;
        if (queue >= 0) {
            node.write_queue = queue;
            node.write_seq = eng.next_queue_seq(queue);
        }
    }
}

template <typename scalar_t>
void BaseMatrix<scalar_t>::updateOrigin(int64_t i, int64_t j)
{
    auto& node = storage_->at(globalIndex(i, j));
    const bool host_is_origin = node.host && node.host->origin();
    const bool dev_is_origin = node.dev && node.dev->origin();
    if (host_is_origin && node.host_state == MOSI::Invalid)
        tileGet(i, j, HostNum, false, -1);
    else if (dev_is_origin && node.dev_state == MOSI::Invalid) {
        tileGet(i, j, 0, false, -1);
        device_copies_pending_ = false;
        Engine::instance().mark_comm();
    }
}

//------------------------------------------------------------------------------
/// Host -> device for tiles in host-memory order.  Maximal runs of neighbours (same host
/// array: a block column's tiles stacked in rows, or a block row's side by side) are moved by
/// one host-to-device copy into a panel on the comm stream and spread to the tiles' blocks by
/// te_panel_pieces; single tiles, tiles already valid on the device and tiles without a valid
/// host instance take tileGet.  Why: each host<->device copy costs the enqueueing thread ~21 us
/// whatever its size, which made the host the bottleneck of the host-origin gemm.
template <typename scalar_t>
void BaseMatrix<scalar_t>::tileGetRun(std::vector<std::pair<int64_t, int64_t>> const& tiles,
                                      bool modify, int queue)
{
    Engine& eng = Engine::instance();
    const int dtype = is_complex ? TE_C64 : TE_F64;
    const size_t es = sizeof(scalar_t);
    const size_t max_panel = (size_t)512 << 20;
    auto node_of = [&](size_t t) { return storage_->find(globalIndex(tiles[t].first, tiles[t].second)); };
    auto eligible = [&](size_t t) {
        auto* nd = node_of(t);
        return tile_runs_enabled() && nd && nd->host && nd->host_state != MOSI::Invalid && nd->host->op() == Op::NoTrans
               && !(nd->dev && nd->dev_state != MOSI::Invalid)
               && !(nd->dev && nd->dev->kind() == TileKind::UserOwned);
    };
    size_t t = 0;
    while (t < tiles.size()) {
        if (!eligible(t)) {
            tileGet(tiles[t].first, tiles[t].second, 0, modify, queue);
            ++t;
            continue;
        }
        // extend the run while the next tile is the same kind of neighbour
        size_t e = t + 1;
        int kind = 0;
        size_t bytes = node_of(t)->host->bytes();
        while (e < tiles.size() && e - t < TE_MAX_PANEL_PIECES && eligible(e)) {
            const int k = hostNeighbours(node_of(e - 1)->host, node_of(e)->host);
            if (k == 0 || (kind != 0 && k != kind)) break;
            if (bytes + node_of(e)->host->bytes() > max_panel) break;
            kind = k;
            bytes += node_of(e)->host->bytes();
            ++e;
        }
        if (e - t < 2) {
            tileGet(tiles[t].first, tiles[t].second, 0, modify, queue);
            ++t;
            continue;
        }
        Tile<scalar_t>* first = node_of(t)->host;
        // panel: kind 1 = (sum of mb) x nb, kind 2 = mb x (sum of nb); column major, tight
        int64_t ext = 0;
        for (size_t u = t; u < e; ++u) ext += (kind == 1) ? node_of(u)->host->mb() : node_of(u)->host->nb();
        const int64_t ld_panel = (kind == 1) ? ext : first->mb();
        const int64_t panel_cols = (kind == 1) ? first->nb() : ext;
        Engine::CommSlab& slab = eng.acquire_slab(false, (size_t)ld_panel * panel_cols * es);
        slate_te_call(te_memcpy2d_async(slab.dev, es * ld_panel, first->data(), es * first->stride(),
                                        es * ld_panel, panel_cols, TE_COPY_H2D, eng.comm_stream()));
        std::vector<te_panel_piece> pieces;
        int64_t off = 0;
        for (size_t u = t; u < e; ++u) {
            const ij_tuple gij = globalIndex(tiles[u].first, tiles[u].second);
            auto& node = storage_->at(gij);
            if (!node.dev) storage_->tileInsert(gij, 0, TileKind::Workspace);
            Tile<scalar_t>* d = node.dev;
            if (d->layout() != Layout::ColMajor) d->setLayout(Layout::ColMajor, d->mb());
            te_panel_piece pc;
            pc.tile = d->data();
            pc.panel_offset = (kind == 1) ? off : off * ld_panel;
            pc.m = (int32_t)d->mb();
            pc.n = (int32_t)d->nb();
            pc.ld = (int32_t)d->stride();
            pc.reserved = 0;
            pieces.push_back(pc);
            off += (kind == 1) ? d->mb() : d->nb();
            // bookkeeping exactly as tileGet
            node.dev_state = MOSI::Shared;
            node.write_queue = -1;
            node.comm_seq = eng.next_comm_seq();
            if (node.comm_seq > needed_comm_seq_) needed_comm_seq_ = node.comm_seq;
            if (node.host_state == MOSI::Modified) node.host_state = MOSI::Shared;
            if (modify) {
                node.dev_state = MOSI::Modified;
                node.host_state = MOSI::Invalid;
                if (queue >= 0) {
                    node.write_queue = queue;
                    node.write_seq = eng.next_queue_seq(queue);
                }
            }
        }
        slate_te_call(te_panel_pieces(eng.device(), eng.comm_stream(), dtype, 1, slab.dev, ld_panel,
                                      pieces.data(), (int)pieces.size()));
        eng.retire_slab(false, slab);
        device_copies_pending_ = true;
        t = e;
    }
}

//------------------------------------------------------------------------------
/// Device -> host (updateOrigin) for tiles in host-memory order: neighbours are gathered into a
/// panel on the home stream and go home as one copy.
template <typename scalar_t>
void BaseMatrix<scalar_t>::updateOriginRun(std::vector<std::pair<int64_t, int64_t>> const& tiles)
{
    Engine& eng = Engine::instance();
    const int dtype = is_complex ? TE_C64 : TE_F64;
    const size_t es = sizeof(scalar_t);
    const size_t max_panel = (size_t)512 << 20;
    auto node_of = [&](size_t t) { return storage_->find(globalIndex(tiles[t].first, tiles[t].second)); };
    auto eligible = [&](size_t t) {
        auto* nd = node_of(t);
        return tile_runs_enabled() && nd && nd->host && nd->host->origin() && nd->host_state == MOSI::Invalid
               && nd->host->op() == Op::NoTrans && nd->dev && nd->dev_state != MOSI::Invalid
               && nd->dev->layout() == Layout::ColMajor;
    };
    size_t t = 0;
    while (t < tiles.size()) {
        if (!eligible(t)) {
            updateOrigin(tiles[t].first, tiles[t].second);
            ++t;
            continue;
        }
        size_t e = t + 1;
        int kind = 0;
        size_t bytes = node_of(t)->host->bytes();
        while (e < tiles.size() && e - t < TE_MAX_PANEL_PIECES && eligible(e)) {
            const int k = hostNeighbours(node_of(e - 1)->host, node_of(e)->host);
            if (k == 0 || (kind != 0 && k != kind)) break;
            if (bytes + node_of(e)->host->bytes() > max_panel) break;
            kind = k;
            bytes += node_of(e)->host->bytes();
            ++e;
        }
        if (e - t < 2) {
            updateOrigin(tiles[t].first, tiles[t].second);
            ++t;
            continue;
        }
        Tile<scalar_t>* first = node_of(t)->host;
        int64_t ext = 0;
        for (size_t u = t; u < e; ++u) ext += (kind == 1) ? node_of(u)->host->mb() : node_of(u)->host->nb();
        const int64_t ld_panel = (kind == 1) ? ext : first->mb();
        const int64_t panel_cols = (kind == 1) ? first->nb() : ext;
        std::vector<te_panel_piece> pieces;
        int64_t off = 0;
        int last_q = -2;
        uint64_t last_qseq = 0, last_cseq = 0;
        for (size_t u = t; u < e; ++u) {
            auto& node = *node_of(u);
            // the home stream behind the launch that produced the tile and the comm work that
            // filled it (one wait per distinct mark)
            if (node.write_queue != last_q || node.write_seq > last_qseq) {
                eng.home_wait_queue_seq(node.write_queue, node.write_seq);
                last_q = node.write_queue;
                last_qseq = node.write_seq;
            }
            if (node.comm_seq > last_cseq) {
                eng.home_wait_comm_seq(node.comm_seq);
                last_cseq = node.comm_seq;
            }
            Tile<scalar_t>* d = node.dev;
            te_panel_piece pc;
            pc.tile = d->data();
            pc.panel_offset = (kind == 1) ? off : off * ld_panel;
            pc.m = (int32_t)d->mb();
            pc.n = (int32_t)d->nb();
            pc.ld = (int32_t)d->stride();
            pc.reserved = 0;
            pieces.push_back(pc);
            off += (kind == 1) ? d->mb() : d->nb();
            node.host_state = MOSI::Shared;
            if (node.dev_state == MOSI::Modified) node.dev_state = MOSI::Shared;
        }
        Engine::CommSlab& slab = eng.acquire_slab(true, (size_t)ld_panel * panel_cols * es);
        slate_te_call(te_panel_pieces(eng.device(), eng.home_stream(), dtype, 0, slab.dev, ld_panel,
                                      pieces.data(), (int)pieces.size()));
        slate_te_call(te_memcpy2d_async(first->data(), es * first->stride(), slab.dev, es * ld_panel,
                                        es * ld_panel, panel_cols, TE_COPY_D2H, eng.home_stream()));
        eng.retire_slab(true, slab);
        host_copies_pending_ = true;
        t = e;
    }
}

//------------------------------------------------------------------------------
/// The tiles this rank is the root of, in plan order: consecutive ones are usually a block
/// column (A(:, k)) or a block row (B(k, :)) of a host-origin matrix.
template <typename scalar_t>
void BaseMatrix<scalar_t>::fetchRootRuns(std::vector<BcastOp> const& plan, int me)
{
    std::vector<std::pair<int64_t, int64_t>> run;
    for (auto const& op : plan)
        if (op.root == me) run.push_back({op.i, op.j});
    if (run.size() >= 2 && op_ == Op::NoTrans) tileGetRun(run, false, -1);
}

//------------------------------------------------------------------------------
template <typename scalar_t>
template <Target target>
void BaseMatrix<scalar_t>::listBcastMT(BcastListTag& bcast_list, Layout layout,
                                       int64_t life_factor, bool is_shared)
{
    Engine& eng = Engine::instance();
    Comm* comm = storage_->comm_;
    const int me = comm->rank;

    if (comm->size == 1) {
        // BaseMatrix.hh:2057-2097: with one rank the "broadcast" is the fetch onto the device
        if (target == Target::Devices && num_devices() > 0) {
            std::vector<std::pair<int64_t, int64_t>> run;
            for (auto& e : bcast_list) {
                int64_t i = std::get<0>(e), j = std::get<1>(e);
                if (tileIsLocal(i, j)) run.push_back({i, j});
            }
            if (op_ == Op::NoTrans) tileGetRun(run, false, -1);
            else for (auto& ij : run) tileGet(ij.first, ij.second, 0, false, -1);
            finishSet(0, -1);
        }
        return;
    }

    slate_assert(comm->nccl != nullptr);
    // Data plane (slate::initialize, SLATE_B200_BCAST): copy engines over CUDA IPC by default,
    // else NCCL send/recv -- one message per tile, or per (root, destination) pair.
    if (comm->peer_plane) {
        bcastPeer(bcastPlan(bcast_list));
        return;
    }
    if (comm->nccl_coalesce) {
        bcastCoalesced(bcastPlan(bcast_list));
        return;
    }
    void* stream = eng.comm_stream();
    auto plan = bcastPlan(bcast_list);
    fetchRootRuns(plan, me);
    // Roots first: fetching / converting a tile may itself put work and a mark on the comm
    // stream, and the receivers below are stamped with the mark that FOLLOWS the group.
    for (auto const& op : plan) {
        if (me != op.root) continue;
        tileGet(op.i, op.j, 0, false, -1);
        auto& node = storage_->at(globalIndex(op.i, op.j));
        // tiles travel in ColMajor (the receivers' workspace tiles are)
        if (node.dev->layout() != Layout::ColMajor)
            tileLayoutConvert(op.i, op.j, 0, Layout::ColMajor, -1);
        // (a host-to-device staging copy just enqueued is on the same comm stream, ahead
        //  of the sends: stream order is enough)
        eng.comm_wait_queue_seq(node.write_queue, node.write_seq);
        // device instances made by this library are contiguous (stride == mb)
        slate_assert(node.dev->isContiguous());
    }
    bool any = false;
    for (auto const& op : plan) {
        const int64_t i = op.i, j = op.j;
        if (me == op.root) {
            Tile<scalar_t>* t = storage_->at(globalIndex(i, j)).dev;
            if (!any) { slate_te_call(te_comm_group_start()); any = true; }
            for (int r : op.dests)
                slate_te_call(te_comm_send(comm->nccl, t->data(), t->bytes(), r, stream));
        }
        else if (std::find(op.dests.begin(), op.dests.end(), me) != op.dests.end()) {
            Tile<scalar_t>* t = prepareToReceive(i, j);
            if (!any) { slate_te_call(te_comm_group_start()); any = true; }
            slate_te_call(te_comm_recv(comm->nccl, t->data(), t->bytes(), op.root, stream));
        }
    }
    if (any) {
        eng.trace_begin(-1, "bcast", (int64_t)bcast_list.size());
        slate_te_call(te_comm_group_end());
        eng.trace_end(-1);
    }
    if (any || device_copies_pending_) {
        eng.mark_comm();
        device_copies_pending_ = false;
    }
}

//------------------------------------------------------------------------------
/// tilePrepareToReceive (MatrixStorage.hh:406-420): the workspace tile a broadcast fills, stamped
/// with the comm-stream mark that will follow the transfer.  A device instance that already
/// exists is reused; launches that may still read it were enqueued before this call, so the
/// comm stream is put behind all compute queues first (the drivers release a panel before it
/// is broadcast again, so this is the rare path), and a leftover RowMajor layout is reset.
template <typename scalar_t>
Tile<scalar_t>* BaseMatrix<scalar_t>::prepareToReceive(int64_t i, int64_t j)
{
    Engine& eng = Engine::instance();
    const ij_tuple gij = globalIndex(i, j);
    auto* node = storage_->find(gij);
    if (!node || !node->dev)
        storage_->tileInsert(gij, 0, TileKind::Workspace);
    else {
        eng.comm_wait_all_queues();
        if (node->dev->layout() != Layout::ColMajor && node->dev->kind() != TileKind::UserOwned)
            node->dev->setLayout(Layout::ColMajor, node->dev->mb());
    }
    auto& nd = storage_->at(gij);
    nd.receive_count += 1;
    nd.dev_state = MOSI::Shared;
    nd.write_queue = -1;
    nd.comm_seq = eng.next_comm_seq();
    return nd.dev;
}

//------------------------------------------------------------------------------
/// listBcastMT on the copy engines (tile_engine.h te_peer_*): per (root, destination) pair the
/// tiles of the list travel as messages of at most te_peer_max_message() bytes.  Message c of
/// every pair is handled in round c: this rank first announces a landing zone for each message
/// it expects, then pushes the messages it owes (peer copies + arrival flag), then makes its
/// comm stream wait for the arrivals and copy the tiles out of the landing zones into their
/// workspace blocks.  Pushes are enqueued BEFORE the waits of the same round, and every rank
/// walks the same lists in the same order, so no stream ever waits for a push that sits behind
/// it.  No kernel is launched on either side.
template <typename scalar_t>
void BaseMatrix<scalar_t>::bcastPeer(std::vector<BcastOp> const& plan)
{
    Engine& eng = Engine::instance();
    Comm* comm = storage_->comm_;
    const int me = comm->rank;
    void* stream = eng.comm_stream();
    std::vector<PairMsg> sends, recvs;
    bcastPairs(plan, me, &sends, &recvs);
    if (sends.empty() && recvs.empty()) {
        if (device_copies_pending_) { eng.mark_comm(); device_copies_pending_ = false; }
        return;
    }
    auto align256 = [](size_t b) { return (b + 255) & ~(size_t)255; };
    auto op_bytes = [&](size_t o) {
        return sizeof(scalar_t) * (size_t)tileMb(plan[o].i) * (size_t)tileNb(plan[o].j);
    };
    const size_t max_msg = te_peer_max_message(comm->nccl);

    // messages -> chunks of consecutive tiles, each tile at a 256-byte aligned offset; root and
    // destination cut the same list the same way
    struct Chunk { size_t first, last, bytes; };
    auto cut = [&](PairMsg const& m) {
        std::vector<Chunk> v;
        size_t cur = 0, first = 0;
        for (size_t t = 0; t < m.ops.size(); ++t) {
            const size_t b = align256(op_bytes(m.ops[t]));
            if (b > max_msg)
                slate_error("broadcast: one tile exceeds the peer arena's message limit; "
                            "raise SLATE_B200_ARENA_MB");
            if (cur + b > max_msg) {
                v.push_back({first, t, cur});
                first = t;
                cur = 0;
            }
            cur += b;
        }
        v.push_back({first, m.ops.size(), cur});
        return v;
    };
    std::vector<std::vector<Chunk>> send_chunks, recv_chunks;
    size_t rounds = 0;
    for (auto const& m : sends) { send_chunks.push_back(cut(m)); rounds = std::max(rounds, send_chunks.back().size()); }
    for (auto const& m : recvs) { recv_chunks.push_back(cut(m)); rounds = std::max(rounds, recv_chunks.back().size()); }

    // roots: tiles valid on the device, ColMajor, the comm stream behind their producers
    fetchRootRuns(plan, me);
    std::vector<char> prepared(plan.size(), 0);
    std::map<size_t, void*> packed_src;
    for (auto const& m : sends)
        for (size_t o : m.ops) {
            if (prepared[o]) continue;
            prepared[o] = 1;
            const int64_t i = plan[o].i, j = plan[o].j;
            tileGet(i, j, 0, false, -1);
            auto& node = storage_->at(globalIndex(i, j));
            if (node.dev->layout() != Layout::ColMajor) tileLayoutConvert(i, j, 0, Layout::ColMajor, -1);
            eng.comm_wait_queue_seq(node.write_queue, node.write_seq);
            if (!node.dev->isContiguous()) {
                // a user-owned device tile inside a larger array (Matrix::fromDevices): packed
                // into a pool block first, the peer copies are one-dimensional
                Tile<scalar_t>* t = node.dev;
                void* packed = storage_->allocScratchBlock();
                slate_te_call(te_memcpy2d_async(packed, sizeof(scalar_t) * t->mb(), t->data(),
                                                sizeof(scalar_t) * t->stride(), sizeof(scalar_t) * t->mb(),
                                                t->nb(), TE_COPY_D2D, stream));
                packed_src[o] = packed;
            }
        }
    // destinations: workspace tiles, stamped with the mark at the end of this call
    std::vector<Tile<scalar_t>*> dst_tile(plan.size(), nullptr);
    for (auto const& m : recvs)
        for (size_t o : m.ops) dst_tile[o] = prepareToReceive(plan[o].i, plan[o].j);

    eng.trace_begin(-1, "bcast_peer", (int64_t)plan.size());
    struct Landed { size_t r; uint64_t seq; char* zone; };
    std::vector<Landed> landed;
    for (size_t c = 0; c < rounds; ++c) {
        // 1. announce where this round's incoming messages go
        landed.clear();
        for (size_t r = 0; r < recvs.size(); ++r) {
            if (c >= recv_chunks[r].size()) continue;
            uint64_t seq = 0;
            void* zone = nullptr;
            slate_te_call(te_peer_post(comm->nccl, recvs[r].peer, recv_chunks[r][c].bytes, &seq, &zone));
            landed.push_back({r, seq, (char*)zone});
        }
        // 2. push what this rank owes: runs of tiles adjacent in memory go as one copy
        for (size_t s = 0; s < sends.size(); ++s) {
            if (c >= send_chunks[s].size()) continue;
            Chunk const& ch = send_chunks[s][c];
            uint64_t seq = 0, off = 0;
            slate_te_call(te_peer_fetch(comm->nccl, sends[s].peer, ch.bytes, &seq, &off));
            const char* run_src = nullptr;
            size_t run_bytes = 0, run_off = 0, o_off = 0;
            auto flush = [&]() {
                if (run_bytes)
                    slate_te_call(te_peer_put(comm->nccl, sends[s].peer, off + run_off, run_src, run_bytes, stream));
                run_bytes = 0;
            };
            for (size_t t = ch.first; t < ch.last; ++t) {
                const size_t o = sends[s].ops[t];
                Tile<scalar_t>* tile = storage_->at(globalIndex(plan[o].i, plan[o].j)).dev;
                const size_t b = tile->bytes();
                auto pk = packed_src.find(o);
                const char* src = pk == packed_src.end() ? (const char*)tile->data() : (const char*)pk->second;
                const bool joins = run_bytes && run_src + run_bytes == src && run_off + run_bytes == o_off;
                if (!joins) {
                    flush();
                    run_src = src;
                    run_off = o_off;
                }
                run_bytes += b;
                o_off += align256(b);
            }
            flush();
            slate_te_call(te_peer_signal(comm->nccl, sends[s].peer, seq, stream));
        }
        // 3. arrivals: wait, copy out, give the landing zone back
        for (auto const& l : landed) {
            Chunk const& ch = recv_chunks[l.r][c];
            slate_te_call(te_peer_wait(comm->nccl, recvs[l.r].peer, l.seq, stream));
            size_t o_off = 0;
            for (size_t t = ch.first; t < ch.last; ++t) {
                Tile<scalar_t>* tile = dst_tile[recvs[l.r].ops[t]];
                if (tile->isContiguous())
                    slate_te_call(te_memcpy_async(tile->data(), l.zone + o_off, tile->bytes(), TE_COPY_D2D, stream));
                else
                    slate_te_call(te_memcpy2d_async(tile->data(), sizeof(scalar_t) * tile->stride(), l.zone + o_off,
                                                    sizeof(scalar_t) * tile->mb(), sizeof(scalar_t) * tile->mb(),
                                                    tile->nb(), TE_COPY_D2D, stream));
                o_off += align256(tile->bytes());
            }
            slate_te_call(te_peer_consumed(comm->nccl, recvs[l.r].peer, l.seq, stream));
        }
    }
    eng.trace_end(-1);
    for (auto& kv : packed_src) storage_->freeDeferred(kv.second);
    if (!packed_src.empty()) storage_->flushErased();
    eng.mark_comm();
    device_copies_pending_ = false;
}

//------------------------------------------------------------------------------
/// listBcastMT with one NCCL message per (root, destination) pair (bcastPairs).  A root whose
/// tiles for a destination are adjacent in memory sends them in place; otherwise it packs them
/// into a staging slab (device-to-device copies on the comm stream).  A destination receives a
/// multi-tile message into the slab and copies the tiles out into their workspace blocks.
template <typename scalar_t>
void BaseMatrix<scalar_t>::bcastCoalesced(std::vector<BcastOp> const& plan)
{
    Engine& eng = Engine::instance();
    Comm* comm = storage_->comm_;
    const int me = comm->rank;
    void* stream = eng.comm_stream();
    std::vector<PairMsg> sends, recvs;
    bcastPairs(plan, me, &sends, &recvs);
    if (sends.empty() && recvs.empty()) {
        if (device_copies_pending_) { eng.mark_comm(); device_copies_pending_ = false; }
        return;
    }
    auto tile_of = [&](size_t o) -> Tile<scalar_t>* {
        return storage_->at(globalIndex(plan[o].i, plan[o].j)).dev;
    };
    auto align256 = [](size_t b) { return (b + 255) & ~(size_t)255; };

    // roots: tiles valid on the device, in ColMajor, behind their producers
    std::vector<char> prepared(plan.size(), 0);
    for (auto const& m : sends)
        for (size_t o : m.ops) {
            if (prepared[o]) continue;
            prepared[o] = 1;
            const int64_t i = plan[o].i, j = plan[o].j;
            tileGet(i, j, 0, false, -1);
            auto& node = storage_->at(globalIndex(i, j));
            if (node.dev->layout() != Layout::ColMajor) tileLayoutConvert(i, j, 0, Layout::ColMajor, -1);
            eng.comm_wait_queue_seq(node.write_queue, node.write_seq);
            slate_assert(node.dev->isContiguous());
        }
    // destinations: workspace tiles (tilePrepareToReceive)
    for (auto const& m : recvs)
        for (size_t o : m.ops) prepareToReceive(plan[o].i, plan[o].j);

    // which messages need staging
    auto adjacent = [&](PairMsg const& m) {
        for (size_t t = 0; t + 1 < m.ops.size(); ++t) {
            Tile<scalar_t>* a = tile_of(m.ops[t]);
            if ((char*)a->data() + a->bytes() != (char*)tile_of(m.ops[t + 1])->data()) return false;
        }
        return true;
    };
    std::vector<char> send_staged(sends.size(), 0);
    size_t need = 0;
    for (size_t s = 0; s < sends.size(); ++s)
        if (sends[s].ops.size() > 1 && !adjacent(sends[s])) {
            send_staged[s] = 1;
            need += align256(sends[s].bytes);
        }
    for (auto const& m : recvs)
        if (m.ops.size() > 1) need += align256(m.bytes);
    Engine::CommSlab* slab = need ? &eng.acquire_comm_slab(need) : nullptr;
    char* cursor = slab ? (char*)slab->dev : nullptr;

    // pack
    std::vector<void*> send_ptr(sends.size()), recv_ptr(recvs.size());
    for (size_t s = 0; s < sends.size(); ++s) {
        if (!send_staged[s]) {
            send_ptr[s] = tile_of(sends[s].ops[0])->data();
            continue;
        }
        send_ptr[s] = cursor;
        char* p = cursor;
        for (size_t o : sends[s].ops) {
            Tile<scalar_t>* t = tile_of(o);
            slate_te_call(te_memcpy_async(p, t->data(), t->bytes(), TE_COPY_D2D, stream));
            p += t->bytes();
        }
        cursor += align256(sends[s].bytes);
    }
    for (size_t r = 0; r < recvs.size(); ++r) {
        if (recvs[r].ops.size() == 1) {
            recv_ptr[r] = tile_of(recvs[r].ops[0])->data();
            continue;
        }
        recv_ptr[r] = cursor;
        cursor += align256(recvs[r].bytes);
    }

    // one NCCL group: one send / recv per peer
    eng.trace_begin(-1, "bcast_coalesced", (int64_t)plan.size());
    slate_te_call(te_comm_group_start());
    for (size_t s = 0; s < sends.size(); ++s)
        slate_te_call(te_comm_send(comm->nccl, send_ptr[s], sends[s].bytes, sends[s].peer, stream));
    for (size_t r = 0; r < recvs.size(); ++r)
        slate_te_call(te_comm_recv(comm->nccl, recv_ptr[r], recvs[r].bytes, recvs[r].peer, stream));
    slate_te_call(te_comm_group_end());

    // unpack
    for (size_t r = 0; r < recvs.size(); ++r) {
        if (recvs[r].ops.size() == 1) continue;
        const char* p = (const char*)recv_ptr[r];
        for (size_t o : recvs[r].ops) {
            Tile<scalar_t>* t = tile_of(o);
            slate_te_call(te_memcpy_async(t->data(), p, t->bytes(), TE_COPY_D2D, stream));
            p += t->bytes();
        }
    }
    eng.trace_end(-1);
    if (slab) eng.retire_comm_slab(*slab);
    eng.mark_comm();
    device_copies_pending_ = false;
}

//------------------------------------------------------------------------------
template <typename scalar_t>
void BaseMatrix<scalar_t>::tileLayoutConvert(int64_t i, int64_t j, int device, Layout layout, int queue)
{
    slate_assert(device != HostNum);
    Engine& eng = Engine::instance();
    auto& node = storage_->at(globalIndex(i, j));
    Tile<scalar_t>* t = node.dev;
    slate_assert(t != nullptr && node.dev_state != MOSI::Invalid);
    if (t->layout() == layout) return;
    if (layout == Layout::RowMajor) storage_->any_rowmajor_ = true;
    const int dtype = is_complex ? TE_C64 : TE_F64;
    // stored array now: rows x cols column-major
    const bool col = (t->layout() == Layout::ColMajor);
    const int64_t tm = t->op() == Op::NoTrans ? t->mb() : t->nb();     // extents of the tile proper
    const int64_t tn = t->op() == Op::NoTrans ? t->nb() : t->mb();
    const int64_t rows = col ? tm : tn, cols = col ? tn : tm;

    // order behind the last writer / the transfer that filled the tile
    void* stream;
    if (queue >= 0) {
        if (node.write_queue >= 0 && node.write_queue != queue)
            eng.queue_wait_queue_seq(queue, node.write_queue, node.write_seq);
        if (node.comm_seq) eng.queue_wait_comm_seq(queue, node.comm_seq);
        stream = eng.compute_stream(queue);
    }
    else {
        eng.comm_wait_queue_seq(node.write_queue, node.write_seq);
        stream = eng.comm_stream();
    }

    if (rows == cols && t->isContiguous()) {
        slate_te_call(te_tile_transpose(eng.device(), stream, dtype, 0, (int)rows, (int)cols, t->data(),
                                        t->stride(), t->data(), t->stride()));
        t->setLayout(layout, t->stride());
    }
    else {
        if (t->kind() == TileKind::UserOwned)
            slate_error("tileLayoutConvert: a rectangular or strided user-owned device tile cannot "
                        "be converted in place");
        scalar_t* fresh = (scalar_t*)storage_->allocScratchBlock();
        slate_te_call(te_tile_transpose(eng.device(), stream, dtype, 0, (int)rows, (int)cols, t->data(),
                                        t->stride(), fresh, cols));
        storage_->freeDeferred(t->data());
        storage_->flushErased();
        t->setLayout(layout, cols, fresh);
    }
    if (queue >= 0) {
        node.write_queue = queue;
        node.write_seq = eng.next_queue_seq(queue);
        eng.mark_queue(queue);
    }
    else {
        node.write_queue = -1;
        node.comm_seq = eng.next_comm_seq();
        eng.mark_comm();
    }
}

//------------------------------------------------------------------------------
template <typename scalar_t>
void BaseMatrix<scalar_t>::listReduce(ReduceList& reduce_list, Layout layout, int tag)
{
    Engine& eng = Engine::instance();
    Comm* comm = storage_->comm_;
    if (comm->size == 1) return;
    const int me = comm->rank;
    slate_assert(comm->nccl != nullptr);
    void* stream = eng.comm_stream();
    const int dtype = is_complex ? TE_C64 : TE_F64;

    struct Incoming { ij_tuple gij; std::vector<void*> parts; };
    std::vector<Incoming> incoming;
    std::vector<ij_tuple> sent;
    bool any = false;
    for (auto const& op : reducePlan(reduce_list)) {
        const ij_tuple gij = globalIndex(op.i, op.j);
        if (me == op.root) {
            // the root's own instance: on the device, about to be modified
            tileGet(op.i, op.j, 0, true, -1);
            auto& node = storage_->at(gij);
            eng.comm_wait_queue_seq(node.write_queue, node.write_seq);
            Incoming in;
            in.gij = gij;
            if (!any) { slate_te_call(te_comm_group_start()); any = true; }
            for (int r : op.sources) {
                void* part = storage_->allocScratchBlock();
                slate_te_call(te_comm_recv(comm->nccl, part, node.dev->bytes(), r, stream));
                in.parts.push_back(part);
            }
            incoming.push_back(std::move(in));
        }
        else if (std::find(op.sources.begin(), op.sources.end(), me) != op.sources.end()) {
            auto* node = storage_->find(gij);
            slate_assert(node && node->dev && node->dev_state != MOSI::Invalid);
            slate_assert(node->dev->isContiguous());
            eng.comm_wait_queue_seq(node->write_queue, node->write_seq);
            if (!any) { slate_te_call(te_comm_group_start()); any = true; }
            slate_te_call(te_comm_send(comm->nccl, node->dev->data(), node->dev->bytes(), op.root,
                                       stream));
            sent.push_back(gij);
        }
    }
    if (!any) return;
    eng.trace_begin(-1, "reduce", (int64_t)reduce_list.size());
    slate_te_call(te_comm_group_end());
    for (auto& in : incoming) {
        auto& node = storage_->at(in.gij);
        Tile<scalar_t>* t = node.dev;
        for (void* part : in.parts) {
            slate_te_call(te_tile_add(eng.device(), stream, dtype, (int)t->mb(), (int)t->nb(), 1.0, part,
                                      t->mb(), 1.0, t->data(), t->stride()));
            storage_->freeDeferred(part);
        }
        // consumers order themselves behind the comm stream, as after a broadcast
        node.dev_state = MOSI::Modified;
        if (node.host) node.host_state = MOSI::Invalid;
        node.write_queue = -1;
        node.comm_seq = eng.next_comm_seq();
    }
    eng.trace_end(-1);
    // a partial sum that has been sent is spent (BaseMatrix.hh:2243-2249)
    for (auto const& gij : sent) {
        storage_->tileErase(gij, 0);
        storage_->tileErase(gij, HostNum);
    }
    storage_->flushErased();
    eng.mark_comm();
}

}  // namespace slate

#endif
