// slate/MatrixStorage.hh -- tile map of one distributed matrix (reference:
// include/slate/internal/MatrixStorage.hh:35-143 TileNode, :146-1156 MatrixStorage).
//
// One process drives one GPU, so a tile has at most two instances: on the host and on
// the device.  MOSI states are kept per instance as in the reference
// (include/slate/enums.hh:138-144).  Device instances also remember which compute queue
// last wrote them, so readers on another queue (or the comm stream) can wait for exactly
// that queue instead of synchronising the host.
#ifndef SLATE_B200_MATRIX_STORAGE_HH
#define SLATE_B200_MATRIX_STORAGE_HH

#include "slate/Tile.hh"
#include "slate/engine.hh"

#include <algorithm>
#include <functional>
#include <map>
#include <memory>
#include <tuple>

namespace slate {

using ij_tuple = std::tuple<int64_t, int64_t>;

namespace func {
/// include/slate/func.hh:100-118, 178-186: 2D block-cyclic owner of tile (i, j).
inline std::function<int(ij_tuple)> process_2d_grid(GridOrder order, int p, int q)
{
    if (order == GridOrder::Col)
        return [p, q](ij_tuple ij) { return int(std::get<0>(ij) % p + (std::get<1>(ij) % q) * p); };
    return [p, q](ij_tuple ij) { return int((std::get<0>(ij) % p) * q + std::get<1>(ij) % q); };
}
/// include/slate/func.hh:57-77: uniform tile size with a short last tile.
inline std::function<int64_t(int64_t)> uniform_blocksize(int64_t n, int64_t nb)
{
    return [n, nb](int64_t j) { return (j + 1) * nb > n ? n % nb : nb; };
}
}  // namespace func

template <typename scalar_t>
struct TileNode {
    Tile<scalar_t>* host = nullptr;
    Tile<scalar_t>* dev = nullptr;
    short host_state = MOSI::Invalid;
    short dev_state = MOSI::Invalid;
    int64_t receive_count = 0;     // MatrixStorage.hh:406-420 (tilePrepareToReceive)
    int write_queue = -1;          // compute queue that last wrote the device instance
    uint64_t write_seq = 0;        //   and the mark of that queue which follows the write
    uint64_t comm_seq = 0;         // comm-stream mark after which the device instance is filled
    bool host_owned_buffer = false;
};

template <typename scalar_t>
class MatrixStorage {
public:
    MatrixStorage(int64_t m, int64_t n, int64_t mb, int64_t nb, GridOrder order, int p, int q,
                  Comm* comm)
        : m_(m), n_(n), mb_(mb), nb_(nb), order_(order), p_(p), q_(q), comm_(comm),
          memory_(sizeof(scalar_t) * (size_t)mb * (size_t)nb)
    {
        slate_assert(mb > 0 && nb > 0 && p > 0 && q > 0);
        slate_assert(comm != nullptr);
        slate_assert(p * q == comm->size || comm->size == 1 || p * q <= comm->size);
        mt_ = (m + mb - 1) / mb;
        nt_ = (n + nb - 1) / nb;
        tileMb = func::uniform_blocksize(m, mb);
        tileNb = func::uniform_blocksize(n, nb);
        tileRank = func::process_2d_grid(order, p, q);
        tileDevice = [](ij_tuple) { return 0; };
    }

    /// include/slate/internal/MatrixStorage.hh (lambda constructor, used by Matrix.hh:33-38): tile
    /// extents, owners and devices given by the caller's functions; tile sizes need not be
    /// uniform.  The pool's block holds the largest tile.  One GPU per process: inTileDevice is
    /// kept for the interface and must map local tiles to device 0.
    MatrixStorage(int64_t m, int64_t n, std::function<int64_t(int64_t)>& inTileMb,
                  std::function<int64_t(int64_t)>& inTileNb, std::function<int(ij_tuple)>& inTileRank,
                  std::function<int(ij_tuple)>& inTileDevice, Comm* comm)
        : m_(m), n_(n), mb_(largest_tile(m, inTileMb)), nb_(largest_tile(n, inTileNb)),
          mt_(count_tiles(m, inTileMb)), nt_(count_tiles(n, inTileNb)), order_(GridOrder::Unknown),
          p_(-1), q_(-1), comm_(comm), tileMb(inTileMb), tileNb(inTileNb), tileRank(inTileRank),
          tileDevice(inTileDevice),
          memory_(sizeof(scalar_t) * (size_t)std::max<int64_t>(mb_, 1) * (size_t)std::max<int64_t>(nb_, 1))
    {
        slate_assert(comm != nullptr);
    }

    ~MatrixStorage()
    {
        try { clear(); } catch (...) {}
    }

    static int64_t count_tiles(int64_t extent, std::function<int64_t(int64_t)>& size_of)
    {
        int64_t t = 0;
        for (int64_t ii = 0; ii < extent; ++t) {
            const int64_t s = size_of(t);
            slate_assert(s > 0);
            ii += s;
            slate_assert(ii <= extent);        // the tile sizes must add up to the extent
        }
        return t;
    }
    static int64_t largest_tile(int64_t extent, std::function<int64_t(int64_t)>& size_of)
    {
        int64_t t = 0, big = 0;
        for (int64_t ii = 0; ii < extent; ++t) {
            const int64_t s = size_of(t);
            slate_assert(s > 0);
            big = std::max(big, s);
            ii += s;
        }
        return big;
    }

    // ---- lookups ----
    TileNode<scalar_t>* find(ij_tuple ij)
    {
        auto it = tiles_.find(ij);
        return it == tiles_.end() ? nullptr : &it->second;
    }
    TileNode<scalar_t>& at(ij_tuple ij)
    {
        auto it = tiles_.find(ij);
        if (it == tiles_.end())
            slate_error("tile (" + std::to_string(std::get<0>(ij)) + ", "
                        + std::to_string(std::get<1>(ij)) + ") does not exist");
        return it->second;
    }
    bool tileIsLocal(ij_tuple ij) const { return tileRank(ij) == comm_->rank; }

    // ---- insertion / removal ----
    /// SLATE-owned or workspace tile with memory from the pool (device) or new[] (host);
    /// MatrixStorage.hh:746-790.  stride = tileMb.
    Tile<scalar_t>* tileInsert(ij_tuple ij, int device, TileKind kind)
    {
        int64_t i = std::get<0>(ij), j = std::get<1>(ij);
        int64_t mb = tileMb(i), nb = tileNb(j);
        TileNode<scalar_t>& node = tiles_[ij];
        Tile<scalar_t>*& slot = (device == HostNum) ? node.host : node.dev;
        slate_assert(slot == nullptr);
        scalar_t* data = (scalar_t*)memory_.alloc(
            device, sizeof(scalar_t) * (size_t)mb * (size_t)nb);
        slot = new Tile<scalar_t>(mb, nb, data, mb, device, kind);
        if (device == HostNum) node.host_owned_buffer = true;
        return slot;
    }

    /// User-owned tile over caller memory (fromLAPACK / fromScaLAPACK / fromDevices).
    Tile<scalar_t>* tileInsert(ij_tuple ij, int device, scalar_t* data, int64_t ld)
    {
        int64_t i = std::get<0>(ij), j = std::get<1>(ij);
        TileNode<scalar_t>& node = tiles_[ij];
        Tile<scalar_t>*& slot = (device == HostNum) ? node.host : node.dev;
        slate_assert(slot == nullptr);
        slot = new Tile<scalar_t>(tileMb(i), tileNb(j), data, ld, device, TileKind::UserOwned);
        if (device == HostNum) node.host_state = MOSI::Shared;
        else node.dev_state = MOSI::Shared;
        return slot;
    }

    /// Drops one instance.  A device block still referenced by enqueued kernels goes back
    /// to the pool only when they have finished (Engine::defer).
    void tileErase(ij_tuple ij, int device)
    {
        auto it = tiles_.find(ij);
        if (it == tiles_.end()) return;
        TileNode<scalar_t>& node = it->second;
        if (device == HostNum) {
            if (node.host) {
                if (node.host->kind() != TileKind::UserOwned)
                    memory_.free(node.host->data(), HostNum);
                delete node.host;
                node.host = nullptr;
                node.host_state = MOSI::Invalid;
            }
        }
        else if (node.dev) {
            if (node.dev->kind() != TileKind::UserOwned)
                pending_free_.push_back(node.dev->data());
            delete node.dev;
            node.dev = nullptr;
            node.dev_state = MOSI::Invalid;
            node.write_queue = -1;
        }
        if (!node.host && !node.dev)
            tiles_.erase(it);
    }

    /// A pool block used as scratch by enqueued work (listReduce receive buffers): returned
    /// with the next flushErased().
    /// set once a device tile of this matrix has been converted to RowMajor: until then the
    /// layout checks of tileGetForReading / Writing are skipped (they cost a map lookup per tile)
    bool any_rowmajor_ = false;

    void* allocScratchBlock() { return memory_.alloc(0, memory_.block_size()); }
    void freeDeferred(void* block) { pending_free_.push_back(block); }

    /// Hands the device blocks dropped since the last call to the engine for deferred
    /// return to the pool.
    void flushErased()
    {
        if (pending_free_.empty()) return;
        std::vector<void*> blocks;
        blocks.swap(pending_free_);
        Memory* mem = &memory_;
        std::weak_ptr<int> alive = alive_;
        Engine::instance().defer(~0u, true, [mem, blocks, alive]() {
            if (alive.expired()) return;      // storage already destroyed (pool cleared)
            for (void* b : blocks) mem->free(b, 0);
        });
    }

    void clear()
    {
        // the pool is about to go away: nothing may still be running on its blocks
        if (!tiles_.empty() && Engine::instance().num_devices() > 0)
            Engine::instance().sync_all();
        for (auto& kv : tiles_) {
            TileNode<scalar_t>& node = kv.second;
            if (node.host) {
                if (node.host->kind() != TileKind::UserOwned) memory_.free(node.host->data(), HostNum);
                delete node.host;
            }
            if (node.dev) delete node.dev;     // blocks go with the pool
        }
        tiles_.clear();
        pending_free_.clear();
        alive_.reset();
    }

    int64_t m_, n_, mb_, nb_, mt_, nt_;
    GridOrder order_;
    int p_, q_;
    Comm* comm_;
    std::function<int64_t(int64_t)> tileMb, tileNb;
    std::function<int(ij_tuple)> tileRank, tileDevice;
    std::map<ij_tuple, TileNode<scalar_t>> tiles_;
    Memory memory_;
    std::vector<void*> pending_free_;
    std::shared_ptr<int> alive_ = std::make_shared<int>(1);
};

}  // namespace slate

#endif
