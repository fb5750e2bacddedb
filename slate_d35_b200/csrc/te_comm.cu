// te_comm.cu -- panel tile broadcasts over NCCL (NVLink 5 / NVSwitch).
//
// Replaces the MPI point-to-point hypercube of the reference
//   BaseMatrix::tileBcastToSet / tileIbcastToSet (include/slate/BaseMatrix.hh:2283-2387),
//   internal::cubeBcastPattern (src/internal/internal_comm.cc:72-115),
//   Tile::isend / irecv / bcast (include/slate/Tile.hh:1084-1267)
// and MPI_Allreduce(MIN) of info (src/internal/internal_reduce_info.cc:31).
// One rank per GPU.  A broadcast of a set of tiles is issued as ONE NCCL group of
// send/recv pairs on the caller's stream: the root posts a send per destination rank,
// each destination a recv into its workspace tile.  NVSwitch gives every pair full
// bandwidth, so a flat fan-out needs no tree; device buffers only (no host staging).
#include "te_comm_internal.cuh"

extern "C" int te_comm_unique_id(void* id_out, size_t bytes)
{
    if (!id_out || bytes < sizeof(ncclUniqueId)) return TE_ERR_ARG;
    ncclUniqueId id;
    TE_NCCL_CHECK(ncclGetUniqueId(&id));
    memcpy(id_out, &id, sizeof(id));
    return TE_OK;
}

extern "C" int te_comm_init_rank(const void* id, size_t bytes, int nranks, int rank, int device,
                                 te_comm** comm)
{
    if (!id || !comm || bytes < sizeof(ncclUniqueId) || nranks < 1 || rank < 0 || rank >= nranks)
        return TE_ERR_ARG;
    int st = te::set_device(device);
    if (st != TE_OK) return st;
    ncclUniqueId uid;
    memcpy(&uid, id, sizeof(uid));
    te_comm* c = new te_comm;
    c->rank = rank;
    c->size = nranks;
    c->device = device;
    c->peer = nullptr;
    // every rank derives the same job tag from the id (names the shared-memory mailbox of
    // the copy-engine transport, te_peer.cu)
    uint64_t h = 1469598103934665603ull;
    for (size_t b = 0; b < sizeof(uid); ++b) h = (h ^ (unsigned char)uid.internal[b]) * 1099511628211ull;
    c->job_tag = h;
    ncclResult_t r = ncclCommInitRank(&c->nccl, nranks, uid, rank);
    if (r != ncclSuccess) {
        te::set_last_error("ncclCommInitRank", ncclGetErrorString(r));
        delete c;
        return TE_ERR_NCCL;
    }
    *comm = c;
    return TE_OK;
}

// NCCL sets up each peer-to-peer connection lazily, at the first send/recv between the
// pair (about a second for a new pair on this box).  Exchange one small message between
// every pair now so that no broadcast inside a factorisation pays for it.
extern "C" int te_comm_connect_all(te_comm* c, void* stream)
{
    if (!c) return TE_ERR_ARG;
    if (c->size == 1) return TE_OK;
    int st = te::set_device(c->device);
    if (st != TE_OK) return st;
    char* buf = nullptr;
    TE_CUDA_CHECK(cudaMalloc(&buf, 2 * 256 * (size_t)c->size));
    TE_CUDA_CHECK(cudaMemsetAsync(buf, 0, 2 * 256 * (size_t)c->size, (cudaStream_t)stream));
    TE_NCCL_CHECK(ncclGroupStart());
    for (int r = 0; r < c->size; ++r) {
        if (r == c->rank) continue;
        TE_NCCL_CHECK(ncclSend(buf + 256 * r, 256, ncclChar, r, c->nccl, (cudaStream_t)stream));
        TE_NCCL_CHECK(ncclRecv(buf + 256 * (c->size + r), 256, ncclChar, r, c->nccl,
                               (cudaStream_t)stream));
    }
    TE_NCCL_CHECK(ncclGroupEnd());
    TE_CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream));
    TE_CUDA_CHECK(cudaFree(buf));
    return TE_OK;
}

extern "C" int te_comm_destroy(te_comm* c)
{
    if (!c) return TE_OK;
    if (c->peer) te_peer_close(c);
    ncclCommDestroy(c->nccl);
    delete c;
    return TE_OK;
}

extern "C" int te_comm_rank(const te_comm* c) { return c ? c->rank : 0; }
extern "C" int te_comm_size(const te_comm* c) { return c ? c->size : 1; }

extern "C" int te_comm_group_start(void)
{
    TE_NCCL_CHECK(ncclGroupStart());
    return TE_OK;
}

extern "C" int te_comm_group_end(void)
{
    TE_NCCL_CHECK(ncclGroupEnd());
    return TE_OK;
}

extern "C" int te_comm_send(te_comm* c, const void* buf, size_t bytes, int peer, void* stream)
{
    if (!c || !buf || peer < 0 || peer >= c->size) return TE_ERR_ARG;
    TE_NCCL_CHECK(ncclSend(buf, bytes, ncclChar, peer, c->nccl, (cudaStream_t)stream));
    return TE_OK;
}

extern "C" int te_comm_recv(te_comm* c, void* buf, size_t bytes, int peer, void* stream)
{
    if (!c || !buf || peer < 0 || peer >= c->size) return TE_ERR_ARG;
    TE_NCCL_CHECK(ncclRecv(buf, bytes, ncclChar, peer, c->nccl, (cudaStream_t)stream));
    return TE_OK;
}

extern "C" int te_comm_bcast(te_comm* c, void* buf, size_t bytes, int root, void* stream)
{
    if (!c || !buf || root < 0 || root >= c->size) return TE_ERR_ARG;
    TE_NCCL_CHECK(ncclBroadcast(buf, buf, bytes, ncclChar, root, c->nccl, (cudaStream_t)stream));
    return TE_OK;
}

extern "C" int te_comm_allreduce_i64(te_comm* c, int64_t* buf_dev, size_t count, int op, void* stream)
{
    if (!c || !buf_dev) return TE_ERR_ARG;
    ncclRedOp_t rop = (op == TE_REDUCE_MIN) ? ncclMin : (op == TE_REDUCE_MAX) ? ncclMax : ncclSum;
    TE_NCCL_CHECK(ncclAllReduce(buf_dev, buf_dev, count, ncclInt64, rop, c->nccl,
                                (cudaStream_t)stream));
    return TE_OK;
}

extern "C" int te_comm_allreduce_f64(te_comm* c, double* buf_dev, size_t count, int op, void* stream)
{
    if (!c || !buf_dev) return TE_ERR_ARG;
    ncclRedOp_t rop = (op == TE_REDUCE_MIN) ? ncclMin : (op == TE_REDUCE_MAX) ? ncclMax : ncclSum;
    TE_NCCL_CHECK(ncclAllReduce(buf_dev, buf_dev, count, ncclDouble, rop, c->nccl,
                                (cudaStream_t)stream));
    return TE_OK;
}
