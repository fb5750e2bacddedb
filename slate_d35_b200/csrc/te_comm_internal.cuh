// te_comm_internal.cuh -- the communicator object shared by te_comm.cu (NCCL) and
// te_peer.cu (copy-engine transport over CUDA IPC).
#pragma once
#include "te_common.cuh"
#include <nccl.h>

namespace te { struct PeerState; }

struct te_comm {
    ncclComm_t nccl;
    int rank, size, device;
    uint64_t job_tag;          // hash of the NCCL unique id: the same on every rank of the job
    te::PeerState* peer;       // te_peer_open / te_peer_close
};

#define TE_NCCL_CHECK(expr)                                                  \
    do {                                                                     \
        ncclResult_t r_ = (expr);                                            \
        if (r_ != ncclSuccess) {                                             \
            te::set_last_error(#expr, ncclGetErrorString(r_));               \
            return TE_ERR_NCCL;                                              \
        }                                                                    \
    } while (0)
