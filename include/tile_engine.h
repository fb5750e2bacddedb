/*
 * tile_engine.h -- C ABI of the B200 (sm_100a) Level-3 tile-update engine.
 *
 * This is the drop-in boundary between SLATE's host C++ (the Target::Devices
 * overloads of internal::gemm / herk / syrk / trsm / potrf, the device-tile
 * Memory pool and the panel tile broadcasts) and hand-written CUDA.  Every
 * entry point takes plain pointers and sizes, returns an int status
 * (TE_OK == 0) and never throws.  Paths cited below are relative to the
 * reference tree (SLATE 2023.11.05).
 *
 * Conventions
 *   - all tile data is column-major, FP64 (TE_F64) or complex-FP64 (TE_C64,
 *     interleaved re/im), leading dimensions in elements;
 *   - `stream` is a cudaStream_t passed as void* (NULL = the legacy default
 *     stream); all work is enqueued on it and the call returns without
 *     synchronising unless stated otherwise;
 *   - `device` is a CUDA device ordinal; the call makes it current;
 *   - descriptor arrays live in DEVICE memory unless the name ends in _host.
 */
#ifndef TILE_ENGINE_H
#define TILE_ENGINE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- status codes ------------------------------------------------------ */
enum {
    TE_OK              = 0,
    TE_ERR_ARG         = 1,   /* bad argument (op, size, null pointer)          */
    TE_ERR_CUDA        = 2,   /* a CUDA runtime call failed: te_last_error()    */
    TE_ERR_NOMEM       = 3,   /* pool exhausted / cudaMalloc failed             */
    TE_ERR_UNSUPPORTED = 4,   /* legal in BLAS but not built (e.g. ConjNoTrans) */
    TE_ERR_NCCL        = 5,   /* an NCCL call failed: te_last_error()           */
    TE_ERR_LIMIT       = 6    /* size beyond a documented fixed limit           */
};

/* ---- enums (values match blas::Op / Uplo / Side / Diag chars) ----------- */
enum { TE_F64 = 0, TE_C64 = 1 };
enum { TE_OP_N = 'N', TE_OP_T = 'T', TE_OP_C = 'C' };
enum { TE_UPLO_LOWER = 'L', TE_UPLO_UPPER = 'U', TE_UPLO_GENERAL = 'G' };
enum { TE_SIDE_LEFT = 'L', TE_SIDE_RIGHT = 'R' };
enum { TE_DIAG_NONUNIT = 'N', TE_DIAG_UNIT = 'U' };

/* Thread-local text of the last CUDA / NCCL failure seen by this thread. */
const char* te_last_error(void);

/* Library / build identification: "tile_engine <ver> sm_100a". */
const char* te_version(void);

/* Number of CUDA devices (0 when none; never an error). */
int te_device_count(void);

/* Kernels this library has launched so far in this process (all devices). */
long long te_launch_count(void);

/* ------------------------------------------------------------------------
 * Streams, events, copies.  Stand in for blas::Queue and blas::device_memcpy /
 * device_malloc (blaspp/include/blas/device.hh:445-726,
 * blaspp/src/device_queue.cc), i.e. MatrixStorage::compute_queue / comm_queue
 * (include/slate/internal/MatrixStorage.hh:198-245, 574-626).
 * ---------------------------------------------------------------------- */
enum { TE_COPY_H2D = 1, TE_COPY_D2H = 2, TE_COPY_D2D = 3, TE_COPY_DEFAULT = 4 };

int te_stream_create(int device, int high_priority, void** stream);
int te_stream_destroy(int device, void* stream);
int te_stream_sync(int device, void* stream);
int te_device_sync(int device);
int te_event_create(int device, int timing, void** event);
int te_event_destroy(void* event);
int te_event_record(void* event, void* stream);
int te_stream_wait_event(void* stream, void* event);
int te_event_sync(void* event);
int te_event_query_done(void* event);   /* 1 finished, 0 not yet, -1 error */
int te_event_elapsed_ms(void* start, void* stop, float* ms);

int te_device_malloc(int device, size_t bytes, void** ptr);
int te_device_free(int device, void* ptr);
int te_device_mem_info(int device, size_t* free_bytes, size_t* total_bytes);
int te_host_alloc_pinned(size_t bytes, void** ptr);
int te_host_free_pinned(void* ptr);
int te_host_register(void* ptr, size_t bytes);
int te_host_unregister(void* ptr);
int te_memcpy_async(void* dst, const void* src, size_t bytes, int kind, void* stream);
/* column-major tile copy: `height` columns of `width_bytes`, pitches in bytes
 * (Tile::copyData, include/slate/Tile.hh:971-1042) */
int te_memcpy2d_async(void* dst, size_t dpitch, const void* src, size_t spitch,
                      size_t width_bytes, size_t height, int kind, void* stream);
int te_memset_async(void* dst, int value, size_t bytes, void* stream);
/* Copy of a small PINNED host buffer (te_host_alloc_pinned; both pointers 16-byte aligned,
 * sizes rounded up to 16 bytes) to the device by a kernel on `stream`, so that it does not
 * queue on the copy engine behind bulk transfers of other streams.  For descriptor arrays. */
int te_upload_small(int device, void* stream, void* dst_dev, const void* src_pinned, size_t bytes);

/* ------------------------------------------------------------------------
 * Device-tile memory pool: fixed-size blocks (sizeof(T)*mb*nb, rounded up to
 * 256 bytes) on one device.  Replaces slate::Memory
 * (include/slate/internal/Memory.hh:29-101, src/core/Memory.cc:17-218):
 *   te_pool_reserve  <-> Memory::addDeviceBlocks  (ONE cudaMalloc slab per call)
 *   te_pool_alloc    <-> Memory::alloc  (pops; grows by one block when empty)
 *   te_pool_free     <-> Memory::free
 *   te_pool_stats    <-> Memory::available / capacity (allocated = cap - avail)
 *   te_pool_clear    <-> Memory::clearDeviceBlocks
 * Thread-safe (the reference uses omp critical(slate_memory)).
 * ---------------------------------------------------------------------- */
typedef struct te_pool te_pool;
int te_pool_create(int device, size_t block_bytes, te_pool** pool);
int te_pool_reserve(te_pool* pool, int64_t nblocks);
int te_pool_alloc(te_pool* pool, void** block);
int te_pool_free(te_pool* pool, void* block);
int te_pool_stats(te_pool* pool, size_t* available, size_t* capacity, size_t* block_bytes);
int te_pool_clear(te_pool* pool);
int te_pool_destroy(te_pool* pool);

/* ------------------------------------------------------------------------
 * Batched tile update:  C_i = alpha * op(A_i) * op(B_i) + beta * C_i
 *
 * One launch covers a whole block outer product on one device, ragged edge
 * tiles included (m, n, k are per tile), so there is no grouping by size.
 * Replaces blas::batch::gemm -> cublas{D,Z}gemmBatched as called from
 *   src/internal/internal_gemm.cc:478-509   (internal::gemm, Devices)
 *   src/internal/internal_herk.cc:490-517   (off-diagonal tiles of herk)
 *   src/internal/internal_syrk.cc:480-510   (off-diagonal tiles of syrk)
 * and, through `uplo`, the per-tile cublas{D,Z}syrk / Zherk loop of
 *   blaspp/src/device_batch_herk.cc:57-74 / device_batch_syrk.cc
 * for diagonal tiles (only the named triangle of C_i is read or written;
 * for TE_TILE_HERMITIAN the imaginary part of the diagonal is set to 0).
 * ---------------------------------------------------------------------- */
enum {
    TE_TILE_GENERAL   = 0,   /* full m x n tile                               */
    TE_TILE_LOWER     = 1,   /* diagonal tile: update rows >= cols only       */
    TE_TILE_UPPER     = 2,   /* diagonal tile: update rows <= cols only       */
    TE_TILE_HERMITIAN = 4    /* or-ed with LOWER/UPPER: herk diagonal rule    */
};

typedef struct te_tile_desc {
    const void* a;           /* op(A_i) is m x k; stored m x k (N) or k x m   */
    const void* b;           /* op(B_i) is k x n; stored k x n (N) or n x k   */
    void*       c;           /* m x n                                         */
    int32_t     m, n, k;
    int32_t     lda, ldb, ldc;
    int32_t     flags;       /* TE_TILE_*                                     */
    int32_t     reserved;
} te_tile_desc;              /* 56 bytes                                      */

/*
 * desc:   DEVICE array of `count` descriptors (upload it on `stream` first).
 * max_m/max_n: upper bounds of m and n over the batch (sizes the work grid).
 * alpha/beta: HOST pointers to one double (TE_F64) or two (TE_C64).
 * beta == 0 means C is not read (NaN-safe, as in BLAS).  As in BLAS, C_i must not
 * overlap A_i or B_i.
 * opa/opb: TE_OP_N / T / C.  For TE_F64, T and C are the same.
 */
int te_gemm_batched(int device, void* stream, int dtype, int opa, int opb,
                    const void* alpha, const void* beta,
                    const te_tile_desc* desc, int64_t count,
                    int32_t max_m, int32_t max_n);

/* Same, with `count` (<= TE_MAX_INLINE_DESC) descriptors passed from HOST
 * memory inside the kernel parameters: no upload, for the few-tile calls on
 * the panel's critical path (single-tile herk fast path,
 * src/internal/internal_herk.cc:365-389). */
#define TE_MAX_INLINE_DESC 8
int te_gemm_inline(int device, void* stream, int dtype, int opa, int opb,
                   const void* alpha, const void* beta,
                   const te_tile_desc* desc_host, int count);

/* Scheduling hint, no reference counterpart: upper bound on the number of 128x128 output
 * blocks one CTA of te_gemm_batched works through before it exits (default 16; 0 restores
 * the default; the TE_ITEMS_PER_CTA environment variable overrides both).  Smaller values
 * free SMs sooner for kernels waiting on higher-priority streams.  Process-wide; takes
 * effect at the next launch. */
int te_gemm_set_items_per_cta(int cap);

/* ------------------------------------------------------------------------
 * Cholesky of one tile, in place:  A = L L^H (lower) or U^H U (upper).
 *
 * Replaces lapack::potrf(uplo, n, A, lda, info, queue) -> cusolverDnXpotrf
 *   src/internal/internal_potrf.cc:57-81, lapackpp/src/cuda/cuda_potrf.cc:128-171.
 * Only the `uplo` triangle of A is read or written.  *info_dev (DEVICE int) is
 * set on the stream to 0, or to the 1-based index of the first pivot that is
 * not positive; there is no host synchronisation (the reference copies info
 * back and syncs every step: internal_potrf.cc:75-78).
 * scratch: DEVICE memory of at least te_potrf_scratch_bytes(dtype) bytes,
 * private to this call until the stream has passed it (inverse of the current
 * diagonal block).
 * ---------------------------------------------------------------------- */
size_t te_potrf_scratch_bytes(int dtype);
int te_potrf_tile(int device, void* stream, int dtype, int uplo, int n, void* A, int lda,
                  int* info_dev, void* scratch, size_t scratch_bytes);

/* ------------------------------------------------------------------------
 * LU of one tile without pivoting, in place: A (m x n) = L U, L unit lower (m x min(m,n)),
 * U upper (min(m,n) x n).
 * Replaces: internal::getrf_nopiv (src/internal/internal_getrf_nopiv.cc, the tile routine of
 * slate::getrf_nopiv, src/getrf_nopiv.cc:86-95), which the reference runs on the host.
 * *info_dev = 0, or j if U(j,j) is exactly zero (first such j; the factorisation goes on
 * without scaling that column).  scratch: te_getrf_nopiv_scratch_bytes(dtype) bytes.
 * ---------------------------------------------------------------------- */
size_t te_getrf_nopiv_scratch_bytes(int dtype);
int te_getrf_nopiv_tile(int device, void* stream, int dtype, int m, int n, void* A, int lda,
                        int* info_dev, void* scratch, size_t scratch_bytes);

/* ------------------------------------------------------------------------
 * Batched triangular solve with ONE triangular tile A shared by the batch:
 *   side R:  B_i = alpha B_i op(A)^-1   (B_i is m_i x na)
 *   side L:  B_i = alpha op(A)^-1 B_i   (B_i is na x n_i)
 *
 * Replaces blas::batch::trsm -> cublas{D,Z}trsmBatched with the A pointer
 * replicated (src/internal/internal_trsm.cc:132-261,
 * src/internal/internal_batch.hh:251-258, blaspp/src/device_batch_trsm.cc:71-111).
 * Method: invert the diagonal blocks of A, then block substitution as tile
 * GEMMs on the DMMA kernel.  b_desc is a DEVICE array; scratch is DEVICE
 * memory of te_trsm_scratch_bytes(dtype, na, count) bytes, private to the call
 * until the stream has passed it.
 * ---------------------------------------------------------------------- */
typedef struct te_trsm_desc {
    void*   b;
    int32_t m, n;            /* extents of B_i (the side-determined one must equal na) */
    int32_t ldb;
    int32_t reserved;
} te_trsm_desc;              /* 24 bytes */

size_t te_trsm_scratch_bytes(int dtype, int na, int64_t count);
int te_trsm_batched(int device, void* stream, int dtype, int side, int uplo, int opa, int diag,
                    const void* alpha, const void* A, int lda, int na,
                    const te_trsm_desc* b_desc, int64_t count, int32_t max_m, int32_t max_n,
                    void* scratch, size_t scratch_bytes);

/* Allocation-free variant of te_trsm_batched.  It has no counterpart in the
 * reference (blas::batch::trsm takes no workspace argument; cuBLAS allocates
 * what it needs behind the call).
 *
 * Who needs it: callers that cannot give the engine device scratch and cannot
 * allocate any at this point -- a solver whose tile pool has already claimed
 * all the HBM it may use, or a panel step that must not risk cudaMalloc
 * stalling the device.  The scratch is a fixed 768 KiB block inside the
 * library image on every GPU that loads it (held until the process exits), so
 * the call makes no cudaMalloc / cudaFree and uses no allocating library
 * routine.  Arguments and results are those of te_trsm_batched without the
 * scratch pair.
 *
 * Limits: na <= TE_TRSM_STATIC_MAX_NA and count <= TE_TRSM_STATIC_MAX_COUNT;
 * anything larger returns TE_ERR_LIMIT with the limits in te_last_error().
 * The first call on a device looks up the block's device address
 * (cudaGetSymbolAddress) and sets kernel attributes; make that first call
 * before starting any stream capture on that device.
 * All calls on a device share the one block: two calls whose work can overlap
 * on the GPU (different streams with no ordering between them) race and give
 * wrong results -- keep them on one stream or order them with events.  The
 * block keeps the previous call's contents; nothing in it needs to start at
 * zero because each call overwrites every part before reading it. */
#define TE_TRSM_STATIC_MAX_NA    512
#define TE_TRSM_STATIC_MAX_COUNT 256
int te_trsm_batched_static(int device, void* stream, int dtype, int side, int uplo, int opa,
                           int diag, const void* alpha, const void* A, int lda, int na,
                           const te_trsm_desc* b_desc, int64_t count, int32_t max_m,
                           int32_t max_n);

/* ------------------------------------------------------------------------
 * Tile broadcasts between the box's GPUs over NCCL (one rank per GPU).
 *
 * Replaces the MPI side of BaseMatrix::tileBcast / listBcast / listBcastMT:
 *   tileBcastToSet / tileIbcastToSet   include/slate/BaseMatrix.hh:2283-2387
 *   internal::cubeBcastPattern         src/internal/internal_comm.cc:72-115
 *   Tile::isend / irecv                include/slate/Tile.hh:1084-1200
 *   internal::reduce_info (MPI_MIN)    src/internal/internal_reduce_info.cc:31
 * A broadcast of tile (i, j) to a set of ranks is: between te_comm_group_start
 * and te_comm_group_end, the root calls te_comm_send once per destination and
 * each destination calls te_comm_recv; every rank brackets the same list.
 * Buffers are DEVICE memory; everything is enqueued on `stream`.
 * te_comm_unique_id is called on one rank and its TE_COMM_ID_BYTES bytes are
 * handed to the others by the launcher (torch.distributed, MPI, a file ...).
 * ---------------------------------------------------------------------- */
#define TE_COMM_ID_BYTES 128
enum { TE_REDUCE_SUM = 0, TE_REDUCE_MIN = 1, TE_REDUCE_MAX = 2 };
typedef struct te_comm te_comm;
int te_comm_unique_id(void* id_out, size_t bytes);
int te_comm_init_rank(const void* id, size_t bytes, int nranks, int rank, int device,
                      te_comm** comm);
/* establishes every pairwise connection now (NCCL does it lazily otherwise); synchronous */
int te_comm_connect_all(te_comm* comm, void* stream);
int te_comm_destroy(te_comm* comm);
int te_comm_rank(const te_comm* comm);
int te_comm_size(const te_comm* comm);
int te_comm_group_start(void);
int te_comm_group_end(void);
int te_comm_send(te_comm* comm, const void* buf, size_t bytes, int peer, void* stream);
int te_comm_recv(te_comm* comm, void* buf, size_t bytes, int peer, void* stream);
int te_comm_bcast(te_comm* comm, void* buf, size_t bytes, int root, void* stream);
int te_comm_allreduce_i64(te_comm* comm, int64_t* buf_dev, size_t count, int op, void* stream);
int te_comm_allreduce_f64(te_comm* comm, double* buf_dev, size_t count, int op, void* stream);

/* ------------------------------------------------------------------------
 * Copy-engine tile transport between the box's GPUs (one rank per GPU, every
 * rank a process on the same box).  The default data plane of
 * BaseMatrix::tileBcast / listBcast / listBcastMT when more than one rank runs.
 *
 * Replaces, like the NCCL calls above, the per-tile MPI messages of
 *   tileBcastToSet / tileIbcastToSet   include/slate/BaseMatrix.hh:2283-2387
 *   Tile::isend / irecv                include/slate/Tile.hh:1084-1200
 * and, for several GPUs under one rank, the per-tile device-to-device
 * cudaMemcpyAsync of Tile::copyData (include/slate/Tile.hh:971-1042,
 * BaseMatrix.hh:2052-2097).  A tile moves as ONE cudaMemcpyAsync into memory
 * of the destination GPU that the root's process has mapped with CUDA IPC,
 * i.e. on a copy engine over NVLink; arrival is a sequence number written
 * behind the data and waited for by a stream memory operation.  Neither side
 * launches a kernel, so a broadcast never queues behind (or takes SMs from)
 * the trailing update that is running.
 *
 * te_peer_open (collective over `comm`, synchronous) gives every rank an
 * arena of `arena_bytes` of landing zones, opens every peer's arena and sets
 * up a host mailbox in POSIX shared memory.  For message number s of the
 * ordered pair (root, dest) -- both ranks count them from the same broadcast
 * plan -- the calls are, in this order:
 *   dest  te_peer_post(root, bytes, &s, &landing)   reserve + announce a zone
 *   root  te_peer_fetch(dest, bytes, &s, &offset)   host wait for the zone
 *         te_peer_put(dest, offset + o, tile, n, stream)   one per tile / run
 *         te_peer_signal(dest, s, stream)           after the last put
 *   dest  te_peer_wait(root, s, stream)             stream waits for arrival
 *         ... device-to-device copies out of `landing` on `stream` ...
 *         te_peer_consumed(root, s, stream)         zone reusable after this
 * A message may not exceed te_peer_max_message() = arena / (2 * ranks) bytes:
 * with every rank posting at most one message per peer before it copies the
 * previous ones out, a post can always be satisfied by waiting for zones that
 * are already being drained.  Host waits (post: ring full; fetch: zone not yet
 * announced) give up after TE_PEER_TIMEOUT_S seconds (default 120) with
 * TE_ERR_CUDA instead of hanging.  TE_PEER_MEMOPS=0 replaces the stream
 * memory operations by a one-thread kernel (wait) and an 8-byte copy (signal).
 * ---------------------------------------------------------------------- */
int te_peer_open(te_comm* comm, size_t arena_bytes);
int te_peer_close(te_comm* comm);             /* collective; te_comm_destroy calls it */
int te_peer_is_open(const te_comm* comm);
size_t te_peer_max_message(const te_comm* comm);
int te_peer_post(te_comm* comm, int root, size_t bytes, uint64_t* seq, void** landing);
int te_peer_fetch(te_comm* comm, int dest, size_t bytes, uint64_t* seq, uint64_t* offset);
int te_peer_put(te_comm* comm, int dest, uint64_t offset, const void* src, size_t bytes,
                void* stream);
int te_peer_signal(te_comm* comm, int dest, uint64_t seq, void* stream);
int te_peer_wait(te_comm* comm, int root, uint64_t seq, void* stream);
int te_peer_consumed(te_comm* comm, int root, uint64_t seq, void* stream);
/* counters since te_peer_open; memops: bit 0 = wait, bit 1 = signal by stream memory op */
int te_peer_stats(const te_comm* comm, uint64_t* puts, uint64_t* put_bytes, uint64_t* signals,
                  uint64_t* waits, int* memops);

/* ------------------------------------------------------------------------
 * Harness kernels around the hot path (inputs and residual checks).
 *
 * te_matgen_tile reproduces matgen bit for bit on the device: entry (i, j) of
 * the tile is Philox-2x64-10 keyed by `seed` at counter (i + ioff, j + joff)
 * (matgen/random.cc:53-159; dist 1 = uniform [0,1), 2 = uniform (-1,1));
 * diag_add is added where i + ioff == j + joff ("rand_dominant",
 * matgen/generate_type_rand.hh:70-78).  te_tile_add is B = alpha A + beta B
 * (src/cuda/device_geadd.cu); te_tile_sumsq adds sum |a_ij|^2 over mask
 * (0 full, 1 lower, 2 upper, 3 strictly lower) to *accum_dev;
 * te_tile_keep_triangle zeroes the other strict triangle (device_tzset.cu).
 * te_tile_scale is A = beta A with *beta of the matrix's type, zeros stored
 * when beta == 0 (src/cuda/device_gescale.cu / device_geset.cu, as
 * internal_gemmA.cc:312-320 uses them).
 * ---------------------------------------------------------------------- */
int te_matgen_tile(int device, void* stream, int dtype, int dist, int64_t seed, int m, int n,
                   int64_t ioff, int64_t joff, void* A, int64_t lda, double diag_add);
int te_tile_add(int device, void* stream, int dtype, int m, int n, double alpha, const void* A,
                int64_t lda, double beta, void* B, int64_t ldb);
int te_tile_scale(int device, void* stream, int dtype, int m, int n, const void* beta, void* A,
                  int64_t lda);

/* Tile transposes (src/cuda/device_transpose.cu: device::transpose and its batched forms,
 * used by tileLayoutConvert).  AT (n x m) = A^T, or A^H if conj (complex only).  AT == A
 * transposes in place and needs m == n, ldat == lda.  The batched form takes device arrays
 * of `batch` tile pointers (batch <= 65535); passing the same array twice means in place. */
int te_tile_transpose(int device, void* stream, int dtype, int conj, int m, int n, const void* A,
                      int64_t lda, void* AT, int64_t ldat);
int te_tile_transpose_batch(int device, void* stream, int dtype, int conj, int m, int n,
                            const void* const* A_array_dev, int64_t lda, void* const* AT_array_dev,
                            int64_t ldat, int batch);

/* Fills the triangle of a square tile that is not stored from the one that is (uplo):
 * A(j,i) = A(i,j), or its conjugate with the diagonal made real if hermitian.  Turns the
 * diagonal tile of a Hermitian / symmetric matrix into a plain GEMM operand for slate::hemm /
 * symm (the reference calls a tile hemm on the host there, src/internal/internal_hemm.cc). */
int te_tile_symmetrize(int device, void* stream, int dtype, int uplo, int hermitian, int n, void* A,
                       int64_t lda);

/* Panel <-> tiles.  Host-origin tiles that are neighbours in the caller's array cross PCIe as
 * ONE copy into / out of a device panel (column-major, leading dimension ld_panel elements);
 * this call moves the pieces between the panel and the tiles' own device blocks: to_tiles != 0
 * panel -> tiles, else tiles -> panel.  pieces_host is a HOST array (it travels in the kernel
 * parameters, TE_MAX_PANEL_PIECES per launch).  Stands in for the per-tile copies of
 * Tile::copyData (include/slate/Tile.hh:971-1042) inside BaseMatrix::tileGetAllForWritingOnDevices /
 * tileUpdateAllOrigin (include/slate/BaseMatrix.hh:3081-3243): one host<->device copy costs the
 * enqueueing thread ~21 us whatever its size, so a slab of 2048 tiles cost 43 ms per direction. */
#define TE_MAX_PANEL_PIECES 96
typedef struct te_panel_piece {
    void*   tile;            /* device block of the tile                               */
    int64_t panel_offset;    /* element offset of the piece inside the panel           */
    int32_t m, n;            /* extents of the piece                                   */
    int32_t ld;              /* leading dimension of the tile block, elements          */
    int32_t reserved;
} te_panel_piece;            /* 32 bytes */
int te_panel_pieces(int device, void* stream, int dtype, int to_tiles, void* panel,
                    int64_t ld_panel, const te_panel_piece* pieces_host, int count);

int te_tile_sumsq(int device, void* stream, int dtype, int mask, int m, int n, const void* A,
                  int64_t lda, double* accum_dev);
int te_tile_keep_triangle(int device, void* stream, int dtype, int uplo, int m, int n, void* A,
                          int64_t lda);

/* ------------------------------------------------------------------------
 * FP64 issue-rate microbenchmarks (the roofline denominators).
 *   kind 0: mma.sync.m8n8k4.f64 (DMMA), register resident
 *   kind 1: fma.rn.f64 (DFMA), register resident
 * Runs `iters` inner iterations on every SM, times with CUDA events and
 * returns TFLOP/s in *tflops.  Synchronous.
 * ---------------------------------------------------------------------- */
int te_peak_fp64(int device, int kind, int iters, double* tflops);

/* Same with the occupancy shape explicit (8 warps x 4 CTAs per SM x 16
 * independent accumulator chains per thread is what te_peak_fp64 uses);
 * chains in {1,2,4,8,16,32}.  For latency / occupancy sweeps. */
int te_peak_fp64_ex(int device, int kind, int iters, int warps_per_cta, int ctas_per_sm,
                    int chains, double* tflops);

#ifdef __cplusplus
}
#endif
#endif /* TILE_ENGINE_H */
