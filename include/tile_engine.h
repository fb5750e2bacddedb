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
 * beta == 0 means C is not read (NaN-safe, as in BLAS).
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
