// te_gemm.cu -- batched FP64 tile update  C_i = alpha op(A_i) op(B_i) + beta C_i
//
// One launch per block outer product per device.  Replaces the
// cublasDgemmBatched / cublasDsyrk calls behind SLATE's
// internal::gemm/herk/syrk<Target::Devices>
// (reference: src/internal/internal_gemm.cc:355-518, internal_herk.cc:351-535,
//  blaspp/src/device_batch_gemm.cc:25-142).
//
// Design (B200, sm_100a)
//   * tcgen05/TMEM has no FP64 kind, so the FP64 tensor path is the
//     warp-synchronous DMMA.8x8x4 (mma.sync.m8n8k4.f64).
//   * CTA tile 128x128, 8 consumer warps with warp tile 64(m) x 32(n): 32
//     accumulator fragments (128 registers) per thread, 12 LDS.64 per 32 DMMA;
//     4 producer warps; setmaxnreg moves registers from producers to consumers.
//   * operands staged GMEM -> SMEM by 16-byte cp.async (.cg, L2 only) through
//     a 5-stage ring of K=16 slices guarded by mbarriers; +4 padding per row
//     makes every 64-bit fragment load bank-conflict free for both storage
//     orders.
//   * the flat list of (tile, 128x128 block) work items, in the order the host
//     laid the descriptors out (ordered for L2 reuse of the A/B panels), is cut
//     into runs of up to 8 consecutive items per CTA (one CTA per SM at a time,
//     ~0.5 ms each).  Inside a run the cp.async ring runs ACROSS items, so the
//     next block's first slices are in flight while the current block's epilogue
//     reads/writes C.  Short-lived CTAs let the block scheduler balance the load
//     and let higher-priority streams (NCCL broadcasts, the Cholesky panel) take
//     SMs as CTAs retire; a persistent grid with fixed shares doubled the kernel
//     time whenever an NCCL kernel held a few SMs (measured on 2 GPUs).
//   * ragged edges (m, n, k not multiples of the tile) are zero-filled by
//     cp.async src-size and masked in the epilogue; odd leading dimensions /
//     8-byte aligned tiles fall back to 8-byte cp.async per element.
//   * epilogue in registers: each lane owns two consecutive rows of one
//     column (the MMA is issued on C^T), so C is read/written as 16-byte
//     vectors forming 64-byte column segments per quad.
#include "te_common.cuh"
#include <stdlib.h>
#include <string.h>

namespace te {

constexpr int BM = 128;
constexpr int BN = 128;                            // widest CTA tile (also the complex kernel's)
constexpr int KC = 16;
constexpr int NCONSUMERS = 8;                      // complex kernel: consumer warps
constexpr int NPRODUCERS = 4;                      //                 producer warps
constexpr int NTHREADS = (NCONSUMERS + NPRODUCERS) * 32;
constexpr int REGS_PRODUCER = 56;                  // setmaxnreg budgets: 56*P + 224*C = 168*(P+C)
constexpr int REGS_CONSUMER = 224;                 //   for C = 2P consumer / producer threads

// Shape of the real kernel.  Two are built:
//   Wide      128x128 CTA tile, 8 consumer + 4 producer warps, 5-stage ring, ONE CTA per SM
//             (the default, and the shape every in-place panel GEMM needs);
//   PairFull  128x64 CTA tile, 4 consumer + 2 producer warps (+ 2 idle warps that complete the
//             producer warpgroup), 3-stage ring, TWO CTAs per SM.  TE_DGEMM_SHAPE=pairfull.
// The warp tile (64x32, 128 accumulator registers) and the DMMA / LDS mix are the same.  The idea
// of the second shape: with one CTA per SM all eight consumer warps reach the epilogue of a block
// together and the tensor pipe idles while C is read, scaled and written (ncu: pipe active
// 88.5 %, long-scoreboard stalls 21 %); two independent CTAs per SM drift apart, so one CTA's
// epilogue runs under the other's DMMA stream.  MEASURED (round 2, one B200, n = 65536): it is
// SLOWER, 30.8 against 32.7 TFLOP/s over 8 block steps (29.6 against 31.6 at the 2x4 grid's launch
// shape): 128 registers per thread at launch leave the consumers 200 instead of 224 (3x the
// spill traffic), the ring is 3 deep instead of 5, and A slices cross L2 -> shared 1.5x as
// often.  Kept selectable as the record of that experiment.  (A 192-thread variant whose
// producer "warpgroup" had only two warps hung in setmaxnreg: the instruction is specified per
// group of four warps.)
template <int BN_, int NCONS_, int NPROD_, int STAGES_, int CTAS_PER_SM_, int NIDLE_ = 0,
          int REGS_C_ = REGS_CONSUMER, int REGS_P_ = REGS_PRODUCER, int KC_ = KC>
struct DCfg {
    static constexpr int KCv = KC_;              // K extent of one ring slice
    static constexpr int LD_Kv = KC_ + 4;        // row stride (doubles) of an [mn][k] slice
    static constexpr int BNv = BN_;
    static constexpr int NCONS = NCONS_;
    static constexpr int NPROD = NPROD_;
    static constexpr int NIDLE = NIDLE_;         // warps that only complete the producer warpgroup
    static constexpr int REGS_C = REGS_C_, REGS_P = REGS_P_;
    static constexpr int STAGES = STAGES_;
    static constexpr int CTAS_PER_SM = CTAS_PER_SM_;
    static constexpr int THREADS = (NCONS_ + NPROD_ + NIDLE_) * 32;
    static constexpr int LD_A = BM + 4;      // row stride (doubles) of a [k][m] slice of A
    static constexpr int LD_B = BN_ + 4;     //                         a [k][n] slice of B
    static constexpr int SLICE_A = (KC_ * LD_A > BM * LD_Kv) ? KC_ * LD_A : BM * LD_Kv;
    static constexpr int SLICE_B = (KC_ * LD_B > BN_ * LD_Kv) ? KC_ * LD_B : BN_ * LD_Kv;
    static constexpr int SMEM = STAGES_ * (SLICE_A + SLICE_B) * (int)sizeof(double) + 2 * STAGES_ * 8;
    static_assert(NCONS_ == 2 * (BN_ / 32), "consumer warps tile the block as 2 (m) x BN/32 (n)");
    // registers per thread at launch (multiple of 8) and the pool setmaxnreg redistributes
    static constexpr int LAUNCH_REGS = (65536 / (THREADS * CTAS_PER_SM_)) / 8 * 8;
    static_assert(REGS_P_ * (NPROD_ + NIDLE_) + REGS_C_ * NCONS_ <= LAUNCH_REGS * (NCONS_ + NPROD_ + NIDLE_),
                  "setmaxnreg budget");
};
using CfgWide = DCfg<128, 8, 4, 5, 1>;
// Wide32 (the default): K slices of 32 in a 3-stage ring (221 KB): half as many full / empty
// barrier hand-offs per block (ncu: they were ~4.7 % of the consumers' K-loop samples with
// slices of 16).  Measured on one B200 against Wide: 33.08 vs 32.85 TFLOP/s at n = 65536
// (12 steps), 32.39 vs 31.7 at the 2x4 grid's launch shape.
using CfgWide32 = DCfg<128, 8, 4, 3, 1, 0, REGS_CONSUMER, REGS_PRODUCER, 32>;
// PairFull: two idle warps complete the producer warpgroup (setmaxnreg is specified per
// warpgroup of four warps); they give their registers back and exit.  256 threads per CTA,
// 128 registers per thread at launch, so the consumers get 200 instead of 224.
using CfgPairFull = DCfg<64, 4, 2, 3, 2, 2, 200, 56>;

struct GemmParams {
    const te_tile_desc* desc;      // device array (batched mode)
    int64_t total_work;            // count * blocks_per_tile
    int32_t items_per_cta;         // work items handled by one CTA
    int32_t wave_ctas;             // CTAs that run side by side (= SM count)
    int32_t nbm, nbn;              // 128-blocks per tile in m and n
    double alpha, beta;
    int32_t inline_count;          // > 0: use inl[] instead of desc
    int32_t use_red;               // beta == 1: interior blocks add into C with red.global.add.f64
    te_tile_desc inl[TE_MAX_INLINE_DESC];
};

// One 128x128 block of one tile.
struct Item {
    const double* a;
    const double* b;
    double* c;
    int m, n, k;                   // extents of this block (m, n may exceed 128: remaining)
    int lda, ldb, ldc;
    int flags;
    int m0, n0;                    // offset of the block inside its tile
    bool a16, b16, c16;            // 16-byte path usable
};

template <bool INL, int BMv = 128, int BNv = 128>
__device__ __forceinline__ bool decode(const GemmParams& p, int64_t w, Item& it)
{
    const int bpt = p.nbm * p.nbn;
    const int64_t t = w / bpt;
    const int sub = (int)(w - t * bpt);
    const int sn = sub / p.nbm;
    const int sm = sub - sn * p.nbm;
    const te_tile_desc* d = INL ? &p.inl[t] : &p.desc[t];
    const int m = d->m, n = d->n;
    const int m0 = sm * BMv, n0 = sn * BNv;
    if (m0 >= m || n0 >= n)
        return false;
    const int flags = d->flags;
    if ((flags & TE_TILE_LOWER) && (m0 + BMv - 1 < n0))
        return false;
    if ((flags & TE_TILE_UPPER) && (n0 + BNv - 1 < m0))
        return false;
    it.m = m - m0;
    it.n = n - n0;
    it.k = d->k;
    it.lda = d->lda;
    it.ldb = d->ldb;
    it.ldc = d->ldc;
    it.flags = flags;
    it.m0 = m0;
    it.n0 = n0;
    it.a = (const double*)d->a;
    it.b = (const double*)d->b;
    it.c = (double*)d->c;
    it.a16 = ((((uintptr_t)it.a) & 15) == 0) && ((it.lda & 1) == 0);
    it.b16 = ((((uintptr_t)it.b) & 15) == 0) && ((it.ldb & 1) == 0);
    it.c16 = ((((uintptr_t)it.c) & 15) == 0) && ((it.ldc & 1) == 0);
    return true;
}

// Walks the work list of this CTA.  CTAs are short-lived on purpose (about a millisecond):
// the hardware block scheduler then balances the load by itself and lets kernels on
// higher-priority streams (NCCL broadcasts, the Cholesky panel) take SMs as CTAs retire,
// which a grid of persistent CTAs with fixed shares would block.
// The list is cut into waves of wave_ctas * items_per_cta consecutive items; CTA c of a
// wave takes items c, c + wave_ctas, c + 2 wave_ctas ... of it.  The CTAs of a wave run
// side by side, so at any moment they work on ~wave_ctas CONSECUTIVE items (about 9
// tiles), which keeps the A / B tiles they share resident in L2; giving each CTA a
// contiguous run instead spread the SMs over 148 different tiles and cost 1.6x more
// DRAM reads (ncu, profiles/).
template <bool INL, int BMv = 128, int BNv = 128, int KCv = KC>
struct Walker {
    int64_t w;
    int left;      // items this CTA may still take
    int stride;
    int nk;        // K slices of the current item
    bool valid;
    Item it;

    __device__ __forceinline__ void seek(const GemmParams& p)
    {
        valid = false;
        while (left > 0 && w < p.total_work) {
            if (decode<INL, BMv, BNv>(p, w, it)) {
                valid = true;
                nk = (it.k + KCv - 1) / KCv;
                return;
            }
            w += stride;
            --left;
        }
    }
    __device__ __forceinline__ void start(const GemmParams& p)
    {
        stride = p.wave_ctas;
        const int64_t wave = blockIdx.x / p.wave_ctas;
        w = wave * (int64_t)p.wave_ctas * p.items_per_cta + (blockIdx.x % p.wave_ctas);
        left = p.items_per_cta;
        seek(p);
    }
    __device__ __forceinline__ void next_item(const GemmParams& p)
    {
        w += stride;
        --left;
        seek(p);
    }
};

// Stage one K slice of one operand with the 128 threads of the producer warpgroup.
// `g` points at element (x0, k0) of the operand; x_rem / k_rem are the valid
// extents from there (anything beyond is zero-filled).
//   KMAJOR == false: global (x, kk) at g[x + kk*ld]  -> smem [kk][LD_MN]
//   KMAJOR == true : global (x, kk) at g[kk + x*ld]  -> smem [x][LD_K]
template <bool KMAJOR, int X, int LD_MNv, int NPROD, int KCv>
__device__ __forceinline__ void load_slice(double* s, const double* g, int ld,
                                           int x_rem, int k_rem, bool al16, int ptid)
{
    const uint32_t sbase = smem_u32(s);
    constexpr int LD_K = KCv + 4;
    constexpr int KHALF = KCv / 2;               // chunks per row of an [mn][k] slice
    constexpr int NCHUNK = X * KCv / 2;          // 16-byte chunks per slice
    constexpr int HALF = X / 2;                  // chunks per k row of an [k][mn] slice
#pragma unroll
    for (int i = 0; i < NCHUNK / (NPROD * 32); ++i) {
        const int c = ptid + i * (NPROD * 32);
        int row, col, nv;
        uint32_t soff;
        if (!KMAJOR) {
            row = c / HALF;            // kk
            col = (c % HALF) * 2;      // x
            nv = (row < k_rem) ? min(max(x_rem - col, 0), 2) : 0;
            soff = (uint32_t)(row * LD_MNv + col);
        }
        else {
            row = c / KHALF;           // x
            col = (c % KHALF) * 2;     // kk
            nv = (row < x_rem) ? min(max(k_rem - col, 0), 2) : 0;
            soff = (uint32_t)(row * LD_K + col);
        }
        const double* src = (nv > 0) ? (g + ((int64_t)row * ld + col)) : g;
        const uint32_t dst = sbase + soff * 8u;
        if (al16) {
            cp_async16(dst, src, nv * 8);
        }
        else {
            cp_async8(dst, src, nv > 0 ? 8 : 0);
            cp_async8(dst + 8u, (nv > 1) ? (src + 1) : src, nv > 1 ? 8 : 0);
        }
    }
}

// Warp roles: warps 0..7 (two warpgroups) consume (LDS + DMMA + epilogue), warps
// 8..11 (one warpgroup) produce (cp.async into the ring, L2 prefetch of the C
// block).  setmaxnreg moves registers from the producers to the consumers.
// full[s] completes when the 128 producer threads' copies of stage s have
// landed; empty[s] when the 8 consumer warps have finished reading it.  There
// is no CTA-wide barrier in the loop.
template <bool TA, bool TB, bool INL, typename Cfg>
__global__ void __launch_bounds__(Cfg::THREADS, Cfg::CTAS_PER_SM)
dgemm_tiles_kernel(const __grid_constant__ GemmParams p)
{
    constexpr int STAGES = Cfg::STAGES;
    constexpr int KC = Cfg::KCv;                 // (shadows the file-scope value)
    constexpr int LD_K = Cfg::LD_Kv;
    constexpr int NCONSUMERS = Cfg::NCONS;       // (shadow the file-scope values of the complex kernel)
    constexpr int NPRODUCERS = Cfg::NPROD;
    constexpr int BN = Cfg::BNv;
    constexpr int SLICE_A = Cfg::SLICE_A, SLICE_B = Cfg::SLICE_B;
    constexpr int LD_MN_A = Cfg::LD_A, LD_MN_B = Cfg::LD_B;
    extern __shared__ __align__(16) double smem[];
    double* sA = smem;                       // STAGES slices
    double* sB = smem + STAGES * SLICE_A;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * (SLICE_A + SLICE_B));
    const uint32_t full0 = smem_u32(bars);
    const uint32_t empty0 = smem_u32(bars + STAGES);

    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int warp = tid >> 5;

    if (tid == 0) {
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(full0 + 8 * s, NPRODUCERS * 32);
            mbar_init(empty0 + 8 * s, NCONSUMERS);
        }
    }
    __syncthreads();

    // op(A) = N: stored m x k, m contiguous -> [k][mn].  op(A) = T: k contiguous -> [mn][k].
    // op(B) = N: stored k x n, k contiguous -> [mn][k].  op(B) = T: n contiguous -> [k][mn].
    constexpr bool A_KMAJOR = TA;
    constexpr bool B_KMAJOR = !TB;

    Walker<INL, BM, BN, KC> wk;
    wk.start(p);
    int stage = 0;
    uint32_t phase = 0;

    if (warp >= NCONSUMERS) {
        // ------------------------------ producers -----------------------------
        asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;\n" :: "n"(Cfg::REGS_P));
        if (warp >= NCONSUMERS + NPRODUCERS)
            return;                  // idle warps: only there to complete the warpgroup
        const int ptid = tid - NCONSUMERS * 32;
#pragma unroll 1
        while (wk.valid) {
            const Item& it = wk.it;
            if (wk.nk > 0) {
                if (p.beta != 0.0) {
                    // pull this block of C towards L2 while the main loop runs
                    const int rows = min(it.m, BM), cols = min(it.n, BN);
                    const double* cblk = it.c + (int64_t)it.n0 * it.ldc + it.m0;
                    for (int q = ptid; q < cols * 8; q += NPRODUCERS * 32) {
                        const int col = q >> 3, seg = (q & 7) * 16;
                        if (seg < rows)
                            prefetch_l2(cblk + (int64_t)col * it.ldc + seg);
                    }
                }
#pragma unroll 1
                for (int kc = 0; kc < wk.nk; ++kc) {
                    const int k0 = kc * KC;
                    const int k_rem = it.k - k0;
                    const double* ga = A_KMAJOR
                        ? it.a + (int64_t)it.m0 * it.lda + k0
                        : it.a + (int64_t)k0 * it.lda + it.m0;
                    const double* gb = B_KMAJOR
                        ? it.b + (int64_t)it.n0 * it.ldb + k0
                        : it.b + (int64_t)k0 * it.ldb + it.n0;
                    mbar_wait(empty0 + 8 * stage, phase ^ 1);
                    load_slice<A_KMAJOR, BM, LD_MN_A, NPRODUCERS, KC>(sA + stage * SLICE_A, ga, it.lda, it.m, k_rem, it.a16, ptid);
                    load_slice<B_KMAJOR, BN, LD_MN_B, NPRODUCERS, KC>(sB + stage * SLICE_B, gb, it.ldb, it.n, k_rem, it.b16, ptid);
                    cp_async_mbar_arrive(full0 + 8 * stage);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
            wk.next_item(p);
        }
        cp_async_wait_all();
        return;
    }

    // -------------------------------- consumers --------------------------------
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;\n" :: "n"(Cfg::REGS_C));
    const int g = lane >> 2;                 // 0..7
    const int t = lane & 3;                  // 0..3
    const int wm = (warp & 1) * 64;          // warp tile origin in the block
    const int wn = (warp >> 1) * 32;

#pragma unroll 1
    while (wk.valid) {
        double acc[4][8][2];
#pragma unroll
        for (int in = 0; in < 4; ++in)
#pragma unroll
            for (int im = 0; im < 8; ++im)
                acc[in][im][0] = acc[in][im][1] = 0.0;

        const int nk = wk.nk;
#pragma unroll 1
        for (int kc = 0; kc < nk; ++kc) {
            mbar_wait(full0 + 8 * stage, phase);
            const double* a_s = sA + stage * SLICE_A;
            const double* b_s = sB + stage * SLICE_B;
#pragma unroll
            for (int k4 = 0; k4 < KC / 4; ++k4) {
                double af[8], bf[4];
                const int kk = k4 * 4 + t;
#pragma unroll
                for (int im = 0; im < 8; ++im) {
                    const int x = wm + im * 8 + g;
                    af[im] = A_KMAJOR ? a_s[x * LD_K + kk] : a_s[kk * LD_MN_A + x];
                }
#pragma unroll
                for (int in = 0; in < 4; ++in) {
                    const int x = wn + in * 8 + g;
                    bf[in] = B_KMAJOR ? b_s[x * LD_K + kk] : b_s[kk * LD_MN_B + x];
                }
#pragma unroll
                for (int in = 0; in < 4; ++in)
#pragma unroll
                    for (int im = 0; im < 8; ++im)
                        dmma884(acc[in][im][0], acc[in][im][1], bf[in], af[im]);
            }
            __syncwarp();
            if (lane == 0)
                mbar_arrive(empty0 + 8 * stage);
            if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }

        // ---- epilogue: lane owns rows (2t, 2t+1) of column g in each 8x8 ----
        {
            const Item& it = wk.it;
            const double alpha = p.alpha, beta = p.beta;
            const bool lower = (it.flags & TE_TILE_LOWER) != 0;
            const bool upper = (it.flags & TE_TILE_UPPER) != 0;
            double* cblk = it.c + (int64_t)it.n0 * it.ldc + it.m0;
            const bool interior = (it.m >= wm + 64) && (it.n >= wn + 32) && !lower && !upper && it.c16;
            if (interior) {
                // Whole 64x32 warp tile inside the block, 16-byte aligned: branch-free, with all
                // loads of a half (16 x 16 bytes per lane) in flight before the first use, so the
                // C read costs one memory round trip instead of one per fragment.
                double* c0 = cblk + (int64_t)(wn + g) * it.ldc + wm + 2 * t;
                const int64_t cstep = (int64_t)8 * it.ldc;
                if (p.use_red) {
                    // beta == 1 (every SUMMA step but the first, every trailing update of a
                    // factorisation): C += alpha * acc as fire-and-forget reductions done in L2.
                    // No C load, so no round trip to wait for before the next block's DMMAs.
                    // Each element belongs to exactly one work item of a launch, so the result is
                    // fl(c + fl(alpha * acc)): bit for bit what the load / fma / store path gives.
#pragma unroll
                    for (int in = 0; in < 4; ++in)
#pragma unroll
                        for (int im = 0; im < 8; ++im) {
                            double* dst = c0 + in * cstep + im * 8;
                            asm volatile("red.global.add.f64 [%0], %1;" :: "l"(dst), "d"(alpha * acc[in][im][0]) : "memory");
                            asm volatile("red.global.add.f64 [%0], %1;" :: "l"(dst + 1), "d"(alpha * acc[in][im][1]) : "memory");
                        }
                }
                else
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    double2 old[2][8];
                    if (beta != 0.0) {
#pragma unroll
                        for (int i2 = 0; i2 < 2; ++i2)
#pragma unroll
                            for (int im = 0; im < 8; ++im)
                                old[i2][im] = __ldcs(reinterpret_cast<const double2*>(
                                    c0 + (2 * h + i2) * cstep + im * 8));
                    }
#pragma unroll
                    for (int i2 = 0; i2 < 2; ++i2)
#pragma unroll
                        for (int im = 0; im < 8; ++im) {
                            double v0 = alpha * acc[2 * h + i2][im][0];
                            double v1 = alpha * acc[2 * h + i2][im][1];
                            if (beta != 0.0) {
                                v0 += beta * old[i2][im].x;
                                v1 += beta * old[i2][im].y;
                            }
                            // C is touched once per launch: streaming loads / stores (evict
                            // first) so that it does not push the reused A / B panels out of L2
                            __stcs(reinterpret_cast<double2*>(c0 + (2 * h + i2) * cstep + im * 8),
                                   make_double2(v0, v1));
                        }
                }
            }
            else {
#pragma unroll
                for (int in = 0; in < 4; ++in) {
                    const int col = wn + in * 8 + g;
                    if (col >= it.n)
                        continue;
                    double* ccol = cblk + (int64_t)col * it.ldc;
                    const int gcol = it.n0 + col;          // column inside the tile
#pragma unroll
                    for (int im = 0; im < 8; ++im) {
                        const int row = wm + im * 8 + 2 * t;
                        if (row >= it.m)
                            continue;
                        const int grow = it.m0 + row;
                        bool ok0 = true, ok1 = (row + 1 < it.m);
                        if (lower) { ok0 = ok0 && (grow >= gcol); ok1 = ok1 && (grow + 1 >= gcol); }
                        if (upper) { ok0 = ok0 && (grow <= gcol); ok1 = ok1 && (grow + 1 <= gcol); }
                        double v0 = alpha * acc[in][im][0];
                        double v1 = alpha * acc[in][im][1];
                        double* dst = ccol + row;
                        if (ok0 && ok1 && it.c16) {
                            if (beta != 0.0) {
                                const double2 old = *reinterpret_cast<const double2*>(dst);
                                v0 += beta * old.x;
                                v1 += beta * old.y;
                            }
                            *reinterpret_cast<double2*>(dst) = make_double2(v0, v1);
                        }
                        else {
                            if (ok0) {
                                if (beta != 0.0) v0 += beta * dst[0];
                                dst[0] = v0;
                            }
                            if (ok1) {
                                if (beta != 0.0) v1 += beta * dst[1];
                                dst[1] = v1;
                            }
                        }
                    }
                }
            }
        }
        wk.next_item(p);
    }
}

// =================================================================================
// complex-FP64:  C_i = alpha op(A_i) op(B_i) + beta C_i,  interleaved (re, im) storage
// =================================================================================
// Same structure as the real kernel (producer warpgroup, cp.async ring with mbarriers,
// runs of consecutive work items per CTA); differences:
//   * one complex product is four real DMMAs on de-interleaved fragments
//     (re += ar*br - ai*bi, im += ar*bi + ai*br); a fragment load is one LDS.128 that
//     brings (re, im) of an element, conjugation flips the sign of im in registers;
//   * the CTA tile is 64 (m) x 128 (n), warp tile 32 x 32: 16 complex accumulator
//     fragments = 128 registers, as in the real kernel;
//   * one cp.async (16 bytes) per element, so any 16-byte aligned complex array takes
//     the fast path whatever its leading dimension;
//   * row padding of +2 / +4 complex elements keeps the 128-bit fragment loads
//     conflict free (8 lanes per phase).
constexpr int ZBM = 64;
constexpr int ZSTAGES = 3;
constexpr int ZLD_A_MN = ZBM + 2;          // [k][m] slice of A, complex elements
constexpr int ZLD_B_MN = BN + 2;           // [k][n] slice of B
constexpr int ZLD_K = KC + 4;              // [mn][k] slices
constexpr int ZSLICE_A = (KC * ZLD_A_MN > ZBM * ZLD_K) ? KC * ZLD_A_MN : ZBM * ZLD_K;   // complex elems
constexpr int ZSLICE_B = (KC * ZLD_B_MN > BN * ZLD_K) ? KC * ZLD_B_MN : BN * ZLD_K;
constexpr int ZSMEM_BYTES = ZSTAGES * (ZSLICE_A + ZSLICE_B) * 16 + 2 * ZSTAGES * 8;

struct ZParams {
    GemmParams g;                 // alpha / beta of g are unused
    double ar, ai, br, bi;        // alpha, beta
};

// X = 64 (A) or 128 (B) rows of the "mn" dimension
template <bool KMAJOR, int X, int LD_MNv>
__device__ __forceinline__ void zload_slice(double2* s, const double2* g, int ld, int x_rem, int k_rem,
                                            bool al16, int ptid)
{
    const uint32_t sbase = smem_u32(s);
    constexpr int NELEM = X * KC;
#pragma unroll
    for (int i = 0; i < NELEM / (NPRODUCERS * 32); ++i) {
        const int c = ptid + i * (NPRODUCERS * 32);
        int row, col;
        bool ok;
        uint32_t soff;
        if (!KMAJOR) {
            row = c / X;               // kk
            col = c % X;               // x
            ok = (row < k_rem) && (col < x_rem);
            soff = (uint32_t)(row * LD_MNv + col);
        }
        else {
            row = c / KC;              // x
            col = c % KC;              // kk
            ok = (row < x_rem) && (col < k_rem);
            soff = (uint32_t)(row * ZLD_K + col);
        }
        const double2* src = ok ? (g + ((int64_t)row * ld + col)) : g;
        const uint32_t dst = sbase + soff * 16u;
        if (al16) {
            cp_async16(dst, src, ok ? 16 : 0);
        }
        else {
            cp_async8(dst, src, ok ? 8 : 0);
            cp_async8(dst + 8u, reinterpret_cast<const double*>(src) + 1, ok ? 8 : 0);
        }
    }
}

// OPA / OPB: 0 = N, 1 = T, 2 = C
template <int OPA, int OPB, bool INL>
__global__ void __launch_bounds__(NTHREADS, 1)
zgemm_tiles_kernel(const __grid_constant__ ZParams zp)
{
    const GemmParams& p = zp.g;
    extern __shared__ __align__(16) double smem[];
    double2* sA = reinterpret_cast<double2*>(smem);
    double2* sB = sA + ZSTAGES * ZSLICE_A;
    uint64_t* bars = reinterpret_cast<uint64_t*>(sB + ZSTAGES * ZSLICE_B);
    const uint32_t full0 = smem_u32(bars);
    const uint32_t empty0 = smem_u32(bars + ZSTAGES);

    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int warp = tid >> 5;

    if (tid == 0) {
        for (int s = 0; s < ZSTAGES; ++s) {
            mbar_init(full0 + 8 * s, NPRODUCERS * 32);
            mbar_init(empty0 + 8 * s, NCONSUMERS);
        }
    }
    __syncthreads();

    constexpr bool A_KMAJOR = (OPA != 0);
    constexpr bool B_KMAJOR = (OPB == 0);
    constexpr bool CONJ_A = (OPA == 2);
    constexpr bool CONJ_B = (OPB == 2);

    Walker<INL, ZBM> wk;
    wk.start(p);
    int stage = 0;
    uint32_t phase = 0;

    if (warp >= NCONSUMERS) {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;\n" :: "n"(REGS_PRODUCER));
        const int ptid = tid - NCONSUMERS * 32;
#pragma unroll 1
        while (wk.valid) {
            const Item& it = wk.it;
            const double2* A = reinterpret_cast<const double2*>(it.a);
            const double2* B = reinterpret_cast<const double2*>(it.b);
            const bool a16 = (((uintptr_t)it.a) & 15) == 0, b16 = (((uintptr_t)it.b) & 15) == 0;
#pragma unroll 1
            for (int kc = 0; kc < wk.nk; ++kc) {
                const int k0 = kc * KC;
                const int k_rem = it.k - k0;
                const double2* ga = A_KMAJOR ? A + (int64_t)it.m0 * it.lda + k0
                                             : A + (int64_t)k0 * it.lda + it.m0;
                const double2* gb = B_KMAJOR ? B + (int64_t)it.n0 * it.ldb + k0
                                             : B + (int64_t)k0 * it.ldb + it.n0;
                mbar_wait(empty0 + 8 * stage, phase ^ 1);
                zload_slice<A_KMAJOR, ZBM, ZLD_A_MN>(sA + stage * ZSLICE_A, ga, it.lda, it.m, k_rem, a16, ptid);
                zload_slice<B_KMAJOR, BN, ZLD_B_MN>(sB + stage * ZSLICE_B, gb, it.ldb, it.n, k_rem, b16, ptid);
                cp_async_mbar_arrive(full0 + 8 * stage);
                if (++stage == ZSTAGES) { stage = 0; phase ^= 1; }
            }
            wk.next_item(p);
        }
        cp_async_wait_all();
        return;
    }

    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;\n" :: "n"(REGS_CONSUMER));
    const int g = lane >> 2;
    const int t = lane & 3;
    const int wm = (warp & 1) * 32;
    const int wn = (warp >> 1) * 32;

#pragma unroll 1
    while (wk.valid) {
        double cre[4][4][2], cim[4][4][2];
#pragma unroll
        for (int in = 0; in < 4; ++in)
#pragma unroll
            for (int im = 0; im < 4; ++im) {
                cre[in][im][0] = cre[in][im][1] = 0.0;
                cim[in][im][0] = cim[in][im][1] = 0.0;
            }
        const int nk = wk.nk;
#pragma unroll 1
        for (int kc = 0; kc < nk; ++kc) {
            mbar_wait(full0 + 8 * stage, phase);
            const double2* a_s = sA + stage * ZSLICE_A;
            const double2* b_s = sB + stage * ZSLICE_B;
#pragma unroll
            for (int k4 = 0; k4 < KC / 4; ++k4) {
                double2 af[4], bf[4];
                const int kk = k4 * 4 + t;
#pragma unroll
                for (int im = 0; im < 4; ++im) {
                    const int x = wm + im * 8 + g;
                    af[im] = A_KMAJOR ? a_s[x * ZLD_K + kk] : a_s[kk * ZLD_A_MN + x];
                    if (CONJ_A) af[im].y = -af[im].y;
                }
#pragma unroll
                for (int in = 0; in < 4; ++in) {
                    const int x = wn + in * 8 + g;
                    bf[in] = B_KMAJOR ? b_s[x * ZLD_K + kk] : b_s[kk * ZLD_B_MN + x];
                    if (CONJ_B) bf[in].y = -bf[in].y;
                }
#pragma unroll
                for (int in = 0; in < 4; ++in) {
                    const double nby = -bf[in].y;
#pragma unroll
                    for (int im = 0; im < 4; ++im) {
                        dmma884(cre[in][im][0], cre[in][im][1], bf[in].x, af[im].x);
                        dmma884(cre[in][im][0], cre[in][im][1], nby, af[im].y);
                        dmma884(cim[in][im][0], cim[in][im][1], bf[in].x, af[im].y);
                        dmma884(cim[in][im][0], cim[in][im][1], bf[in].y, af[im].x);
                    }
                }
            }
            __syncwarp();
            if (lane == 0)
                mbar_arrive(empty0 + 8 * stage);
            if (++stage == ZSTAGES) { stage = 0; phase ^= 1; }
        }

        // ---- epilogue: lane owns rows (2t, 2t+1) of column g of each 8 x 8 fragment ----
        {
            const Item& it = wk.it;
            const bool lower = (it.flags & TE_TILE_LOWER) != 0;
            const bool upper = (it.flags & TE_TILE_UPPER) != 0;
            const bool herm = (it.flags & TE_TILE_HERMITIAN) != 0;
            const bool beta_nz = (zp.br != 0.0) || (zp.bi != 0.0);
            const bool c16 = (((uintptr_t)it.c) & 15) == 0;
            double2* cblk = reinterpret_cast<double2*>(it.c) + (int64_t)it.n0 * it.ldc + it.m0;
            const bool interior = p.use_red != 0 && (it.m >= wm + 32) && (it.n >= wn + 32) && !lower
                                  && !upper && !herm && c16;
            if (interior) {
                // Whole 32x32 warp tile inside the block (as in the real kernel): no per-element
                // tests, and no load -> use -> store chain per element.
                double2* c0 = cblk + (int64_t)(wn + g) * it.ldc + wm + 2 * t;
                const int64_t cstep = (int64_t)8 * it.ldc;
                if (zp.br == 1.0 && zp.bi == 0.0) {
                    // beta == 1: C += alpha * acc by reductions performed in L2 (no C load)
#pragma unroll
                    for (int in = 0; in < 4; ++in)
#pragma unroll
                        for (int im = 0; im < 4; ++im)
#pragma unroll
                            for (int e = 0; e < 2; ++e) {
                                const double ar = cre[in][im][e], ai = cim[in][im][e];
                                double* dst = reinterpret_cast<double*>(c0 + in * cstep + im * 8 + e);
                                asm volatile("red.global.add.f64 [%0], %1;" :: "l"(dst), "d"(zp.ar * ar - zp.ai * ai) : "memory");
                                asm volatile("red.global.add.f64 [%0], %1;" :: "l"(dst + 1), "d"(zp.ar * ai + zp.ai * ar) : "memory");
                            }
                }
                else {
                    // half of the warp tile at a time: its 16 loads in flight before the first use
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        double2 old[2][4][2];
                        if (beta_nz) {
#pragma unroll
                            for (int i2 = 0; i2 < 2; ++i2)
#pragma unroll
                                for (int im = 0; im < 4; ++im)
#pragma unroll
                                    for (int e = 0; e < 2; ++e)
                                        old[i2][im][e] = __ldcs(c0 + (2 * h + i2) * cstep + im * 8 + e);
                        }
#pragma unroll
                        for (int i2 = 0; i2 < 2; ++i2)
#pragma unroll
                            for (int im = 0; im < 4; ++im)
#pragma unroll
                                for (int e = 0; e < 2; ++e) {
                                    const double ar = cre[2 * h + i2][im][e], ai = cim[2 * h + i2][im][e];
                                    double vr = zp.ar * ar - zp.ai * ai;
                                    double vi = zp.ar * ai + zp.ai * ar;
                                    if (beta_nz) {
                                        vr += zp.br * old[i2][im][e].x - zp.bi * old[i2][im][e].y;
                                        vi += zp.br * old[i2][im][e].y + zp.bi * old[i2][im][e].x;
                                    }
                                    __stcs(c0 + (2 * h + i2) * cstep + im * 8 + e, make_double2(vr, vi));
                                }
                    }
                }
            }
            else
#pragma unroll
            for (int in = 0; in < 4; ++in) {
                const int col = wn + in * 8 + g;
                if (col >= it.n) continue;
                const int gcol = it.n0 + col;
#pragma unroll
                for (int im = 0; im < 4; ++im) {
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        const int row = wm + im * 8 + 2 * t + e;
                        if (row >= it.m) continue;
                        const int grow = it.m0 + row;
                        if (lower && grow < gcol) continue;
                        if (upper && grow > gcol) continue;
                        const double ar = cre[in][im][e], ai = cim[in][im][e];
                        double vr = zp.ar * ar - zp.ai * ai;
                        double vi = zp.ar * ai + zp.ai * ar;
                        double2* dst = cblk + (int64_t)col * it.ldc + row;
                        if (beta_nz) {
                            double2 old;
                            if (c16) old = *dst;
                            else { old.x = reinterpret_cast<double*>(dst)[0]; old.y = reinterpret_cast<double*>(dst)[1]; }
                            vr += zp.br * old.x - zp.bi * old.y;
                            vi += zp.br * old.y + zp.bi * old.x;
                        }
                        if (herm && grow == gcol) vi = 0.0;
                        if (c16) *dst = make_double2(vr, vi);
                        else { reinterpret_cast<double*>(dst)[0] = vr; reinterpret_cast<double*>(dst)[1] = vi; }
                    }
                }
            }
        }
        wk.next_item(p);
    }
}

// ---- host side ----------------------------------------------------------------
template <bool TA, bool TB, bool INL, typename Cfg>
static int launch(int device, cudaStream_t stream, const GemmParams& p)
{
    static bool configured[64] = {};
    auto kern = dgemm_tiles_kernel<TA, TB, INL, Cfg>;
    if (device < 64 && !configured[device]) {
        TE_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           Cfg::SMEM));
        configured[device] = true;
    }
    const int64_t per_wave = (int64_t)p.wave_ctas * p.items_per_cta;
    const int64_t grid = ((p.total_work + per_wave - 1) / per_wave) * p.wave_ctas;
    kern<<<(unsigned)grid, Cfg::THREADS, Cfg::SMEM, stream>>>(p);
    te::count_launch();
    TE_CUDA_CHECK(cudaGetLastError());
    return TE_OK;
}

// Up to 16 consecutive 128x128 blocks per CTA (about 1 ms at k = 512) once there are
// enough blocks for several waves; fewer when the launch is small, so that it still
// spreads over all SMs.
//
// The cap is also a latency knob: a CTA keeps its SM until its blocks are done, and a
// kernel on a high-priority stream (potrf's panel chain) can only start when a CTA of
// the running update exits.  te_gemm_set_items_per_cta lets a driver that has such a
// chain trade a little locality for shorter CTAs (measured on 4 B200, dpotrf n = 98304:
// cap 16 -> 117.0, 4 -> 121.2, 2 -> 120.0, 1 -> 118.1 TFLOP/s; dgemm unchanged at 4).
static int g_items_cap_override = 0;

extern "C" int te_gemm_set_items_per_cta(int cap)
{
    if (cap < 0 || cap > 64) return TE_ERR_ARG;
    g_items_cap_override = cap;
    return TE_OK;
}

static int items_per_cta(int device, int64_t total_work)
{
    static int env_cap = -1;
    if (env_cap < 0) {
        const char* e = getenv("TE_ITEMS_PER_CTA");
        env_cap = (e && atoi(e) > 0) ? atoi(e) : 0;
    }
    // the environment wins (experiments), then the driver's hint, then the default
    const int cap = env_cap > 0 ? env_cap : (g_items_cap_override > 0 ? g_items_cap_override : 16);
    const int64_t sms = num_sms(device);
    int64_t g = total_work / (4 * sms);
    if (g < 1) g = 1;
    if (g > cap) g = cap;
    return (int)g;
}

template <bool INL, typename Cfg>
static int dispatch(int device, cudaStream_t stream, int opa, int opb, const GemmParams& p)
{
    const bool ta = (opa != TE_OP_N), tb = (opb != TE_OP_N);
    if (!ta && !tb) return launch<false, false, INL, Cfg>(device, stream, p);
    if (!ta &&  tb) return launch<false, true,  INL, Cfg>(device, stream, p);
    if ( ta && !tb) return launch<true,  false, INL, Cfg>(device, stream, p);
    return launch<true, true, INL, Cfg>(device, stream, p);
}

// Which shape the real kernel runs in: 3 wide with K slices of 32 (default), 0 wide with K slices
// of 16, 2 two CTAs per SM (TE_DGEMM_SHAPE=wide32|wide|pairfull; see DCfg).
static int dgemm_shape()
{
    static int choice = -1;
    if (choice < 0) {
        const char* e = getenv("TE_DGEMM_SHAPE");
        choice = (e && strcmp(e, "pairfull") == 0) ? 2 : (e && strcmp(e, "wide") == 0) ? 0 : 3;
    }
    return choice;
}

template <typename Cfg>
static void shape_params(int device, GemmParams& p, int64_t count, int32_t max_m, int32_t max_n)
{
    p.nbm = (max_m + BM - 1) / BM;
    p.nbn = (max_n + Cfg::BNv - 1) / Cfg::BNv;
    p.total_work = count * p.nbm * p.nbn;
    p.wave_ctas = num_sms(device) * Cfg::CTAS_PER_SM;
    // same time per CTA (about a millisecond at k = 512) whatever the block size
    p.items_per_cta = items_per_cta(device, p.total_work * Cfg::BNv / BN) * (BN / Cfg::BNv);
}

// Reductions instead of the read-modify-write epilogue when beta == 1 (TE_DGEMM_RED=0 turns it
// off).  Measured on one B200: 33.51 against 33.06 TFLOP/s at n = 65536 (12 steps, 11 of them
// with beta == 1), 32.75 against 32.30 at the 2x4 grid's launch shape.
static bool dgemm_red_enabled()
{
    static int on = -1;
    if (on < 0) {
        const char* e = getenv("TE_DGEMM_RED");
        on = (e && e[0] == '0') ? 0 : 1;
    }
    return on == 1;
}

template <bool INL>
static int dgemm_launch_shape(int device, cudaStream_t stream, int opa, int opb, GemmParams& p,
                              int64_t count, int32_t max_m, int32_t max_n, bool wide)
{
    p.use_red = (p.beta == 1.0 && !wide && dgemm_red_enabled()) ? 1 : 0;
    const int shape = (wide && dgemm_shape() == 2) ? 0 : dgemm_shape();
    if (shape == 2) {
        shape_params<CfgPairFull>(device, p, count, max_m, max_n);
        return dispatch<INL, CfgPairFull>(device, stream, opa, opb, p);
    }
    if (shape == 3) {
        shape_params<CfgWide32>(device, p, count, max_m, max_n);
        return dispatch<INL, CfgWide32>(device, stream, opa, opb, p);
    }
    shape_params<CfgWide>(device, p, count, max_m, max_n);
    return dispatch<INL, CfgWide>(device, stream, opa, opb, p);
}

static bool op_ok(int op) { return op == TE_OP_N || op == TE_OP_T || op == TE_OP_C; }

// `wide`: the call updates op(A) in place (C aliases A) and relies on ONE CTA producing all
// <= 128 columns of a block row; it must run in the 128-wide shape.
int dgemm_batched(int device, cudaStream_t stream, int opa, int opb, double alpha, double beta,
                  const te_tile_desc* desc, int64_t count, int32_t max_m, int32_t max_n, bool wide)
{
    GemmParams p;
    p.desc = desc;
    p.alpha = alpha;
    p.beta = beta;
    p.inline_count = 0;
    return dgemm_launch_shape<false>(device, stream, opa, opb, p, count, max_m, max_n, wide);
}

int dgemm_inline(int device, cudaStream_t stream, int opa, int opb, double alpha, double beta,
                 const te_tile_desc* desc_host, int count, bool wide)
{
    GemmParams p;
    p.desc = nullptr;
    int32_t max_m = 0, max_n = 0;
    for (int i = 0; i < count; ++i) {
        p.inl[i] = desc_host[i];
        if (desc_host[i].m > max_m) max_m = desc_host[i].m;
        if (desc_host[i].n > max_n) max_n = desc_host[i].n;
    }
    p.alpha = alpha;
    p.beta = beta;
    p.inline_count = count;
    return dgemm_launch_shape<true>(device, stream, opa, opb, p, count, max_m, max_n, wide);
}


// Interior warp tiles of the complex kernel take the branch-free epilogue the real kernel has
// (batched C loads, reductions for beta == 1); TE_ZGEMM_EPILOGUE=0: element by element, as in
// round 1.  Measured on one B200, zgemm n = 24576 nb = 384: 34.05 against 31.94 TFLOP/s.
static bool zgemm_fast_epilogue()
{
    static int on = -1;
    if (on < 0) {
        const char* e = getenv("TE_ZGEMM_EPILOGUE");
        on = (e && e[0] == '0') ? 0 : 1;
    }
    return on == 1;
}

template <int OPA, int OPB, bool INL>
static int zlaunch(int device, cudaStream_t stream, const ZParams& zp)
{
    static bool configured[64] = {};
    auto kern = zgemm_tiles_kernel<OPA, OPB, INL>;
    if (device < 64 && !configured[device]) {
        TE_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           ZSMEM_BYTES));
        configured[device] = true;
    }
    const GemmParams& p = zp.g;
    const int64_t per_wave = (int64_t)p.wave_ctas * p.items_per_cta;
    const int64_t grid = ((p.total_work + per_wave - 1) / per_wave) * p.wave_ctas;
    kern<<<(unsigned)grid, NTHREADS, ZSMEM_BYTES, stream>>>(zp);
    te::count_launch();
    TE_CUDA_CHECK(cudaGetLastError());
    return TE_OK;
}

static int op_index(int op) { return op == TE_OP_N ? 0 : (op == TE_OP_T ? 1 : 2); }

template <bool INL>
static int zdispatch(int device, cudaStream_t stream, int opa, int opb, const ZParams& zp)
{
    switch (op_index(opa) * 3 + op_index(opb)) {
        case 0: return zlaunch<0, 0, INL>(device, stream, zp);
        case 1: return zlaunch<0, 1, INL>(device, stream, zp);
        case 2: return zlaunch<0, 2, INL>(device, stream, zp);
        case 3: return zlaunch<1, 0, INL>(device, stream, zp);
        case 4: return zlaunch<1, 1, INL>(device, stream, zp);
        case 5: return zlaunch<1, 2, INL>(device, stream, zp);
        case 6: return zlaunch<2, 0, INL>(device, stream, zp);
        case 7: return zlaunch<2, 1, INL>(device, stream, zp);
        default: return zlaunch<2, 2, INL>(device, stream, zp);
    }
}

int zgemm_batched(int device, cudaStream_t stream, int opa, int opb, const double* alpha,
                  const double* beta, const te_tile_desc* desc, int64_t count, int32_t max_m,
                  int32_t max_n)
{
    ZParams zp;
    GemmParams& p = zp.g;
    p.desc = desc;
    p.nbm = (max_m + ZBM - 1) / ZBM;
    p.nbn = (max_n + BN - 1) / BN;
    p.total_work = count * p.nbm * p.nbn;
    p.items_per_cta = items_per_cta(device, p.total_work);
    p.wave_ctas = num_sms(device);
    p.alpha = p.beta = 0.0;
    p.inline_count = 0;
    p.use_red = zgemm_fast_epilogue() ? 1 : 0;
    zp.ar = alpha[0];  zp.ai = alpha[1];  zp.br = beta[0];  zp.bi = beta[1];
    return zdispatch<false>(device, stream, opa, opb, zp);
}

int zgemm_inline(int device, cudaStream_t stream, int opa, int opb, const double* alpha,
                 const double* beta, const te_tile_desc* desc_host, int count)
{
    ZParams zp;
    GemmParams& p = zp.g;
    p.desc = nullptr;
    int32_t max_m = 0, max_n = 0;
    for (int i = 0; i < count; ++i) {
        p.inl[i] = desc_host[i];
        if (desc_host[i].m > max_m) max_m = desc_host[i].m;
        if (desc_host[i].n > max_n) max_n = desc_host[i].n;
    }
    p.nbm = (max_m + ZBM - 1) / ZBM;
    p.nbn = (max_n + BN - 1) / BN;
    p.total_work = (int64_t)count * p.nbm * p.nbn;
    p.items_per_cta = items_per_cta(device, p.total_work);
    p.wave_ctas = num_sms(device);
    p.alpha = p.beta = 0.0;
    p.inline_count = count;
    p.use_red = zgemm_fast_epilogue() ? 1 : 0;
    zp.ar = alpha[0];  zp.ai = alpha[1];  zp.br = beta[0];  zp.bi = beta[1];
    return zdispatch<true>(device, stream, opa, opb, zp);
}

// dtype dispatch used by the C ABI and by the panel code (te_panel.cu)
int gemm_batched_any(int device, cudaStream_t stream, int dtype, int opa, int opb,
                     const void* alpha, const void* beta, const te_tile_desc* desc,
                     int64_t count, int32_t max_m, int32_t max_n, bool inplace_a)
{
    if (count == 0 || max_m == 0 || max_n == 0)
        return TE_OK;
    if (dtype == TE_F64)
        return dgemm_batched(device, stream, opa, opb, *(const double*)alpha,
                             *(const double*)beta, desc, count, max_m, max_n, inplace_a);
    if (dtype == TE_C64)
        return zgemm_batched(device, stream, opa, opb, (const double*)alpha, (const double*)beta,
                             desc, count, max_m, max_n);
    return TE_ERR_ARG;
}

int gemm_inline_any(int device, cudaStream_t stream, int dtype, int opa, int opb,
                    const void* alpha, const void* beta, const te_tile_desc* desc_host, int count,
                    bool inplace_a)
{
    if (count == 0)
        return TE_OK;
    if (dtype == TE_F64)
        return dgemm_inline(device, stream, opa, opb, *(const double*)alpha,
                            *(const double*)beta, desc_host, count, inplace_a);
    if (dtype == TE_C64)
        return zgemm_inline(device, stream, opa, opb, (const double*)alpha, (const double*)beta,
                            desc_host, count);
    return TE_ERR_ARG;
}

}  // namespace te

extern "C" int te_gemm_batched(int device, void* stream, int dtype, int opa, int opb,
                               const void* alpha, const void* beta,
                               const te_tile_desc* desc, int64_t count,
                               int32_t max_m, int32_t max_n)
{
    if (!alpha || !beta || count < 0 || max_m < 0 || max_n < 0 ||
        !te::op_ok(opa) || !te::op_ok(opb) || (count > 0 && !desc))
        return TE_ERR_ARG;
    int st = te::set_device(device);
    if (st != TE_OK) return st;
    return te::gemm_batched_any(device, (cudaStream_t)stream, dtype, opa, opb, alpha, beta, desc,
                                count, max_m, max_n, false);
}

extern "C" int te_gemm_inline(int device, void* stream, int dtype, int opa, int opb,
                              const void* alpha, const void* beta,
                              const te_tile_desc* desc_host, int count)
{
    if (!alpha || !beta || count < 0 || count > TE_MAX_INLINE_DESC ||
        !te::op_ok(opa) || !te::op_ok(opb) || (count > 0 && !desc_host))
        return TE_ERR_ARG;
    int st = te::set_device(device);
    if (st != TE_OK) return st;
    return te::gemm_inline_any(device, (cudaStream_t)stream, dtype, opa, opb, alpha, beta,
                               desc_host, count, false);
}
