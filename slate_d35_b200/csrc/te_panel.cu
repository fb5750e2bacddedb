// te_panel.cu -- the panel (critical-path) kernels of the tiled Cholesky:
// single-tile potrf and batched triangular solve, FP64 and complex-FP64.
//
// Replaces, on the Target::Devices path,
//   lapack::potrf(uplo, n, A, lda, info, queue) -> cusolverDnXpotrf
//       (reference: src/internal/internal_potrf.cc:57-81,
//        lapackpp/src/cuda/cuda_potrf.cc:128-171)
//   blas::batch::trsm -> cublas{D,Z}trsmBatched
//       (reference: src/internal/internal_trsm.cc:132-261,
//        blaspp/src/device_batch_trsm.cc:71-111).
//
// Structure: a tile is processed in diagonal blocks of IB columns.  One CTA
// factors an IB x IB diagonal block in shared memory (right-looking Cholesky)
// and also forms the inverse of its triangular factor; everything else -- the
// solve of the block column below it, the symmetric trailing update, and all of
// the batched triangular solve -- is then expressed as tile GEMMs and runs on
// the DMMA kernel of te_gemm.cu (in place where the CTA tile covers the whole
// block, which holds because IB <= 128).  `info` stays on the device: nothing
// here synchronises with the host.
#include "te_common.cuh"
#include <mutex>

namespace te {

// implemented in te_gemm.cu
// inplace_a: C aliases op(A) (X <- X * M with n <= 128): one CTA must own whole block rows,
// so the real kernel is run in its 128-column shape (te_gemm.cu, DCfg)
int gemm_batched_any(int device, cudaStream_t stream, int dtype, int opa, int opb,
                     const void* alpha, const void* beta, const te_tile_desc* desc,
                     int64_t count, int32_t max_m, int32_t max_n, bool inplace_a = false);
int gemm_inline_any(int device, cudaStream_t stream, int dtype, int opa, int opb,
                    const void* alpha, const void* beta, const te_tile_desc* desc_host, int count,
                    bool inplace_a = false);

constexpr int PANEL_THREADS = 512;

template <typename T> struct PanelCfg;
template <> struct PanelCfg<double> { static constexpr int IB = 128; };
template <> struct PanelCfg<zd>     { static constexpr int IB = 64; };

// One CTA.  Works on the n x n (n <= IB) triangular block at A (uplo given by
// `upper`), held in shared memory as a LOWER matrix S (an upper block is loaded
// conjugate-transposed and stored back the same way).
//   do_potf2 == 2: as 1 but the inverse is not formed (last diagonal block of a tile:
//                  nothing below it needs it)
//   do_potf2 != 0: S <- chol(S); failure at local pivot j sets *info = info_base + j + 1
//                  (first failure wins; a non-zero *info on entry skips the block).
//   always:        inv <- inverse of the triangular factor, dense IB-strided block in
//                  the same orientation as A's storage (zeros in the other triangle).
//   unit_diag:     the diagonal of the factor is taken as 1 (trtri-only mode).
template <typename T, int IB>
__global__ void __launch_bounds__(PANEL_THREADS)
potf2_trtri_kernel(T* A, int lda, int upper, T* inv, int* info, int info_base,
                   int do_potf2, int unit_diag, int64_t a_block_stride, int64_t inv_block_stride,
                   int n_total)
{
    extern __shared__ __align__(16) unsigned char panel_smem[];
    T* S = reinterpret_cast<T*>(panel_smem);               // [IB][IB+1]
    T* xd = S + IB * (IB + 1);                             // [IB] diagonal of the inverse
    __shared__ int s_fail;
    constexpr int LD = IB + 1;
    const int tid = threadIdx.x;

    // batched over diagonal blocks (trtri-only mode): block b starts at (b*IB, b*IB)
    A += (int64_t)blockIdx.x * a_block_stride;
    inv += (int64_t)blockIdx.x * inv_block_stride;
    const int n = min(IB, n_total - (int)blockIdx.x * IB);

    if (do_potf2 && *info != 0)
        return;
    if (tid == 0) s_fail = 0;

    for (int idx = tid; idx < n * n; idx += PANEL_THREADS) {
        const int r = idx % n, c = idx / n;
        const T v = A[r + (int64_t)c * lda];
        if (!upper) { if (r >= c) S[r * LD + c] = v; }
        else        { if (r <= c) S[c * LD + r] = conj_(v); }
    }
    __syncthreads();

    if (do_potf2) {
        // Right-looking Cholesky with the matrix held in REGISTERS: the 512 threads form a
        // 32 x 16 grid, thread (tr, tc) owns the elements (tr + 32 m, tc + 16 q).  Per column:
        // the owners publish the column through a double-buffered shared vector, one
        // block-wide barrier, and every thread applies the rank-1 update to its own patch.
        constexpr int RP = IB / 32;                 // rows per thread
        constexpr int CP = IB / 16;                 // columns per thread
        T* colbuf = xd + IB;                        // [2][IB]
        const int tr = tid & 31, tc = tid >> 5;
        T a[RP][CP];
#pragma unroll
        for (int m = 0; m < RP; ++m)
#pragma unroll
            for (int q = 0; q < CP; ++q) {
                const int i = tr + 32 * m, c = tc + 16 * q;
                a[m][q] = (i < n && c <= i) ? S[i * LD + c] : zero_<T>();
            }
        // publish column 0
        if (tc == 0) {
#pragma unroll
            for (int m = 0; m < RP; ++m) colbuf[tr + 32 * m] = a[m][0];
        }
        __syncthreads();
        for (int j = 0; j < n; ++j) {
            const T* col = colbuf + (j & 1) * IB;
            const double d = real_(col[j]);
            if (!(d > 0.0)) {                              // also catches NaN; uniform
                if (tid == 0) s_fail = j + 1;
                break;
            }
            // rsqrt is a few instructions (sqrt + divide are ~40, executed by every warp in
            // every step); d * rsqrt(d) is within 2 ulp of sqrt(d)
            const double rinv = rsqrt(d);
            const double rdiag = d * rinv;
            const int qj = j >> 4;
            const bool own = (tc == (j & 15));
            T li[RP], lc[CP];
#pragma unroll
            for (int m = 0; m < RP; ++m) li[m] = col[tr + 32 * m] * rinv;
#pragma unroll
            for (int q = 0; q < CP; ++q) lc[q] = conj_(col[tc + 16 * q] * rinv);
            // branch-free rank-1 update of the columns right of j (select, not branch: the
            // unrolled 4 x 8 patch must stay a handful of instructions per element)
#pragma unroll
            for (int q = 0; q < CP; ++q) {
                // tc is the warp index, so this test is warp-uniform: a real branch that
                // skips the column groups already factored
                if (tc + 16 * q > j) {
#pragma unroll
                    for (int m = 0; m < RP; ++m) a[m][q] = a[m][q] - li[m] * lc[q];
                }
            }
            if (own) {
                // this thread holds part of column j itself: keep the scaled column
#pragma unroll
                for (int q = 0; q < CP; ++q)
                    if (q == qj) {
#pragma unroll
                        for (int m = 0; m < RP; ++m) {
                            const int i = tr + 32 * m;
                            if (i == j) set_real(a[m][q], rdiag);
                            else if (i > j) a[m][q] = li[m];
                        }
                    }
            }
            // publish column j + 1 (now final) into the other buffer
            if (j + 1 < n && tc == ((j + 1) & 15)) {
                T* nxt = colbuf + ((j + 1) & 1) * IB;
                const int qn = (j + 1) >> 4;
#pragma unroll
                for (int q = 0; q < CP; ++q)
                    if (q == qn) {
#pragma unroll
                        for (int m = 0; m < RP; ++m) nxt[tr + 32 * m] = a[m][q];
                    }
            }
            __syncthreads();
        }
        __syncthreads();
        if (s_fail != 0) {
            if (tid == 0) *info = info_base + s_fail;
            return;                                        // LAPACK leaves the tile partly factored
        }
#pragma unroll
        for (int m = 0; m < RP; ++m)
#pragma unroll
            for (int q = 0; q < CP; ++q) {
                const int i = tr + 32 * m, c = tc + 16 * q;
                if (i < n && c <= i) S[i * LD + c] = a[m][q];
            }
        __syncthreads();
    }

    // ---- inverse of the lower factor: X = S^-1, X^T kept in the strict upper part ----
    // Column j of X belongs to 4 adjacent lanes; lane `part` keeps the running sums
    //   s_i = sum_{l < current} S[i][l] X[l][j]   for the rows i = part + 4 m
    // in registers.  Step l: x_l = -s_l / S[l][l] comes from the lane that owns row l
    // (shuffle), then every lane adds S[i][l] x_l to its rows below l: 32 independent
    // LDS + FMA per step and no barrier at all.
    const bool want_inv = (do_potf2 != 2);
    if (want_inv)
    for (int i = tid; i < n; i += PANEL_THREADS) {
        if (unit_diag) xd[i] = from_real<T>(1.0);
        else {
            // complex diagonal of a Cholesky factor is real; a general triangular block may not be
            const T d = S[i * LD + i];
            const double den = real_(d * conj_(d));
            xd[i] = conj_(d) * (1.0 / den);
        }
    }
    __syncthreads();
    if (want_inv) {
        constexpr int MP = IB / 4;                  // rows per lane
        const int j = tid >> 2, part = tid & 3;
        if (j < IB) {                               // warps past the last column are idle
            T sums[MP];
#pragma unroll
            for (int m = 0; m < MP; ++m) sums[m] = zero_<T>();
            const bool active = (j < n);
            // the 4 lanes of a column move together; columns in one warp differ by < 8, so
            // all start at the first column of the warp (earlier steps are no-ops for them)
            const int l_begin = ((tid & ~31) >> 2) & ~3;      // multiple of 4
            const int gbase = threadIdx.x & 31 & ~3;
            for (int l4 = l_begin; l4 < n; l4 += 4) {
                // the running sums of rows l4 .. l4+3 sit in sums[l4 / 4] of lanes 0 .. 3
                T own_sum = zero_<T>();
                const int m0 = l4 >> 2;
#pragma unroll
                for (int m = 0; m < MP; ++m)
                    if (m == m0) own_sum = sums[m];
#pragma unroll
                for (int pp = 0; pp < 4; ++pp) {
                    const int l = l4 + pp;
                    // rows l4 .. l4+pp-1 were finished in this group of four: fold their
                    // contribution to row l in before it is read (done below via own_sum)
                    T xl;
                    if constexpr (sizeof(T) == 8) {
                        xl = __shfl_sync(0xffffffffu, own_sum, gbase | pp);
                    }
                    else {
                        xl.x = __shfl_sync(0xffffffffu, own_sum.x, gbase | pp);
                        xl.y = __shfl_sync(0xffffffffu, own_sum.y, gbase | pp);
                    }
                    if (l >= n || l < j) continue;             // no shuffles below this point
                    xl = (l == j) ? xd[l] : -(xl * xd[l]);
                    if (active && part == 0 && l > j) S[j * LD + l] = xl;   // X[l][j], transposed
                    // the lanes that own rows l+1 .. l4+3 need this x_l before their turn
                    if (part > pp) own_sum = own_sum + S[l * LD + (part - pp) * LD + l] * xl;
                    // rows below l: m >= mth.  Rows >= n only hold padding and sums[m0] is not
                    // read again, so neither needs a test of its own.
                    const T* Lcol = S + part * LD + l;          // row part, column l
                    const int mth = (l - part + 4) >> 2;
#pragma unroll
                    for (int m = 0; m < MP; ++m)
                        if (m >= mth) sums[m] = sums[m] + Lcol[m * 4 * LD] * xl;
                }
            }
        }
    }
    __syncthreads();

    // ---- write back ----
    for (int idx = tid; idx < n * n; idx += PANEL_THREADS) {
        const int r = idx % n, c = idx / n;
        // factor, in A's own triangle only
        if (do_potf2) {
            if (!upper) { if (r >= c) A[r + (int64_t)c * lda] = S[r * LD + c]; }
            else        { if (r <= c) A[r + (int64_t)c * lda] = conj_(S[c * LD + r]); }
        }
        if (!want_inv) continue;
        // inverse, dense block with leading dimension IB
        T v = zero_<T>();
        if (!upper) {
            if (r == c) v = xd[r];
            else if (r > c) v = S[c * LD + r];                     // X[r][c]
        }
        else {
            if (r == c) v = conj_(xd[r]);
            else if (r < c) v = conj_(S[r * LD + c]);              // conj(X[c][r])
        }
        inv[r + c * IB] = v;
    }
}

template <typename T>
static int panel_smem_bytes()
{
    constexpr int IB = PanelCfg<T>::IB;
    return (IB * (IB + 1) + 3 * IB) * (int)sizeof(T);
}

template <typename T>
static int configure_panel_kernel(int device)
{
    static bool done[64] = {};
    constexpr int IB = PanelCfg<T>::IB;
    if (device < 64 && done[device]) return TE_OK;
    TE_CUDA_CHECK(cudaFuncSetAttribute(potf2_trtri_kernel<T, IB>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       panel_smem_bytes<T>()));
    if (device < 64) done[device] = true;
    return TE_OK;
}

// ------------------------------- tile potrf ----------------------------------
template <typename T>
static int potrf_tile(int device, cudaStream_t stream, int dtype, int uplo, int n, T* A, int lda,
                      int* info, T* scratch)
{
    constexpr int IB = PanelCfg<T>::IB;
    int st = configure_panel_kernel<T>(device);
    if (st != TE_OK) return st;
    const int upper = (uplo == TE_UPLO_UPPER);
    TE_CUDA_CHECK(cudaMemsetAsync(info, 0, sizeof(int), stream));
    const double one[2] = {1.0, 0.0}, zero[2] = {0.0, 0.0}, mone[2] = {-1.0, 0.0};
    for (int jb = 0; jb < n; jb += IB) {
        const int ib = (n - jb < IB) ? n - jb : IB;
        T* Ajj = A + jb + (int64_t)jb * lda;
        const int rem = n - jb - ib;
        potf2_trtri_kernel<T, IB><<<1, PANEL_THREADS, panel_smem_bytes<T>(), stream>>>(
            Ajj, lda, upper, scratch, info, jb, rem > 0 ? 1 : 2, 0, 0, 0, ib);
        te::count_launch();
        TE_CUDA_CHECK(cudaGetLastError());
        if (rem <= 0) break;
        te_tile_desc d;
        d.reserved = 0;
        if (!upper) {
            T* P = A + (jb + ib) + (int64_t)jb * lda;            // rem x ib, below the block
            T* A22 = A + (jb + ib) + (int64_t)(jb + ib) * lda;
            // P <- P * inv(L)^H   (in place: the CTA tile spans all ib columns)
            d.a = P; d.b = scratch; d.c = P; d.m = rem; d.n = ib; d.k = ib;
            d.lda = lda; d.ldb = IB; d.ldc = lda; d.flags = TE_TILE_GENERAL;
            st = gemm_inline_any(device, stream, dtype, TE_OP_N, TE_OP_C, one, zero, &d, 1, true);
            if (st != TE_OK) return st;
            // A22 <- A22 - P P^H on the lower triangle
            d.a = P; d.b = P; d.c = A22; d.m = rem; d.n = rem; d.k = ib;
            d.lda = lda; d.ldb = lda; d.ldc = lda; d.flags = TE_TILE_LOWER | TE_TILE_HERMITIAN;
            st = gemm_inline_any(device, stream, dtype, TE_OP_N, TE_OP_C, mone, one, &d, 1);
            if (st != TE_OK) return st;
        }
        else {
            T* P = A + jb + (int64_t)(jb + ib) * lda;            // ib x rem, right of the block
            T* A22 = A + (jb + ib) + (int64_t)(jb + ib) * lda;
            // P <- inv(U)^H * P   (in place: the CTA tile spans all ib rows)
            d.a = scratch; d.b = P; d.c = P; d.m = ib; d.n = rem; d.k = ib;
            d.lda = IB; d.ldb = lda; d.ldc = lda; d.flags = TE_TILE_GENERAL;
            st = gemm_inline_any(device, stream, dtype, TE_OP_C, TE_OP_N, one, zero, &d, 1);
            if (st != TE_OK) return st;
            // A22 <- A22 - P^H P on the upper triangle
            d.a = P; d.b = P; d.c = A22; d.m = rem; d.n = rem; d.k = ib;
            d.lda = lda; d.ldb = lda; d.ldc = lda; d.flags = TE_TILE_UPPER | TE_TILE_HERMITIAN;
            st = gemm_inline_any(device, stream, dtype, TE_OP_C, TE_OP_N, mone, one, &d, 1);
            if (st != TE_OK) return st;
        }
    }
    return TE_OK;
}

// ------------------------------- batched trsm --------------------------------
struct TrsmPlan {
    const te_trsm_desc* b;     // device, count entries
    te_tile_desc* upd;         // device, nblk * count
    te_tile_desc* mul;         // device, nblk * count
    const void* A;
    const void* inv;
    int64_t count;
    int lda, na, nblk, IB, esize;
    int right, eff_lower, op_is_n;
};

// One thread per (step s, batch entry i): writes the update and multiply
// descriptors of diagonal block blk(s), in processing order.
__global__ void trsm_build_descs_kernel(TrsmPlan p)
{
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (int64_t)p.nblk * p.count) return;
    const int s = (int)(idx / p.count);
    const int64_t i = idx - (int64_t)s * p.count;
    // right: ascending when op(A) is upper; left: ascending when op(A) is lower
    const bool ascending = p.right ? !p.eff_lower : p.eff_lower;
    const int blk = ascending ? s : (p.nblk - 1 - s);
    const int j0 = blk * p.IB;
    const int ib = min(p.IB, p.na - j0);
    const int l0 = ascending ? 0 : j0 + ib;          // start of the already-solved range
    const int kk = ascending ? j0 : p.na - (j0 + ib);
    const te_trsm_desc bd = p.b[i];
    char* B = (char*)bd.b;
    const char* A = (const char*)p.A;
    const int64_t es = p.esize;
    te_tile_desc u, m;
    u.flags = m.flags = TE_TILE_GENERAL;
    u.reserved = m.reserved = 0;
    const char* invb = (const char*)p.inv + (int64_t)blk * p.IB * p.IB * es;
    if (p.right) {
        // B is m x na.  T = alpha B(:, blk) - X(:, solved) * M(solved, blk);  X(:, blk) = T * inv
        u.c = B + (int64_t)j0 * bd.ldb * es;  u.ldc = bd.ldb;  u.m = bd.m;  u.n = ib;  u.k = kk;
        u.a = B + (int64_t)l0 * bd.ldb * es;  u.lda = bd.ldb;
        u.b = p.op_is_n ? A + ((int64_t)l0 + (int64_t)j0 * p.lda) * es
                        : A + ((int64_t)j0 + (int64_t)l0 * p.lda) * es;
        u.ldb = p.lda;
        m.c = (void*)u.c;  m.ldc = bd.ldb;  m.a = u.c;  m.lda = bd.ldb;
        m.b = invb;  m.ldb = p.IB;  m.m = bd.m;  m.n = ib;  m.k = ib;
    }
    else {
        // B is na x n.  T = alpha B(blk, :) - M(blk, solved) * X(solved, :);  X(blk, :) = inv * T
        u.c = B + (int64_t)j0 * es;  u.ldc = bd.ldb;  u.m = ib;  u.n = bd.n;  u.k = kk;
        u.a = p.op_is_n ? A + ((int64_t)j0 + (int64_t)l0 * p.lda) * es
                        : A + ((int64_t)l0 + (int64_t)j0 * p.lda) * es;
        u.lda = p.lda;
        u.b = B + (int64_t)l0 * es;  u.ldb = bd.ldb;
        m.c = (void*)u.c;  m.ldc = bd.ldb;  m.b = u.c;  m.ldb = bd.ldb;
        m.a = invb;  m.lda = p.IB;  m.m = ib;  m.n = bd.n;  m.k = ib;
    }
    p.upd[idx] = u;
    p.mul[idx] = m;
}

template <typename T>
static size_t trsm_scratch_bytes(int na, int64_t count)
{
    constexpr int IB = PanelCfg<T>::IB;
    const int nblk = (na + IB - 1) / IB;
    size_t inv = (size_t)nblk * IB * IB * sizeof(T);
    size_t descs = 2 * (size_t)nblk * (size_t)count * sizeof(te_tile_desc);
    return inv + descs + 256;
}

template <typename T>
static int trsm_batched(int device, cudaStream_t stream, int dtype, int side, int uplo, int opa,
                        int diag, const void* alpha, const T* A, int lda, int na,
                        const te_trsm_desc* b_desc, int64_t count, int max_m, int max_n,
                        void* scratch, size_t scratch_bytes)
{
    constexpr int IB = PanelCfg<T>::IB;
    if (scratch_bytes < trsm_scratch_bytes<T>(na, count)) return TE_ERR_ARG;
    int st = configure_panel_kernel<T>(device);
    if (st != TE_OK) return st;
    const int nblk = (na + IB - 1) / IB;
    T* inv = (T*)scratch;
    size_t off = ((size_t)nblk * IB * IB * sizeof(T) + 255) & ~(size_t)255;
    te_tile_desc* upd = (te_tile_desc*)((char*)scratch + off);
    te_tile_desc* mul = upd + (size_t)nblk * count;

    // 1. inverses of the diagonal blocks (one CTA each); info is unused in this mode
    potf2_trtri_kernel<T, IB><<<nblk, PANEL_THREADS, panel_smem_bytes<T>(), stream>>>(
        const_cast<T*>(A), lda, uplo == TE_UPLO_UPPER, inv, nullptr, 0, 0,
        diag == TE_DIAG_UNIT, (int64_t)IB + (int64_t)IB * lda, (int64_t)IB * IB, na);
    te::count_launch();
    TE_CUDA_CHECK(cudaGetLastError());

    // 2. descriptors of every step
    TrsmPlan p;
    p.b = b_desc; p.upd = upd; p.mul = mul; p.A = A; p.inv = inv; p.count = count;
    p.lda = lda; p.na = na; p.nblk = nblk; p.IB = IB; p.esize = (int)sizeof(T);
    p.right = (side == TE_SIDE_RIGHT);
    p.op_is_n = (opa == TE_OP_N);
    p.eff_lower = ((uplo == TE_UPLO_LOWER) == (opa == TE_OP_N));
    const int64_t total = (int64_t)nblk * count;
    trsm_build_descs_kernel<<<(unsigned)((total + 127) / 128), 128, 0, stream>>>(p);
    te::count_launch();
    TE_CUDA_CHECK(cudaGetLastError());

    // 3. block substitution, two GEMM launches per diagonal block
    const double one[2] = {1.0, 0.0}, zero[2] = {0.0, 0.0}, mone[2] = {-1.0, 0.0};
    for (int s = 0; s < nblk; ++s) {
        const bool ascending = p.right ? !p.eff_lower : p.eff_lower;
        const int blk = ascending ? s : (nblk - 1 - s);
        const int ib = (na - blk * IB < IB) ? na - blk * IB : IB;
        const int mm = p.right ? max_m : ib;
        const int nn = p.right ? ib : max_n;
        if (s > 0) {
            // right: (X, op(A) block);  left: (op(A) block, X)
            st = gemm_batched_any(device, stream, dtype, p.right ? TE_OP_N : opa,
                                  p.right ? opa : TE_OP_N, mone, alpha,
                                  upd + (size_t)s * count, count, mm, nn);
            if (st != TE_OK) return st;
        }
        // in place: X(:, blk) <- T * inv (right: C aliases A), X(blk, :) <- inv * T (left: C
        // aliases B, safe in any shape because a CTA reads only its own columns of B)
        st = gemm_batched_any(device, stream, dtype, p.right ? TE_OP_N : opa,
                              p.right ? opa : TE_OP_N, (s == 0) ? alpha : (const void*)one, zero,
                              mul + (size_t)s * count, count, mm, nn, p.right != 0);
        if (st != TE_OK) return st;
    }
    return TE_OK;
}

// ------------------------------- tile LU without pivoting --------------------
// Reference: src/internal/internal_getrf_nopiv.cc (Tile_getrf_nopiv.hh: recursive LU of the
// diagonal tile on the host).  Here, like the tile Cholesky: diagonal blocks of IB columns are
// factored by one CTA in shared memory, everything else is GEMM on the DMMA kernel
// (L21 = A21 inv(U11), U12 = inv(L11) A12, A22 -= L21 U12, the inverses by the trtri part of
// potf2_trtri_kernel).
__device__ __forceinline__ double recip_(double a) { return 1.0 / a; }
__device__ __forceinline__ zd recip_(zd a)
{
    // Smith's formula: no overflow of |a|^2
    if (fabs(a.x) >= fabs(a.y)) {
        const double r = a.y / a.x, d = a.x + a.y * r;
        return make_zd(1.0 / d, -r / d);
    }
    const double r = a.x / a.y, d = a.x * r + a.y;
    return make_zd(r / d, -1.0 / d);
}
__device__ __forceinline__ bool is_zero_(double a) { return a == 0.0; }
__device__ __forceinline__ bool is_zero_(zd a) { return a.x == 0.0 && a.y == 0.0; }

// One CTA: A (n x n, n <= IB) <- L \ U with unit-diagonal L, right-looking, no pivoting.
// A zero pivot at local column j sets *info = info_base + j + 1 (first one wins) and that
// column is left unscaled, as LAPACK's unblocked getrf does.
template <typename T, int IB>
__global__ void __launch_bounds__(PANEL_THREADS)
lu_nopiv_block_kernel(T* A, int lda, int n, int* info, int info_base)
{
    extern __shared__ __align__(16) unsigned char panel_smem[];
    T* S = reinterpret_cast<T*>(panel_smem);               // column-major, leading dimension LD
    constexpr int LD = IB + 1;
    const int tid = threadIdx.x;
    for (int idx = tid; idx < n * n; idx += PANEL_THREADS) {
        const int r = idx % n, c = idx / n;
        S[r + c * LD] = A[r + (int64_t)c * lda];
    }
    __syncthreads();
    for (int k = 0; k < n; ++k) {
        const T pivot = S[k + k * LD];
        const bool singular = is_zero_(pivot);
        if (singular) {
            if (tid == 0 && *info == 0) *info = info_base + k + 1;
        }
        else {
            const T rp = recip_(pivot);
            for (int i = k + 1 + tid; i < n; i += PANEL_THREADS)
                S[i + k * LD] = S[i + k * LD] * rp;
        }
        __syncthreads();
        const int r = n - k - 1;
        for (int idx = tid; idx < r * r; idx += PANEL_THREADS) {
            const int i = k + 1 + idx % r, j = k + 1 + idx / r;
            S[i + j * LD] = S[i + j * LD] - S[i + k * LD] * S[k + j * LD];
        }
        __syncthreads();
    }
    for (int idx = tid; idx < n * n; idx += PANEL_THREADS) {
        const int r = idx % n, c = idx / n;
        A[r + (int64_t)c * lda] = S[r + c * LD];
    }
}

template <typename T>
static int getrf_nopiv_tile(int device, cudaStream_t stream, int dtype, int m, int n, T* A, int lda,
                            int* info, T* scratch)
{
    constexpr int IB = PanelCfg<T>::IB;
    int st = configure_panel_kernel<T>(device);
    if (st != TE_OK) return st;
    static bool lu_done[64] = {};
    if (!(device < 64 && lu_done[device])) {
        TE_CUDA_CHECK(cudaFuncSetAttribute(lu_nopiv_block_kernel<T, IB>,
                                           cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           panel_smem_bytes<T>()));
        if (device < 64) lu_done[device] = true;
    }
    TE_CUDA_CHECK(cudaMemsetAsync(info, 0, sizeof(int), stream));
    T* invL = scratch;
    T* invU = scratch + (size_t)IB * IB;
    const double one[2] = {1.0, 0.0}, zero[2] = {0.0, 0.0}, mone[2] = {-1.0, 0.0};
    const int mn = m < n ? m : n;
    for (int jb = 0; jb < mn; jb += IB) {
        const int ib = (mn - jb < IB) ? mn - jb : IB;
        T* A11 = A + jb + (int64_t)jb * lda;
        lu_nopiv_block_kernel<T, IB><<<1, PANEL_THREADS, panel_smem_bytes<T>(), stream>>>(
            A11, lda, ib, info, jb);
        te::count_launch();
        TE_CUDA_CHECK(cudaGetLastError());
        const int rem_m = m - jb - ib, rem_n = n - jb - ib;
        if (rem_m <= 0 && rem_n <= 0) break;
        // inverses of the two factors of the block (trtri-only mode, one CTA each)
        if (rem_n > 0) {
            potf2_trtri_kernel<T, IB><<<1, PANEL_THREADS, panel_smem_bytes<T>(), stream>>>(
                A11, lda, 0, invL, nullptr, 0, 0, 1, 0, 0, ib);
            te::count_launch();
        }
        if (rem_m > 0) {
            potf2_trtri_kernel<T, IB><<<1, PANEL_THREADS, panel_smem_bytes<T>(), stream>>>(
                A11, lda, 1, invU, nullptr, 0, 0, 0, 0, 0, ib);
            te::count_launch();
        }
        TE_CUDA_CHECK(cudaGetLastError());
        te_tile_desc d;
        d.reserved = 0;
        d.flags = TE_TILE_GENERAL;
        T* A21 = A + (jb + ib) + (int64_t)jb * lda;
        T* A12 = A + jb + (int64_t)(jb + ib) * lda;
        T* A22 = A + (jb + ib) + (int64_t)(jb + ib) * lda;
        if (rem_m > 0) {
            // L21 <- A21 inv(U11)   (in place: a CTA tile spans all ib columns)
            d.a = A21; d.b = invU; d.c = A21; d.m = rem_m; d.n = ib; d.k = ib;
            d.lda = lda; d.ldb = IB; d.ldc = lda;
            st = gemm_inline_any(device, stream, dtype, TE_OP_N, TE_OP_N, one, zero, &d, 1, true);
            if (st != TE_OK) return st;
        }
        if (rem_n > 0) {
            // U12 <- inv(L11) A12   (in place: a CTA tile spans all ib rows)
            d.a = invL; d.b = A12; d.c = A12; d.m = ib; d.n = rem_n; d.k = ib;
            d.lda = IB; d.ldb = lda; d.ldc = lda;
            st = gemm_inline_any(device, stream, dtype, TE_OP_N, TE_OP_N, one, zero, &d, 1);
            if (st != TE_OK) return st;
        }
        if (rem_m > 0 && rem_n > 0) {
            d.a = A21; d.b = A12; d.c = A22; d.m = rem_m; d.n = rem_n; d.k = ib;
            d.lda = lda; d.ldb = lda; d.ldc = lda;
            st = gemm_inline_any(device, stream, dtype, TE_OP_N, TE_OP_N, mone, one, &d, 1);
            if (st != TE_OK) return st;
        }
    }
    return TE_OK;
}

}  // namespace te

// ------------------------------------ C ABI ------------------------------------
extern "C" size_t te_getrf_nopiv_scratch_bytes(int dtype)
{
    return 2 * (dtype == TE_C64 ? sizeof(te::zd) * 64 * 64 : sizeof(double) * 128 * 128);
}

extern "C" int te_getrf_nopiv_tile(int device, void* stream, int dtype, int m, int n, void* A, int lda,
                                   int* info_dev, void* scratch, size_t scratch_bytes)
{
    if (!A || !info_dev || !scratch || m < 0 || n < 0 || lda < (m > 1 ? m : 1) ||
        scratch_bytes < te_getrf_nopiv_scratch_bytes(dtype))
        return TE_ERR_ARG;
    int st = te::set_device(device);
    if (st != TE_OK) return st;
    if (dtype == TE_F64)
        return te::getrf_nopiv_tile<double>(device, (cudaStream_t)stream, dtype, m, n, (double*)A, lda,
                                            info_dev, (double*)scratch);
    if (dtype == TE_C64)
        return te::getrf_nopiv_tile<te::zd>(device, (cudaStream_t)stream, dtype, m, n, (te::zd*)A, lda,
                                            info_dev, (te::zd*)scratch);
    return TE_ERR_ARG;
}

extern "C" size_t te_potrf_scratch_bytes(int dtype)
{
    return dtype == TE_C64 ? sizeof(te::zd) * 64 * 64 : sizeof(double) * 128 * 128;
}

extern "C" int te_potrf_tile(int device, void* stream, int dtype, int uplo, int n, void* A,
                             int lda, int* info_dev, void* scratch, size_t scratch_bytes)
{
    if (!A || !info_dev || !scratch || n < 0 || lda < (n > 1 ? n : 1) ||
        (uplo != TE_UPLO_LOWER && uplo != TE_UPLO_UPPER) ||
        scratch_bytes < te_potrf_scratch_bytes(dtype))
        return TE_ERR_ARG;
    int st = te::set_device(device);
    if (st != TE_OK) return st;
    if (dtype == TE_F64)
        return te::potrf_tile<double>(device, (cudaStream_t)stream, dtype, uplo, n, (double*)A, lda,
                                      info_dev, (double*)scratch);
    if (dtype == TE_C64)
        return te::potrf_tile<te::zd>(device, (cudaStream_t)stream, dtype, uplo, n, (te::zd*)A,
                                      lda, info_dev, (te::zd*)scratch);
    return TE_ERR_ARG;
}

extern "C" size_t te_trsm_scratch_bytes(int dtype, int na, int64_t count)
{
    return dtype == TE_C64 ? te::trsm_scratch_bytes<te::zd>(na, count)
                           : te::trsm_scratch_bytes<double>(na, count);
}

extern "C" int te_trsm_batched(int device, void* stream, int dtype, int side, int uplo, int opa,
                               int diag, const void* alpha, const void* A, int lda, int na,
                               const te_trsm_desc* b_desc, int64_t count, int32_t max_m,
                               int32_t max_n, void* scratch, size_t scratch_bytes)
{
    if (!alpha || !A || na < 0 || count < 0 || (count > 0 && !b_desc) || !scratch ||
        (side != TE_SIDE_LEFT && side != TE_SIDE_RIGHT) ||
        (uplo != TE_UPLO_LOWER && uplo != TE_UPLO_UPPER) ||
        (opa != TE_OP_N && opa != TE_OP_T && opa != TE_OP_C) ||
        (diag != TE_DIAG_NONUNIT && diag != TE_DIAG_UNIT))
        return TE_ERR_ARG;
    if (count == 0 || na == 0 || max_m == 0 || max_n == 0)
        return TE_OK;
    int st = te::set_device(device);
    if (st != TE_OK) return st;
    if (dtype == TE_F64)
        return te::trsm_batched<double>(device, (cudaStream_t)stream, dtype, side, uplo, opa, diag,
                                        alpha, (const double*)A, lda, na, b_desc, count, max_m,
                                        max_n, scratch, scratch_bytes);
    if (dtype == TE_C64)
        return te::trsm_batched<te::zd>(device, (cudaStream_t)stream, dtype, side, uplo, opa, diag,
                                        alpha, (const te::zd*)A, lda, na, b_desc, count, max_m,
                                        max_n, scratch, scratch_bytes);
    return TE_ERR_ARG;
}

// ---------------------- allocation-free batched trsm ---------------------------
// Same computation as te_trsm_batched, for callers that have no device memory to
// hand over as scratch and are not in a position to obtain any: a solver that has
// already committed its whole HBM budget to the tile pool, or a panel step on the
// critical path that must not run into cudaMalloc (which can stall every stream of
// the device while it runs).  The scratch lives in the library image instead.
//
// te_trsm_static_scratch is that storage: one fixed block per GPU that loads the
// library, held until the process exits.  In host code the symbol below is only a
// handle, not something a kernel may dereference; the device address is looked up
// once per device with cudaGetSymbolAddress and that pointer is what the kernels get.
#define TE_TRSM_STATIC_SCRATCH_BYTES (768u * 1024u)
static __device__ __align__(256) unsigned char te_trsm_static_scratch[TE_TRSM_STATIC_SCRATCH_BYTES];

namespace te {

// the block must hold the scratch of the largest call the two limits allow, for both types
// (keeps a change of te_tile_desc, IB or the limits from silently overrunning it)
template <typename T>
constexpr size_t trsm_static_need()
{
    constexpr size_t nblk = (TE_TRSM_STATIC_MAX_NA + PanelCfg<T>::IB - 1) / PanelCfg<T>::IB;
    return nblk * PanelCfg<T>::IB * PanelCfg<T>::IB * sizeof(T)
           + 2 * nblk * TE_TRSM_STATIC_MAX_COUNT * sizeof(te_tile_desc) + 256;
}
static_assert(trsm_static_need<double>() <= TE_TRSM_STATIC_SCRATCH_BYTES, "static trsm scratch too small (real)");
static_assert(trsm_static_need<zd>() <= TE_TRSM_STATIC_SCRATCH_BYTES, "static trsm scratch too small (complex)");

static std::mutex g_static_scratch_mu;
static void* g_static_scratch_ptr[64] = {};

// Device address of te_trsm_static_scratch on `device` (which must be current).
static int static_scratch_address(int device, void** out)
{
    if (device < 0 || device >= 64) return TE_ERR_ARG;
    std::lock_guard<std::mutex> lock(g_static_scratch_mu);
    if (!g_static_scratch_ptr[device]) {
        void* ptr = nullptr;
        TE_CUDA_CHECK(cudaGetSymbolAddress((void**)&ptr, te_trsm_static_scratch));
        g_static_scratch_ptr[device] = ptr;
    }
    *out = g_static_scratch_ptr[device];
    return TE_OK;
}

}  // namespace te

extern "C" int te_trsm_batched_static(int device, void* stream, int dtype, int side, int uplo,
                                      int opa, int diag, const void* alpha, const void* A, int lda,
                                      int na, const te_trsm_desc* b_desc, int64_t count,
                                      int32_t max_m, int32_t max_n)
{
    if (!alpha || !A || na < 0 || count < 0 || (count > 0 && !b_desc) ||
        (dtype != TE_F64 && dtype != TE_C64) ||
        (side != TE_SIDE_LEFT && side != TE_SIDE_RIGHT) ||
        (uplo != TE_UPLO_LOWER && uplo != TE_UPLO_UPPER) ||
        (opa != TE_OP_N && opa != TE_OP_T && opa != TE_OP_C) ||
        (diag != TE_DIAG_NONUNIT && diag != TE_DIAG_UNIT))
        return TE_ERR_ARG;
    if (na > TE_TRSM_STATIC_MAX_NA || count > TE_TRSM_STATIC_MAX_COUNT) {
        char msg[200];
        snprintf(msg, sizeof(msg),
                 "na = %d, count = %lld exceed the built-in scratch (na <= %d, count <= %d); "
                 "use te_trsm_batched with caller scratch",
                 na, (long long)count, TE_TRSM_STATIC_MAX_NA, TE_TRSM_STATIC_MAX_COUNT);
        te::set_last_error("te_trsm_batched_static", msg);
        return TE_ERR_LIMIT;
    }
    if (count == 0 || na == 0 || max_m == 0 || max_n == 0)
        return TE_OK;
    int st = te::set_device(device);
    if (st != TE_OK) return st;
    void* scratch = nullptr;
    st = te::static_scratch_address(device, &scratch);
    if (st != TE_OK) return st;
    const size_t need = te_trsm_scratch_bytes(dtype, na, count);
    if (need > TE_TRSM_STATIC_SCRATCH_BYTES) {      // cannot happen within the two limits
        te::set_last_error("te_trsm_batched_static", "internal: limits exceed the static block");
        return TE_ERR_LIMIT;
    }
    // Every byte of the block that this call reads (the inverted diagonal blocks and the
    // per-step descriptors) is written by this call first, so whatever an earlier call
    // left behind needs no clearing.
    if (dtype == TE_F64)
        return te::trsm_batched<double>(device, (cudaStream_t)stream, dtype, side, uplo, opa, diag,
                                        alpha, (const double*)A, lda, na, b_desc, count, max_m,
                                        max_n, scratch, TE_TRSM_STATIC_SCRATCH_BYTES);
    return te::trsm_batched<te::zd>(device, (cudaStream_t)stream, dtype, side, uplo, opa, diag,
                                    alpha, (const te::zd*)A, lda, na, b_desc, count, max_m,
                                    max_n, scratch, TE_TRSM_STATIC_SCRATCH_BYTES);
}
