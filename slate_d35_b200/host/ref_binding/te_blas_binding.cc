// te_blas_binding.cc -- the reference-side binding of the tile engine at the level where the
// reference's Target::Devices path reaches the vendor library: the BLAS++ device entry points
//
//   blas::batch::gemm (queue)   blaspp/src/device_batch_gemm.cc   -> cublas{D,Z}gemmBatched
//   blas::batch::herk (queue)   blaspp/src/device_batch_herk.cc   -> loop of cublas{D,Z}syrk/Zherk
//   blas::batch::syrk (queue)   blaspp/src/device_batch_syrk.cc   -> loop of cublas{D,Z}syrk
//   blas::batch::trsm (queue)   blaspp/src/device_batch_trsm.cc   -> cublas{D,Z}trsmBatched
//   blas::herk / blas::syrk (queue)  blaspp/src/device_herk.cc, device_syrk.cc
//       (single-tile fast path of internal::herk / syrk, src/internal/internal_herk.cc:365-389)
//
// This file is compiled IN PLACE OF those six BLAS++ sources (see Makefile); everything above
// it -- slate::gemm / slate::potrf, gemmC.cc, potrf.cc, internal_{gemm,herk,syrk,trsm,potrf}.cc,
// internal_batch.hh, BaseMatrix / MatrixStorage / Tile -- is the reference's own source, compiled
// unmodified.  Each entry point builds tile descriptors from the pointer arrays the reference
// hands over and makes ONE launch of the sm_100a kernels through include/tile_engine.h; no
// cuBLAS routine is called.  FP64 and complex-FP64 are built; single precision throws.
#include "blas/batch_common.hh"
#include "blas/device_blas.hh"

#include "tile_engine.h"

#include <complex>
#include <string>
#include <vector>

namespace {

template <typename T> struct te_scalar { static constexpr bool built = false; static constexpr int dtype = -1; };
template <> struct te_scalar<double> { static constexpr bool built = true; static constexpr int dtype = TE_F64; };
template <> struct te_scalar<std::complex<double>> { static constexpr bool built = true; static constexpr int dtype = TE_C64; };

void te_ok(int status, const char* what)
{
    if (status != TE_OK)
        throw blas::Error((std::string(what) + " failed: " + te_last_error()).c_str(), what);
}

[[noreturn]] void not_built(const char* func)
{
    throw blas::Error("single precision is not built on the B200 tile engine", func);
}

inline int op_code(blas::Op op) { return op == blas::Op::NoTrans ? TE_OP_N : op == blas::Op::Trans ? TE_OP_T : TE_OP_C; }

/// One launch over `d` with uniform ops and scalars.  A handful of tiles travel inside the
/// kernel parameters; larger batches are uploaded into the queue's device workspace first
/// (stream ordered, so the previous launch on this queue has finished reading it).
template <typename T>
void launch(std::vector<te_tile_desc>& d, int opa, int opb, T alpha, T beta, int max_m, int max_n,
            blas::Queue& queue)
{
    if (d.empty()) return;
    const int dtype = te_scalar<T>::dtype;
    if (d.size() <= TE_MAX_INLINE_DESC) {
        te_ok(te_gemm_inline(queue.device(), queue.stream(), dtype, opa, opb, &alpha, &beta, d.data(),
                             (int)d.size()), "te_gemm_inline");
        return;
    }
    const size_t bytes = d.size() * sizeof(te_tile_desc);
    queue.work_ensure_size<char>(bytes);
    te_ok(te_memcpy_async(queue.work(), d.data(), bytes, TE_COPY_H2D, queue.stream()), "te_memcpy_async");
    te_ok(te_gemm_batched(queue.device(), queue.stream(), dtype, opa, opb, &alpha, &beta,
                          (const te_tile_desc*)queue.work(), (int64_t)d.size(), max_m, max_n),
          "te_gemm_batched");
}

//------------------------------------------------------------------------------
template <typename T>
void batch_gemm(blas::Layout layout, std::vector<blas::Op> const& transA, std::vector<blas::Op> const& transB,
                std::vector<int64_t> const& m, std::vector<int64_t> const& n, std::vector<int64_t> const& k,
                std::vector<T> const& alpha, std::vector<T*> const& Aarray, std::vector<int64_t> const& lda,
                std::vector<T*> const& Barray, std::vector<int64_t> const& ldb, std::vector<T> const& beta,
                std::vector<T*> const& Carray, std::vector<int64_t> const& ldc, size_t batch_size,
                std::vector<int64_t>& info, blas::Queue& queue)
{
    if constexpr (!te_scalar<T>::built) not_built(__func__);
    else {
        using blas::batch::extract;
        blas_error_if(layout != blas::Layout::ColMajor && layout != blas::Layout::RowMajor);
        blas_error_if(info.size() != 0 && info.size() != 1 && info.size() != batch_size);
        if (info.size() > 0)
            blas::batch::gemm_check(layout, transA, transB, m, n, k, alpha, Aarray, lda, Barray, ldb, beta,
                                    Carray, ldc, batch_size, info);
        // tiles that share (opA, opB, alpha, beta) go into one launch; SLATE passes one value of
        // each per call (src/internal/internal_gemm.cc:478-509), i.e. one launch
        std::vector<char> done(batch_size, 0);
        for (size_t first = 0; first < batch_size; ++first) {
            if (done[first]) continue;
            const blas::Op ta = extract(transA, first), tb = extract(transB, first);
            const T al = extract(alpha, first), be = extract(beta, first);
            std::vector<te_tile_desc> d;
            int max_m = 0, max_n = 0;
            for (size_t i = first; i < batch_size; ++i) {
                if (done[i] || extract(transA, i) != ta || extract(transB, i) != tb
                    || extract(alpha, i) != al || extract(beta, i) != be)
                    continue;
                done[i] = 1;
                te_tile_desc t{};
                if (layout == blas::Layout::ColMajor) {
                    t.a = extract(Aarray, i);  t.lda = (int32_t)extract(lda, i);
                    t.b = extract(Barray, i);  t.ldb = (int32_t)extract(ldb, i);
                    t.m = (int32_t)extract(m, i);  t.n = (int32_t)extract(n, i);
                }
                else {      // row major: C^T = op(B)^T op(A)^T in column-major terms
                    t.a = extract(Barray, i);  t.lda = (int32_t)extract(ldb, i);
                    t.b = extract(Aarray, i);  t.ldb = (int32_t)extract(lda, i);
                    t.m = (int32_t)extract(n, i);  t.n = (int32_t)extract(m, i);
                }
                t.c = extract(Carray, i);  t.ldc = (int32_t)extract(ldc, i);
                t.k = (int32_t)extract(k, i);
                t.flags = TE_TILE_GENERAL;
                if (t.m == 0 || t.n == 0) continue;
                max_m = std::max(max_m, t.m);
                max_n = std::max(max_n, t.n);
                d.push_back(t);
            }
            if (layout == blas::Layout::ColMajor)
                launch<T>(d, op_code(ta), op_code(tb), al, be, max_m, max_n, queue);
            else
                launch<T>(d, op_code(tb), op_code(ta), al, be, max_m, max_n, queue);
        }
    }
}

//------------------------------------------------------------------------------
/// herk (hermitian == true) and syrk: C = alpha op(A) op(A)^{H|T} + beta C on one triangle,
/// as a tile update with B = A and the triangle mask (TE_TILE_LOWER / UPPER [| HERMITIAN]).
template <typename T, typename scalar_ab>
void batch_rank_k(bool hermitian, blas::Layout layout, std::vector<blas::Uplo> const& uplo,
                  std::vector<blas::Op> const& trans, std::vector<int64_t> const& n, std::vector<int64_t> const& k,
                  std::vector<scalar_ab> const& alpha, std::vector<T*> const& Aarray, std::vector<int64_t> const& lda,
                  std::vector<scalar_ab> const& beta, std::vector<T*> const& Carray, std::vector<int64_t> const& ldc,
                  size_t batch_size, blas::Queue& queue)
{
    if constexpr (!te_scalar<T>::built) not_built(__func__);
    else {
        using blas::batch::extract;
        blas_error_if(layout != blas::Layout::ColMajor);      // SLATE's tiles on this path are ColMajor
        std::vector<char> done(batch_size, 0);
        for (size_t first = 0; first < batch_size; ++first) {
            if (done[first]) continue;
            const blas::Op tr = extract(trans, first);
            const blas::Uplo ul = extract(uplo, first);
            const scalar_ab al = extract(alpha, first), be = extract(beta, first);
            std::vector<te_tile_desc> d;
            int max_n = 0;
            for (size_t i = first; i < batch_size; ++i) {
                if (done[i] || extract(trans, i) != tr || extract(uplo, i) != ul || extract(alpha, i) != al
                    || extract(beta, i) != be)
                    continue;
                done[i] = 1;
                te_tile_desc t{};
                t.a = t.b = extract(Aarray, i);
                t.lda = t.ldb = (int32_t)extract(lda, i);
                t.c = extract(Carray, i);  t.ldc = (int32_t)extract(ldc, i);
                t.m = t.n = (int32_t)extract(n, i);
                t.k = (int32_t)extract(k, i);
                t.flags = (ul == blas::Uplo::Lower ? TE_TILE_LOWER : TE_TILE_UPPER) | (hermitian ? TE_TILE_HERMITIAN : 0);
                if (t.n == 0) continue;
                max_n = std::max(max_n, t.n);
                d.push_back(t);
            }
            // NoTrans: A is n x k, C += A A^H;  (Conj)Trans: A is k x n, C += A^H A
            const int other = hermitian ? TE_OP_C : TE_OP_T;
            const int opa = tr == blas::Op::NoTrans ? TE_OP_N : other;
            const int opb = tr == blas::Op::NoTrans ? other : TE_OP_N;
            launch<T>(d, opa, opb, T(al), T(be), max_n, max_n, queue);
        }
    }
}

//------------------------------------------------------------------------------
/// SLATE's panel solve replicates ONE triangular tile through Aarray
/// (src/internal/internal_batch.hh:251-258): runs of equal A go to te_trsm_batched together.
template <typename T>
void batch_trsm(blas::Layout layout, std::vector<blas::Side> const& side, std::vector<blas::Uplo> const& uplo,
                std::vector<blas::Op> const& trans, std::vector<blas::Diag> const& diag,
                std::vector<int64_t> const& m, std::vector<int64_t> const& n, std::vector<T> const& alpha,
                std::vector<T*> const& Aarray, std::vector<int64_t> const& lda, std::vector<T*> const& Barray,
                std::vector<int64_t> const& ldb, size_t batch_size, blas::Queue& queue)
{
    if constexpr (!te_scalar<T>::built) not_built(__func__);
    else {
        using blas::batch::extract;
        blas_error_if(layout != blas::Layout::ColMajor);
        const int dtype = te_scalar<T>::dtype;
        size_t first = 0;
        while (first < batch_size) {
            const blas::Side sd = extract(side, first);
            const blas::Uplo ul = extract(uplo, first);
            const blas::Op tr = extract(trans, first);
            const blas::Diag dg = extract(diag, first);
            const T al = extract(alpha, first);
            T* const A = extract(Aarray, first);
            const int64_t la = extract(lda, first);
            const int64_t na = sd == blas::Side::Left ? extract(m, first) : extract(n, first);
            std::vector<te_trsm_desc> d;
            int max_m = 0, max_n = 0;
            size_t i = first;
            for (; i < batch_size; ++i) {
                const int64_t na_i = extract(side, i) == blas::Side::Left ? extract(m, i) : extract(n, i);
                if (extract(Aarray, i) != A || extract(lda, i) != la || extract(side, i) != sd
                    || extract(uplo, i) != ul || extract(trans, i) != tr || extract(diag, i) != dg
                    || extract(alpha, i) != al || na_i != na)
                    break;
                te_trsm_desc t{};
                t.b = extract(Barray, i);
                t.m = (int32_t)extract(m, i);
                t.n = (int32_t)extract(n, i);
                t.ldb = (int32_t)extract(ldb, i);
                if (t.m == 0 || t.n == 0) continue;
                max_m = std::max(max_m, t.m);
                max_n = std::max(max_n, t.n);
                d.push_back(t);
            }
            first = i;
            if (d.empty() || na == 0) continue;
            // device workspace of the queue: [descriptors | solver scratch]
            const size_t dbytes = (d.size() * sizeof(te_trsm_desc) + 255) & ~(size_t)255;
            const size_t sbytes = te_trsm_scratch_bytes(dtype, (int)na, (int64_t)d.size());
            queue.work_ensure_size<char>(dbytes + sbytes);
            char* work = (char*)queue.work();
            te_ok(te_memcpy_async(work, d.data(), d.size() * sizeof(te_trsm_desc), TE_COPY_H2D, queue.stream()),
                  "te_memcpy_async");
            te_ok(te_trsm_batched(queue.device(), queue.stream(), dtype, sd == blas::Side::Left ? TE_SIDE_LEFT : TE_SIDE_RIGHT,
                                  ul == blas::Uplo::Lower ? TE_UPLO_LOWER : TE_UPLO_UPPER, op_code(tr),
                                  dg == blas::Diag::Unit ? TE_DIAG_UNIT : TE_DIAG_NONUNIT, &al, A, (int)la, (int)na,
                                  (const te_trsm_desc*)work, (int64_t)d.size(), max_m, max_n, work + dbytes, sbytes),
                  "te_trsm_batched");
        }
    }
}

//------------------------------------------------------------------------------
template <typename T, typename scalar_ab>
void single_rank_k(bool hermitian, blas::Layout layout, blas::Uplo uplo, blas::Op trans, int64_t n, int64_t k,
                   scalar_ab alpha, T const* A, int64_t lda, scalar_ab beta, T* C, int64_t ldc, blas::Queue& queue)
{
    if constexpr (!te_scalar<T>::built) not_built(__func__);
    else {
        std::vector<blas::Uplo> ul{uplo};
        std::vector<blas::Op> tr{trans};
        std::vector<int64_t> nn{n}, kk{k}, la{lda}, lc{ldc};
        std::vector<scalar_ab> al{alpha}, be{beta};
        std::vector<T*> Aa{const_cast<T*>(A)}, Ca{C};
        batch_rank_k<T, scalar_ab>(hermitian, layout, ul, tr, nn, kk, al, Aa, la, be, Ca, lc, 1, queue);
    }
}

}  // namespace

//==============================================================================
// The overload sets BLAS++ declares in blas/device_blas.hh.
namespace blas {

#define TE_SINGLE_RANK_K(name, hermitian, T, AB)                                                          \
    void name(blas::Layout layout, blas::Uplo uplo, blas::Op trans, int64_t n, int64_t k, AB alpha,        \
              T const* A, int64_t lda, AB beta, T* C, int64_t ldc, blas::Queue& queue)                     \
    { single_rank_k<T, AB>(hermitian, layout, uplo, trans, n, k, alpha, A, lda, beta, C, ldc, queue); }

TE_SINGLE_RANK_K(herk, true, float, float)
TE_SINGLE_RANK_K(herk, true, double, double)
TE_SINGLE_RANK_K(herk, true, std::complex<float>, float)
TE_SINGLE_RANK_K(herk, true, std::complex<double>, double)
TE_SINGLE_RANK_K(syrk, false, float, float)
TE_SINGLE_RANK_K(syrk, false, double, double)
TE_SINGLE_RANK_K(syrk, false, std::complex<float>, std::complex<float>)
TE_SINGLE_RANK_K(syrk, false, std::complex<double>, std::complex<double>)

namespace batch {

#define TE_BATCH_GEMM(T)                                                                                   \
    void gemm(blas::Layout layout, std::vector<blas::Op> const& transA, std::vector<blas::Op> const& transB, \
              std::vector<int64_t> const& m, std::vector<int64_t> const& n, std::vector<int64_t> const& k, \
              std::vector<T> const& alpha, std::vector<T*> const& Aarray, std::vector<int64_t> const& lda,  \
              std::vector<T*> const& Barray, std::vector<int64_t> const& ldb, std::vector<T> const& beta,   \
              std::vector<T*> const& Carray, std::vector<int64_t> const& ldc, size_t batch_size,            \
              std::vector<int64_t>& info, blas::Queue& queue)                                              \
    { batch_gemm<T>(layout, transA, transB, m, n, k, alpha, Aarray, lda, Barray, ldb, beta, Carray, ldc,   \
                    batch_size, info, queue); }

TE_BATCH_GEMM(float)
TE_BATCH_GEMM(double)
TE_BATCH_GEMM(std::complex<float>)
TE_BATCH_GEMM(std::complex<double>)

#define TE_BATCH_RANK_K(name, hermitian, T, AB)                                                            \
    void name(blas::Layout layout, std::vector<blas::Uplo> const& uplo, std::vector<blas::Op> const& trans, \
              std::vector<int64_t> const& n, std::vector<int64_t> const& k, std::vector<AB> const& alpha,   \
              std::vector<T*> const& Aarray, std::vector<int64_t> const& lda, std::vector<AB> const& beta,  \
              std::vector<T*> const& Carray, std::vector<int64_t> const& ldc, size_t batch_size,            \
              std::vector<int64_t>& info, blas::Queue& queue)                                              \
    {                                                                                                      \
        blas_error_if(info.size() != 0 && info.size() != 1 && info.size() != batch_size);                  \
        batch_rank_k<T, AB>(hermitian, layout, uplo, trans, n, k, alpha, Aarray, lda, beta, Carray, ldc,   \
                            batch_size, queue);                                                            \
    }

TE_BATCH_RANK_K(herk, true, float, float)
TE_BATCH_RANK_K(herk, true, double, double)
TE_BATCH_RANK_K(herk, true, std::complex<float>, float)
TE_BATCH_RANK_K(herk, true, std::complex<double>, double)
TE_BATCH_RANK_K(syrk, false, float, float)
TE_BATCH_RANK_K(syrk, false, double, double)
TE_BATCH_RANK_K(syrk, false, std::complex<float>, std::complex<float>)
TE_BATCH_RANK_K(syrk, false, std::complex<double>, std::complex<double>)

#define TE_BATCH_TRSM(T)                                                                                   \
    void trsm(blas::Layout layout, std::vector<blas::Side> const& side, std::vector<blas::Uplo> const& uplo, \
              std::vector<blas::Op> const& trans, std::vector<blas::Diag> const& diag,                      \
              std::vector<int64_t> const& m, std::vector<int64_t> const& n, std::vector<T> const& alpha,    \
              std::vector<T*> const& Aarray, std::vector<int64_t> const& lda, std::vector<T*> const& Barray, \
              std::vector<int64_t> const& ldb, size_t batch_size, std::vector<int64_t>& info,               \
              blas::Queue& queue)                                                                          \
    {                                                                                                      \
        blas_error_if(info.size() != 0 && info.size() != 1 && info.size() != batch_size);                  \
        batch_trsm<T>(layout, side, uplo, trans, diag, m, n, alpha, Aarray, lda, Barray, ldb, batch_size,  \
                      queue);                                                                              \
    }

TE_BATCH_TRSM(float)
TE_BATCH_TRSM(double)
TE_BATCH_TRSM(std::complex<float>)
TE_BATCH_TRSM(std::complex<double>)

}  // namespace batch
}  // namespace blas
