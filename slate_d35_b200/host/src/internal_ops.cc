// internal_ops.cc -- Target::Devices overloads of the internal Level-3 operations.
// See slate/internal.hh.  Reference files: src/internal/internal_gemm.cc:355-518,
// internal_herk.cc:351-535, internal_syrk.cc:346-527, internal_trsm.cc:132-261,
// internal_potrf.cc:57-81, internal_batch.hh:169-394.
#include "slate/internal.hh"

#include <cstring>
#include <limits>

namespace slate {
namespace internal {

namespace {

// Tiles are handed to the engine in 8 x 8 super-blocks of the local (row, column) index
// space: the CTAs that run concurrently then share at most 8 A tiles and 8 B tiles
// (32 MiB at nb = 512), which stay resident in the 126 MB L2 while 64 C tiles stream.
const int64_t SuperBlock = 8;

struct DescBatch {
    std::vector<te_tile_desc> descs;
    int32_t max_m = 0, max_n = 0;
    void add(const void* a, int64_t lda, const void* b, int64_t ldb, void* c, int64_t ldc,
             int64_t m, int64_t n, int64_t k, int flags)
    {
        te_tile_desc d;
        d.a = a;  d.b = b;  d.c = c;
        d.m = (int32_t)m;  d.n = (int32_t)n;  d.k = (int32_t)k;
        d.lda = (int32_t)lda;  d.ldb = (int32_t)ldb;  d.ldc = (int32_t)ldc;
        d.flags = flags;  d.reserved = 0;
        descs.push_back(d);
        if (d.m > max_m) max_m = d.m;
        if (d.n > max_n) max_n = d.n;
    }
};

template <typename scalar_t>
void launch_batch(DescBatch& batch, int q, Op opa, Op opb, scalar_t alpha, scalar_t beta)
{
    if (batch.descs.empty()) return;
    Engine& eng = Engine::instance();
    const int dtype = te_dtype_of<scalar_t>::value;
    void* stream = eng.compute_stream(q);
    eng.trace_begin(q, "gemm_tiles", (int64_t)batch.descs.size());
    if ((int)batch.descs.size() <= TE_MAX_INLINE_DESC) {
        // few tiles (panel / lookahead calls): descriptors ride in the kernel parameters
        slate_te_call(te_gemm_inline(eng.device(), stream, dtype, int(opa), int(opb), &alpha, &beta,
                                     batch.descs.data(), (int)batch.descs.size()));
    }
    else {
        size_t bytes = batch.descs.size() * sizeof(te_tile_desc);
        Engine::DescSlot& slot = eng.acquire_desc(q, bytes);
        std::memcpy(slot.host, batch.descs.data(), bytes);
        eng.upload_desc(q, slot, bytes);
        slate_te_call(te_gemm_batched(eng.device(), stream, dtype, int(opa), int(opb), &alpha, &beta,
                                      (const te_tile_desc*)slot.dev, (int64_t)batch.descs.size(),
                                      batch.max_m, batch.max_n));
        eng.retire_desc(q, slot);
    }
    eng.trace_end(q);
    eng.mark_queue(q);
}

/// op(X)^H expressed as an op on the stored tile; ConjNoTrans is not representable.
template <typename scalar_t>
Op conj_trans_of(Op op, bool* ok)
{
    *ok = true;
    if (op == Op::NoTrans) return Op::ConjTrans;
    if (op == Op::ConjTrans || !is_complex<scalar_t>::value) return Op::NoTrans;
    *ok = false;
    return Op::NoTrans;
}
template <typename scalar_t>
Op trans_of(Op op, bool* ok)
{
    *ok = true;
    if (op == Op::NoTrans) return Op::Trans;
    if (op == Op::Trans || !is_complex<scalar_t>::value) return Op::NoTrans;
    *ok = false;
    return Op::NoTrans;
}

}  // namespace

namespace {
/// Fetches the local tiles of one block outer product onto the device (queue q) and appends
/// their descriptors, in 8 x 8 super-blocks, to `batch`.
template <typename scalar_t>
void append_gemm_tiles(DescBatch& batch, Matrix<scalar_t>& A, Matrix<scalar_t>& B, Matrix<scalar_t>& C,
                       Layout layout, int q)
{
    const int device = 0;
    // local tiles of C and the A / B tiles they need
    std::vector<int64_t> rows, cols;
    std::vector<char> row_used(C.mt(), 0), col_used(C.nt(), 0);
    std::set<ij_tuple> A_set, B_set, C_set;
    for (int64_t j = 0; j < C.nt(); ++j)
        for (int64_t i = 0; i < C.mt(); ++i)
            if (C.tileIsLocal(i, j)) {
                row_used[i] = col_used[j] = 1;
                C_set.insert({i, j});
            }
    for (int64_t i = 0; i < C.mt(); ++i)
        if (row_used[i]) { rows.push_back(i); A_set.insert({i, 0}); }
    for (int64_t j = 0; j < C.nt(); ++j)
        if (col_used[j]) { cols.push_back(j); B_set.insert({0, j}); }
    if (C_set.empty()) return;

    A.tileGetForReading(A_set, device, LayoutConvert(layout), q);
    B.tileGetForReading(B_set, device, LayoutConvert(layout), q);
    C.tileGetForWriting(C_set, device, LayoutConvert(layout), q);

    const int64_t k = A.tileNb(0);
    batch.descs.reserve(batch.descs.size() + C_set.size());
    const bool swap_ab = (C.op() != Op::NoTrans);
    for (size_t jb = 0; jb < cols.size(); jb += SuperBlock)
        for (size_t ib = 0; ib < rows.size(); ib += SuperBlock)
            for (size_t jj = jb; jj < std::min(cols.size(), jb + (size_t)SuperBlock); ++jj)
                for (size_t ii = ib; ii < std::min(rows.size(), ib + (size_t)SuperBlock); ++ii) {
                    const int64_t i = rows[ii], j = cols[jj];
                    if (!C.tileIsLocal(i, j)) continue;
                    auto Ai = A(i, 0, device);
                    auto Bj = B(0, j, device);
                    auto Cij = C(i, j, device);
                    if (!swap_ab)
                        batch.add(Ai.data(), Ai.stride(), Bj.data(), Bj.stride(), Cij.data(),
                                  Cij.stride(), C.tileMb(i), C.tileNb(j), k, TE_TILE_GENERAL);
                    else   // the stored tile is op(C(i,j)): (n x m) = op'(B_j) op'(A_i)
                        batch.add(Bj.data(), Bj.stride(), Ai.data(), Ai.stride(), Cij.data(),
                                  Cij.stride(), C.tileNb(j), C.tileMb(i), k, TE_TILE_GENERAL);
                }
}
}  // namespace

//------------------------------------------------------------------------------
template <typename scalar_t>
void gemm(TargetType<Target::Devices>, scalar_t alpha, Matrix<scalar_t>& A, Matrix<scalar_t>& B,
          scalar_t beta, Matrix<scalar_t>& C, Layout layout, int priority, int64_t queue_index)
{
    slate_assert(layout == Layout::ColMajor);
    slate_assert(C.mt() > 0 && C.nt() > 0);
    slate_assert(A.nt() == 1 && B.mt() == 1);
    slate_assert(A.mt() == C.mt() && B.nt() == C.nt());
    slate_assert(C.num_devices() > 0);
    const int q = (int)queue_index;

    // src/internal/internal_gemm.cc:383-414: if op(C) is not NoTrans, compute the
    // transposed product on the stored tiles.
    Op opA = A.op(), opB = B.op();
    if (C.op() != Op::NoTrans) {
        if (opA == Op::NoTrans) opA = C.op();
        else if (A.op() == C.op() || C.is_real) opA = Op::NoTrans;
        else slate_error("internal::gemm: ConjNoTrans is not supported");
        if (opB == Op::NoTrans) opB = C.op();
        else if (B.op() == C.op() || C.is_real) opB = Op::NoTrans;
        else slate_error("internal::gemm: ConjNoTrans is not supported");
    }
    if (C.op() == Op::ConjTrans) {
        alpha = conj(alpha);
        beta = conj(beta);
    }

    DescBatch batch;
    append_gemm_tiles(batch, A, B, C, layout, q);
    const bool swap_ab = (C.op() != Op::NoTrans);
    if (swap_ab) std::swap(opA, opB);
    launch_batch<scalar_t>(batch, q, opA, opB, alpha, beta);
}

//------------------------------------------------------------------------------
/// Several block outer products C_p += alpha A_p B_p (+ beta C_p) in ONE launch: the parts may
/// belong to different block steps k (a skewed schedule over column slabs of C, slate::gemm with
/// host-origin C).  All parts share op(A), op(B) and the scalars; C must be NoTrans.
template <typename scalar_t>
void gemm_multi(TargetType<Target::Devices>, scalar_t alpha, std::vector<GemmPart<scalar_t>>& parts,
                scalar_t beta, Layout layout, int64_t queue_index)
{
    if (parts.empty()) return;
    const int q = (int)queue_index;
    DescBatch batch;
    for (auto& part : parts) {
        slate_assert(part.C.op() == Op::NoTrans);
        slate_assert(part.A.op() == parts[0].A.op() && part.B.op() == parts[0].B.op());
        slate_assert(part.A.nt() == 1 && part.B.mt() == 1);
        append_gemm_tiles(batch, part.A, part.B, part.C, layout, q);
    }
    launch_batch<scalar_t>(batch, q, parts[0].A.op(), parts[0].B.op(), alpha, beta);
}

//------------------------------------------------------------------------------
/// C(:, 0) = alpha A B(:, 0) + beta C(:, 0) with the work placed where the tiles of A are
/// (src/internal/internal_gemmA.cc:224-560): every rank multiplies its local A(i, j) into
/// C(i, 0) -- a workspace tile, zeroed first, when C(i, 0) lives elsewhere -- and the caller
/// then sums the partial tiles with listReduce.  One batched launch per block column j of A
/// (the rows of one launch write different C tiles; consecutive launches accumulate in stream
/// order).  beta is applied up front by a scaling pass over the C tiles this rank touches
/// (internal_gemmA.cc:286-330 does so for untouched tiles only and passes beta to the first
/// product; one launch has one beta here, so all tiles take the same route).
template <typename scalar_t>
void gemmA(TargetType<Target::Devices>, scalar_t alpha, Matrix<scalar_t>& A, Matrix<scalar_t>& B,
           scalar_t beta, Matrix<scalar_t>& C, Layout layout, int priority, int64_t queue_index)
{
    slate_assert(layout == Layout::ColMajor);
    slate_assert(A.mt() == C.mt() && A.nt() == B.mt());
    slate_assert(B.nt() == 1 && C.nt() == 1);
    slate_assert(C.num_devices() > 0);
    Engine& eng = Engine::instance();
    const int device = 0;
    const int q = (int)queue_index;
    const int dtype = te_dtype_of<scalar_t>::value;
    const scalar_t one = 1.0, zero = 0.0;

    Op opA = A.op(), opB = B.op();
    if (C.op() != Op::NoTrans) {
        if (opA == Op::NoTrans) opA = C.op();
        else if (A.op() == C.op() || C.is_real) opA = Op::NoTrans;
        else slate_error("internal::gemmA: ConjNoTrans is not supported");
        if (opB == Op::NoTrans) opB = C.op();
        else if (B.op() == C.op() || C.is_real) opB = Op::NoTrans;
        else slate_error("internal::gemmA: ConjNoTrans is not supported");
    }
    if (C.op() == Op::ConjTrans) {
        alpha = conj(alpha);
        beta = conj(beta);
    }

    // rows with local work, the B tiles they need, C tiles to create / scale
    std::set<ij_tuple> A_set, B_set, C_local_set;
    std::vector<int64_t> new_rows, scale_rows;
    for (int64_t i = 0; i < A.mt(); ++i) {
        bool touched = false;
        for (int64_t j = 0; j < A.nt(); ++j)
            if (A.tileIsLocal(i, j)) {
                A_set.insert({i, j});
                B_set.insert({j, 0});
                touched = true;
            }
        if (C.tileIsLocal(i, 0)) {
            C_local_set.insert({i, 0});
            scale_rows.push_back(i);
        }
        else if (touched)
            new_rows.push_back(i);
    }
    if (A_set.empty() && C_local_set.empty()) return;

    A.tileGetForReading(A_set, device, LayoutConvert(layout), q);
    B.tileGetForReading(B_set, device, LayoutConvert(layout), q);
    C.tileGetForWriting(C_local_set, device, LayoutConvert(layout), q);
    void* stream = eng.compute_stream(q);
    for (int64_t i : new_rows) {
        // partial sum for a C tile owned elsewhere
        slate_assert(!C.tileExists(i, 0, device));
        C.tileInsertWorkspace(i, 0, device);
        C.tileModified(i, 0, device, q);
        auto T = C(i, 0, device);
        slate_te_call(te_tile_scale(eng.device(), stream, dtype, (int)T.mb(), (int)T.nb(), &zero,
                                    T.data(), T.stride()));
    }
    if (beta != one)
        for (int64_t i : scale_rows) {
            auto T = C(i, 0, device);
            slate_te_call(te_tile_scale(eng.device(), stream, dtype, (int)T.mb(), (int)T.nb(), &beta,
                                        T.data(), T.stride()));
        }

    const bool swap_ab = (C.op() != Op::NoTrans);
    if (swap_ab) std::swap(opA, opB);
    for (int64_t j = 0; j < A.nt(); ++j) {
        DescBatch batch;
        for (int64_t i = 0; i < A.mt(); ++i) {
            if (!A.tileIsLocal(i, j)) continue;
            auto Aij = A(i, j, device);
            auto Bj = B(j, 0, device);
            auto Ci = C(i, 0, device);
            if (!swap_ab)
                batch.add(Aij.data(), Aij.stride(), Bj.data(), Bj.stride(), Ci.data(), Ci.stride(),
                          C.tileMb(i), C.tileNb(0), A.tileNb(j), TE_TILE_GENERAL);
            else
                batch.add(Bj.data(), Bj.stride(), Aij.data(), Aij.stride(), Ci.data(), Ci.stride(),
                          C.tileNb(0), C.tileMb(i), A.tileNb(j), TE_TILE_GENERAL);
        }
        launch_batch<scalar_t>(batch, q, opA, opB, alpha, one);
    }
    // the C tiles are final only after the last launch: stamp them with the mark that follows
    // it (tileGetForWriting stamped them with the first one)
    for (int64_t i : scale_rows) C.tileModified(i, 0, device, q);
    for (int64_t i : new_rows) C.tileModified(i, 0, device, q);
    eng.mark_queue(q);
}

//------------------------------------------------------------------------------
namespace {
/// Shared body of hemm (hermitian = true) and symm: C(0, :) = alpha A B(0, :) + beta C(0, :)
/// with A one diagonal tile of which only the uplo triangle is stored.  The reference runs a
/// tile hemm on the host here even for Target::Devices (src/internal/internal_hemm.cc, called
/// from src/hemmC.cc:146-151); on this engine the tile is copied into a pool block, its other
/// triangle filled in (te_tile_symmetrize), and the copy is an ordinary operand of the batched
/// GEMM kernel.
template <typename scalar_t, typename AMatrix>
void diag_tile_mm(bool hermitian, Side side, scalar_t alpha, AMatrix& A, Matrix<scalar_t>& B,
                  scalar_t beta, Matrix<scalar_t>& C, int queue_index)
{
    slate_assert(side == Side::Left);               // drivers turn Right into Left by views
    slate_assert(A.mt() == 1 && A.nt() == 1);
    slate_assert(B.mt() == 1 && C.mt() == 1 && B.nt() == C.nt());
    slate_assert(C.num_devices() > 0);
    Engine& eng = Engine::instance();
    const int device = 0;
    const int q = queue_index;
    const int dtype = te_dtype_of<scalar_t>::value;

    // op(A) of the filled tile: A itself if the view is untransposed, and also if the view's op
    // matches the matrix kind (A^H = A, A^T = A); anything else would be conj(A)
    const Op same_op = hermitian ? Op::ConjTrans : Op::Trans;
    Op opA = Op::NoTrans;
    if (A.op() != Op::NoTrans && A.op() != same_op && is_complex<scalar_t>::value)
        opA = (A.op() == Op::Trans) ? Op::Trans : Op::ConjTrans;   // conj(A) = (A^H)^T = (A^T)^H
    Op opB = B.op();
    if (C.op() != Op::NoTrans) {
        // stored P = op(C): P = op(B) op(A), as in internal::gemm
        if (opA == Op::NoTrans) opA = (C.op() == same_op || C.is_real) ? Op::NoTrans : C.op();
        else if (opA == C.op() || C.is_real) opA = Op::NoTrans;
        else slate_error("internal::hemm/symm: ConjNoTrans is not supported");
        if (opB == Op::NoTrans) opB = C.op();
        else if (B.op() == C.op() || C.is_real) opB = Op::NoTrans;
        else slate_error("internal::hemm/symm: ConjNoTrans is not supported");
    }
    if (C.op() == Op::ConjTrans) {
        alpha = conj(alpha);
        beta = conj(beta);
    }

    std::set<ij_tuple> A_set{{0, 0}}, B_set, C_set;
    for (int64_t j = 0; j < C.nt(); ++j)
        if (C.tileIsLocal(0, j)) {
            C_set.insert({0, j});
            B_set.insert({0, j});
        }
    if (C_set.empty()) return;
    A.tileGetForReading(A_set, device, LayoutConvert::ColMajor, q);
    B.tileGetForReading(B_set, device, LayoutConvert::ColMajor, q);
    C.tileGetForWriting(C_set, device, LayoutConvert::ColMajor, q);

    // full copy of the diagonal tile
    auto A00 = A(0, 0, device);
    const int64_t n = A00.mb();
    slate_assert(A00.nb() == n);
    void* stream = eng.compute_stream(q);
    scalar_t* full = (scalar_t*)A.storage()->allocScratchBlock();
    slate_te_call(te_memcpy2d_async(full, sizeof(scalar_t) * n, A00.data(), sizeof(scalar_t) * A00.stride(),
                                    sizeof(scalar_t) * n, n, TE_COPY_D2D, stream));
    slate_te_call(te_tile_symmetrize(eng.device(), stream, dtype,
                                     A00.uploPhysical() == Uplo::Lower ? TE_UPLO_LOWER : TE_UPLO_UPPER,
                                     hermitian ? 1 : 0, (int)n, full, n));

    DescBatch batch;
    const bool swap_ab = (C.op() != Op::NoTrans);
    for (int64_t j = 0; j < C.nt(); ++j) {
        if (!C.tileIsLocal(0, j)) continue;
        auto Bj = B(0, j, device);
        auto Cj = C(0, j, device);
        if (!swap_ab)
            batch.add(full, n, Bj.data(), Bj.stride(), Cj.data(), Cj.stride(), C.tileMb(0),
                      C.tileNb(j), n, TE_TILE_GENERAL);
        else
            batch.add(Bj.data(), Bj.stride(), full, n, Cj.data(), Cj.stride(), C.tileNb(j),
                      C.tileMb(0), n, TE_TILE_GENERAL);
    }
    if (swap_ab) std::swap(opA, opB);
    launch_batch<scalar_t>(batch, q, opA, opB, alpha, beta);
    A.storage()->freeDeferred(full);
    A.storage()->flushErased();
}
}  // namespace

template <typename scalar_t>
void hemm(TargetType<Target::Devices>, Side side, scalar_t alpha, HermitianMatrix<scalar_t>& A,
          Matrix<scalar_t>& B, scalar_t beta, Matrix<scalar_t>& C, int priority, int queue_index)
{
    diag_tile_mm<scalar_t>(true, side, alpha, A, B, beta, C, queue_index);
}

template <typename scalar_t>
void symm(TargetType<Target::Devices>, Side side, scalar_t alpha, SymmetricMatrix<scalar_t>& A,
          Matrix<scalar_t>& B, scalar_t beta, Matrix<scalar_t>& C, int priority, int queue_index)
{
    diag_tile_mm<scalar_t>(false, side, alpha, A, B, beta, C, queue_index);
}

//------------------------------------------------------------------------------
namespace {
/// Shared body of herk (conjugate = true) and syrk: lower-logical C, A one block column.
template <typename scalar_t, typename CMatrix>
void rank_k_update(bool conjugate, scalar_t alpha, Matrix<scalar_t>& A, scalar_t beta, CMatrix& C,
                   int queue_index, Layout layout)
{
    slate_assert(layout == Layout::ColMajor);
    slate_assert(C.num_devices() > 0);
    slate_assert(A.nt() == 1 && A.mt() == C.mt() && C.mt() == C.nt());
    slate_assert(C.uploLogical() == Uplo::Lower);      // drivers flip upper to lower first
    const int device = 0;
    const int q = queue_index;

    // C(i,j) = alpha op(A_i) op(A_j)^{H|T} + beta C(i,j), i >= j.
    bool ok = true;
    Op opa = A.op();
    Op opb = conjugate ? conj_trans_of<scalar_t>(A.op(), &ok) : trans_of<scalar_t>(A.op(), &ok);
    if (!ok) slate_error("internal::herk/syrk: op(A) mixes Trans and ConjTrans");
    // stored C tiles are transposed when op(C) != NoTrans:
    //   P = C(i,j)^{H|T} = alpha op(A_j) op(A_i)^{H|T} + beta P
    const bool swap_ij = (C.op() != Op::NoTrans);
    if (swap_ij && is_complex<scalar_t>::value && (C.op() == Op::ConjTrans) != conjugate)
        slate_error("internal::herk/syrk: op(C) does not match the update");
    if (swap_ij && conjugate) { alpha = conj(alpha); beta = conj(beta); }

    std::set<ij_tuple> A_set, C_set;
    for (int64_t j = 0; j < C.nt(); ++j)
        for (int64_t i = j; i < C.mt(); ++i)
            if (C.tileIsLocal(i, j)) {
                C_set.insert({i, j});
                A_set.insert({i, 0});
                A_set.insert({j, 0});
            }
    if (C_set.empty()) return;
    A.tileGetForReading(A_set, device, LayoutConvert(layout), q);
    C.tileGetForWriting(C_set, device, LayoutConvert(layout), q);

    const int64_t k = A.tileNb(0);
    const int diag_flag = (C.uploPhysical() == Uplo::Lower ? TE_TILE_LOWER : TE_TILE_UPPER)
                          | (conjugate ? TE_TILE_HERMITIAN : 0);
    DescBatch batch;
    batch.descs.reserve(C_set.size());
    const int64_t nt = C.nt();
    for (int64_t jb = 0; jb < nt; jb += SuperBlock)
        for (int64_t ib = jb; ib < nt; ib += SuperBlock)
            for (int64_t j = jb; j < std::min(nt, jb + SuperBlock); ++j)
                for (int64_t i = std::max(ib, j); i < std::min(nt, ib + SuperBlock); ++i) {
                    if (!C.tileIsLocal(i, j)) continue;
                    auto Ai = A(i, 0, device);
                    auto Aj = A(j, 0, device);
                    auto Cij = C(i, j, device);
                    const int flags = (i == j) ? diag_flag : TE_TILE_GENERAL;
                    if (!swap_ij)
                        batch.add(Ai.data(), Ai.stride(), Aj.data(), Aj.stride(), Cij.data(),
                                  Cij.stride(), C.tileMb(i), C.tileNb(j), k, flags);
                    else
                        batch.add(Aj.data(), Aj.stride(), Ai.data(), Ai.stride(), Cij.data(),
                                  Cij.stride(), C.tileNb(j), C.tileMb(i), k, flags);
                }
    launch_batch<scalar_t>(batch, q, opa, opb, alpha, beta);
}
}  // namespace

template <typename scalar_t>
void herk(TargetType<Target::Devices>, real_type<scalar_t> alpha, Matrix<scalar_t>& A,
          real_type<scalar_t> beta, HermitianMatrix<scalar_t>& C, int priority, int queue_index,
          Layout layout)
{
    rank_k_update<scalar_t>(true, scalar_t(alpha), A, scalar_t(beta), C, queue_index, layout);
}

template <typename scalar_t>
void syrk(TargetType<Target::Devices>, scalar_t alpha, Matrix<scalar_t>& A, scalar_t beta,
          SymmetricMatrix<scalar_t>& C, int priority, int queue_index, Layout layout)
{
    rank_k_update<scalar_t>(false, alpha, A, beta, C, queue_index, layout);
}

//------------------------------------------------------------------------------
namespace {
/// Shared body of her2k (conjugate = true) and syr2k: lower-logical C, A and B one block
/// column each.  Two launches on the same queue, as the reference issues two batched gemms
/// for the off-diagonal tiles (src/internal/internal_her2k.cc:520-600):
///   1.  C(i,j) = alpha  op(A_i) op(B_j)^{H|T} + beta C(i,j)
///   2.  C(i,j) = alpha' op(B_i) op(A_j)^{H|T} +      C(i,j),  alpha' = conj(alpha) | alpha
/// Diagonal tiles take the same two products restricted to their stored triangle (the sum of
/// the two triangles is the triangle of the sum); the second launch carries the Hermitian
/// flag, which leaves the diagonal exactly real as blas::her2k does.
template <typename scalar_t, typename CMatrix>
void rank_2k_update(bool conjugate, scalar_t alpha, Matrix<scalar_t>& A, Matrix<scalar_t>& B,
                    scalar_t beta, CMatrix& C, int queue_index, Layout layout)
{
    slate_assert(layout == Layout::ColMajor);
    slate_assert(C.num_devices() > 0);
    slate_assert(A.nt() == 1 && B.nt() == 1);
    slate_assert(A.mt() == C.mt() && B.mt() == C.mt() && C.mt() == C.nt());
    slate_assert(A.op() == B.op());
    slate_assert(C.uploLogical() == Uplo::Lower);
    const int device = 0;
    const int q = queue_index;

    bool ok = true;
    Op op_left = A.op();
    Op op_right = conjugate ? conj_trans_of<scalar_t>(A.op(), &ok) : trans_of<scalar_t>(A.op(), &ok);
    if (!ok) slate_error("internal::her2k/syr2k: op(A) mixes Trans and ConjTrans");
    const bool swap_ij = (C.op() != Op::NoTrans);
    if (swap_ij && is_complex<scalar_t>::value && (C.op() == Op::ConjTrans) != conjugate)
        slate_error("internal::her2k/syr2k: op(C) does not match the update");
    scalar_t alpha1 = alpha, alpha2 = conjugate ? conj(alpha) : alpha;
    if (swap_ij && conjugate) {
        // stored P = C(i,j)^H: P = conj(alpha) op(B_j) op(A_i)^H + alpha op(A_j) op(B_i)^H + conj(beta) P
        alpha1 = conj(alpha);
        alpha2 = alpha;
        beta = conj(beta);
    }

    std::set<ij_tuple> AB_set, C_set;
    for (int64_t j = 0; j < C.nt(); ++j)
        for (int64_t i = j; i < C.mt(); ++i)
            if (C.tileIsLocal(i, j)) {
                C_set.insert({i, j});
                AB_set.insert({i, 0});
                AB_set.insert({j, 0});
            }
    if (C_set.empty()) return;
    A.tileGetForReading(AB_set, device, LayoutConvert(layout), q);
    B.tileGetForReading(AB_set, device, LayoutConvert(layout), q);
    C.tileGetForWriting(C_set, device, LayoutConvert(layout), q);

    const int64_t k = A.tileNb(0);
    const int tri_flag = (C.uploPhysical() == Uplo::Lower ? TE_TILE_LOWER : TE_TILE_UPPER);
    const int64_t nt = C.nt();
    for (int pass = 0; pass < 2; ++pass) {
        Matrix<scalar_t>& L = (pass == 0) ? A : B;      // left factor of this pass
        Matrix<scalar_t>& R = (pass == 0) ? B : A;
        const int diag_flag = tri_flag | ((pass == 1 && conjugate) ? TE_TILE_HERMITIAN : 0);
        DescBatch batch;
        batch.descs.reserve(C_set.size());
        for (int64_t jb = 0; jb < nt; jb += SuperBlock)
            for (int64_t ib = jb; ib < nt; ib += SuperBlock)
                for (int64_t j = jb; j < std::min(nt, jb + SuperBlock); ++j)
                    for (int64_t i = std::max(ib, j); i < std::min(nt, ib + SuperBlock); ++i) {
                        if (!C.tileIsLocal(i, j)) continue;
                        auto Li = L(i, 0, device), Lj = L(j, 0, device);
                        auto Ri = R(i, 0, device), Rj = R(j, 0, device);
                        auto Cij = C(i, j, device);
                        const int flags = (i == j) ? diag_flag : TE_TILE_GENERAL;
                        if (!swap_ij)      // C(i,j) += a op(L_i) op(R_j)^{H|T}
                            batch.add(Li.data(), Li.stride(), Rj.data(), Rj.stride(), Cij.data(),
                                      Cij.stride(), C.tileMb(i), C.tileNb(j), k, flags);
                        else               // P(j,i) += a' op(R_j) op(L_i)^{H|T}
                            batch.add(Rj.data(), Rj.stride(), Li.data(), Li.stride(), Cij.data(),
                                      Cij.stride(), C.tileNb(j), C.tileMb(i), k, flags);
                        (void)Lj; (void)Ri;
                    }
        launch_batch<scalar_t>(batch, q, op_left, op_right, pass == 0 ? alpha1 : alpha2,
                               pass == 0 ? beta : scalar_t(1.0));
    }
    // tileGetForWriting stamped the C tiles with the mark of the first launch; they are final
    // after the second one (a device-to-host copy of a host-origin C waits for this mark)
    for (auto const& ij : C_set) C.tileModified(std::get<0>(ij), std::get<1>(ij), device, q);
    Engine::instance().mark_queue(q);
}
}  // namespace

template <typename scalar_t>
void her2k(TargetType<Target::Devices>, scalar_t alpha, Matrix<scalar_t>& A, Matrix<scalar_t>& B,
           real_type<scalar_t> beta, HermitianMatrix<scalar_t>& C, int priority, int queue_index,
           Layout layout)
{
    rank_2k_update<scalar_t>(true, alpha, A, B, scalar_t(beta), C, queue_index, layout);
}

template <typename scalar_t>
void syr2k(TargetType<Target::Devices>, scalar_t alpha, Matrix<scalar_t>& A, Matrix<scalar_t>& B,
           scalar_t beta, SymmetricMatrix<scalar_t>& C, int priority, int queue_index, Layout layout)
{
    rank_2k_update<scalar_t>(false, alpha, A, B, beta, C, queue_index, layout);
}

//------------------------------------------------------------------------------
template <typename scalar_t>
void trsm(TargetType<Target::Devices>, Side side, scalar_t alpha, TriangularMatrix<scalar_t>& A,
          Matrix<scalar_t>& B, int priority, Layout layout, int64_t queue_index)
{
    slate_assert(layout == Layout::ColMajor);
    slate_assert(B.num_devices() > 0);
    slate_assert(A.mt() == 1 && A.nt() == 1);
    slate_assert(side == Side::Left ? A.mt() == B.mt() : A.mt() == B.nt());
    const int device = 0;
    const int q = (int)queue_index;
    Engine& eng = Engine::instance();

    // src/internal/internal_trsm.cc:148-170
    Uplo uploA = A.uploPhysical();
    Diag diagA = A.diag();
    Op opA = A.op();
    Side sideA = side;
    if (B.op() != Op::NoTrans) {
        if (A.is_complex && A.op() != Op::NoTrans && A.op() != B.op())
            slate_error("internal::trsm: op(A) and op(B) mix Trans and ConjTrans");
        sideA = (side == Side::Left ? Side::Right : Side::Left);
        if (A.op() == Op::NoTrans) opA = B.op();
        else opA = Op::NoTrans;
        if (B.op() == Op::ConjTrans) alpha = conj(alpha);
    }

    std::set<ij_tuple> B_set;
    if (side == Side::Right) {
        for (int64_t i = 0; i < B.mt(); ++i)
            if (B.tileIsLocal(i, 0)) B_set.insert({i, 0});
    }
    else {
        for (int64_t j = 0; j < B.nt(); ++j)
            if (B.tileIsLocal(0, j)) B_set.insert({0, j});
    }
    if (B_set.empty()) return;

    {
        std::set<ij_tuple> one = {{0, 0}};
        A.tileGetForReading(one, device, LayoutConvert(layout), q);
    }
    B.tileGetForWriting(B_set, device, LayoutConvert(layout), q);

    auto A00 = A(0, 0, device);
    std::vector<te_trsm_desc> descs;
    int32_t max_m = 0, max_n = 0;
    for (auto const& ij : B_set) {
        auto t = B(std::get<0>(ij), std::get<1>(ij), device);
        t.op(Op::NoTrans);                       // stored extents
        te_trsm_desc d;
        d.b = t.data();  d.m = (int32_t)t.mb();  d.n = (int32_t)t.nb();
        d.ldb = (int32_t)t.stride();  d.reserved = 0;
        descs.push_back(d);
        max_m = std::max(max_m, d.m);
        max_n = std::max(max_n, d.n);
    }
    // a triangle cut from a rectangular corner tile (getrf_nopiv at a ragged edge) is the
    // leading square part of that tile
    const int na = (int)std::min(A.tileMb(0), A.tileNb(0));
    const int dtype = te_dtype_of<scalar_t>::value;
    size_t bytes = descs.size() * sizeof(te_trsm_desc);
    Engine::DescSlot& slot = eng.acquire_desc(q, bytes);
    std::memcpy(slot.host, descs.data(), bytes);
    eng.upload_desc(q, slot, bytes);
    size_t sb = te_trsm_scratch_bytes(dtype, na, (int64_t)descs.size());
    void* scratch = eng.scratch(q, sb);
    eng.trace_begin(q, "trsm", (int64_t)descs.size());
    slate_te_call(te_trsm_batched(eng.device(), eng.compute_stream(q), dtype, int(sideA), int(uploA),
                                  int(opA), int(diagA), &alpha, A00.data(), (int)A00.stride(), na,
                                  (const te_trsm_desc*)slot.dev, (int64_t)descs.size(), max_m, max_n,
                                  scratch, sb));
    eng.trace_end(q);
    eng.retire_desc(q, slot);
    eng.mark_queue(q);
}

//------------------------------------------------------------------------------
template <typename scalar_t>
int64_t potrf(TargetType<Target::Devices>, HermitianMatrix<scalar_t>& A, int priority,
              int64_t queue_index, device_info_int* device_info)
{
    slate_assert(A.mt() == 1 && A.nt() == 1);
    if (!A.tileIsLocal(0, 0)) return 0;
    const int device = 0;
    const int q = (int)queue_index;
    Engine& eng = Engine::instance();
    std::set<ij_tuple> one = {{0, 0}};
    A.tileGetForWriting(one, device, LayoutConvert::ColMajor, q);
    auto A00 = A(0, 0, device);
    const int dtype = te_dtype_of<scalar_t>::value;
    int* info_dev = device_info ? device_info : eng.info_scratch_slot();
    size_t sb = te_potrf_scratch_bytes(dtype);
    // the trsm that follows on the same queue uses the same scratch: stream order keeps them apart
    void* scratch = eng.scratch(q, sb);
    eng.trace_begin(q, "potrf_tile", 1);
    slate_te_call(te_potrf_tile(eng.device(), eng.compute_stream(q), dtype,
                                int(A00.uploPhysical()), (int)A.tileMb(0), A00.data(),
                                (int)A00.stride(), info_dev, scratch, sb));
    eng.trace_end(q);
    eng.mark_queue(q);
    if (device_info) return 0;
    int info = 0;
    slate_te_call(te_memcpy_async(&info, info_dev, sizeof(int), TE_COPY_D2H, eng.compute_stream(q)));
    eng.sync_queue(q);
    return info;
}

//------------------------------------------------------------------------------
/// LU without pivoting of one (diagonal) tile, on the device (the reference runs
/// Tile_getrf_nopiv on the host: src/internal/internal_getrf_nopiv.cc:20-60).  `ib` is
/// accepted for the signature's sake; the device routine has its own blocking.
template <typename scalar_t>
int64_t getrf_nopiv(TargetType<Target::Devices>, Matrix<scalar_t>& A, int64_t ib, int priority,
                    int64_t queue_index, device_info_int* device_info)
{
    slate_assert(A.mt() == 1 && A.nt() == 1);
    slate_assert(A.op() == Op::NoTrans);
    if (!A.tileIsLocal(0, 0)) return 0;
    const int device = 0;
    const int q = (int)queue_index;
    Engine& eng = Engine::instance();
    std::set<ij_tuple> one = {{0, 0}};
    A.tileGetForWriting(one, device, LayoutConvert::ColMajor, q);
    auto A00 = A(0, 0, device);
    const int dtype = te_dtype_of<scalar_t>::value;
    int* info_dev = device_info ? device_info : eng.info_scratch_slot();
    size_t sb = te_getrf_nopiv_scratch_bytes(dtype);
    void* scratch = eng.scratch(q, sb);
    eng.trace_begin(q, "getrf_nopiv_tile", 1);
    slate_te_call(te_getrf_nopiv_tile(eng.device(), eng.compute_stream(q), dtype, (int)A.tileMb(0),
                                      (int)A.tileNb(0), A00.data(), (int)A00.stride(), info_dev,
                                      scratch, sb));
    eng.trace_end(q);
    eng.mark_queue(q);
    if (device_info) return 0;
    int info = 0;
    slate_te_call(te_memcpy_async(&info, info_dev, sizeof(int), TE_COPY_D2H, eng.compute_stream(q)));
    eng.sync_queue(q);
    return info;
}

//------------------------------------------------------------------------------
void reduce_info(int64_t* info, MPI_Comm comm)
{
    if (!comm || comm->size == 1) return;
    Engine& eng = Engine::instance();
    int64_t send = (*info == 0) ? std::numeric_limits<int64_t>::max() : *info;
    int64_t* buf = (int64_t*)eng.scratch(Engine::MaxQueues - 1, sizeof(int64_t));
    void* stream = eng.comm_stream();
    slate_te_call(te_memcpy_async(buf, &send, sizeof(send), TE_COPY_H2D, stream));
    slate_te_call(te_comm_allreduce_i64(comm->nccl, buf, 1, TE_REDUCE_MIN, stream));
    slate_te_call(te_memcpy_async(&send, buf, sizeof(send), TE_COPY_D2H, stream));
    eng.sync_comm();
    *info = (send == std::numeric_limits<int64_t>::max()) ? 0 : send;
}

//------------------------------------------------------------------------------
// explicit instantiations (src/internal/internal_gemm.cc:523-636 lists float too; this
// engine is FP64 / complex-FP64 only)
#define INSTANTIATE(T)                                                                          \
    template void gemm<T>(TargetType<Target::Devices>, T, Matrix<T>&, Matrix<T>&, T, Matrix<T>&, \
                          Layout, int, int64_t);                                                \
    template void gemm_multi<T>(TargetType<Target::Devices>, T, std::vector<GemmPart<T>>&, T,  \
                                Layout, int64_t);                                               \
    template void gemmA<T>(TargetType<Target::Devices>, T, Matrix<T>&, Matrix<T>&, T,           \
                           Matrix<T>&, Layout, int, int64_t);                                   \
    template void hemm<T>(TargetType<Target::Devices>, Side, T, HermitianMatrix<T>&, Matrix<T>&, \
                          T, Matrix<T>&, int, int);                                             \
    template void symm<T>(TargetType<Target::Devices>, Side, T, SymmetricMatrix<T>&, Matrix<T>&, \
                          T, Matrix<T>&, int, int);                                             \
    template void herk<T>(TargetType<Target::Devices>, real_type<T>, Matrix<T>&, real_type<T>,  \
                          HermitianMatrix<T>&, int, int, Layout);                               \
    template void syrk<T>(TargetType<Target::Devices>, T, Matrix<T>&, T, SymmetricMatrix<T>&,   \
                          int, int, Layout);                                                    \
    template void her2k<T>(TargetType<Target::Devices>, T, Matrix<T>&, Matrix<T>&,              \
                           real_type<T>, HermitianMatrix<T>&, int, int, Layout);                \
    template void syr2k<T>(TargetType<Target::Devices>, T, Matrix<T>&, Matrix<T>&, T,           \
                           SymmetricMatrix<T>&, int, int, Layout);                              \
    template void trsm<T>(TargetType<Target::Devices>, Side, T, TriangularMatrix<T>&,           \
                          Matrix<T>&, int, Layout, int64_t);                                    \
    template int64_t potrf<T>(TargetType<Target::Devices>, HermitianMatrix<T>&, int, int64_t,   \
                              device_info_int*);                                                \
    template int64_t getrf_nopiv<T>(TargetType<Target::Devices>, Matrix<T>&, int64_t, int,      \
                                    int64_t, device_info_int*);
INSTANTIATE(double)
INSTANTIATE(std::complex<double>)

}  // namespace internal
}  // namespace slate
