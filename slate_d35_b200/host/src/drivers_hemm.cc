// drivers_hemm.cc -- slate::hemm and slate::symm on the tile engine (SURVEY.md section 8f,
// item 3: the hemm / symm neighbours of the trailing update on the same batched GEMM kernel).
//
// Reference: src/hemm.cc (dispatch), src/hemmC.cc:28-420, src/symm.cc.  C stationary:
//
//   side L:  C = alpha A B + beta C        side R:  C = alpha B A + beta C
//
// with A Hermitian (symmetric), one triangle stored.  Right is turned into Left by
// (conj-)transposing the views of A, B and C (src/hemmC.cc:54-60).  Step k multiplies block
// column k of A -- A(i, k) where it is stored, A(k, i)^H (A(k, i)^T) where it is not -- into C:
//
//   C(i, :) += alpha A(i, k) B(k, :)        i != k   internal::gemm
//   C(k, :) += alpha A(k, k) B(k, :)                 internal::hemm / symm on the diagonal tile
//
// beta is applied at k = 0.  Broadcasts as src/hemmC.cc:98-137: the tile standing for A(i, k)
// goes to the owners of block row i of C, B(k, j) to the owners of block column j of C.
#include "slate/slate.hh"

#include <type_traits>

namespace slate {

/// Broadcasts of step k of slate::hemm / symm, side Left (src/hemmC.cc:98-137):
/// which = 0: the tile standing for A(i, k) -- (i, k) where the triangle stores it, (k, i)
/// where it does not -- to the owners of block row i of C; which = 1: B(k, j) to the owners of
/// block column j of C.
template <typename scalar_t, typename AMatrix>
typename Matrix<scalar_t>::BcastListTag hemm_bcast_list(int which, AMatrix& A, Matrix<scalar_t>& B,
                                                        Matrix<scalar_t>& C, int64_t k)
{
    typename Matrix<scalar_t>::BcastListTag list;
    const int64_t mt = C.mt(), nt = C.nt();
    const bool lower = (A.uploLogical() == Uplo::Lower);
    if (which == 0) {
        for (int64_t i = 0; i < mt; ++i) {
            const bool direct = (i == k) || (lower ? i > k : i < k);
            list.push_back({direct ? i : k, direct ? k : i, {C.sub(i, i, 0, nt - 1)}, i});
        }
    }
    else {
        for (int64_t j = 0; j < nt; ++j)
            list.push_back({k, j, {C.sub(0, mt - 1, j, j)}, j});
    }
    return list;
}

namespace {

template <typename scalar_t, typename AMatrix>
void hemm_symm(bool hermitian, Side side, scalar_t alpha, AMatrix A, Matrix<scalar_t> B,
               scalar_t beta, Matrix<scalar_t> C, Options const& opts)
{
    if (get_target(opts) != Target::Devices)
        slate_not_implemented("slate::hemm/symm: this build implements Target::Devices only "
                              "(there is no CPU fallback)");
    using BcastListTag = typename Matrix<scalar_t>::BcastListTag;
    const int64_t lookahead = get_option<int64_t>(opts, Option::Lookahead, 1);
    const Layout layout = Layout::ColMajor;
    const scalar_t one = 1.0;
    const int queue_0 = 0;

    if (side == Side::Right) {
        if (hermitian) {
            A = conj_transpose(A);
            B = conj_transpose(B);
            C = conj_transpose(C);
            alpha = conj(alpha);
            beta = conj(beta);
        }
        else {
            A = transpose(A);
            B = transpose(B);
            C = transpose(C);
        }
    }
    slate_assert(A.mt() == B.mt() && A.nt() == B.mt());
    slate_assert(B.mt() == C.mt() && B.nt() == C.nt());
    const int64_t mt = C.mt(), nt = C.nt();
    if (mt == 0 || nt == 0) return;
    const bool lower = (A.uploLogical() == Uplo::Lower);

    C.allocateBatchArrays();
    C.reserveDeviceWorkspace(0);
    A.ensureDeviceWorkspace((lookahead + 3) * mt + 2);
    B.ensureDeviceWorkspace((lookahead + 3) * nt);
    C.tileGetAllForWritingOnDevices(LayoutConvert(layout), -1);

    // logical A(i, k), i != k, as a view on the stored tile
    auto stored = [&](int64_t i, int64_t k) { return lower ? i > k : i < k; };

    auto bcast = [&](int64_t k) {
        BcastListTag list_A = hemm_bcast_list<scalar_t>(0, A, B, C, k);
        A.template listBcastMT<Target::Devices>(list_A, layout);
        BcastListTag list_B = hemm_bcast_list<scalar_t>(1, A, B, C, k);
        B.template listBcastMT<Target::Devices>(list_B, layout);
    };
    for (int64_t k = 0; k <= lookahead && k < mt; ++k)
        bcast(k);

    for (int64_t k = 0; k < mt; ++k) {
        if (k > 0 && k + lookahead < mt)
            bcast(k + lookahead);
        const scalar_t bet = (k == 0) ? beta : one;
        auto B_row = B.sub(k, k, 0, nt - 1);

        // rows above the diagonal tile, then rows below: one of the two ranges reads stored
        // tiles of block column k, the other (conj-)transposed tiles of block row k
        for (int part = 0; part < 2; ++part) {
            const int64_t i1 = part == 0 ? 0 : k + 1;
            const int64_t i2 = part == 0 ? k - 1 : mt - 1;
            if (i1 > i2) continue;
            const bool direct = stored(i1, k);
            if (direct) {
                internal::gemm<Target::Devices>(alpha, A.sub(i1, i2, k, k), B.sub(k, k, 0, nt - 1), bet,
                                                C.sub(i1, i2, 0, nt - 1), layout, 0, queue_0);
            }
            else {
                auto A_row = A.sub(k, k, i1, i2);
                auto A_rowT = hermitian ? conj_transpose(A_row) : transpose(A_row);
                internal::gemm<Target::Devices>(alpha, std::move(A_rowT), B.sub(k, k, 0, nt - 1), bet,
                                                C.sub(i1, i2, 0, nt - 1), layout, 0, queue_0);
            }
        }
        if constexpr (std::is_same<AMatrix, HermitianMatrix<scalar_t>>::value)
            internal::hemm<Target::Devices>(Side::Left, alpha, A.sub(k, k), std::move(B_row), bet,
                                            C.sub(k, k, 0, nt - 1), 0, queue_0);
        else
            internal::symm<Target::Devices>(Side::Left, alpha, A.sub(k, k), std::move(B_row), bet,
                                            C.sub(k, k, 0, nt - 1), 0, queue_0);

        // drop what this step received.  A stored tile (a, b) travels twice -- in step b as
        // A(a, b) to the owners of block row a of C and in step a, transposed, as A(b, a) to the
        // owners of block row b -- so a rank gives back only the copy it was a destination of in
        // THIS step (it owns part of block row i of C); otherwise, with lookahead, it would drop
        // the copy a later step has already delivered (seen on a 2 x 1 grid).
        for (int64_t i = 0; i < mt; ++i) {
            const bool direct = (i == k) || stored(i, k);
            const int64_t ti = direct ? i : k, tj = direct ? k : i;
            if (!A.tileIsLocal(ti, tj)) {
                bool destination = false;
                for (int64_t j = 0; j < nt && !destination; ++j)
                    destination = C.tileIsLocal(i, j);
                if (destination) A.releaseRemoteWorkspaceTile(ti, tj);
            }
            else A.releaseLocalWorkspaceTile(ti, tj);
        }
        A.storage()->flushErased();
        B_row.releaseRemoteWorkspace();
        B_row.releaseLocalWorkspace();
        Engine::instance().poll_deferred();
    }
    C.tileUpdateAllOrigin();
    C.releaseWorkspace();
    Engine::instance().sync_all();
}

}  // namespace

template <typename scalar_t>
void hemm(Side side, scalar_t alpha, HermitianMatrix<scalar_t>& A, Matrix<scalar_t>& B,
          scalar_t beta, Matrix<scalar_t>& C, Options const& opts)
{
    hemm_symm<scalar_t>(true, side, alpha, A, B, beta, C, opts);
}

template <typename scalar_t>
void symm(Side side, scalar_t alpha, SymmetricMatrix<scalar_t>& A, Matrix<scalar_t>& B,
          scalar_t beta, Matrix<scalar_t>& C, Options const& opts)
{
    hemm_symm<scalar_t>(false, side, alpha, A, B, beta, C, opts);
}

#define INSTANTIATE(T)                                                                          \
    template typename Matrix<T>::BcastListTag hemm_bcast_list<T, HermitianMatrix<T>>(             \
        int, HermitianMatrix<T>&, Matrix<T>&, Matrix<T>&, int64_t);                             \
    template typename Matrix<T>::BcastListTag hemm_bcast_list<T, SymmetricMatrix<T>>(             \
        int, SymmetricMatrix<T>&, Matrix<T>&, Matrix<T>&, int64_t);                             \
    template void hemm<T>(Side, T, HermitianMatrix<T>&, Matrix<T>&, T, Matrix<T>&, Options const&); \
    template void symm<T>(Side, T, SymmetricMatrix<T>&, Matrix<T>&, T, Matrix<T>&, Options const&);
INSTANTIATE(double)
INSTANTIATE(std::complex<double>)

}  // namespace slate
