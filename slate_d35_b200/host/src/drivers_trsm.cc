// drivers_trsm.cc -- slate::trsm, slate::potrs and slate::posv on the tile engine
// (the drivers that complete the Cholesky solve, so that the
// reference tester's own potrf check -- solve, then ||B - A X|| -- runs on the GPU path).
//
// Loop structure of the reference: src/work/work_trsm.cc:40-300 (block forward / backward
// substitution over block rows), src/potrs.cc:41-62, src/posv.cc.
//
//   side L:  op(A) X = alpha B          side R:  X op(A) = alpha B   (B overwritten by X)
//
// Right is turned into left by (conj-)transposing the views of A and B, as the reference
// does (work_trsm.cc:73-83); the internal ops work on the stored tiles with the op flags.
// Per block row k: broadcast A(k,k) along block row k of B, internal::trsm on that row
// (panel queue), broadcast the row and the block column A(:,k), one internal::gemm update of
// the remaining rows.  Everything is enqueued on streams; the per-tile write / comm marks
// order the panel queue against the update queue.
#include "slate/slate.hh"

namespace slate {

template <typename scalar_t>
void trsm(Side side, scalar_t alpha, TriangularMatrix<scalar_t>& A_in, Matrix<scalar_t>& B_in,
          Options const& opts)
{
    if (get_target(opts) != Target::Devices)
        slate_not_implemented("slate::trsm: this build implements Target::Devices only "
                              "(there is no CPU fallback)");
    using BcastListTag = typename Matrix<scalar_t>::BcastListTag;
    const Layout layout = Layout::ColMajor;
    const scalar_t one = 1.0;
    const int queue_update = 0, queue_panel = 1;

    TriangularMatrix<scalar_t> A = A_in;
    Matrix<scalar_t> B = B_in;
    if (side == Side::Right) {
        if (A.op() == Op::ConjTrans || B.op() == Op::ConjTrans) {
            A = conj_transpose(A);
            B = conj_transpose(B);
            alpha = conj(alpha);
        }
        else {
            A = transpose(A);
            B = transpose(B);
        }
    }
    slate_assert(A.mt() == B.mt() && A.nt() == B.mt());
    const int64_t mt = B.mt(), nt = B.nt();
    if (mt == 0 || nt == 0) return;

    B.allocateBatchArrays(0, 2);
    B.reserveDeviceWorkspace(0);
    B.ensureDeviceWorkspace(4 * (mt + nt));
    A.ensureDeviceWorkspace(4 * mt);
    B.tileGetAllForWritingOnDevices(LayoutConvert(layout), -1);

    const bool lower = (A.uplo() == Uplo::Lower);
    for (int64_t step = 0; step < mt; ++step) {
        // forward sweep for a lower op(A), backward for an upper one (work_trsm.cc:98, 200)
        const int64_t k = lower ? step : mt - 1 - step;
        const scalar_t alph = (step == 0) ? alpha : one;
        const int64_t i1 = lower ? k + 1 : 0;          // rows still to be updated
        const int64_t i2 = lower ? mt - 1 : k - 1;

        // send A(k, k) to ranks owning block row B(k, :), solve that row
        {
            BcastListTag diag;
            diag.push_back({k, k, {B.sub(k, k, 0, nt - 1)}, k});
            A.template listBcastMT<Target::Devices>(diag, layout);
        }
        internal::trsm<Target::Devices>(Side::Left, alph, A.sub(k, k), B.sub(k, k, 0, nt - 1), 1,
                                        layout, queue_panel);
        if (i1 <= i2) {
            // send A(i, k) to ranks owning block row B(i, :), B(k, j) to those owning
            // block column B(i1:i2, j)
            BcastListTag list_A, list_B;
            for (int64_t i = i1; i <= i2; ++i)
                list_A.push_back({i, k, {B.sub(i, i, 0, nt - 1)}, i});
            A.template listBcastMT<Target::Devices>(list_A, layout);
            for (int64_t j = 0; j < nt; ++j)
                list_B.push_back({k, j, {B.sub(i1, i2, j, j)}, j});
            B.template listBcastMT<Target::Devices>(list_B, layout);

            // B(i1:i2, :) = alph B(i1:i2, :) - A(i1:i2, k) B(k, :)
            internal::gemm<Target::Devices>(-one, A.sub(i1, i2, k, k), B.sub(k, k, 0, nt - 1), alph,
                                            B.sub(i1, i2, 0, nt - 1), layout, 0, queue_update);
        }
        // drop what this step received
        auto A_col = A.sub(0, mt - 1, k, k);
        A_col.releaseRemoteWorkspace();
        A_col.releaseLocalWorkspace();
        B.sub(k, k, 0, nt - 1).releaseRemoteWorkspace();
        Engine::instance().poll_deferred();
    }
    B.tileUpdateAllOrigin();
    B.releaseWorkspace();
    Engine::instance().sync_all();
}

//------------------------------------------------------------------------------
/// Solves A X = B with the Cholesky factor held in A (src/potrs.cc:41-62).
template <typename scalar_t>
void potrs(HermitianMatrix<scalar_t>& A, Matrix<scalar_t>& B, Options const& opts)
{
    const scalar_t one = 1;
    slate_assert(B.mt() == A.mt());
    auto A_ = A;
    if (A_.uplo() == Uplo::Upper)
        A_ = conj_transpose(A_);
    auto L = TriangularMatrix<scalar_t>(Diag::NonUnit, A_);
    auto LT = conj_transpose(L);
    trsm(Side::Left, one, L, B, opts);
    trsm(Side::Left, one, LT, B, opts);
}

/// Factor and solve (src/posv.cc); returns potrf's info, B is only touched when info == 0.
template <typename scalar_t>
int64_t posv(HermitianMatrix<scalar_t>& A, Matrix<scalar_t>& B, Options const& opts)
{
    int64_t info = potrf(A, opts);
    if (info == 0)
        potrs(A, B, opts);
    return info;
}

#define INSTANTIATE(T)                                                                         \
    template void trsm<T>(Side, T, TriangularMatrix<T>&, Matrix<T>&, Options const&);           \
    template void potrs<T>(HermitianMatrix<T>&, Matrix<T>&, Options const&);                    \
    template int64_t posv<T>(HermitianMatrix<T>&, Matrix<T>&, Options const&);
INSTANTIATE(double)
INSTANTIATE(std::complex<double>)

}  // namespace slate
