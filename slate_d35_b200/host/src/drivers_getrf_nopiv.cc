// drivers_getrf_nopiv.cc -- slate::getrf_nopiv on the tile engine (SURVEY.md section 8f, item
// 4: the LU trailing update, internal::trsm + internal::gemm on Target::Devices; this is the
// reference's pivot-free LU, src/getrf_nopiv.cc:26-260, whose panel needs no row exchange).
//
// Step k:  A(k,k) = L U on the diagonal tile (internal::getrf_nopiv, on the device here);
//          A(k+1:, k) <- A(k+1:, k) U^-1   and   A(k, k+1:) <- L^-1 A(k, k+1:)   (internal::trsm);
//          A(k+1:, k+1:) -= A(k+1:, k) A(k, k+1:)                                   (internal::gemm).
// Queues as in slate::potrf: the panel on a high-priority queue, lookahead columns on their
// own high-priority queues, the trailing update on queue 0; per-tile write / comm marks order
// them.  info stays on the device until the end.
#include "slate/slate.hh"

namespace slate {

/// The three broadcasts of step k (src/getrf_nopiv.cc:96-100, 113-120, 178-188):
/// which = 0: A(k, k) to the owners of block column A(k+1:, k) and block row A(k, k+1:);
/// which = 1: A(i, k), i > k, to the owners of block row A(i, k+1:);
/// which = 2: A(k, j), j > k, to the owners of block column A(k+1:, j).
template <typename scalar_t>
typename Matrix<scalar_t>::BcastListTag getrf_nopiv_bcast_list(int which, Matrix<scalar_t>& A, int64_t k)
{
    typename Matrix<scalar_t>::BcastListTag list;
    const int64_t A_mt = A.mt(), A_nt = A.nt();
    if (which == 0) {
        std::list<BaseMatrix<scalar_t>> dests;
        if (k + 1 < A_mt) dests.push_back(A.sub(k + 1, A_mt - 1, k, k));
        if (k + 1 < A_nt) dests.push_back(A.sub(k, k, k + 1, A_nt - 1));
        if (!dests.empty()) list.push_back({k, k, dests, k});
    }
    else if (which == 1) {
        if (k + 1 < A_nt)
            for (int64_t i = k + 1; i < A_mt; ++i)
                list.push_back({i, k, {A.sub(i, i, k + 1, A_nt - 1)}, i});
    }
    else {
        if (k + 1 < A_mt)
            for (int64_t j = k + 1; j < A_nt; ++j)
                list.push_back({k, j, {A.sub(k + 1, A_mt - 1, j, j)}, j + A_mt});
    }
    return list;
}

template <typename scalar_t>
int64_t getrf_nopiv(Matrix<scalar_t>& A, Options const& opts)
{
    if (get_target(opts) != Target::Devices)
        slate_not_implemented("slate::getrf_nopiv: this build implements Target::Devices only "
                              "(there is no CPU fallback)");
    using BcastListTag = typename Matrix<scalar_t>::BcastListTag;
    const scalar_t one = 1.0;
    const Layout layout = Layout::ColMajor;
    const int64_t lookahead = get_option<int64_t>(opts, Option::Lookahead, 1);
    const int64_t ib = get_option<int64_t>(opts, Option::InnerBlocking, 16);
    Engine& eng = Engine::instance();
    const int queue_trail = 0, queue_panel = 1, queue_la0 = 2;
    slate_assert(lookahead >= 0 && queue_la0 + lookahead <= Engine::MaxQueues - 1);
    slate_assert(A.op() == Op::NoTrans);

    const int64_t A_mt = A.mt(), A_nt = A.nt();
    const int64_t min_mt_nt = std::min(A_mt, A_nt);
    Engine::instance().reserve_info_slots(min_mt_nt);
    if (min_mt_nt == 0) return 0;

    // src/getrf_nopiv.cc:43-48
    A.allocateBatchArrays(0, 2 + lookahead);
    A.reserveDeviceWorkspace();

    slate_te_call(te_memset_async(eng.info_slot(0), 0, sizeof(int) * min_mt_nt,
                                  eng.compute_stream(queue_panel)));
    eng.mark_queue(queue_panel);

    for (int64_t k = 0; k < min_mt_nt; ++k) {
        // ---- panel: factor A(k, k) and send it down its column and along its row
        internal::getrf_nopiv<Target::Devices>(A.sub(k, k, k, k), ib, 1, queue_panel,
                                               eng.info_slot((int)k));
        {
            BcastListTag diag = getrf_nopiv_bcast_list(0, A, k);
            if (!diag.empty()) A.template listBcastMT<Target::Devices>(diag, layout);
        }
        auto Akk = A.sub(k, k, k, k);

        // ---- block column: A(k+1:, k) <- A(k+1:, k) U(k,k)^-1, then along the rows
        if (k + 1 < A_mt) {
            auto Ukk = TriangularMatrix<scalar_t>(Uplo::Upper, Diag::NonUnit, Akk);
            internal::trsm<Target::Devices>(Side::Right, one, std::move(Ukk),
                                            A.sub(k + 1, A_mt - 1, k, k), 1, layout, queue_panel);
            BcastListTag list = getrf_nopiv_bcast_list(1, A, k);
            if (!list.empty()) A.template listBcastMT<Target::Devices>(list, layout);
        }
        // ---- block row: A(k, k+1:) <- L(k,k)^-1 A(k, k+1:), then down the columns
        if (k + 1 < A_nt) {
            auto Lkk = TriangularMatrix<scalar_t>(Uplo::Lower, Diag::Unit, Akk);
            internal::trsm<Target::Devices>(Side::Left, one, std::move(Lkk),
                                            A.sub(k, k, k + 1, A_nt - 1), 1, layout, queue_panel);
            BcastListTag list = getrf_nopiv_bcast_list(2, A, k);
            if (!list.empty()) A.template listBcastMT<Target::Devices>(list, layout);
        }

        if (k + 1 < A_mt && k + 1 < A_nt) {
            // ---- lookahead columns first: the next panel waits for them
            for (int64_t j = k + 1; j < k + 1 + lookahead && j < A_nt; ++j) {
                const int queue_j = queue_la0 + (int)(j - k - 1);
                internal::gemm<Target::Devices>(-one, A.sub(k + 1, A_mt - 1, k, k), A.sub(k, k, j, j),
                                                one, A.sub(k + 1, A_mt - 1, j, j), layout, 1, queue_j);
            }
            // ---- trailing submatrix
            if (k + 1 + lookahead < A_nt) {
                internal::gemm<Target::Devices>(-one, A.sub(k + 1, A_mt - 1, k, k),
                                                A.sub(k, k, k + 1 + lookahead, A_nt - 1), one,
                                                A.sub(k + 1, A_mt - 1, k + 1 + lookahead, A_nt - 1),
                                                layout, 0, queue_trail);
            }
        }

        // ---- release what this step received; host-origin tiles of the finished panel go home
        {
            auto col = A.sub(k, A_mt - 1, k, k);
            auto row = A.sub(k, k, k, A_nt - 1);
            col.releaseRemoteWorkspace();
            row.releaseRemoteWorkspace();
            if (A.origin() != Target::Devices) {
                col.tileUpdateAllOriginAsync();
                col.releaseLocalWorkspace();
                row.tileUpdateAllOriginAsync();
                row.releaseLocalWorkspace();
            }
        }
        eng.poll_deferred();
    }

    A.tileUpdateAllOrigin();
    A.releaseWorkspace();
    eng.sync_all();
    eng.trace_dump("slate::getrf_nopiv");

    std::vector<int> step_info(min_mt_nt, 0);
    slate_te_call(te_memcpy_async(step_info.data(), eng.info_slot(0), sizeof(int) * min_mt_nt,
                                  TE_COPY_D2H, eng.comm_stream()));
    eng.sync_comm();
    int64_t info = 0, kk = 0;
    for (int64_t k = 0; k < min_mt_nt; ++k) {
        if (step_info[k] != 0 && A.tileIsLocal(k, k)) {
            info = kk + step_info[k];
            break;
        }
        kk += A.tileNb(k);
    }
    internal::reduce_info(&info, A.mpiComm());
    return info;
}

#define INSTANTIATE(T)                                                                    \
    template typename Matrix<T>::BcastListTag getrf_nopiv_bcast_list<T>(int, Matrix<T>&,   \
                                                                        int64_t);         \
    template int64_t getrf_nopiv<T>(Matrix<T>&, Options const&);
INSTANTIATE(double)
INSTANTIATE(std::complex<double>)

}  // namespace slate
