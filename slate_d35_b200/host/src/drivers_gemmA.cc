// drivers_gemmA.cc -- slate::gemmA and the slate::gemm method dispatch (SURVEY.md section 8f,
// item 3: internal::gemmA, "used when B has one tile column").
//
// Reference: src/gemm.cc:66-88 (dispatch, MethodGemm::select_algo in include/slate/method.hh:
// 86-98), src/gemmA.cc:32-210.  gemmA leaves A where it is: for every block column k of B,
// B(i, k) is broadcast to the ranks owning block column i of A, each rank multiplies its
// local tiles of A into (partial) tiles of C(:, k), and the partial tiles are summed onto the
// owner of each C(i, k) (listReduce).  The tester's residual checks (C X with nrhs = 10) take
// this route: gemmC would broadcast all of A for ten columns of output.
#include "slate/slate.hh"

namespace slate {

/// src/gemmA.cc:72-78: B(i, k) goes to the ranks owning block column A(:, i)
template <typename scalar_t>
typename Matrix<scalar_t>::BcastListTag gemmA_bcast_list(Matrix<scalar_t>& A, Matrix<scalar_t>& B,
                                                         int64_t k)
{
    typename Matrix<scalar_t>::BcastListTag list;
    for (int64_t i = 0; i < B.mt(); ++i)
        list.push_back({i, k, {A.sub(0, A.mt() - 1, i, i)}, i});
    return list;
}

/// src/gemmA.cc:110-118: C(i, k) is summed over the ranks owning block row A(i, :)
template <typename scalar_t>
typename Matrix<scalar_t>::ReduceList gemmA_reduce_list(Matrix<scalar_t>& A, Matrix<scalar_t>& C,
                                                        int64_t k)
{
    typename Matrix<scalar_t>::ReduceList list;
    for (int64_t i = 0; i < C.mt(); ++i)
        list.push_back({i, k, C.sub(i, i, k, k), {A.sub(i, i, 0, A.nt() - 1)}});
    return list;
}

template <typename scalar_t>
void gemmA(scalar_t alpha, Matrix<scalar_t>& A, Matrix<scalar_t>& B, scalar_t beta,
           Matrix<scalar_t>& C, Options const& opts)
{
    if (get_target(opts) != Target::Devices)
        slate_not_implemented("slate::gemmA: this build implements Target::Devices only "
                              "(there is no CPU fallback)");
    using BcastListTag = typename Matrix<scalar_t>::BcastListTag;
    using ReduceList = typename Matrix<scalar_t>::ReduceList;
    const int64_t lookahead = get_option<int64_t>(opts, Option::Lookahead, 1);
    const Layout layout = Layout::ColMajor;
    const int queue_0 = 0;

    slate_assert(A.mt() == C.mt());
    slate_assert(B.nt() == C.nt());
    slate_assert(A.nt() == B.mt());
    if (C.mt() == 0 || C.nt() == 0) return;

    // src/gemmA.cc:56-62
    A.allocateBatchArrays();
    A.reserveDeviceWorkspace(0);
    B.ensureDeviceWorkspace((lookahead + 2) * B.mt());
    C.ensureDeviceWorkspace(2 * C.mt());

    auto bcast = [&](int64_t k) {
        BcastListTag list = gemmA_bcast_list(A, B, k);
        B.template listBcastMT<Target::Devices>(list, layout);
    };
    for (int64_t k = 0; k <= lookahead && k < B.nt(); ++k)
        bcast(k);

    for (int64_t k = 0; k < B.nt(); ++k) {
        if (k > 0 && k + lookahead < B.nt())
            bcast(k + lookahead);

        // src/gemmA.cc:97-121: multiply where A is, then sum the partial tiles of C(:, k)
        internal::gemmA<Target::Devices>(alpha, std::move(A), B.sub(0, B.mt() - 1, k, k), beta,
                                         C.sub(0, C.mt() - 1, k, k), layout, 0, queue_0);
        ReduceList reduce_list = gemmA_reduce_list(A, C, k);
        C.listReduce(reduce_list, layout);

        // src/gemmA.cc:124-137
        auto B_col = B.sub(0, B.mt() - 1, k, k);
        B_col.releaseRemoteWorkspace();
        B_col.releaseLocalWorkspace();
        Engine::instance().poll_deferred();
    }
    // src/gemmA.cc:196-199
    C.tileUpdateAllOrigin();
    C.releaseWorkspace();
    A.releaseLocalWorkspace();
    Engine::instance().sync_all();
}

//------------------------------------------------------------------------------
template <typename scalar_t>
void gemm(scalar_t alpha, Matrix<scalar_t>& A, Matrix<scalar_t>& B, scalar_t beta,
          Matrix<scalar_t>& C, Options const& opts)
{
    // src/gemm.cc:70-87 with MethodGemm::select_algo (method.hh:86-98): one process drives one
    // GPU here, so the reference's "gemmA does not support several devices" exception never applies
    Method method = (Method)get_option<int64_t>(opts, Option::MethodGemm, MethodGemm::Auto);
    if (method == MethodGemm::Auto)
        method = (B.nt() < 2) ? MethodGemm::GemmA : MethodGemm::GemmC;
    // gemmA as written needs C in its stored orientation
    if (method == MethodGemm::GemmA && C.op() == Op::NoTrans)
        gemmA(alpha, A, B, beta, C, opts);
    else if (method == MethodGemm::GemmA || method == MethodGemm::GemmC)
        gemmC(alpha, A, B, beta, C, opts);
    else
        throw Exception("slate::gemm: unknown MethodGemm");
}

#define INSTANTIATE(T)                                                                     \
    template typename Matrix<T>::BcastListTag gemmA_bcast_list<T>(Matrix<T>&, Matrix<T>&,   \
                                                                  int64_t);                \
    template typename Matrix<T>::ReduceList gemmA_reduce_list<T>(Matrix<T>&, Matrix<T>&,    \
                                                                 int64_t);                 \
    template void gemmA<T>(T, Matrix<T>&, Matrix<T>&, T, Matrix<T>&, Options const&);      \
    template void gemm<T>(T, Matrix<T>&, Matrix<T>&, T, Matrix<T>&, Options const&);
INSTANTIATE(double)
INSTANTIATE(std::complex<double>)

}  // namespace slate
