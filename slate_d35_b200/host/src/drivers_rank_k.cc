// drivers_rank_k.cc -- slate::herk, syrk, her2k and syr2k on the tile engine (SURVEY.md
// section 8f, items 1 and 3: the callers that drive the same internal::herk / syrk as potrf's
// trailing update, and their rank-2k neighbours on the same kernel).  Loop structure of the
// reference: src/herk.cc:15-160, src/syrk.cc:15-163, src/her2k.cc:25-190, src/syr2k.cc.
//
//   C = alpha A A^H + beta C                          (herk; alpha, beta real, C Hermitian)
//   C = alpha A A^T + beta C                          (syrk; C symmetric)
//   C = alpha A B^H + conj(alpha) B A^H + beta C      (her2k; beta real)
//   C = alpha A B^T + alpha B A^T + beta C            (syr2k)
//
// For every block column k of A: broadcast A(i, k) to the owners of row i and of column i
// of C (src/herk.cc:58-63), then one internal::herk / syrk launch on the local tiles, beta
// on the first column only.  As in slate::gemm everything is enqueued on streams; broadcasts
// run `lookahead` columns ahead of the updates.
#include "slate/slate.hh"

namespace slate {

namespace {

template <typename scalar_t, typename CMatrix>
typename Matrix<scalar_t>::BcastListTag rank_k_bcast_list(Matrix<scalar_t>& A, CMatrix& C, int64_t k)
{
    typename Matrix<scalar_t>::BcastListTag list;
    for (int64_t i = 0; i < A.mt(); ++i)
        list.push_back({i, k, {C.sub(i, i, 0, i), C.sub(i, C.mt() - 1, i, i)}, i});
    return list;
}

template <typename scalar_t, typename CMatrix, typename Update>
void rank_k_driver(Matrix<scalar_t>& A, CMatrix& C, Options const& opts, const char* name,
                   Update&& update, Matrix<scalar_t>* B = nullptr)
{
    if (get_target(opts) != Target::Devices)
        slate_not_implemented("this build implements Target::Devices only (there is no CPU fallback)");
    const int64_t lookahead = get_option<int64_t>(opts, Option::Lookahead, 1);
    const Layout layout = Layout::ColMajor;
    slate_assert(A.mt() == C.mt());
    (void)name;
    if (B) {
        // src/her2k.cc:46-48
        slate_assert(B->mt() == C.mt());
        slate_assert(B->nt() == A.nt());
    }

    C.allocateBatchArrays();
    C.reserveDeviceWorkspace(0);
    A.ensureDeviceWorkspace((lookahead + 3) * A.mt());
    if (B) B->ensureDeviceWorkspace((lookahead + 3) * A.mt());
    C.tileGetAllForWritingOnDevices(LayoutConvert(layout), -1);

    const int64_t kt = A.nt();
    auto bcast = [&](int64_t k) {
        auto list = rank_k_bcast_list(A, C, k);
        A.template listBcastMT<Target::Devices>(list, layout);
        if (B) {
            // B(i, k) goes where A(i, k) goes (src/her2k.cc:77-87)
            auto list_B = rank_k_bcast_list(*B, C, k);
            B->template listBcastMT<Target::Devices>(list_B, layout);
        }
    };
    for (int64_t k = 0; k <= lookahead && k < kt; ++k)
        bcast(k);
    for (int64_t k = 0; k < kt; ++k) {
        if (k > 0 && k + lookahead < kt)
            bcast(k + lookahead);
        update(k);
        auto A_colblock = A.sub(0, A.mt() - 1, k, k);
        A_colblock.releaseRemoteWorkspace();
        A_colblock.releaseLocalWorkspace();
        if (B) {
            auto B_colblock = B->sub(0, B->mt() - 1, k, k);
            B_colblock.releaseRemoteWorkspace();
            B_colblock.releaseLocalWorkspace();
        }
        Engine::instance().poll_deferred();
    }
    C.tileUpdateAllOrigin();
    C.releaseWorkspace();
    Engine::instance().sync_all();
}

}  // namespace

//------------------------------------------------------------------------------
template <typename scalar_t>
void herk(real_type<scalar_t> alpha, Matrix<scalar_t>& A, real_type<scalar_t> beta,
          HermitianMatrix<scalar_t>& C_in, Options const& opts)
{
    using real_t = real_type<scalar_t>;
    // if upper, change to lower (src/herk.cc:29-30)
    HermitianMatrix<scalar_t> C = C_in;
    if (C.uplo() == Uplo::Upper)
        C = conj_transpose(C);
    rank_k_driver(A, C, opts, "herk", [&](int64_t k) {
        internal::herk<Target::Devices>(alpha, A.sub(0, A.mt() - 1, k, k),
                                        k == 0 ? beta : real_t(1.0), std::move(C), 0, 0,
                                        Layout::ColMajor);
    });
}

template <typename scalar_t>
void syrk(scalar_t alpha, Matrix<scalar_t>& A, scalar_t beta, SymmetricMatrix<scalar_t>& C_in,
          Options const& opts)
{
    SymmetricMatrix<scalar_t> C = C_in;
    if (C.uplo() == Uplo::Upper)
        C = transpose(C);
    rank_k_driver(A, C, opts, "syrk", [&](int64_t k) {
        internal::syrk<Target::Devices>(alpha, A.sub(0, A.mt() - 1, k, k),
                                        k == 0 ? beta : scalar_t(1.0), std::move(C), 0, 0,
                                        Layout::ColMajor);
    });
}

//------------------------------------------------------------------------------
template <typename scalar_t>
void her2k(scalar_t alpha, Matrix<scalar_t>& A, Matrix<scalar_t>& B, real_type<scalar_t> beta,
           HermitianMatrix<scalar_t>& C_in, Options const& opts)
{
    using real_t = real_type<scalar_t>;
    // if upper, change to lower (src/her2k.cc:41-42)
    HermitianMatrix<scalar_t> C = C_in;
    if (C.uplo() == Uplo::Upper)
        C = conj_transpose(C);
    rank_k_driver(A, C, opts, "her2k", [&](int64_t k) {
        internal::her2k<Target::Devices>(alpha, A.sub(0, A.mt() - 1, k, k), B.sub(0, B.mt() - 1, k, k),
                                         k == 0 ? beta : real_t(1.0), std::move(C), 0, 0,
                                         Layout::ColMajor);
    }, &B);
}

template <typename scalar_t>
void syr2k(scalar_t alpha, Matrix<scalar_t>& A, Matrix<scalar_t>& B, scalar_t beta,
           SymmetricMatrix<scalar_t>& C_in, Options const& opts)
{
    SymmetricMatrix<scalar_t> C = C_in;
    if (C.uplo() == Uplo::Upper)
        C = transpose(C);
    rank_k_driver(A, C, opts, "syr2k", [&](int64_t k) {
        internal::syr2k<Target::Devices>(alpha, A.sub(0, A.mt() - 1, k, k), B.sub(0, B.mt() - 1, k, k),
                                         k == 0 ? beta : scalar_t(1.0), std::move(C), 0, 0,
                                         Layout::ColMajor);
    }, &B);
}

#define INSTANTIATE(T)                                                                         \
    template void herk<T>(real_type<T>, Matrix<T>&, real_type<T>, HermitianMatrix<T>&,          \
                          Options const&);                                                     \
    template void syrk<T>(T, Matrix<T>&, T, SymmetricMatrix<T>&, Options const&);               \
    template void her2k<T>(T, Matrix<T>&, Matrix<T>&, real_type<T>, HermitianMatrix<T>&,        \
                           Options const&);                                                    \
    template void syr2k<T>(T, Matrix<T>&, Matrix<T>&, T, SymmetricMatrix<T>&, Options const&);
INSTANTIATE(double)
INSTANTIATE(std::complex<double>)

}  // namespace slate
