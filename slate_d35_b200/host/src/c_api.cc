// c_api.cc -- C entry points of the host library for the Python harness (tests, bench,
// smoke).  They play the role of the reference's generated C API
// (include/slate/c_api/slate.h, src/c_api/wrappers.cc): opaque matrix handles, the two
// drivers, and -- for unit-level parity -- the internal ops the reference's
// unit_test/test_internal_blas.cc calls directly.  Every call returns 0 or a negative
// status; sb_last_error() holds the exception text.
#include "slate/slate.hh"

#include <cstring>
#include <functional>
#include <vector>
#include <memory>
#include <string>

using namespace slate;
typedef std::complex<double> zdouble;

namespace {

thread_local std::string g_error;

enum Kind { General = 0, Hermitian = 1, Symmetric = 2, Triangular = 3 };

struct Handle {
    int dtype;     // TE_F64 / TE_C64
    int kind;
    std::shared_ptr<void> obj;
};

template <typename T> Matrix<T>& as_general(Handle* h) { return *static_cast<Matrix<T>*>(h->obj.get()); }
template <typename T> HermitianMatrix<T>& as_herm(Handle* h) { return *static_cast<HermitianMatrix<T>*>(h->obj.get()); }
template <typename T> SymmetricMatrix<T>& as_symm(Handle* h) { return *static_cast<SymmetricMatrix<T>*>(h->obj.get()); }
template <typename T> TriangularMatrix<T>& as_tri(Handle* h) { return *static_cast<TriangularMatrix<T>*>(h->obj.get()); }
template <typename T> BaseMatrix<T>& as_base(Handle* h)
{
    switch (h->kind) {
        case Hermitian: return as_herm<T>(h);
        case Symmetric: return as_symm<T>(h);
        case Triangular: return as_tri<T>(h);
        default: return as_general<T>(h);
    }
}

template <typename F>
int guarded(F&& f)
{
    try {
        f();
        return 0;
    }
    catch (std::exception const& e) {
        g_error = e.what();
        return -1;
    }
    catch (...) {
        g_error = "unknown exception";
        return -2;
    }
}

template <typename T>
T scalar_from(const double* p)
{
    if constexpr (is_complex<T>::value) return T(p[0], p[1]);
    else return p[0];
}

template <typename M>
M view_op(M& A, int op)
{
    if (op == 'T') return transpose(A);
    if (op == 'C') return conj_transpose(A);
    return A;
}

Options make_opts(int target, int lookahead)
{
    return {{Option::Target, Target(char(target))}, {Option::Lookahead, (int64_t)lookahead}};
}

template <typename T>
void create(Handle* h, int kind, int uplo, int diag, int64_t m, int64_t n, int64_t nb, int p, int q,
            int origin, T* host, int64_t ld)
{
    Uplo u = (uplo == 'U') ? Uplo::Upper : Uplo::Lower;
    Diag d = (diag == 'U') ? Diag::Unit : Diag::NonUnit;
    Target org = (origin == 'D') ? Target::Devices : Target::Host;
    const bool no_tiles = (origin == 'N');        // index map only (planning)
    MPI_Comm comm = MPI_COMM_WORLD;
    if (p * q == 1 && comm->size > 1) comm = MPI_COMM_SELF;
    switch (kind) {
        case General: {
            auto M = host ? std::make_shared<Matrix<T>>(
                                origin == 'S' ? Matrix<T>::fromScaLAPACK(m, n, host, ld, nb, p, q, comm)
                                              : Matrix<T>::fromLAPACK(m, n, host, ld, nb, p, q, comm))
                          : std::make_shared<Matrix<T>>(m, n, nb, p, q, comm);
            if (!host && !no_tiles) M->insertLocalTiles(org);
            h->obj = M;
            break;
        }
        case Hermitian: {
            auto M = host ? std::make_shared<HermitianMatrix<T>>(
                                HermitianMatrix<T>::fromLAPACK(u, n, host, ld, nb, p, q, comm))
                          : std::make_shared<HermitianMatrix<T>>(u, n, nb, p, q, comm);
            if (!host && !no_tiles) M->insertLocalTiles(org);
            h->obj = M;
            break;
        }
        case Symmetric: {
            auto M = host ? std::make_shared<SymmetricMatrix<T>>(
                                SymmetricMatrix<T>::fromLAPACK(u, n, host, ld, nb, p, q, comm))
                          : std::make_shared<SymmetricMatrix<T>>(u, n, nb, p, q, comm);
            if (!host && !no_tiles) M->insertLocalTiles(org);
            h->obj = M;
            break;
        }
        case Triangular: {
            auto M = host ? std::make_shared<TriangularMatrix<T>>(
                                TriangularMatrix<T>::fromLAPACK(u, d, n, host, ld, nb, p, q, comm))
                          : std::make_shared<TriangularMatrix<T>>(u, d, n, nb, p, q, comm);
            if (!host && !no_tiles) M->insertLocalTiles(org);
            h->obj = M;
            break;
        }
        default: slate_error("unknown matrix kind");
    }
}

/// local tiles <-> a LAPACK-layout host array of the whole matrix (harness only)
template <typename T>
void copy_all(Handle* h, T* host, int64_t ld, bool to_host)
{
    BaseMatrix<T>& A = as_base<T>(h);
    Engine& eng = Engine::instance();
    auto st = A.storage();
    const bool trap = A.uploPhysical() != Uplo::General;
    std::vector<int64_t> roff(A.mt() + 1, 0), coff(A.nt() + 1, 0);
    for (int64_t i = 0; i < A.mt(); ++i) roff[i + 1] = roff[i] + A.tileMb(i);
    for (int64_t j = 0; j < A.nt(); ++j) coff[j + 1] = coff[j] + A.tileNb(j);
    for (int64_t j = 0; j < A.nt(); ++j)
        for (int64_t i = 0; i < A.mt(); ++i) {
            if (!A.tileIsLocal(i, j)) continue;
            if (trap && (A.uploLogical() == Uplo::Lower ? i < j : i > j)) continue;
            T* hp = host + roff[i] + coff[j] * ld;
            if (to_host) {
                A.tileGetForReading(i, j, HostNum);
                auto t = A(i, j, HostNum);
                t.op(Op::NoTrans);
                for (int64_t c = 0; c < t.nb(); ++c)
                    std::memcpy(hp + c * ld, t.data() + c * t.stride(), sizeof(T) * t.mb());
            }
            else {
                A.tileGetForWriting(i, j, HostNum);
                auto t = A(i, j, HostNum);
                t.op(Op::NoTrans);
                for (int64_t c = 0; c < t.nb(); ++c)
                    std::memcpy(t.data() + c * t.stride(), hp + c * ld, sizeof(T) * t.mb());
            }
        }
    if (!to_host && A.origin() == Target::Devices) {
        A.tileUpdateAllOrigin();
        eng.sync_all();
        // host staging copies are not needed any more
        for (int64_t j = 0; j < A.nt(); ++j)
            for (int64_t i = 0; i < A.mt(); ++i)
                if (A.tileIsLocal(i, j) && A.tileExists(i, j, HostNum) && A.tileExists(i, j, 0))
                    A.tileErase(i, j, HostNum);
    }
    if (to_host && A.origin() == Target::Devices) {
        for (int64_t j = 0; j < A.nt(); ++j)
            for (int64_t i = 0; i < A.mt(); ++i)
                if (A.tileIsLocal(i, j) && A.tileExists(i, j, HostNum) && A.tileExists(i, j, 0))
                    A.tileErase(i, j, HostNum);
    }
}

}  // namespace

namespace {
template <typename M>
int64_t flatten_plan(M& mat, typename M::BcastListTag& list, int which, int64_t* out, int64_t cap,
                     int64_t pos)
{
    for (auto const& op : mat.bcastPlan(list)) {
        int64_t need = 5 + (int64_t)op.dests.size();
        if (pos + need > cap) slate_error("plan buffer too small");
        out[pos++] = which;  out[pos++] = op.i;  out[pos++] = op.j;  out[pos++] = op.root;
        out[pos++] = (int64_t)op.dests.size();
        for (int d : op.dests) out[pos++] = d;
    }
    return pos;
}
}  // namespace


#define DISPATCH(h, call_d, call_z)          \
    do {                                     \
        if ((h)->dtype == TE_C64) { call_z; } \
        else { call_d; }                     \
    } while (0)

extern "C" {

const char* sb_last_error(void) { return g_error.c_str(); }

int sb_initialize(int device, int rank, int size, const void* nccl_id)
{
    return guarded([&] { slate::initialize(device, rank, size, nccl_id); });
}

int sb_finalize(void)
{
    return guarded([&] { slate::finalize(); });
}

int sb_sync(void)
{
    return guarded([&] { Engine::instance().sync_all(); });
}

int sb_matrix_create(int dtype, int kind, int uplo, int diag, int64_t m, int64_t n, int64_t nb,
                     int p, int q, int origin, void* host, int64_t ld, void** out)
{
    return guarded([&] {
        auto h = std::make_unique<Handle>();
        h->dtype = dtype;
        h->kind = kind;
        if (dtype == TE_C64) create<zdouble>(h.get(), kind, uplo, diag, m, n, nb, p, q, origin, (zdouble*)host, ld);
        else create<double>(h.get(), kind, uplo, diag, m, n, nb, p, q, origin, (double*)host, ld);
        *out = h.release();
    });
}

/// Lambda-constructed matrix (include/slate/Matrix.hh:33-38, HermitianMatrix.hh): tile sizes from
/// the arrays `row_sizes` (mt entries, general matrices only) and `col_sizes` (nt entries; both
/// dimensions of a Hermitian matrix), owners 2D block-cyclic over p x q (column-major ranks),
/// tiles SLATE-owned on the device ('D') or the host ('H').  kind: General or Hermitian.
int sb_matrix_create_tiled(int dtype, int kind, int uplo, int64_t m, int64_t n, const int64_t* row_sizes,
                           int64_t mt, const int64_t* col_sizes, int64_t nt, int p, int q, int origin,
                           void** out)
{
    return guarded([&] {
        auto h = std::make_unique<Handle>();
        h->dtype = dtype;
        h->kind = kind;
        std::vector<int64_t> rs(row_sizes ? row_sizes : col_sizes, (row_sizes ? row_sizes + mt : col_sizes + nt));
        std::vector<int64_t> cs(col_sizes, col_sizes + nt);
        std::function<int64_t(int64_t)> tmb = [rs](int64_t i) { return i < (int64_t)rs.size() ? rs[i] : 1; };
        std::function<int64_t(int64_t)> tnb = [cs](int64_t j) { return j < (int64_t)cs.size() ? cs[j] : 1; };
        std::function<int(ij_tuple)> trank = func::process_2d_grid(GridOrder::Col, p, q);
        std::function<int(ij_tuple)> tdev = [](ij_tuple) { return 0; };
        MPI_Comm comm = MPI_COMM_WORLD;
        if (p * q == 1 && comm->size > 1) comm = MPI_COMM_SELF;
        const Target org = (origin == 'D') ? Target::Devices : Target::Host;
        auto build = [&](auto tag) {
            using T = decltype(tag);
            if (kind == Hermitian) {
                auto M = std::make_shared<HermitianMatrix<T>>(uplo == 'U' ? Uplo::Upper : Uplo::Lower, n, tnb,
                                                              trank, tdev, comm);
                M->insertLocalTiles(org);
                h->obj = M;
            }
            else if (kind == General) {
                auto M = std::make_shared<Matrix<T>>(m, n, tmb, tnb, trank, tdev, comm);
                M->insertLocalTiles(org);
                h->obj = M;
            }
            else slate_error("sb_matrix_create_tiled: General or Hermitian only");
        };
        if (dtype == TE_C64) build(zdouble());
        else build(double());
        *out = h.release();
    });
}

/// Matrix::fromDevices / HermitianMatrix::fromDevices (include/slate/Matrix.hh:102-112): tiles are
/// views into the caller's device array `dev` (this rank's 2D block-cyclic local matrix, column
/// major, leading dimension lda).
int sb_matrix_from_devices(int dtype, int kind, int uplo, int64_t m, int64_t n, void* dev, int64_t lda,
                           int64_t nb, int p, int q, void** out)
{
    return guarded([&] {
        auto h = std::make_unique<Handle>();
        h->dtype = dtype;
        h->kind = kind;
        MPI_Comm comm = MPI_COMM_WORLD;
        if (p * q == 1 && comm->size > 1) comm = MPI_COMM_SELF;
        auto build = [&](auto tag) {
            using T = decltype(tag);
            T* arr[1] = {(T*)dev};
            if (kind == Hermitian)
                h->obj = std::make_shared<HermitianMatrix<T>>(HermitianMatrix<T>::fromDevices(
                    uplo == 'U' ? Uplo::Upper : Uplo::Lower, n, arr, 1, lda, nb, p, q, comm));
            else if (kind == General)
                h->obj = std::make_shared<Matrix<T>>(Matrix<T>::fromDevices(m, n, arr, 1, lda, nb, p, q, comm));
            else slate_error("sb_matrix_from_devices: General or Hermitian only");
        };
        if (dtype == TE_C64) build(zdouble());
        else build(double());
        *out = h.release();
    });
}

int sb_matrix_destroy(void* handle)
{
    return guarded([&] {
        Engine::instance().sync_all();
        delete static_cast<Handle*>(handle);
    });
}

int sb_matrix_generate(void* handle, int64_t seed, int dominant, int dist)
{
    Handle* h = static_cast<Handle*>(handle);
    return guarded([&] {
        DISPATCH(h, generate_rand(as_base<double>(h), seed, dominant != 0, dist),
                 generate_rand(as_base<zdouble>(h), seed, dominant != 0, dist));
    });
}

int sb_matrix_to_host(void* handle, void* out, int64_t ld)
{
    Handle* h = static_cast<Handle*>(handle);
    return guarded([&] {
        DISPATCH(h, copy_all<double>(h, (double*)out, ld, true),
                 copy_all<zdouble>(h, (zdouble*)out, ld, true));
    });
}

int sb_matrix_from_host(void* handle, const void* in, int64_t ld)
{
    Handle* h = static_cast<Handle*>(handle);
    return guarded([&] {
        DISPATCH(h, copy_all<double>(h, (double*)in, ld, false),
                 copy_all<zdouble>(h, (zdouble*)in, ld, false));
    });
}

int sb_matrix_sumsq(void* handle, double* out)
{
    Handle* h = static_cast<Handle*>(handle);
    return guarded([&] {
        DISPATCH(h, *out = local_sumsq(as_base<double>(h)), *out = local_sumsq(as_base<zdouble>(h)));
    });
}

int sb_pool_stats(void* handle, int64_t* available, int64_t* capacity)
{
    Handle* h = static_cast<Handle*>(handle);
    return guarded([&] {
        if (h->dtype == TE_C64) {
            auto& mem = as_base<zdouble>(h).storage()->memory_;
            *available = (int64_t)mem.available(0);  *capacity = (int64_t)mem.capacity(0);
        }
        else {
            auto& mem = as_base<double>(h).storage()->memory_;
            *available = (int64_t)mem.available(0);  *capacity = (int64_t)mem.capacity(0);
        }
    });
}

/// C = alpha op(A) op(B) + beta C through slate::gemm; *ms = device time of the call.
/// method: 0 = auto (MethodGemm::select_algo), 1 = gemmA, 2 = gemmC
int sb_gemm(int opa, int opb, const double* alpha, void* A, void* B, const double* beta, void* C,
            int lookahead, int target, float* ms, int method)
{
    Handle *ha = (Handle*)A, *hb = (Handle*)B, *hc = (Handle*)C;
    return guarded([&] {
        Engine& eng = Engine::instance();
        Options opts = make_opts(target, lookahead);
        opts[Option::MethodGemm] = (int64_t)method;
        void* t0 = ms ? eng.timer_begin() : nullptr;
        if (hc->dtype == TE_C64) {
            auto opA = view_op(as_general<zdouble>(ha), opa);
            auto opB = view_op(as_general<zdouble>(hb), opb);
            gemm(scalar_from<zdouble>(alpha), opA, opB, scalar_from<zdouble>(beta),
                 as_general<zdouble>(hc), opts);
        }
        else {
            auto opA = view_op(as_general<double>(ha), opa);
            auto opB = view_op(as_general<double>(hb), opb);
            gemm(alpha[0], opA, opB, beta[0], as_general<double>(hc), opts);
        }
        if (ms) *ms = eng.timer_end(t0);
    });
}

int sb_potrf(void* A, int lookahead, int target, int64_t* info, float* ms)
{
    Handle* h = (Handle*)A;
    return guarded([&] {
        Engine& eng = Engine::instance();
        Options opts = make_opts(target, lookahead);
        void* t0 = ms ? eng.timer_begin() : nullptr;
        DISPATCH(h, *info = potrf(as_herm<double>(h), opts), *info = potrf(as_herm<zdouble>(h), opts));
        if (ms) *ms = eng.timer_end(t0);
    });
}

int sb_getrf_nopiv(void* A, int lookahead, int target, int64_t* info)
{
    Handle* h = (Handle*)A;
    return guarded([&] {
        Options opts = make_opts(target, lookahead);
        DISPATCH(h, *info = getrf_nopiv(as_general<double>(h), opts),
                 *info = getrf_nopiv(as_general<zdouble>(h), opts));
    });
}

/// slate::herk (kind = 0; alpha, beta real) or slate::syrk (kind = 1) on op(A)
int sb_rank_k(int kind, int opa, const double* alpha, void* A, const double* beta, void* C,
              int lookahead, int target)
{
    Handle *ha = (Handle*)A, *hc = (Handle*)C;
    return guarded([&] {
        Options opts = make_opts(target, lookahead);
        if (hc->dtype == TE_C64) {
            auto opA = view_op(as_general<zdouble>(ha), opa);
            if (kind == 0) herk(alpha[0], opA, beta[0], as_herm<zdouble>(hc), opts);
            else syrk(scalar_from<zdouble>(alpha), opA, scalar_from<zdouble>(beta), as_symm<zdouble>(hc), opts);
        }
        else {
            auto opA = view_op(as_general<double>(ha), opa);
            if (kind == 0) herk(alpha[0], opA, beta[0], as_herm<double>(hc), opts);
            else syrk(alpha[0], opA, beta[0], as_symm<double>(hc), opts);
        }
    });
}

/// slate::her2k (kind = 0; beta real) or slate::syr2k (kind = 1) on op(A), op(B)
int sb_rank_2k(int kind, int opa, const double* alpha, void* A, void* B, const double* beta, void* C,
               int lookahead, int target)
{
    Handle *ha = (Handle*)A, *hb = (Handle*)B, *hc = (Handle*)C;
    return guarded([&] {
        Options opts = make_opts(target, lookahead);
        if (hc->dtype == TE_C64) {
            auto opA = view_op(as_general<zdouble>(ha), opa);
            auto opB = view_op(as_general<zdouble>(hb), opa);
            if (kind == 0) her2k(scalar_from<zdouble>(alpha), opA, opB, beta[0], as_herm<zdouble>(hc), opts);
            else syr2k(scalar_from<zdouble>(alpha), opA, opB, scalar_from<zdouble>(beta), as_symm<zdouble>(hc), opts);
        }
        else {
            auto opA = view_op(as_general<double>(ha), opa);
            auto opB = view_op(as_general<double>(hb), opa);
            if (kind == 0) her2k(alpha[0], opA, opB, beta[0], as_herm<double>(hc), opts);
            else syr2k(alpha[0], opA, opB, beta[0], as_symm<double>(hc), opts);
        }
    });
}

/// slate::hemm (A Hermitian handle) or slate::symm (A symmetric handle)
int sb_hemm(int side, const double* alpha, void* A, void* B, const double* beta, void* C,
            int lookahead, int target)
{
    Handle *ha = (Handle*)A, *hb = (Handle*)B, *hc = (Handle*)C;
    return guarded([&] {
        Options opts = make_opts(target, lookahead);
        const Side sd = (side == 'R') ? Side::Right : Side::Left;
        const bool herm = (ha->kind == 1);
        if (hc->dtype == TE_C64) {
            auto al = scalar_from<zdouble>(alpha), be = scalar_from<zdouble>(beta);
            if (herm) hemm(sd, al, as_herm<zdouble>(ha), as_general<zdouble>(hb), be, as_general<zdouble>(hc), opts);
            else symm(sd, al, as_symm<zdouble>(ha), as_general<zdouble>(hb), be, as_general<zdouble>(hc), opts);
        }
        else {
            if (herm) hemm(sd, alpha[0], as_herm<double>(ha), as_general<double>(hb), beta[0], as_general<double>(hc), opts);
            else symm(sd, alpha[0], as_symm<double>(ha), as_general<double>(hb), beta[0], as_general<double>(hc), opts);
        }
    });
}

int sb_trsm(int side, int opa, const double* alpha, void* A, void* B, int lookahead, int target)
{
    Handle *ha = (Handle*)A, *hb = (Handle*)B;
    return guarded([&] {
        Options opts = make_opts(target, lookahead);
        Side s = (side == 'L') ? Side::Left : Side::Right;
        if (hb->dtype == TE_C64) {
            auto opA = view_op(as_tri<zdouble>(ha), opa);
            trsm(s, scalar_from<zdouble>(alpha), opA, as_general<zdouble>(hb), opts);
        }
        else {
            auto opA = view_op(as_tri<double>(ha), opa);
            trsm(s, alpha[0], opA, as_general<double>(hb), opts);
        }
    });
}

/// factor != 0: slate::posv (factor + solve, *info = potrf's info); else slate::potrs
int sb_potrs(void* A, void* B, int factor, int lookahead, int target, int64_t* info)
{
    Handle *ha = (Handle*)A, *hb = (Handle*)B;
    return guarded([&] {
        Options opts = make_opts(target, lookahead);
        *info = 0;
        if (hb->dtype == TE_C64) {
            if (factor) *info = posv(as_herm<zdouble>(ha), as_general<zdouble>(hb), opts);
            else potrs(as_herm<zdouble>(ha), as_general<zdouble>(hb), opts);
        }
        else {
            if (factor) *info = posv(as_herm<double>(ha), as_general<double>(hb), opts);
            else potrs(as_herm<double>(ha), as_general<double>(hb), opts);
        }
    });
}

int sb_potrf_residual(void* A, int64_t seed, int dominant, double* resid_sumsq, double* a0_sumsq)
{
    Handle* h = (Handle*)A;
    return guarded([&] {
        DISPATCH(h,
                 *resid_sumsq = potrf_residual_sumsq(as_herm<double>(h), seed, dominant != 0, a0_sumsq),
                 *resid_sumsq = potrf_residual_sumsq(as_herm<zdouble>(h), seed, dominant != 0, a0_sumsq));
    });
}

// ---- broadcast plans (host logic only; runs without a GPU) ---------------------------------
/// Broadcasts of SUMMA step k as slate::gemm issues them: records
/// [which (0 = A, 1 = B), i, j, root, ndest, dest...], A's list first.
int sb_plan_gemm_step(void* A, void* B, void* C, int64_t k, int64_t* out, int64_t cap, int64_t* used)
{
    Handle *ha = (Handle*)A, *hb = (Handle*)B, *hc = (Handle*)C;
    return guarded([&] {
        auto& a = as_general<double>(ha);  auto& b = as_general<double>(hb);
        auto& c = as_general<double>(hc);
        auto la = gemm_bcast_list(0, a, b, c, k);
        auto lb = gemm_bcast_list(1, a, b, c, k);
        int64_t pos = flatten_plan(a, la, 0, out, cap, 0);
        *used = flatten_plan(b, lb, 1, out, cap, pos);
    });
}

/// Column k of slate::gemmA: records [0, i, j, root, ndest, dest...] for the broadcast of
/// B(:, k), then [2, i, j, root, nsource, source...] for the reduction of C(:, k).
int sb_plan_gemmA_step(void* A, void* B, void* C, int64_t k, int64_t* out, int64_t cap, int64_t* used)
{
    Handle *ha = (Handle*)A, *hb = (Handle*)B, *hc = (Handle*)C;
    return guarded([&] {
        auto& a = as_general<double>(ha);  auto& b = as_general<double>(hb);
        auto& c = as_general<double>(hc);
        auto lb = gemmA_bcast_list(a, b, k);
        int64_t pos = flatten_plan(b, lb, 0, out, cap, 0);
        auto lr = gemmA_reduce_list(a, c, k);
        for (auto const& op : c.reducePlan(lr)) {
            int64_t need = 5 + (int64_t)op.sources.size();
            if (pos + need > cap) slate_error("plan buffer too small");
            out[pos++] = 2;  out[pos++] = op.i;  out[pos++] = op.j;  out[pos++] = op.root;
            out[pos++] = (int64_t)op.sources.size();
            for (int r : op.sources) out[pos++] = r;
        }
        *used = pos;
    });
}

/// Broadcasts of step k of slate::hemm (side Left): which = 0 tiles of A, 1 block row k of B.
int sb_plan_hemm_step(void* A, void* B, void* C, int64_t k, int64_t* out, int64_t cap, int64_t* used)
{
    Handle *ha = (Handle*)A, *hb = (Handle*)B, *hc = (Handle*)C;
    return guarded([&] {
        auto& a = as_herm<double>(ha);  auto& b = as_general<double>(hb);
        auto& c = as_general<double>(hc);
        auto la = hemm_bcast_list<double>(0, a, b, c, k);
        auto lb = hemm_bcast_list<double>(1, a, b, c, k);
        int64_t pos = flatten_plan(a, la, 0, out, cap, 0);
        *used = flatten_plan(b, lb, 1, out, cap, pos);
    });
}

/// Broadcasts of LU step k as slate::getrf_nopiv issues them: which = 0 the diagonal tile,
/// 1 the column tiles along their rows, 2 the row tiles down their columns.
int sb_plan_getrf_nopiv_step(void* A, int64_t k, int64_t* out, int64_t cap, int64_t* used)
{
    Handle* h = (Handle*)A;
    return guarded([&] {
        auto& a = as_general<double>(h);
        int64_t pos = 0;
        for (int which = 0; which < 3; ++which) {
            auto l = getrf_nopiv_bcast_list(which, a, k);
            pos = flatten_plan(a, l, which, out, cap, pos);
        }
        *used = pos;
    });
}

/// Broadcasts of Cholesky step k as slate::potrf issues them (lower storage): which = 0 is
/// A(k,k) down its column (before the trsm), which = 1 the panel tiles A(i,k) (after it).
int sb_plan_potrf_step(void* A, int64_t k, int64_t* out, int64_t cap, int64_t* used)
{
    Handle* h = (Handle*)A;
    return guarded([&] {
        auto& a = as_herm<double>(h);
        auto l0 = potrf_bcast_list(0, a, k);
        auto l1 = potrf_bcast_list(1, a, k);
        int64_t pos = flatten_plan(a, l0, 0, out, cap, 0);
        *used = flatten_plan(a, l1, 1, out, cap, pos);
    });
}

/// Layout of the device instance of tile (i, j): 'C' or 'R'; 0 if there is none.
int sb_tile_layout(void* A, int64_t i, int64_t j)
{
    Handle* h = (Handle*)A;
    int out = 0;
    guarded([&] {
        DISPATCH(h,
                 out = as_base<double>(h).tileExists(i, j, 0) ? int(as_base<double>(h).tileLayout(i, j, 0)) : 0,
                 out = as_base<zdouble>(h).tileExists(i, j, 0) ? int(as_base<zdouble>(h).tileLayout(i, j, 0)) : 0);
    });
    return out;
}

/// BaseMatrix::tileLayoutConvert on the device instance of tile (i, j), compute queue 0.
int sb_tile_layout_convert(void* A, int64_t i, int64_t j, int layout)
{
    Handle* h = (Handle*)A;
    return guarded([&] {
        const Layout l = (layout == 'R') ? Layout::RowMajor : Layout::ColMajor;
        DISPATCH(h, as_base<double>(h).tileLayoutConvert(i, j, 0, l, 0),
                 as_base<zdouble>(h).tileLayoutConvert(i, j, 0, l, 0));
    });
}

/// owner of tile (i, j) on a p x q grid with ranks laid out column-major ('C') or row-major
/// ('R') over the grid (include/slate/func.hh:100-118, process_2d_grid)
int sb_grid_tile_rank(int order, int p, int q, int64_t i, int64_t j)
{
    int out = -1;
    guarded([&] {
        auto f = func::process_2d_grid(order == 'R' ? GridOrder::Row : GridOrder::Col, p, q);
        out = f(ij_tuple{i, j});
    });
    return out;
}

/// The (root, destination) messages of one broadcast list as rank `rank` sees them
/// (BaseMatrix::bcastPairs; SLATE_B200_BCAST_COALESCE): kind 0 = list `which` of gemm step k
/// (A, B, C given), kind 1 = list `which` of potrf step k (A given).  Records
/// [direction (0 send, 1 receive), peer, ntiles, bytes, i0, j0, i1, j1, ...].
int sb_plan_pairs(int kind, void* A, void* B, void* C, int64_t k, int which, int rank, int64_t* out,
                  int64_t cap, int64_t* used)
{
    Handle *ha = (Handle*)A, *hb = (Handle*)B, *hc = (Handle*)C;
    return guarded([&] {
        typedef BaseMatrix<double> Base;
        std::vector<Base::BcastOp> plan;
        std::vector<Base::PairMsg> sends, recvs;
        Base* M = nullptr;
        if (kind == 0) {
            auto& a = as_general<double>(ha);  auto& b = as_general<double>(hb);
            auto& c = as_general<double>(hc);
            auto list = gemm_bcast_list(which, a, b, c, k);
            M = (which == 0) ? (Base*)&a : (Base*)&b;
            plan = M->bcastPlan(list);
        }
        else {
            auto& a = as_herm<double>(ha);
            auto list = potrf_bcast_list(which, a, k);
            M = &a;
            plan = M->bcastPlan(list);
        }
        M->bcastPairs(plan, rank, &sends, &recvs);
        int64_t pos = 0;
        for (int dir = 0; dir < 2; ++dir)
            for (auto const& m : (dir == 0 ? sends : recvs)) {
                if (pos + 4 + 2 * (int64_t)m.ops.size() > cap) slate_error("plan buffer too small");
                out[pos++] = dir;  out[pos++] = m.peer;  out[pos++] = (int64_t)m.ops.size();
                out[pos++] = (int64_t)m.bytes;
                for (size_t o : m.ops) { out[pos++] = plan[o].i;  out[pos++] = plan[o].j; }
            }
        *used = pos;
    });
}

int sb_tile_rank(void* A, int64_t i, int64_t j)
{
    Handle* h = (Handle*)A;
    return h->dtype == TE_C64 ? as_base<zdouble>(h).tileRank(i, j) : as_base<double>(h).tileRank(i, j);
}

// ---- internal ops, as unit_test/test_internal_blas.cc drives them ------------------------
int sb_internal_gemm(int opa, int opb, int opc, const double* alpha, void* A, void* B,
                     const double* beta, void* C, int target, int queue)
{
    Handle *ha = (Handle*)A, *hb = (Handle*)B, *hc = (Handle*)C;
    return guarded([&] {
        if (target != 'D') slate_not_implemented("only Target::Devices is built");
        if (hc->dtype == TE_C64) {
            auto opA = view_op(as_general<zdouble>(ha), opa);
            auto opB = view_op(as_general<zdouble>(hb), opb);
            auto opC = view_op(as_general<zdouble>(hc), opc);
            internal::gemm<Target::Devices>(scalar_from<zdouble>(alpha), std::move(opA), std::move(opB),
                                            scalar_from<zdouble>(beta), std::move(opC),
                                            Layout::ColMajor, 0, queue);
        }
        else {
            auto opA = view_op(as_general<double>(ha), opa);
            auto opB = view_op(as_general<double>(hb), opb);
            auto opC = view_op(as_general<double>(hc), opc);
            internal::gemm<Target::Devices>(alpha[0], std::move(opA), std::move(opB), beta[0],
                                            std::move(opC), Layout::ColMajor, 0, queue);
        }
    });
}

int sb_internal_herk(int opa, int opc, double alpha, void* A, double beta, void* C, int queue)
{
    Handle *ha = (Handle*)A, *hc = (Handle*)C;
    return guarded([&] {
        if (hc->dtype == TE_C64) {
            auto opA = view_op(as_general<zdouble>(ha), opa);
            auto opC = view_op(as_herm<zdouble>(hc), opc);
            if (opC.uplo() == Uplo::Upper) opC = conj_transpose(opC);     // src/herk.cc
            internal::herk<Target::Devices>(alpha, std::move(opA), beta, std::move(opC), 0, queue);
        }
        else {
            auto opA = view_op(as_general<double>(ha), opa);
            auto opC = view_op(as_herm<double>(hc), opc);
            if (opC.uplo() == Uplo::Upper) opC = conj_transpose(opC);
            internal::herk<Target::Devices>(alpha, std::move(opA), beta, std::move(opC), 0, queue);
        }
    });
}

int sb_internal_syrk(int opa, int opc, const double* alpha, void* A, const double* beta, void* C,
                     int queue)
{
    Handle *ha = (Handle*)A, *hc = (Handle*)C;
    return guarded([&] {
        if (hc->dtype == TE_C64) {
            auto opA = view_op(as_general<zdouble>(ha), opa);
            auto opC = view_op(as_symm<zdouble>(hc), opc);
            if (opC.uplo() == Uplo::Upper) opC = transpose(opC);          // src/syrk.cc
            internal::syrk<Target::Devices>(scalar_from<zdouble>(alpha), std::move(opA),
                                            scalar_from<zdouble>(beta), std::move(opC), 0, queue);
        }
        else {
            auto opA = view_op(as_general<double>(ha), opa);
            auto opC = view_op(as_symm<double>(hc), opc);
            if (opC.uplo() == Uplo::Upper) opC = transpose(opC);
            internal::syrk<Target::Devices>(alpha[0], std::move(opA), beta[0], std::move(opC), 0, queue);
        }
    });
}

int sb_internal_trsm(int side, int opa, int opb, const double* alpha, void* A, void* B, int queue)
{
    Handle *ha = (Handle*)A, *hb = (Handle*)B;
    return guarded([&] {
        Side s = (side == 'L') ? Side::Left : Side::Right;
        if (hb->dtype == TE_C64) {
            auto opA = view_op(as_tri<zdouble>(ha), opa);
            auto opB = view_op(as_general<zdouble>(hb), opb);
            internal::trsm<Target::Devices>(s, scalar_from<zdouble>(alpha), std::move(opA),
                                            std::move(opB), 0, Layout::ColMajor, queue);
        }
        else {
            auto opA = view_op(as_tri<double>(ha), opa);
            auto opB = view_op(as_general<double>(hb), opb);
            internal::trsm<Target::Devices>(s, alpha[0], std::move(opA), std::move(opB), 0,
                                            Layout::ColMajor, queue);
        }
    });
}

int sb_internal_potrf(void* A, int queue, int64_t* info)
{
    Handle* h = (Handle*)A;
    return guarded([&] {
        DISPATCH(h, *info = internal::potrf<Target::Devices>(as_herm<double>(h).sub(0, 0), 0, queue),
                 *info = internal::potrf<Target::Devices>(as_herm<zdouble>(h).sub(0, 0), 0, queue));
    });
}

}  // extern "C"
