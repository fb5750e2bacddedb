// scalapack_api.cc -- ScaLAPACK-convention entry points on the tile engine
// (include/slate_b200_scalapack.h).
//
// What the reference does in scalapack_api/scalapack_{gemm,potrf,potrs,posv,trsm,syrk,herk,hemm,
// symm,her2k,syr2k}.cc:
// read the process grid of the descriptor's BLACS context, wrap the caller's local array
// with fromScaLAPACK, cut the (ia, ja) sub-matrix at tile boundaries
// (scalapack_api/scalapack_slate.hh:81-95), apply the op flags as views, call the driver.
// The same here, with Target::Devices as the only target: origin tiles are user-owned host
// memory, staged over PCIe by the drivers (slate::gemm's slab schedule, tileUpdateAllOrigin).
#include "slate/slate.hh"
#include "slate_b200_scalapack.h"

#include <cctype>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>

// BLACS, if the process has one (weak: a null address means "not linked").
extern "C" {
void Cblacs_gridinfo(int context, int* np_row, int* np_col, int* my_row, int* my_col)
    __attribute__((weak));
void Cblacs_pinfo(int* mypnum, int* nprocs) __attribute__((weak));
void Cblacs_pcoord(int icontxt, int pnum, int* prow, int* pcol) __attribute__((weak));
void Cblacs_get(int icontxt, int what, int* val) __attribute__((weak));
}

namespace {

using zdouble = std::complex<double>;

std::mutex g_mu;
std::string g_error;
int g_status = 0;

struct Grid { slate::GridOrder order; int p, q; };
std::map<int, Grid> g_grids;

// descriptor fields (ScaLAPACK array descriptor, DTYPE 1; DTYPE 2 carries IMB/INB first)
enum { DTYPE_ = 0, CTXT_ = 1, M_ = 2, N_ = 3 };
struct Desc {
    int ctxt;
    int64_t m, n, mb, nb, lld;
    int rsrc, csrc;
    explicit Desc(const int* d)
    {
        slate_assert(d != nullptr);
        const int o = (d[DTYPE_] == 1) ? 0 : 2;      // BLOCK_CYCLIC_2D : BLOCK_CYCLIC_2D_INB
        slate_assert(d[DTYPE_] == 1 || d[DTYPE_] == 2);
        ctxt = d[CTXT_];
        m = d[M_];
        n = d[N_];
        mb = d[4 + o];
        nb = d[5 + o];
        rsrc = d[6 + o];
        csrc = d[7 + o];
        lld = d[8 + o];
        if (rsrc != 0 || csrc != 0)
            throw slate::Exception("scalapack_api: descriptors with RSRC/CSRC != 0 are not supported");
    }
};

Grid grid_of(int ctxt)
{
    if (Cblacs_gridinfo && Cblacs_pinfo && Cblacs_pcoord && Cblacs_get) {
        // scalapack_api/scalapack_slate.hh:58-78: where process number 1 sits tells the order
        Grid g{slate::GridOrder::Col, 1, 1};
        int myrow, mycol, me, nprocs;
        Cblacs_gridinfo(ctxt, &g.p, &g.q, &myrow, &mycol);
        Cblacs_pinfo(&me, &nprocs);
        if (nprocs > 1) {
            int sys = -1, prow, pcol;
            Cblacs_get(-1, 0, &sys);
            Cblacs_pcoord(sys, 1, &prow, &pcol);
            g.order = (prow == 0 && pcol == 1) ? slate::GridOrder::Col : slate::GridOrder::Row;
            // the reference's test reads (0,1) as column-major; kept as is
        }
        return g;
    }
    std::lock_guard<std::mutex> lock(g_mu);
    auto it = g_grids.find(ctxt);
    if (it == g_grids.end())
        throw slate::Exception("scalapack_api: no BLACS linked and no grid registered for context "
                               + std::to_string(ctxt) + " (slate_scalapack_grid_set)");
    return it->second;
}

slate::Options call_options()
{
    // SLATE_SCALAPACK_TARGET: 5th character of hostTask/hostNest/hostBatch/deviCes
    // (scalapack_slate.hh:144-158).  Unset means Devices here.
    if (const char* t = std::getenv("SLATE_SCALAPACK_TARGET")) {
        if (std::strlen(t) < 5 || std::toupper((unsigned char)t[4]) != 'C')
            slate_not_implemented("scalapack_api: SLATE_SCALAPACK_TARGET names a host target; this "
                                  "library computes on the GPU only (no CPU fallback)");
    }
    int64_t la = 1;
    if (const char* s = std::getenv("SLATE_SCALAPACK_LOOKAHEAD")) {
        long v = std::strtol(s, nullptr, 0);
        if (v != 0) la = v;
    }
    return {{slate::Option::Lookahead, la}, {slate::Option::Target, slate::Target::Devices}};
}

bool verbose()
{
    const char* s = std::getenv("SLATE_SCALAPACK_VERBOSE");
    return s && s[0] == '1';
}

void say(const char* what)
{
    if (verbose() && slate::comm_world()->rank == 0)
        std::fprintf(stderr, "slate_b200 scalapack_api: %s\n", what);
}

slate::Op op_of(const char* s)
{
    switch (std::toupper((unsigned char)s[0])) {
        case 'N': return slate::Op::NoTrans;
        case 'T': return slate::Op::Trans;
        case 'C': return slate::Op::ConjTrans;
    }
    throw slate::Exception("scalapack_api: bad trans character");
}
slate::Uplo uplo_of(const char* s)
{
    switch (std::toupper((unsigned char)s[0])) {
        case 'L': return slate::Uplo::Lower;
        case 'U': return slate::Uplo::Upper;
    }
    throw slate::Exception("scalapack_api: bad uplo character");
}
slate::Side side_of(const char* s)
{
    switch (std::toupper((unsigned char)s[0])) {
        case 'L': return slate::Side::Left;
        case 'R': return slate::Side::Right;
    }
    throw slate::Exception("scalapack_api: bad side character");
}
slate::Diag diag_of(const char* s)
{
    switch (std::toupper((unsigned char)s[0])) {
        case 'N': return slate::Diag::NonUnit;
        case 'U': return slate::Diag::Unit;
    }
    throw slate::Exception("scalapack_api: bad diag character");
}

// ---- wrapping and sub-matrices (scalapack_slate.hh:81-141) --------------------------------
template <typename T>
slate::Matrix<T> general(int64_t m, int64_t n, T* a, int ia, int ja, const int* desc)
{
    Desc d(desc);
    Grid g = grid_of(d.ctxt);
    auto A = slate::Matrix<T>::fromScaLAPACK(d.m, d.n, a, d.lld, d.mb, d.nb, g.order, g.p, g.q,
                                             MPI_COMM_WORLD);
    if (ia == 1 && ja == 1 && m == A.m() && n == A.n())
        return A;
    if ((ia - 1) % d.mb || (ja - 1) % d.nb || m % d.mb || n % d.nb)
        throw slate::Exception("scalapack_api: sub-matrix offsets and sizes must be multiples of "
                               "the block size");
    const int64_t i1 = (ia - 1) / d.mb, j1 = (ja - 1) / d.nb;
    return A.sub(i1, i1 + m / d.mb - 1, j1, j1 + n / d.nb - 1);
}

template <typename MatrixType, typename T, typename... Extra>
MatrixType square(int64_t n, T* a, int ia, int ja, const int* desc, Extra... extra)
{
    Desc d(desc);
    Grid g = grid_of(d.ctxt);
    auto A = MatrixType::fromScaLAPACK(extra..., d.n, a, d.lld, d.nb, g.order, g.p, g.q,
                                       MPI_COMM_WORLD);
    if (ia == 1 && ja == 1 && n == A.n())
        return A;
    if (ia != ja || (ia - 1) % d.nb || n % d.nb)
        throw slate::Exception("scalapack_api: a diagonal sub-matrix must start and end on block "
                               "boundaries with ia == ja");
    const int64_t i1 = (ia - 1) / d.nb;
    return A.sub(i1, i1 + n / d.nb - 1);
}

template <typename M>
void apply_op(M& A, slate::Op op)
{
    if (op == slate::Op::Trans) A = transpose(A);
    else if (op == slate::Op::ConjTrans) A = conj_transpose(A);
}

// ---- the type-generic bodies ---------------------------------------------------------------
template <typename T>
void pgemm(const char* ta, const char* tb, int m, int n, int k, T alpha, T* a, int ia, int ja,
           const int* desca, T* b, int ib, int jb, const int* descb, T beta, T* c, int ic, int jc,
           const int* descc)
{
    const slate::Op opA = op_of(ta), opB = op_of(tb);
    const bool na = (opA == slate::Op::NoTrans), nb = (opB == slate::Op::NoTrans);
    auto A = general<T>(na ? m : k, na ? k : m, a, ia, ja, desca);
    auto B = general<T>(nb ? k : n, nb ? n : k, b, ib, jb, descb);
    auto C = general<T>(m, n, c, ic, jc, descc);
    apply_op(A, opA);
    apply_op(B, opB);
    say("gemm");
    slate::gemm(alpha, A, B, beta, C, call_options());
}

template <typename T>
void ppotrf(const char* uplo, int n, T* a, int ia, int ja, const int* desca, int* info)
{
    auto A = square<slate::HermitianMatrix<T>>(n, a, ia, ja, desca, uplo_of(uplo));
    say("potrf");
    *info = (int)slate::potrf(A, call_options());
}

template <typename T>
void ppotrs(const char* uplo, int n, int nrhs, T* a, int ia, int ja, const int* desca, T* b, int ib,
            int jb, const int* descb, int* info)
{
    auto A = square<slate::HermitianMatrix<T>>(n, a, ia, ja, desca, uplo_of(uplo));
    auto B = general<T>(n, nrhs, b, ib, jb, descb);
    say("potrs");
    slate::potrs(A, B, call_options());
    *info = 0;
}

template <typename T>
void pposv(const char* uplo, int n, int nrhs, T* a, int ia, int ja, const int* desca, T* b, int ib,
           int jb, const int* descb, int* info)
{
    auto A = square<slate::HermitianMatrix<T>>(n, a, ia, ja, desca, uplo_of(uplo));
    auto B = general<T>(n, nrhs, b, ib, jb, descb);
    say("posv");
    *info = (int)slate::posv(A, B, call_options());
}

template <typename T>
void ptrsm(const char* side, const char* uplo, const char* ta, const char* diag, int m, int n,
           T alpha, T* a, int ia, int ja, const int* desca, T* b, int ib, int jb, const int* descb)
{
    const slate::Side sd = side_of(side);
    const int64_t an = (sd == slate::Side::Left) ? m : n;
    auto A = square<slate::TriangularMatrix<T>>(an, a, ia, ja, desca, uplo_of(uplo), diag_of(diag));
    auto B = general<T>(m, n, b, ib, jb, descb);
    apply_op(A, op_of(ta));
    say("trsm");
    slate::trsm(sd, alpha, A, B, call_options());
}

template <typename T>
void psyrk(const char* uplo, const char* trans, int n, int k, T alpha, T* a, int ia, int ja,
           const int* desca, T beta, T* c, int ic, int jc, const int* descc)
{
    const slate::Op op = op_of(trans);
    const bool nt = (op == slate::Op::NoTrans);
    auto A = general<T>(nt ? n : k, nt ? k : n, a, ia, ja, desca);
    auto Cm = square<slate::SymmetricMatrix<T>>(n, c, ic, jc, descc, uplo_of(uplo));
    apply_op(A, op);
    slate_assert(A.mt() == Cm.mt());
    say("syrk");
    slate::syrk(alpha, A, beta, Cm, call_options());
}

template <typename T>
void pherk(const char* uplo, const char* trans, int n, int k, double alpha, T* a, int ia, int ja,
           const int* desca, double beta, T* c, int ic, int jc, const int* descc)
{
    const slate::Op op = op_of(trans);
    const bool nt = (op == slate::Op::NoTrans);
    auto A = general<T>(nt ? n : k, nt ? k : n, a, ia, ja, desca);
    auto Cm = square<slate::HermitianMatrix<T>>(n, c, ic, jc, descc, uplo_of(uplo));
    apply_op(A, op);
    slate_assert(A.mt() == Cm.mt());
    say("herk");
    slate::herk(alpha, A, beta, Cm, call_options());
}

template <typename T>
void phemm(bool hermitian, const char* side, const char* uplo, int m, int n, T alpha, T* a, int ia,
           int ja, const int* desca, T* b, int ib, int jb, const int* descb, T beta, T* c, int ic,
           int jc, const int* descc)
{
    const slate::Side sd = side_of(side);
    const int64_t an = (sd == slate::Side::Left) ? m : n;
    auto B = general<T>(m, n, b, ib, jb, descb);
    auto Cm = general<T>(m, n, c, ic, jc, descc);
    if (hermitian) {
        auto A = square<slate::HermitianMatrix<T>>(an, a, ia, ja, desca, uplo_of(uplo));
        slate_assert(A.mt() == (sd == slate::Side::Left ? Cm.mt() : Cm.nt()));
        say("hemm");
        slate::hemm(sd, alpha, A, B, beta, Cm, call_options());
    }
    else {
        auto A = square<slate::SymmetricMatrix<T>>(an, a, ia, ja, desca, uplo_of(uplo));
        slate_assert(A.mt() == (sd == slate::Side::Left ? Cm.mt() : Cm.nt()));
        say("symm");
        slate::symm(sd, alpha, A, B, beta, Cm, call_options());
    }
}

template <typename T>
void psyr2k(const char* uplo, const char* trans, int n, int k, T alpha, T* a, int ia, int ja,
            const int* desca, T* b, int ib, int jb, const int* descb, T beta, T* c, int ic, int jc,
            const int* descc)
{
    const slate::Op op = op_of(trans);
    const bool nt = (op == slate::Op::NoTrans);
    auto A = general<T>(nt ? n : k, nt ? k : n, a, ia, ja, desca);
    auto B = general<T>(nt ? n : k, nt ? k : n, b, ib, jb, descb);
    auto Cm = square<slate::SymmetricMatrix<T>>(n, c, ic, jc, descc, uplo_of(uplo));
    apply_op(A, op);
    apply_op(B, op);
    slate_assert(A.mt() == Cm.mt() && B.mt() == Cm.mt() && A.nt() == B.nt());
    say("syr2k");
    slate::syr2k(alpha, A, B, beta, Cm, call_options());
}

template <typename T>
void pher2k(const char* uplo, const char* trans, int n, int k, T alpha, T* a, int ia, int ja,
            const int* desca, T* b, int ib, int jb, const int* descb, double beta, T* c, int ic,
            int jc, const int* descc)
{
    const slate::Op op = op_of(trans);
    const bool nt = (op == slate::Op::NoTrans);
    auto A = general<T>(nt ? n : k, nt ? k : n, a, ia, ja, desca);
    auto B = general<T>(nt ? n : k, nt ? k : n, b, ib, jb, descb);
    auto Cm = square<slate::HermitianMatrix<T>>(n, c, ic, jc, descc, uplo_of(uplo));
    apply_op(A, op);
    apply_op(B, op);
    slate_assert(A.mt() == Cm.mt() && B.mt() == Cm.mt() && A.nt() == B.nt());
    say("her2k");
    slate::her2k(alpha, A, B, beta, Cm, call_options());
}

template <typename F>
void fortran_call(const char* name, F&& body)
{
    try {
        body();
        std::lock_guard<std::mutex> lock(g_mu);
        g_status = 0;
        g_error.clear();
    }
    catch (std::exception const& e) {
        std::lock_guard<std::mutex> lock(g_mu);
        g_status = 1;
        g_error = std::string(name) + ": " + e.what();
        std::fprintf(stderr, "slate_b200 scalapack_api: %s\n", g_error.c_str());
    }
}

}  // namespace

#define Z(p) (*(zdouble*)(p))
#define ZP(p) ((zdouble*)(p))

extern "C" {

int slate_scalapack_grid_set(int ctxt, char order, int nprow, int npcol)
{
    const char o = (char)std::toupper((unsigned char)order);
    if ((o != 'C' && o != 'R') || nprow < 1 || npcol < 1) return -1;
    std::lock_guard<std::mutex> lock(g_mu);
    g_grids[ctxt] = Grid{o == 'C' ? slate::GridOrder::Col : slate::GridOrder::Row, nprow, npcol};
    return 0;
}

int slate_scalapack_status(void) { return g_status; }
const char* slate_scalapack_last_error(void) { return g_error.c_str(); }

void slate_pdgemm(const char* ta, const char* tb, int* m, int* n, int* k, double* alpha, double* a,
                  int* ia, int* ja, int* desca, double* b, int* ib, int* jb, int* descb,
                  double* beta, double* c, int* ic, int* jc, int* descc)
{
    fortran_call("pdgemm", [&] {
        pgemm<double>(ta, tb, *m, *n, *k, *alpha, a, *ia, *ja, desca, b, *ib, *jb, descb, *beta, c,
                      *ic, *jc, descc);
    });
}
void slate_pzgemm(const char* ta, const char* tb, int* m, int* n, int* k, void* alpha, void* a,
                  int* ia, int* ja, int* desca, void* b, int* ib, int* jb, int* descb, void* beta,
                  void* c, int* ic, int* jc, int* descc)
{
    fortran_call("pzgemm", [&] {
        pgemm<zdouble>(ta, tb, *m, *n, *k, Z(alpha), ZP(a), *ia, *ja, desca, ZP(b), *ib, *jb, descb,
                       Z(beta), ZP(c), *ic, *jc, descc);
    });
}

void slate_pdpotrf(const char* uplo, int* n, double* a, int* ia, int* ja, int* desca, int* info)
{
    fortran_call("pdpotrf", [&] { ppotrf<double>(uplo, *n, a, *ia, *ja, desca, info); });
}
void slate_pzpotrf(const char* uplo, int* n, void* a, int* ia, int* ja, int* desca, int* info)
{
    fortran_call("pzpotrf", [&] { ppotrf<zdouble>(uplo, *n, ZP(a), *ia, *ja, desca, info); });
}

void slate_pdpotrs(const char* uplo, int* n, int* nrhs, double* a, int* ia, int* ja, int* desca,
                   double* b, int* ib, int* jb, int* descb, int* info)
{
    fortran_call("pdpotrs", [&] {
        ppotrs<double>(uplo, *n, *nrhs, a, *ia, *ja, desca, b, *ib, *jb, descb, info);
    });
}
void slate_pzpotrs(const char* uplo, int* n, int* nrhs, void* a, int* ia, int* ja, int* desca,
                   void* b, int* ib, int* jb, int* descb, int* info)
{
    fortran_call("pzpotrs", [&] {
        ppotrs<zdouble>(uplo, *n, *nrhs, ZP(a), *ia, *ja, desca, ZP(b), *ib, *jb, descb, info);
    });
}

void slate_pdposv(const char* uplo, int* n, int* nrhs, double* a, int* ia, int* ja, int* desca,
                  double* b, int* ib, int* jb, int* descb, int* info)
{
    fortran_call("pdposv", [&] {
        pposv<double>(uplo, *n, *nrhs, a, *ia, *ja, desca, b, *ib, *jb, descb, info);
    });
}
void slate_pzposv(const char* uplo, int* n, int* nrhs, void* a, int* ia, int* ja, int* desca,
                  void* b, int* ib, int* jb, int* descb, int* info)
{
    fortran_call("pzposv", [&] {
        pposv<zdouble>(uplo, *n, *nrhs, ZP(a), *ia, *ja, desca, ZP(b), *ib, *jb, descb, info);
    });
}

void slate_pdtrsm(const char* side, const char* uplo, const char* ta, const char* diag, int* m,
                  int* n, double* alpha, double* a, int* ia, int* ja, int* desca, double* b,
                  int* ib, int* jb, int* descb)
{
    fortran_call("pdtrsm", [&] {
        ptrsm<double>(side, uplo, ta, diag, *m, *n, *alpha, a, *ia, *ja, desca, b, *ib, *jb, descb);
    });
}
void slate_pztrsm(const char* side, const char* uplo, const char* ta, const char* diag, int* m,
                  int* n, void* alpha, void* a, int* ia, int* ja, int* desca, void* b, int* ib,
                  int* jb, int* descb)
{
    fortran_call("pztrsm", [&] {
        ptrsm<zdouble>(side, uplo, ta, diag, *m, *n, Z(alpha), ZP(a), *ia, *ja, desca, ZP(b), *ib,
                       *jb, descb);
    });
}

void slate_pdsyrk(const char* uplo, const char* trans, int* n, int* k, double* alpha, double* a,
                  int* ia, int* ja, int* desca, double* beta, double* c, int* ic, int* jc,
                  int* descc)
{
    fortran_call("pdsyrk", [&] {
        psyrk<double>(uplo, trans, *n, *k, *alpha, a, *ia, *ja, desca, *beta, c, *ic, *jc, descc);
    });
}
void slate_pzsyrk(const char* uplo, const char* trans, int* n, int* k, void* alpha, void* a,
                  int* ia, int* ja, int* desca, void* beta, void* c, int* ic, int* jc, int* descc)
{
    fortran_call("pzsyrk", [&] {
        psyrk<zdouble>(uplo, trans, *n, *k, Z(alpha), ZP(a), *ia, *ja, desca, Z(beta), ZP(c), *ic,
                       *jc, descc);
    });
}
void slate_pzherk(const char* uplo, const char* trans, int* n, int* k, double* alpha, void* a,
                  int* ia, int* ja, int* desca, double* beta, void* c, int* ic, int* jc,
                  int* descc)
{
    fortran_call("pzherk", [&] {
        pherk<zdouble>(uplo, trans, *n, *k, *alpha, ZP(a), *ia, *ja, desca, *beta, ZP(c), *ic, *jc,
                       descc);
    });
}

void slate_pdsymm(const char* side, const char* uplo, int* m, int* n, double* alpha, double* a,
                  int* ia, int* ja, int* desca, double* b, int* ib, int* jb, int* descb, double* beta,
                  double* c, int* ic, int* jc, int* descc)
{
    fortran_call("pdsymm", [&] {
        phemm<double>(false, side, uplo, *m, *n, *alpha, a, *ia, *ja, desca, b, *ib, *jb, descb, *beta,
                      c, *ic, *jc, descc);
    });
}
void slate_pzsymm(const char* side, const char* uplo, int* m, int* n, void* alpha, void* a, int* ia,
                  int* ja, int* desca, void* b, int* ib, int* jb, int* descb, void* beta, void* c,
                  int* ic, int* jc, int* descc)
{
    fortran_call("pzsymm", [&] {
        phemm<zdouble>(false, side, uplo, *m, *n, Z(alpha), ZP(a), *ia, *ja, desca, ZP(b), *ib, *jb,
                       descb, Z(beta), ZP(c), *ic, *jc, descc);
    });
}
void slate_pzhemm(const char* side, const char* uplo, int* m, int* n, void* alpha, void* a, int* ia,
                  int* ja, int* desca, void* b, int* ib, int* jb, int* descb, void* beta, void* c,
                  int* ic, int* jc, int* descc)
{
    fortran_call("pzhemm", [&] {
        phemm<zdouble>(true, side, uplo, *m, *n, Z(alpha), ZP(a), *ia, *ja, desca, ZP(b), *ib, *jb,
                       descb, Z(beta), ZP(c), *ic, *jc, descc);
    });
}

void slate_pdsyr2k(const char* uplo, const char* trans, int* n, int* k, double* alpha, double* a,
                   int* ia, int* ja, int* desca, double* b, int* ib, int* jb, int* descb,
                   double* beta, double* c, int* ic, int* jc, int* descc)
{
    fortran_call("pdsyr2k", [&] {
        psyr2k<double>(uplo, trans, *n, *k, *alpha, a, *ia, *ja, desca, b, *ib, *jb, descb, *beta, c,
                       *ic, *jc, descc);
    });
}
void slate_pzsyr2k(const char* uplo, const char* trans, int* n, int* k, void* alpha, void* a, int* ia,
                   int* ja, int* desca, void* b, int* ib, int* jb, int* descb, void* beta, void* c,
                   int* ic, int* jc, int* descc)
{
    fortran_call("pzsyr2k", [&] {
        psyr2k<zdouble>(uplo, trans, *n, *k, Z(alpha), ZP(a), *ia, *ja, desca, ZP(b), *ib, *jb, descb,
                        Z(beta), ZP(c), *ic, *jc, descc);
    });
}
void slate_pzher2k(const char* uplo, const char* trans, int* n, int* k, void* alpha, void* a, int* ia,
                   int* ja, int* desca, void* b, int* ib, int* jb, int* descb, double* beta, void* c,
                   int* ic, int* jc, int* descc)
{
    fortran_call("pzher2k", [&] {
        pher2k<zdouble>(uplo, trans, *n, *k, Z(alpha), ZP(a), *ia, *ja, desca, ZP(b), *ib, *jb, descb,
                        *beta, ZP(c), *ic, *jc, descc);
    });
}

}  // extern "C"
