"""Parity of the host C++ mirror (slate::gemm, slate::potrf, internal::gemm/herk/syrk/
trsm/potrf on Target::Devices) against the CPU oracle and, when present, the unmodified
reference (oracle/_ref), on the same matgen inputs.  Mirrors the reference's
unit_test/test_internal_blas.cc (27 op combinations, ld = m + 1 LAPACK views,
3*sqrt(k)*eps) and the tester metrics of test/test_gemm.cc / test_posv.cc.  Needs a GPU.
"""
import numpy as np
import pytest

import oracle_lib as ol

pytestmark = pytest.mark.gpu
EPS = np.finfo(np.float64).eps
ALPHA, BETA = 3.141592653589793, 2.718281828459045


@pytest.fixture(scope="module")
def sl():
    from slate_d35_b200 import host
    host.initialize(0)
    yield host
    host.sync()


def opn(a, op):
    return {"N": a, "T": a.T, "C": a.conj().T}[op]


@pytest.mark.parametrize("opa", "NTC")
@pytest.mark.parametrize("opb", "NTC")
@pytest.mark.parametrize("opc", "NTC")
def test_internal_gemm_27_ops(sl, opa, opb, opc):
    """unit_test/test_internal_blas.cc:282-450: m=80, n=64, k=16, nb=16, ld = m+1."""
    m, n, k, nb = 80, 64, 16, 16
    ash = (m, k) if opa == "N" else (k, m)
    bsh = (k, n) if opb == "N" else (n, k)
    csh = (m, n) if opc == "N" else (n, m)
    Ab = ol.matgen(1, ash[0] + 1, ash[1], dist=2)
    Bb = ol.matgen(2, bsh[0] + 1, bsh[1], dist=2)
    Cb = ol.matgen(3, csh[0] + 1, csh[1], dist=2)
    A0, B0, C0 = Ab[:ash[0]].copy(), Bb[:bsh[0]].copy(), Cb[:csh[0]].copy()
    A = sl.Matrix.fromLAPACK(Ab, nb, m=ash[0])
    B = sl.Matrix.fromLAPACK(Bb, nb, m=bsh[0])
    Cm = sl.Matrix.fromLAPACK(Cb, nb, m=csh[0])
    sl.internal_gemm(ALPHA, A, B, BETA, Cm, opa, opb, opc)
    got = Cm.to_numpy()
    want = ALPHA * opn(A0, opa) @ opn(B0, opb) + BETA * opn(C0, opc)
    want = opn(want, opc)              # back to storage orientation
    assert np.abs(got - want).max() <= 3 * np.sqrt(k) * EPS * max(1.0, np.abs(want).max())
    assert np.array_equal(Cb[csh[0]:], ol.matgen(3, csh[0] + 1, csh[1], dist=2)[csh[0]:])


@pytest.mark.parametrize("kind", ["herk", "syrk"])
@pytest.mark.parametrize("uplo", "LU")
@pytest.mark.parametrize("opa", "NT")
def test_internal_herk_syrk(sl, kind, uplo, opa):
    """unit_test/test_internal_blas.cc:454-796, vs the oracle's tile herk/syrk."""
    n, k, nb = 80, 16, 16
    ash = (n, k) if opa == "N" else (k, n)
    A0 = ol.matgen(4, *ash, dist=2)
    C0 = ol.matgen(5, n, n, dist=2)
    A = sl.Matrix(ash[0], ash[1], nb).from_numpy(A0)
    cls = sl.HermitianMatrix if kind == "herk" else sl.SymmetricMatrix
    Cm = cls(uplo, n, nb).from_numpy(C0)
    if kind == "herk":
        sl.internal_herk(-1.0, A, 1.0, Cm, opa)
    else:
        sl.internal_syrk(-1.0, A, 1.0, Cm, opa)
    got = Cm.to_numpy()
    full = -opn(A0, opa) @ opn(A0, opa).T + C0
    tri = np.tril if uplo == "L" else np.triu
    assert np.abs(tri(got) - tri(full)).max() <= 3 * np.sqrt(k) * EPS * np.abs(full).max()


def test_internal_trsm_and_potrf(sl):
    nb = 96
    T0 = ol.matgen(16, nb, nb, dist=2)
    T0[np.arange(nb), np.arange(nb)] += 20
    B0 = ol.matgen(17, 250, nb, dist=2)
    T = sl.TriangularMatrix("L", "N", nb, nb).from_numpy(T0)
    B = sl.Matrix(250, nb, nb).from_numpy(B0)
    sl.internal_trsm("R", 1.0, T, B, opa="C")
    want = B0.copy(order="F")
    ol.oracle().oracle_tile_dtrsm(ord("R"), ord("L"), ord("C"), ord("N"), 250, nb, 1.0, ol.P(T0), nb,
                                  ol.P(want), 250)
    assert np.abs(B.to_numpy() - want).max() <= 50 * EPS * np.sqrt(nb) * np.abs(want).max()
    H0 = ol.matgen(18, nb, nb)
    H0[np.arange(nb), np.arange(nb)] += nb
    H = sl.HermitianMatrix("L", nb, nb).from_numpy(H0)
    assert sl.internal_potrf(H) == 0
    want = H0.copy(order="F")
    assert ol.oracle().oracle_tile_dpotrf(ord("L"), nb, ol.P(want), nb) == 0
    assert np.abs(np.tril(H.to_numpy()) - np.tril(want)).max() <= 50 * EPS * np.abs(want).max()


@pytest.mark.parametrize("origin", ["D", "H"])
@pytest.mark.parametrize("ops", ["NN", "TN", "NT", "CC"])
def test_gemm_driver_vs_oracle(sl, origin, ops):
    """config 1 scaled down: ragged 5x5 tiles, tester alpha/beta (test/test.cc:392-393)."""
    opa, opb = ops
    m, n, k, nb = 600, 520, 450, 128
    ash = (m, k) if opa == "N" else (k, m)
    bsh = (k, n) if opb == "N" else (n, k)
    A0, B0, C0 = ol.matgen(1, *ash), ol.matgen(2, *bsh), ol.matgen(3, m, n)
    if origin == "D":
        A = sl.Matrix(ash[0], ash[1], nb).generate(1)
        B = sl.Matrix(bsh[0], bsh[1], nb).generate(2)
        Cm = sl.Matrix(m, n, nb).generate(3)
        assert np.array_equal(A.to_numpy(), A0)          # device matgen == host matgen
    else:
        Ah, Bh, Ch = A0.copy(order="F"), B0.copy(order="F"), C0.copy(order="F")
        A, B, Cm = sl.Matrix.fromLAPACK(Ah, nb), sl.Matrix.fromLAPACK(Bh, nb), sl.Matrix.fromLAPACK(Ch, nb)
    sl.gemm(ALPHA, A, B, BETA, Cm, opa, opb)
    got = Cm.to_numpy() if origin == "D" else Ch
    want = C0.copy(order="F")
    ol.oracle().oracle_dgemm(ord(opa), ord(opb), m, n, k, nb, ALPHA, ol.P(A0), ash[0], ol.P(B0),
                             bsh[0], BETA, ol.P(want), m)
    err = np.abs(got - want).max() / np.abs(want).max()
    assert err <= 3 * np.sqrt(k) * EPS, err
    if ol.have_ref() and ops == "NN":
        ref = C0.copy(order="F")
        ol.ref().ref_dgemm(ord(opa), ord(opb), m, n, k, nb, ALPHA, ol.P(A0), ash[0], ol.P(B0),
                           bsh[0], BETA, ol.P(ref), m, 1)
        # north star: relative error against the HostTask path within 50 eps
        assert np.linalg.norm(got - ref) / np.linalg.norm(ref) <= 50 * EPS


def test_gemm_tester_residual(sl):
    """test/test_gemm.cc:136-158, 196-212: ||C X - (alpha A (B X) + beta C0 X)||_1 / ||Y||_1 <= 3 eps
    with X of kind rand (uniform [0,1)) as the tester generates it."""
    n, nb, nrhs = 1000, 256, 10
    A = sl.Matrix(n, n, nb).generate(1)
    B = sl.Matrix(n, n, nb).generate(2)
    Cm = sl.Matrix(n, n, nb).generate(3)
    X = sl.Matrix(n, nrhs, nb).generate(4)
    W = sl.Matrix(n, nrhs, nb)
    Y = sl.Matrix(n, nrhs, nb)
    sl.gemm(1.0, B, X, 0.0, W)                 # W = B X
    sl.gemm(BETA, Cm, X, 0.0, Y)               # Y = beta C0 X
    sl.gemm(ALPHA, A, W, 1.0, Y)               # Y += alpha A W
    sl.gemm(ALPHA, A, B, BETA, Cm)
    Z = sl.Matrix(n, nrhs, nb)
    sl.gemm(1.0, Cm, X, 0.0, Z)
    y, z = Y.to_numpy(), Z.to_numpy()
    assert np.abs(z - y).sum(axis=0).max() / np.abs(y).sum(axis=0).max() <= 3 * EPS


@pytest.mark.parametrize("origin", ["D", "H"])
@pytest.mark.parametrize("uplo", "LU")
@pytest.mark.parametrize("la", [0, 1, 2])
def test_potrf_driver_vs_oracle(sl, origin, uplo, la):
    n, nb = 1000, 128
    A0 = ol.matgen(7, n, n)
    A0[np.arange(n), np.arange(n)] += n                   # rand_dominant, test_posv.cc:124
    if origin == "D":
        A = sl.HermitianMatrix(uplo, n, nb).from_numpy(A0)
    else:
        Ah = A0.copy(order="F")
        A = sl.HermitianMatrix.fromLAPACK(uplo, Ah, nb)
    info = sl.potrf(A, lookahead=la)
    assert info == 0
    got = A.to_numpy() if origin == "D" else Ah
    want = A0.copy(order="F")
    assert ol.oracle().oracle_dpotrf(ord(uplo), n, nb, ol.P(want), n) == 0
    tri = np.tril if uplo == "L" else np.triu
    assert np.abs(tri(got) - tri(want)).max() / np.abs(tri(want)).max() <= 50 * EPS
    # north-star metric, dense evaluation: ||A - L L^H|| / (n ||A||) <= 50 eps
    F = tri(got)
    S = np.tril(A0) + np.tril(A0, -1).T if uplo == "L" else np.triu(A0) + np.triu(A0, 1).T
    R = S - (F @ F.T if uplo == "L" else F.T @ F)
    assert np.linalg.norm(R) / (n * np.linalg.norm(S)) <= 50 * EPS
    if origin == "H":
        other = (lambda x: np.triu(x, 1)) if uplo == "L" else (lambda x: np.tril(x, -1))
        assert np.array_equal(other(got), other(A0)), "the other triangle is not referenced"
    if ol.have_ref() and la == 1:
        # north star: against the reference's own HostTask path on the same matgen matrix
        ref = A0.copy(order="F")
        assert ol.ref().ref_dpotrf(ord(uplo), n, nb, ol.P(ref), n, 1) == 0
        assert np.linalg.norm(tri(got) - tri(ref)) / np.linalg.norm(tri(ref)) <= 50 * EPS


def test_potrf_info_not_positive_definite(sl):
    n, nb = 700, 128
    A0 = ol.matgen(7, n, n)
    A0[np.arange(n), np.arange(n)] += n
    A0[300, 300] = -5.0
    A = sl.HermitianMatrix("L", n, nb).from_numpy(A0)
    want = A0.copy(order="F")
    assert sl.potrf(A) == ol.oracle().oracle_dpotrf(ord("L"), n, nb, ol.P(want), n) == 301


def test_potrf_device_generate_and_residual(sl):
    """device-origin, generated on the GPU, checked with the engine's own residual."""
    n, nb = 2048, 256
    A = sl.HermitianMatrix("L", n, nb).generate(11, dominant=True)
    assert sl.potrf(A) == 0
    r, a0 = sl.potrf_residual(A, 11, True)
    assert np.sqrt(r) / (n * np.sqrt(a0)) <= 50 * EPS
    # and the same number evaluated densely on the host
    F = np.tril(A.to_numpy())
    A0 = ol.matgen(11, n, n)
    A0[np.arange(n), np.arange(n)] += n
    S = np.tril(A0) + np.tril(A0, -1).T
    dense = np.linalg.norm(S - F @ F.T)
    # both are pure rounding noise, accumulated in different orders: same magnitude
    assert 0.5 * dense <= np.sqrt(r) <= 2.0 * dense
    assert abs(np.sqrt(a0) - np.linalg.norm(S)) <= 1e-9 * np.linalg.norm(S)


def test_host_targets_are_not_built(sl):
    A = sl.Matrix(64, 64, 32).generate(1)
    with pytest.raises(sl.SlateError, match="Target::Devices only"):
        sl.gemm(1.0, A, A, 0.0, A, target="T")
    H = sl.HermitianMatrix("L", 64, 32).generate(1, dominant=True)
    with pytest.raises(sl.SlateError, match="Target::Devices only"):
        sl.potrf(H, target="T")


def test_pool_returns_to_level_after_gemm(sl):
    """workspace blocks taken for host-origin staging all come back (deferred release)."""
    n, nb = 512, 128
    Ah, Bh, Ch = ol.matgen(1, n, n), ol.matgen(2, n, n), ol.matgen(3, n, n)
    A, B, Cm = sl.Matrix.fromLAPACK(Ah, nb), sl.Matrix.fromLAPACK(Bh, nb), sl.Matrix.fromLAPACK(Ch, nb)
    sl.gemm(1.0, A, B, 1.0, Cm)
    for M in (A, B, Cm):
        avail, cap = M.pool_stats()
        assert avail == cap, (avail, cap)


# ------------------------------ complex-FP64 drivers ------------------------------
ZALPHA, ZBETA = ALPHA + 1.4142135623730951j, BETA + 1.7320508075688772j    # test/test.cc:392-393


@pytest.mark.parametrize("ops", ["NN", "CN", "NT", "CC"])
def test_zgemm_driver_vs_oracle(sl, ops):
    """config 5 scaled down: nb = 96 with a ragged last tile in m, n and k."""
    opa, opb = ops
    m, n, k, nb = 500, 420, 330, 96
    ash = (m, k) if opa == "N" else (k, m)
    bsh = (k, n) if opb == "N" else (n, k)
    A0 = ol.matgen(1, *ash, np.complex128)
    B0 = ol.matgen(2, *bsh, np.complex128)
    C0 = ol.matgen(3, m, n, np.complex128)
    A = sl.Matrix(ash[0], ash[1], nb, dtype=np.complex128).generate(1)
    B = sl.Matrix(bsh[0], bsh[1], nb, dtype=np.complex128).generate(2)
    Cm = sl.Matrix(m, n, nb, dtype=np.complex128).generate(3)
    assert np.array_equal(A.to_numpy(), A0)
    sl.gemm(ZALPHA, A, B, ZBETA, Cm, opa, opb)
    want = C0.copy(order="F")
    ol.oracle().oracle_zgemm(ord(opa), ord(opb), m, n, k, nb, ol.Z(ZALPHA), ol.P(A0), ash[0], ol.P(B0),
                             bsh[0], ol.Z(ZBETA), ol.P(want), m)
    err = np.abs(Cm.to_numpy() - want).max() / np.abs(want).max()
    assert err <= 3 * np.sqrt(4 * k) * EPS, err


@pytest.mark.parametrize("uplo", "LU")
def test_zpotrf_driver_vs_oracle(sl, uplo):
    n, nb = 700, 96
    A0 = ol.matgen(7, n, n, np.complex128)
    A0[np.arange(n), np.arange(n)] += n
    A = sl.HermitianMatrix(uplo, n, nb, dtype=np.complex128).from_numpy(A0)
    assert sl.potrf(A) == 0
    want = A0.copy(order="F")
    assert ol.oracle().oracle_zpotrf(ord(uplo), n, nb, ol.P(want), n) == 0
    tri = np.tril if uplo == "L" else np.triu
    got = A.to_numpy()
    assert np.abs(tri(got) - tri(want)).max() / np.abs(tri(want)).max() <= 50 * EPS
    F = tri(got)
    D = np.diag(A0).real
    S = (np.tril(A0, -1) + np.tril(A0, -1).conj().T if uplo == "L"
         else np.triu(A0, 1) + np.triu(A0, 1).conj().T) + np.diag(D)
    R = S - (F @ F.conj().T if uplo == "L" else F.conj().T @ F)
    assert np.linalg.norm(R) / (n * np.linalg.norm(S)) <= 50 * EPS


# ------------------------------ herk / syrk drivers ------------------------------
@pytest.mark.parametrize("kind", ["herk", "syrk"])
@pytest.mark.parametrize("uplo", "LU")
@pytest.mark.parametrize("dt", [np.float64, np.complex128])
def test_herk_syrk_driver(sl, kind, uplo, dt):
    """test/test_herk.cc / test_syrk.cc shapes scaled down: n = 520, k = 330, nb = 128 (ragged)."""
    n, k, nb = 520, 330, 128
    A0 = ol.matgen(1, n, k, dt, dist=2)
    C0 = ol.matgen(2, n, n, dt, dist=2)
    if kind == "herk":
        C0[np.arange(n), np.arange(n)] = C0[np.arange(n), np.arange(n)].real
    A = sl.Matrix(n, k, nb, dtype=dt).from_numpy(A0)
    cls = sl.HermitianMatrix if kind == "herk" else sl.SymmetricMatrix
    Cm = cls(uplo, n, nb, dtype=dt).from_numpy(C0)
    if kind == "herk":
        alpha, beta = -1.25, 0.5
        sl.herk(alpha, A, beta, Cm)
        full = alpha * A0 @ A0.conj().T + beta * C0
    else:
        alpha, beta = (-1.25 + 0.5j, 0.5 - 0.25j) if dt == np.complex128 else (-1.25, 0.5)
        sl.syrk(alpha, A, beta, Cm)
        full = alpha * A0 @ A0.T + beta * C0
    got = Cm.to_numpy()
    tri = np.tril if uplo == "L" else np.triu
    kk = k * (4 if dt == np.complex128 else 1)
    assert np.abs(tri(got) - tri(full)).max() <= 3 * np.sqrt(kk) * EPS * np.abs(full).max()


# ------------------------------ gemmA ------------------------------
@pytest.mark.parametrize("dt", [np.float64, np.complex128])
@pytest.mark.parametrize("opa,opb", [("N", "N"), ("T", "N"), ("C", "N"), ("N", "T")])
@pytest.mark.parametrize("beta", [0.0, 1.0, 2.718281828459045])
def test_gemmA_driver(sl, dt, opa, opb, beta):
    """slate::gemmA (src/gemmA.cc) against gemmC and numpy: several block columns of B, ragged
    tiles, beta = 0 must not read C (it is filled with NaN)."""
    m, n, k, nb = 700, 300, 520, 128
    ash = (m, k) if opa == "N" else (k, m)
    bsh = (k, n) if opb == "N" else (n, k)
    A0, B0, C0 = ol.matgen(1, *ash, dt, dist=2), ol.matgen(2, *bsh, dt, dist=2), ol.matgen(3, m, n, dt, dist=2)
    alpha = (1.5 - 0.5j) if dt == np.complex128 else 3.141592653589793
    b = (beta * (1 + 0.5j)) if dt == np.complex128 else beta
    Cstart = np.full_like(C0, np.nan) if beta == 0.0 else C0
    A = sl.Matrix(*ash, nb, dtype=dt).from_numpy(A0)
    B = sl.Matrix(*bsh, nb, dtype=dt).from_numpy(B0)
    out = {}
    for method in "AC":
        Cm = sl.Matrix(m, n, nb, dtype=dt).from_numpy(Cstart)
        sl.gemm(alpha, A, B, b, Cm, opa=opa, opb=opb, method=method)
        out[method] = Cm.to_numpy()
    want = alpha * opn(A0, opa) @ opn(B0, opb) + (b * C0 if beta != 0.0 else 0)
    tol = 3 * np.sqrt(k) * EPS * np.abs(want).max() * (4 if dt == np.complex128 else 1)
    assert np.abs(out["A"] - want).max() <= tol
    assert np.abs(out["C"] - want).max() <= tol


def test_gemm_auto_selects_gemmA_for_one_block_column(sl):
    """method.hh:86-98: B.nt() < 2 -> gemmA.  Both routes must agree with the tester's check."""
    m, k, nb, nrhs = 900, 640, 128, 10
    A0, X0 = ol.matgen(1, m, k), ol.matgen(6, k, nrhs)
    A = sl.Matrix(m, k, nb).from_numpy(A0)
    X = sl.Matrix(k, nrhs, nb).from_numpy(X0)
    Y = sl.Matrix(m, nrhs, nb).from_numpy(np.zeros((m, nrhs)))
    sl.gemm(1.0, A, X, 0.0, Y)
    assert np.abs(Y.to_numpy() - A0 @ X0).max() <= 3 * np.sqrt(k) * EPS * np.abs(A0 @ X0).max()


# ------------------------------ her2k / syr2k ------------------------------
@pytest.mark.parametrize("kind", ["her2k", "syr2k"])
@pytest.mark.parametrize("uplo", "LU")
@pytest.mark.parametrize("opa", "NC")
@pytest.mark.parametrize("dt", [np.float64, np.complex128])
@pytest.mark.parametrize("origin", "DH")
def test_her2k_syr2k_driver(sl, kind, uplo, opa, dt, origin):
    """test/test_her2k.cc / test_syr2k.cc scaled down (ragged tiles, complex alpha, both
    triangles, op(A) = A or its (conjugate) transpose); the diagonal of a Hermitian C comes
    back exactly real."""
    n, k, nb = 520, 330, 128
    if kind == "syr2k" and opa == "C":
        opa = "T"
    ash = (n, k) if opa == "N" else (k, n)
    A0 = ol.matgen(1, *ash, dt, dist=2)
    B0 = ol.matgen(5, *ash, dt, dist=2)
    C0 = ol.matgen(2, n, n, dt, dist=2)
    if kind == "her2k":
        C0[np.arange(n), np.arange(n)] = C0[np.arange(n), np.arange(n)].real
    cls = sl.HermitianMatrix if kind == "her2k" else sl.SymmetricMatrix
    if origin == "D":
        A = sl.Matrix(*ash, nb, dtype=dt).from_numpy(A0)
        B = sl.Matrix(*ash, nb, dtype=dt).from_numpy(B0)
        Cm = cls(uplo, n, nb, dtype=dt).from_numpy(C0)
    else:
        # user-owned host arrays: the result is copied back behind the LAST launch of the update
        a_h, b_h, c_h = A0.copy(order="F"), B0.copy(order="F"), C0.copy(order="F")
        A, B = sl.Matrix.fromLAPACK(a_h, nb), sl.Matrix.fromLAPACK(b_h, nb)
        Cm = cls.fromLAPACK(uplo, c_h, nb)
    Aop, Bop = opn(A0, opa), opn(B0, opa)
    alpha = (-1.25 + 0.5j) if dt == np.complex128 else -1.25
    if kind == "her2k":
        beta = 0.5
        sl.her2k(alpha, A, B, beta, Cm, opa=opa)
        full = alpha * Aop @ Bop.conj().T + np.conj(alpha) * Bop @ Aop.conj().T + beta * C0
    else:
        beta = (0.5 - 0.25j) if dt == np.complex128 else 0.5
        sl.syr2k(alpha, A, B, beta, Cm, opa=opa)
        full = alpha * Aop @ Bop.T + alpha * Bop @ Aop.T + beta * C0
    got = Cm.to_numpy() if origin == "D" else c_h
    tri = np.tril if uplo == "L" else np.triu
    kk = 2 * k * (4 if dt == np.complex128 else 1)
    assert np.abs(tri(got) - tri(full)).max() <= 3 * np.sqrt(kk) * EPS * np.abs(full).max()
    if kind == "her2k" and dt == np.complex128:
        assert not np.diag(got).imag.any()


# ------------------------------ RowMajor tiles ------------------------------
@pytest.mark.parametrize("dt", [np.float64, np.complex128])
def test_tile_layout_convert_round_trip_and_gemm(sl, dt):
    """tileLayoutConvert (src/cuda/device_transpose.cu behind it): square tiles in place,
    ragged edge tiles out of place; the matrix reads back unchanged, and a gemm on RowMajor
    tiles (converted back by the internal op, LayoutConvert::ColMajor) gives the same bits as
    on untouched tiles."""
    m, n, k, nb = 300, 200, 280, 128                 # last tile row / column ragged
    A0, B0, C0 = ol.matgen(1, m, k, dt, dist=2), ol.matgen(2, k, n, dt, dist=2), ol.matgen(3, m, n, dt, dist=2)
    A = sl.Matrix(m, k, nb, dtype=dt).from_numpy(A0)
    mt, kt = -(-m // nb), -(-k // nb)
    for i in range(mt):
        for j in range(kt):
            assert A.tile_layout(i, j) == "C"
            A.tile_layout_convert(i, j, "R")
            assert A.tile_layout(i, j) == "R"
    assert np.array_equal(A.to_numpy(), A0)            # reading converts back
    assert A.tile_layout(0, 0) == "C"
    # twice in a row, through both layouts
    A.tile_layout_convert(mt - 1, kt - 1, "R").tile_layout_convert(mt - 1, kt - 1, "C")
    assert np.array_equal(A.to_numpy(), A0)

    out = []
    for rowmajor in (False, True):
        A = sl.Matrix(m, k, nb, dtype=dt).from_numpy(A0)
        B = sl.Matrix(k, n, nb, dtype=dt).from_numpy(B0)
        Cm = sl.Matrix(m, n, nb, dtype=dt).from_numpy(C0)
        if rowmajor:
            for M, rows, cols in ((A, m, k), (B, k, n), (Cm, m, n)):
                for i in range(-(-rows // nb)):
                    for j in range(-(-cols // nb)):
                        M.tile_layout_convert(i, j, "R")
        sl.gemm(ALPHA, A, B, BETA, Cm)
        out.append(Cm.to_numpy())
    assert np.array_equal(out[0], out[1])
    want = ALPHA * A0 @ B0 + BETA * C0
    assert np.abs(out[1] - want).max() <= 12 * np.sqrt(k) * EPS * np.abs(want).max()


# ------------------------------ getrf_nopiv ------------------------------
@pytest.mark.parametrize("dt", [np.float64, np.complex128])
@pytest.mark.parametrize("m,n", [(900, 900), (700, 500), (500, 700), (100, 100)])
@pytest.mark.parametrize("la", [0, 1, 2])
def test_getrf_nopiv_driver(sl, dt, m, n, la):
    """test/test_gesv.cc with the nopiv method, scaled down: ragged tiles, rectangular shapes,
    several lookahead depths; backward error || A - L U || / (n ||A||)."""
    nb = 128
    A0 = ol.matgen(61, m, n, dt, dist=2)
    mn = min(m, n)
    A0[np.arange(mn), np.arange(mn)] += max(m, n)
    A = sl.Matrix(m, n, nb, dtype=dt).from_numpy(A0)
    assert sl.getrf_nopiv(A, lookahead=la) == 0
    got = A.to_numpy()
    Lf = np.tril(got[:, :mn], -1) + np.eye(m, mn)
    Uf = np.triu(got[:mn, :])
    assert np.linalg.norm(Lf @ Uf - A0) / (max(m, n) * np.linalg.norm(A0)) <= 50 * EPS


def test_getrf_nopiv_info(sl):
    n, nb = 600, 128
    A0 = np.asfortranarray(np.eye(n) * 3.0 + np.triu(np.ones((n, n)), 1))
    A0[333, 333] = 0.0
    A = sl.Matrix(n, n, nb).from_numpy(A0)
    assert sl.getrf_nopiv(A) == 334


# ------------------------------ hemm / symm ------------------------------
@pytest.mark.parametrize("kind", ["hemm", "symm"])
@pytest.mark.parametrize("side", "LR")
@pytest.mark.parametrize("uplo", "LU")
@pytest.mark.parametrize("dt", [np.float64, np.complex128])
def test_hemm_symm_driver(sl, kind, side, uplo, dt):
    """test/test_hemm.cc / test_symm.cc scaled down: both sides and triangles, ragged tiles,
    complex alpha and beta; the triangle of A that is not stored holds NaN."""
    na, other, nb = 520, 330, 128
    m, n = (na, other) if side == "L" else (other, na)
    A0 = ol.matgen(41, na, na, dt, dist=2)
    if kind == "hemm":
        S = (A0 + A0.conj().T) / 2
    else:
        S = (A0 + A0.T) / 2
    tri = np.tril if uplo == "L" else np.triu
    other_strict = (lambda x: np.triu(x, 1)) if uplo == "L" else (lambda x: np.tril(x, -1))
    stored = tri(S).astype(dt) + other_strict(np.full((na, na), np.nan))
    B0, C0 = ol.matgen(42, m, n, dt, dist=2), ol.matgen(43, m, n, dt, dist=2)
    cls = sl.HermitianMatrix if kind == "hemm" else sl.SymmetricMatrix
    A = cls(uplo, na, nb, dtype=dt).from_numpy(stored)
    B = sl.Matrix(m, n, nb, dtype=dt).from_numpy(B0)
    Cm = sl.Matrix(m, n, nb, dtype=dt).from_numpy(C0)
    alpha, beta = ((1.5 - 0.5j), (0.25 + 2j)) if dt == np.complex128 else (3.141592653589793, 2.718281828459045)
    sl.hemm(side, alpha, A, B, beta, Cm)
    want = alpha * (S @ B0 if side == "L" else B0 @ S) + beta * C0
    got = Cm.to_numpy()
    kk = na * (4 if dt == np.complex128 else 1)
    assert np.abs(got - want).max() <= 3 * np.sqrt(kk) * EPS * np.abs(want).max()


# ------------------------------ trsm / potrs / posv drivers ------------------------------
@pytest.mark.parametrize("side", "LR")
@pytest.mark.parametrize("uplo", "LU")
@pytest.mark.parametrize("op", "NTC")
@pytest.mark.parametrize("dt", [np.float64, np.complex128])
def test_trsm_driver(sl, side, uplo, op, dt):
    """test/test_trsm.cc scaled down: every side / uplo / op, ragged tiles, alpha != 1."""
    nb = 96
    na, other = 420, 330
    m, n = (na, other) if side == "L" else (other, na)
    A0 = ol.matgen(21, na, na, dt, dist=2)
    A0[np.arange(na), np.arange(na)] += 2 * np.sqrt(na)
    B0 = ol.matgen(22, m, n, dt, dist=2)
    A = sl.TriangularMatrix(uplo, "N", na, nb, dtype=dt).from_numpy(A0)
    B = sl.Matrix(m, n, nb, dtype=dt).from_numpy(B0)
    alpha = 0.75 - 0.5j if dt == np.complex128 else 0.75
    sl.trsm(side, alpha, A, B, opa=op)
    X = B.to_numpy()
    T = np.tril(A0) if uplo == "L" else np.triu(A0)
    opT = {"N": T, "T": T.T, "C": T.conj().T}[op]
    lhs = opT @ X if side == "L" else X @ opT
    # residual form of the tester: || op(A) X - alpha B || / (||A|| ||X|| n)
    err = np.linalg.norm(lhs - alpha * B0, 1) / (np.linalg.norm(T, 1) * np.linalg.norm(X, 1) * na)
    assert err <= 3 * EPS, err


@pytest.mark.parametrize("uplo", "LU")
@pytest.mark.parametrize("dt", [np.float64, np.complex128])
def test_posv_tester_check(sl, uplo, dt):
    """the reference tester's own potrf / posv check (test/test_posv.cc:267-343): factor, solve
    with nrhs = 10, || B - A X ||_1 / (n ||A||_1 ||X||_1) <= tol * eps / 2 with tol = 50."""
    n, nb, nrhs = 1000, 128, 10
    A0 = ol.matgen(7, n, n, dt)
    A0[np.arange(n), np.arange(n)] += n
    tri = np.tril if uplo == "L" else np.triu
    strict = (lambda x: np.tril(x, -1)) if uplo == "L" else (lambda x: np.triu(x, 1))
    S = strict(A0) + strict(A0).conj().T + np.diag(np.diag(A0).real)
    B0 = ol.matgen(8, n, nrhs, dt)
    A = sl.HermitianMatrix(uplo, n, nb, dtype=dt).from_numpy(A0)
    B = sl.Matrix(n, nrhs, nb, dtype=dt).from_numpy(B0)
    assert sl.posv(A, B) == 0
    X = B.to_numpy()
    resid = np.linalg.norm(B0 - S @ X, 1) / (n * np.linalg.norm(S, 1) * np.linalg.norm(X, 1))
    assert resid <= 50 * EPS / 2, resid
    # potrs on an already factored matrix gives the same solution
    B2 = sl.Matrix(n, nrhs, nb, dtype=dt).from_numpy(B0)
    sl.potrs(A, B2)
    assert np.abs(B2.to_numpy() - X).max() <= 50 * EPS * np.abs(X).max()


def test_host_origin_tiles_through_device_panels():
    """SLATE_B200_TILE_RUNS=1 (read once per process, hence the subprocess): host-origin tiles that
    are neighbours in the caller's array cross PCIe as one copy through a device panel
    (BaseMatrix::tileGetRun / updateOriginRun, te_panel_pieces).  gemm and potrf on fromLAPACK
    arrays with padding rows, ragged tiles, against numpy."""
    import os
    import subprocess
    import sys
    code = r'''
import numpy as np, sys, os
sys.path.insert(0, os.path.join(os.getcwd(), "tests"))
import oracle_lib as ol
from slate_d35_b200 import host as sl
eps = np.finfo(np.float64).eps
for dtype in (np.float64, np.complex128):
    m, n, k, nb, pad = 1100, 900, 700, 256, 7
    bufs = [np.asfortranarray(np.full((r + pad, c), np.nan, dtype=dtype)) for r, c in ((m, k), (k, n), (m, n))]
    A0, B0, C0 = ol.matgen(1, m, k, dtype), ol.matgen(2, k, n, dtype), ol.matgen(3, m, n, dtype)
    for b, src in zip(bufs, (A0, B0, C0)):
        b[:src.shape[0], :] = src
    A = sl.Matrix.fromLAPACK(bufs[0], nb, m=m)
    B = sl.Matrix.fromLAPACK(bufs[1], nb, m=k)
    Cm = sl.Matrix.fromLAPACK(bufs[2], nb, m=m)
    sl.gemm(1.5, A, B, -0.5, Cm, method="C")
    sl.sync()
    want = 1.5 * A0 @ B0 - 0.5 * C0
    assert np.abs(bufs[2][:m] - want).max() / np.abs(want).max() <= 3 * np.sqrt(4 * k) * eps
    assert np.isnan(bufs[2][m:].real).all()          # padding rows untouched
n, nb = 1300, 256
S0 = ol.matgen(7, n, n); S = (S0 + S0.T) / 2 + n * np.eye(n)
buf = np.asfortranarray(np.tril(S))
H = sl.HermitianMatrix.fromLAPACK("L", buf, nb)
assert sl.potrf(H) == 0
sl.sync()
L = np.linalg.cholesky(S)
assert np.abs(np.tril(buf) - L).max() / np.abs(L).max() <= 50 * eps
print("PANEL PATH OK")
'''
    env = dict(os.environ, SLATE_B200_TILE_RUNS="1")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, "-c", code], cwd=root, env=env, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and "PANEL PATH OK" in out.stdout, out.stdout[-2000:] + out.stderr[-3000:]
