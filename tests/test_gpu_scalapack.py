"""The ScaLAPACK-convention entry points (include/slate_b200_scalapack.h) on one GPU, 1 x 1
grid: every routine called the way a ScaLAPACK program calls it (by reference, 1-based
offsets, 9-integer descriptors, local arrays with lld > rows), checked against the oracle's
matgen inputs and numpy.  Mirrors what the reference's scalapack_api/ promises: same
arguments in, same numbers out as the slate:: driver.  The multi-rank version of this test
is scripts/mp_parity.py (2, 4, 8 GPUs).  Needs a GPU.
"""
import numpy as np
import pytest

import oracle_lib as ol

pytestmark = pytest.mark.gpu
EPS = np.finfo(np.float64).eps
CTXT = 0


@pytest.fixture(scope="module")
def sc():
    from slate_d35_b200 import host, scalapack
    host.initialize(0)
    scalapack.grid_set(CTXT, "C", 1, 1)
    yield scalapack
    host.sync()


def op(a, t):
    return {"N": a, "T": a.T, "C": a.conj().T}[t]


def padded(a, extra=3):
    """local array with lld = rows + extra, and the view of its first rows"""
    buf = np.zeros((a.shape[0] + extra, a.shape[1]), dtype=a.dtype, order="F")
    buf[:a.shape[0]] = a
    return buf


@pytest.mark.parametrize("dt", [np.float64, np.complex128])
@pytest.mark.parametrize("ta,tb", [("N", "N"), ("T", "N"), ("N", "C"), ("C", "T")])
def test_pgemm(sc, dt, ta, tb):
    m, n, k, nb = 700, 520, 300, 128
    ash = (m, k) if ta == "N" else (k, m)
    bsh = (k, n) if tb == "N" else (n, k)
    A0, B0, C0 = ol.matgen(1, *ash, dt), ol.matgen(2, *bsh, dt), ol.matgen(3, m, n, dt)
    a, b, c = padded(A0), padded(B0), padded(C0)
    alpha, beta = (1.5 - 0.5j, 0.25 + 2j) if dt == np.complex128 else (3.141592653589793, 2.718281828459045)
    sc.pgemm(ta, tb, m, n, k, alpha, a, 1, 1, sc.descinit(*ash, nb, nb, CTXT, a.shape[0]),
             b, 1, 1, sc.descinit(*bsh, nb, nb, CTXT, b.shape[0]), beta,
             c, 1, 1, sc.descinit(m, n, nb, nb, CTXT, c.shape[0]))
    want = alpha * op(A0, ta) @ op(B0, tb) + beta * C0
    assert np.abs(c[:m] - want).max() <= 3 * np.sqrt(k) * EPS * np.abs(want).max() * (4 if dt == np.complex128 else 1)
    assert not c[m:].any() and np.array_equal(a[:ash[0]], A0)     # padding and inputs untouched


def test_pgemm_submatrix_offsets(sc):
    """C(ic:, jc:) = alpha A(ia:, ja:) B(ib:, jb:) + beta C(ic:, jc:) on tile-aligned offsets
    (scalapack_api/scalapack_slate.hh:81-95); everything outside the sub-matrix stays."""
    nb = 64
    A0, B0, C0 = ol.matgen(1, 512, 448), ol.matgen(2, 448, 384), ol.matgen(3, 512, 384)
    a, b, c = A0.copy(order="F"), B0.copy(order="F"), C0.copy(order="F")
    m, n, k = 256, 192, 128
    ia, ja, ib, jb, ic, jc = 65, 129, 193, 65, 129, 1
    sc.pgemm("N", "N", m, n, k, 2.0, a, ia, ja, sc.descinit(512, 448, nb, nb, CTXT, 512),
             b, ib, jb, sc.descinit(448, 384, nb, nb, CTXT, 448), -1.0,
             c, ic, jc, sc.descinit(512, 384, nb, nb, CTXT, 512))
    want = C0.copy()
    want[ic - 1:ic - 1 + m, jc - 1:jc - 1 + n] = (
        2.0 * A0[ia - 1:ia - 1 + m, ja - 1:ja - 1 + k] @ B0[ib - 1:ib - 1 + k, jb - 1:jb - 1 + n]
        - C0[ic - 1:ic - 1 + m, jc - 1:jc - 1 + n])
    assert np.abs(c - want).max() <= 3 * np.sqrt(k) * EPS * np.abs(want).max()
    outside = np.ones_like(c, dtype=bool)
    outside[ic - 1:ic - 1 + m, jc - 1:jc - 1 + n] = False
    assert np.array_equal(c[outside], C0[outside])
    with pytest.raises(sc.ScalapackError):          # offset off a tile boundary
        sc.pgemm("N", "N", m, n, k, 2.0, a, 2, 1, sc.descinit(512, 448, nb, nb, CTXT, 512),
                 b, 1, 1, sc.descinit(448, 384, nb, nb, CTXT, 448), -1.0,
                 c, 1, 1, sc.descinit(512, 384, nb, nb, CTXT, 512))


def spd(n, dt, seed=7):
    A0 = ol.matgen(seed, n, n, dt)
    S = (A0 + A0.conj().T) / 2
    S[np.arange(n), np.arange(n)] = S.diagonal().real + n
    return S


@pytest.mark.parametrize("dt", [np.float64, np.complex128])
@pytest.mark.parametrize("uplo", "LU")
def test_ppotrf_pposv_ppotrs(sc, dt, uplo):
    n, nb, nrhs = 900, 128, 10
    S = spd(n, dt)
    tri = np.tril if uplo == "L" else np.triu
    other = (lambda x: np.triu(x, 1)) if uplo == "L" else (lambda x: np.tril(x, -1))
    junk = other(ol.matgen(9, n, n, dt))             # the other triangle must never be read
    a = padded(tri(S) + junk)
    desca = sc.descinit(n, n, nb, nb, CTXT, a.shape[0])
    assert sc.ppotrf(uplo, n, a, 1, 1, desca) == 0
    F = tri(a[:n])
    back = F @ F.conj().T if uplo == "L" else F.conj().T @ F
    assert np.linalg.norm(back - S) / (n * np.linalg.norm(S)) <= 50 * EPS
    assert np.array_equal(other(a[:n]), junk)

    B0 = ol.matgen(8, n, nrhs, dt)
    b = padded(B0)
    descb = sc.descinit(n, nrhs, nb, nb, CTXT, b.shape[0])
    assert sc.ppotrs(uplo, n, nrhs, a, 1, 1, desca, b, 1, 1, descb) == 0
    X = b[:n]
    resid = np.linalg.norm(B0 - S @ X, 1) / (n * np.linalg.norm(S, 1) * np.linalg.norm(X, 1))
    assert resid <= 50 * EPS / 2

    a2, b2 = padded(tri(S)), padded(B0)
    assert sc.pposv(uplo, n, nrhs, a2, 1, 1, desca, b2, 1, 1, descb) == 0
    assert np.array_equal(tri(a2[:n]), F) and np.array_equal(b2[:n], X)


def test_ppotrf_reports_info(sc):
    """the reference returns info = 0 unconditionally ("todo"); here it is the LAPACK info"""
    n, nb = 600, 128
    S = spd(n, np.float64)
    S[333, 333] = -1.0
    a = np.asfortranarray(np.tril(S))
    assert sc.ppotrf("L", n, a, 1, 1, sc.descinit(n, n, nb, nb, CTXT, n)) == 334


@pytest.mark.parametrize("dt", [np.float64, np.complex128])
@pytest.mark.parametrize("side,uplo,ta,diag", [("L", "L", "N", "N"), ("L", "U", "C", "N"),
                                               ("R", "L", "T", "U"), ("R", "U", "N", "N")])
def test_ptrsm(sc, dt, side, uplo, ta, diag):
    m, n, nb = 640, 384, 128
    an = m if side == "L" else n
    T0 = ol.matgen(11, an, an, dt)
    T0[np.arange(an), np.arange(an)] += an
    tri = np.tril if uplo == "L" else np.triu
    T = tri(T0)
    if diag == "U":
        T = T / an                                  # keep the unit-diagonal system well conditioned
        np.fill_diagonal(T, 1.0)
    stored = np.asfortranarray(tri(T0) if diag == "N" else tri(T0) / an)   # diagonal ignored for 'U'
    B0 = ol.matgen(12, m, n, dt)
    b = padded(B0)
    alpha = 0.75
    sc.ptrsm(side, uplo, ta, diag, m, n, alpha, stored, 1, 1, sc.descinit(an, an, nb, nb, CTXT, an),
             b, 1, 1, sc.descinit(m, n, nb, nb, CTXT, b.shape[0]))
    X = b[:m]
    back = op(T, ta) @ X if side == "L" else X @ op(T, ta)
    assert np.abs(back - alpha * B0).max() <= 200 * EPS * np.abs(B0).max() * np.linalg.cond(T, 1)


@pytest.mark.parametrize("uplo", "LU")
@pytest.mark.parametrize("trans", "NT")
def test_psyrk_pherk(sc, uplo, trans):
    n, k, nb = 512, 384, 128
    tri = np.tril if uplo == "L" else np.triu
    ash = (n, k) if trans == "N" else (k, n)
    # syrk, real
    A0, C0 = ol.matgen(13, *ash), ol.matgen(14, n, n)
    c = np.asfortranarray(C0.copy())
    sc.psyrk(uplo, trans, n, k, -1.0, np.asfortranarray(A0), 1, 1, sc.descinit(*ash, nb, nb, CTXT, ash[0]),
             0.5, c, 1, 1, sc.descinit(n, n, nb, nb, CTXT, n))
    Aop = A0 if trans == "N" else A0.T
    want = -1.0 * Aop @ Aop.T + 0.5 * C0
    assert np.abs(tri(c) - tri(want)).max() <= 3 * np.sqrt(k) * EPS * np.abs(want).max()
    strict_other = (lambda x: np.triu(x, 1)) if uplo == "L" else (lambda x: np.tril(x, -1))
    assert np.array_equal(strict_other(c), strict_other(C0))
    # herk, complex ('T' is not a herk op: use 'C')
    th = "N" if trans == "N" else "C"
    Z0 = ol.matgen(15, *ash, np.complex128)
    H0 = ol.matgen(16, n, n, np.complex128)
    H0 = (H0 + H0.conj().T) / 2
    h = np.asfortranarray(H0.copy())
    sc.pherk(uplo, th, n, k, 2.0, np.asfortranarray(Z0), 1, 1, sc.descinit(*ash, nb, nb, CTXT, ash[0]),
             -1.0, h, 1, 1, sc.descinit(n, n, nb, nb, CTXT, n))
    Zop = Z0 if th == "N" else Z0.conj().T
    wanth = 2.0 * Zop @ Zop.conj().T - H0
    assert np.abs(tri(h) - tri(wanth)).max() <= 12 * np.sqrt(k) * EPS * np.abs(wanth).max()
    assert not np.diag(h).imag.any()


@pytest.mark.parametrize("side", "LR")
@pytest.mark.parametrize("uplo", "LU")
def test_psymm_phemm(sc, side, uplo):
    m, n, nb = 512, 384, 128
    an = m if side == "L" else n
    tri = np.tril if uplo == "L" else np.triu
    for herm, dt in ((False, np.float64), (False, np.complex128), (True, np.complex128)):
        A0 = ol.matgen(21, an, an, dt)
        S = (A0 + (A0.conj().T if herm else A0.T)) / 2
        B0, C0 = ol.matgen(22, m, n, dt), ol.matgen(23, m, n, dt)
        c = padded(C0)
        alpha, beta = ((1.5 - 0.5j), (0.25 + 2j)) if dt == np.complex128 else (3.0, -0.5)
        sc.phemm(side, uplo, m, n, alpha, np.asfortranarray(tri(S)), 1, 1, sc.descinit(an, an, nb, nb, CTXT, an),
                 np.asfortranarray(B0), 1, 1, sc.descinit(m, n, nb, nb, CTXT, m), beta,
                 c, 1, 1, sc.descinit(m, n, nb, nb, CTXT, c.shape[0]), hermitian=herm)
        want = alpha * (S @ B0 if side == "L" else B0 @ S) + beta * C0
        assert np.abs(c[:m] - want).max() <= 12 * np.sqrt(an) * EPS * np.abs(want).max()
        assert not c[m:].any()


@pytest.mark.parametrize("uplo", "LU")
@pytest.mark.parametrize("trans", "NT")
def test_psyr2k_pher2k(sc, uplo, trans):
    n, k, nb = 512, 256, 128
    tri = np.tril if uplo == "L" else np.triu
    ash = (n, k) if trans == "N" else (k, n)
    for herm, dt in ((False, np.float64), (False, np.complex128), (True, np.complex128)):
        tr = ("N" if trans == "N" else "C") if herm else trans
        A0, B0, C0 = ol.matgen(24, *ash, dt), ol.matgen(25, *ash, dt), ol.matgen(26, n, n, dt)
        if herm:
            C0 = (C0 + C0.conj().T) / 2
        c = np.asfortranarray(C0.copy())
        alpha = (1.5 - 0.5j) if dt == np.complex128 else 3.0
        da = sc.descinit(*ash, nb, nb, CTXT, ash[0])
        dc = sc.descinit(n, n, nb, nb, CTXT, n)
        Aop, Bop = op(A0, tr), op(B0, tr)
        if herm:
            sc.pher2k(uplo, tr, n, k, alpha, np.asfortranarray(A0), 1, 1, da, np.asfortranarray(B0), 1, 1, da,
                      -0.5, c, 1, 1, dc)
            want = alpha * Aop @ Bop.conj().T + np.conj(alpha) * Bop @ Aop.conj().T - 0.5 * C0
        else:
            beta = (0.25 + 2j) if dt == np.complex128 else -0.5
            sc.psyr2k(uplo, tr, n, k, alpha, np.asfortranarray(A0), 1, 1, da, np.asfortranarray(B0), 1, 1, da,
                      beta, c, 1, 1, dc)
            want = alpha * Aop @ Bop.T + alpha * Bop @ Aop.T + beta * C0
        assert np.abs(tri(c) - tri(want)).max() <= 12 * np.sqrt(2 * k) * EPS * np.abs(want).max()
        other = (lambda x: np.triu(x, 1)) if uplo == "L" else (lambda x: np.tril(x, -1))
        assert np.array_equal(other(c), other(C0))


def test_errors_are_loud(sc, monkeypatch):
    nb = 64
    a = np.zeros((128, 128), order="F")
    d_ok = sc.descinit(128, 128, nb, nb, CTXT, 128)
    with pytest.raises(sc.ScalapackError, match="no grid registered"):
        sc.pgemm("N", "N", 128, 128, 128, 1.0, a, 1, 1, sc.descinit(128, 128, nb, nb, 77, 128),
                 a, 1, 1, d_ok, 0.0, a, 1, 1, d_ok)
    with pytest.raises(sc.ScalapackError, match="RSRC"):
        sc.pgemm("N", "N", 128, 128, 128, 1.0, a, 1, 1, sc.descinit(128, 128, nb, nb, CTXT, 128, rsrc=1),
                 a, 1, 1, d_ok, 0.0, a, 1, 1, d_ok)
    # a host target is refused, never silently computed on the CPU
    monkeypatch.setenv("SLATE_SCALAPACK_TARGET", "HostTask")
    with pytest.raises(sc.ScalapackError, match="no CPU fallback"):
        sc.pgemm("N", "N", 128, 128, 128, 1.0, a, 1, 1, d_ok, a, 1, 1, d_ok, 0.0, a, 1, 1, d_ok)
    monkeypatch.setenv("SLATE_SCALAPACK_TARGET", "Devices")
    sc.pgemm("N", "N", 128, 128, 128, 1.0, a, 1, 1, d_ok, a, 1, 1, d_ok, 0.0, a, 1, 1, d_ok)
