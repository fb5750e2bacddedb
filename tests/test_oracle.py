"""Pin the CPU oracle (oracle/*.c) against the UNMODIFIED reference.

The reference has no golden vectors for this path, so the pin is
the reference itself: tests/golden/ref_hosttask.npz was produced by
tests/golden/make_golden.py from oracle/_ref/libslate_ref.so (reference HostTask path,
compiled from /root/reference).  When _ref is present (authoring container) the oracle
is additionally compared with it directly on fresh inputs.

Tolerances: matgen is integer arithmetic -> bit exact.  One block update: 3*sqrt(k)*eps
elementwise (unit_test/test_internal_blas.cc:446-447) relative to the row scale; full
drivers: 3*sqrt(K)*eps with K the whole inner dimension; potrf/trsm: 50*eps*n.
"""
import os

import numpy as np
import pytest

import oracle_lib as ol

EPS = np.finfo(np.float64).eps
G = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_hosttask.npz"))
O = ol.oracle()
P, Z = ol.P, ol.Z
ALPHA, BETA = 3.141592653589793, 2.718281828459045
ZA, ZB = ALPHA + 1.4142135623730951j, BETA + 1.7320508075688772j


def close(got, want, k, scale=None):
    scale = scale if scale is not None else max(1.0, np.abs(want).max())
    err = np.abs(got - want).max() / scale
    assert err <= 3 * np.sqrt(max(k, 1)) * EPS, err


def test_matgen_bit_exact_vs_golden():
    assert np.array_equal(ol.matgen(7, 9, 11, np.float64, 1, 3, 5), G["matgen_d_seed7_off_3_5"])
    assert np.array_equal(ol.matgen(7, 9, 11, np.complex128, 1, 3, 5), G["matgen_z_seed7_off_3_5"])
    assert np.array_equal(ol.matgen(42, 6, 4, np.float64, 2), G["matgen_d_signed_seed42"])
    assert np.array_equal(ol.matgen(1, 4, 4, np.float64, 1, 98300, 65530), G["matgen_d_seed1_big_off"])


def test_matgen_is_pure_function_of_index():
    whole = ol.matgen(9, 40, 30)
    part = ol.matgen(9, 7, 5, ioff=11, joff=13)
    assert np.array_equal(whole[11:18, 13:18], part)
    assert whole.min() >= 0.0 and whole.max() < 1.0


@pytest.mark.parametrize("opa", "NTC")
@pytest.mark.parametrize("opb", "NTC")
def test_dgemm_vs_golden(opa, opb):
    m, n, k, nb = 70, 53, 45, 16
    A = ol.matgen(1, *((m, k) if opa == "N" else (k, m)))
    B = ol.matgen(2, *((k, n) if opb == "N" else (n, k)))
    Cm = ol.matgen(3, m, n)
    O.oracle_dgemm(ord(opa), ord(opb), m, n, k, nb, ALPHA, P(A), A.shape[0], P(B), B.shape[0],
                   BETA, P(Cm), m)
    close(Cm, G[f"dgemm_{opa}{opb}"], k)


@pytest.mark.parametrize("opa", "NTC")
@pytest.mark.parametrize("opb", "NTC")
def test_zgemm_vs_golden(opa, opb):
    m, n, k, nb = 70, 53, 45, 16
    A = ol.matgen(1, *((m, k) if opa == "N" else (k, m)), np.complex128)
    B = ol.matgen(2, *((k, n) if opb == "N" else (n, k)), np.complex128)
    Cm = ol.matgen(3, m, n, np.complex128)
    O.oracle_zgemm(ord(opa), ord(opb), m, n, k, nb, Z(ZA), P(A), A.shape[0], P(B), B.shape[0],
                   Z(ZB), P(Cm), m)
    close(Cm, G[f"zgemm_{opa}{opb}"], 4 * k)


@pytest.mark.parametrize("pre,dt", [("d", np.float64), ("z", np.complex128)])
@pytest.mark.parametrize("uplo", "LU")
def test_potrf_vs_golden(pre, dt, uplo):
    n, nb = 90, 32
    A = ol.matgen(5, n, n, dt)
    A[np.arange(n), np.arange(n)] += n
    info = getattr(O, f"oracle_{pre}potrf")(ord(uplo), n, nb, P(A), n)
    assert info == int(G[f"{pre}potrf_{uplo}_info"][0]) == 0
    tri = np.tril if uplo == "L" else np.triu
    want = tri(G[f"{pre}potrf_{uplo}"])
    err = np.abs(tri(A) - want).max() / np.abs(want).max()
    assert err <= 50 * EPS, err


def test_potrf_info_names_first_bad_minor():
    n, nb = 90, 32
    A = ol.matgen(5, n, n)
    A[np.arange(n), np.arange(n)] += n
    A[40, 40] = -1.0
    assert O.oracle_dpotrf(ord("L"), n, nb, P(A), n) == int(G["dpotrf_L_bad_info"][0]) == 41


@pytest.mark.parametrize("opa", "NT")
@pytest.mark.parametrize("opb", "NT")
def test_tile_gemm_vs_internal_golden(opa, opb):
    m, n, k = 80, 64, 16
    A = ol.matgen(11, *((m, k) if opa == "N" else (k, m)))
    B = ol.matgen(12, *((k, n) if opb == "N" else (n, k)))
    Cm = ol.matgen(13, m, n)
    O.oracle_tile_dgemm(ord(opa), ord(opb), m, n, k, ALPHA, P(A), A.shape[0], P(B), B.shape[0],
                        BETA, P(Cm), m)
    close(Cm, G[f"internal_dgemm_{opa}{opb}"], k)


@pytest.mark.parametrize("pre,dt", [("d", np.float64), ("z", np.complex128)])
@pytest.mark.parametrize("uplo", "LU")
def test_tile_herk_vs_internal_golden(pre, dt, uplo):
    n, k = 64, 16
    A = ol.matgen(14, n, k, dt)
    Cm = ol.matgen(15, n, n, dt)
    if pre == "z":
        Cm[np.arange(n), np.arange(n)] = Cm[np.arange(n), np.arange(n)].real
    before = Cm.copy()
    fn = O.oracle_tile_dsyrk if pre == "d" else O.oracle_tile_zherk
    fn(ord(uplo), ord("N"), n, k, -1.0, P(A), n, 1.0, P(Cm), n)
    tri = np.tril if uplo == "L" else np.triu
    close(tri(Cm), tri(G[f"internal_{pre}herk_{uplo}"]), 4 * k)
    other = (lambda x: np.triu(x, 1)) if uplo == "L" else (lambda x: np.tril(x, -1))
    assert np.array_equal(other(Cm), other(before)), "other triangle must not be touched"


@pytest.mark.parametrize("pre,dt", [("d", np.float64), ("z", np.complex128)])
def test_tile_trsm_and_potrf_vs_internal_golden(pre, dt):
    nb = 32
    T = ol.matgen(16, nb, nb, dt)
    T[np.arange(nb), np.arange(nb)] += nb
    Bm = ol.matgen(17, 75, nb, dt)
    if pre == "d":
        O.oracle_tile_dtrsm(ord("R"), ord("L"), ord("C"), ord("N"), 75, nb, 1.0, P(T), nb, P(Bm), 75)
    else:
        O.oracle_tile_ztrsm(ord("R"), ord("L"), ord("C"), ord("N"), 75, nb, Z(1.0), P(T), nb, P(Bm), 75)
    want = G[f"internal_{pre}trsm_RLC"]
    assert np.abs(Bm - want).max() / np.abs(want).max() <= 50 * EPS
    T2 = ol.matgen(18, nb, nb, dt)
    T2[np.arange(nb), np.arange(nb)] += nb
    info = getattr(O, f"oracle_tile_{pre}potrf")(ord("L"), nb, P(T2), nb)
    assert info == int(G[f"internal_{pre}potrf_L_info"][0]) == 0
    want = np.tril(G[f"internal_{pre}potrf_L"])
    assert np.abs(np.tril(T2) - want).max() / np.abs(want).max() <= 50 * EPS


@pytest.mark.parametrize("side", "LR")
@pytest.mark.parametrize("uplo", "LU")
@pytest.mark.parametrize("op", "NTC")
@pytest.mark.parametrize("diag", "NU")
def test_tile_trsm_solves(side, uplo, op, diag):
    """property: op(A) X = alpha B (or X op(A) = alpha B) holds for every variant."""
    m, n = 13, 9
    na = m if side == "L" else n
    A = ol.matgen(21, na, na, np.complex128)
    A[np.arange(na), np.arange(na)] += na
    B0 = ol.matgen(22, m, n, np.complex128)
    X = B0.copy(order="F")
    alpha = 0.7 - 0.2j
    O.oracle_tile_ztrsm(ord(side), ord(uplo), ord(op), ord(diag), m, n, Z(alpha), P(A), na, P(X), m)
    T = np.tril(A) if uplo == "L" else np.triu(A)
    if diag == "U":
        T[np.arange(na), np.arange(na)] = 1.0
    opT = {"N": T, "T": T.T, "C": T.conj().T}[op]
    lhs = opT @ X if side == "L" else X @ opT
    assert np.abs(lhs - alpha * B0).max() <= 200 * EPS


@pytest.mark.skipif(not ol.have_ref(), reason="oracle/_ref not built (needs /root/reference)")
def test_oracle_vs_reference_direct():
    """fresh shapes, not in the fixtures: config 1's shape scaled down (ragged 8x8 tiles)."""
    R = ol.ref()
    n, nb = 250, 32
    A, B = ol.matgen(1, n, n), ol.matgen(2, n, n)
    assert np.array_equal(A, ol.matgen(1, n, n, lib=R))
    C1, C2 = ol.matgen(3, n, n), ol.matgen(3, n, n)
    O.oracle_dgemm(ord("N"), ord("N"), n, n, n, nb, ALPHA, P(A), n, P(B), n, BETA, P(C1), n)
    R.ref_dgemm(ord("N"), ord("N"), n, n, n, nb, ALPHA, P(A), n, P(B), n, BETA, P(C2), n, 1)
    close(C1, C2, n)
    H1 = ol.matgen(4, n, n)
    H1[np.arange(n), np.arange(n)] += n
    H2 = H1.copy(order="F")
    assert O.oracle_dpotrf(ord("L"), n, nb, P(H1), n) == R.ref_dpotrf(ord("L"), n, nb, P(H2), n, 1) == 0
    assert np.abs(np.tril(H1) - np.tril(H2)).max() / np.abs(np.tril(H2)).max() <= 50 * EPS
