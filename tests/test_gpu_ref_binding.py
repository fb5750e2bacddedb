"""The drop-in, demonstrated: the reference's OWN slate::gemm and slate::potrf -- src/gemm.cc,
src/gemmC.cc, src/potrf.cc, the Target::Devices overloads of internal::gemm / herk / trsm / potrf,
internal_batch.hh, BaseMatrix, MatrixStorage, Tile, all compiled unmodified from /root/reference by
slate_d35_b200/host/ref_binding/Makefile -- running on the B200 tile engine, which is bound in at the
BLAS++ / LAPACK++ device entry points (blas::batch::gemm / herk / syrk / trsm, lapack::potrf(queue))
and at slate::Memory.  Checks, on one GPU:
  * Target::Devices agrees with Target::HostTask of the SAME binary on the same inputs within
    50 eps (north star), host-origin and device-origin matrices, ragged tiles, real and complex;
  * the work really went through the engine's kernels (te_launch_count moves) -- and does NOT
    move in the sibling build that keeps BLAS++'s cuBLAS wrappers, whose results must agree too.
The libraries are built where /root/reference exists and travel to the GPU box prebuilt."""
import ctypes as C
import os

import numpy as np
import pytest

import oracle_lib as ol

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BUILD = os.path.join(ROOT, "slate_d35_b200", "host", "ref_binding", "_build")
EPS = np.finfo(np.float64).eps
i64, dp = C.c_int64, C.POINTER(C.c_double)
_libs = {}


def binding(flavour):
    if flavour not in _libs:
        path = os.path.join(BUILD, f"libslate_ref_{flavour}.so")
        if not os.path.exists(path):
            pytest.skip(f"{path} not built (needs /root/reference at build time)")
        from slate_d35_b200 import lib
        lib()                                     # the engine first: one shared instance
        L = C.CDLL(path)
        L.rb_flavour.restype = C.c_char_p
        L.rb_dgemm.argtypes = [C.c_int, C.c_int, i64, i64, i64, i64, C.c_double, dp, i64, dp, i64,
                               C.c_double, dp, i64, C.c_int, dp]
        L.rb_zgemm.argtypes = [C.c_int, C.c_int, i64, i64, i64, i64, dp, dp, i64, dp, i64, dp, dp, i64,
                               C.c_int, dp]
        for f in (L.rb_dpotrf, L.rb_zpotrf):
            f.argtypes = [C.c_int, C.c_int, C.c_int, i64, i64, dp, i64, C.c_int, dp]
            f.restype = i64
        _libs[flavour] = L
    return _libs[flavour]


def launches():
    from slate_d35_b200 import lib
    te = lib()
    te.te_launch_count.restype = C.c_longlong
    return te.te_launch_count()


def test_flavours_are_what_they_say():
    assert binding("te").rb_flavour() == b"tile_engine"
    assert binding("te").rb_num_devices() >= 1
    assert binding("cublas").rb_flavour() == b"cublas"


@pytest.mark.parametrize("origin", "HD")
@pytest.mark.parametrize("dtype", [np.float64, np.complex128])
def test_reference_gemm_devices_on_the_engine(origin, dtype):
    L = binding("te")
    m, n, k, nb = 1000, 900, 700, 256                     # ragged in every dimension
    z = np.dtype(dtype) == np.complex128
    A0, B0, C0 = ol.matgen(1, m, k, dtype), ol.matgen(2, k, n, dtype), ol.matgen(3, m, n, dtype)
    alpha, beta = (3.141592653589793 + 1.4142135623730951j, 2.718281828459045 + 1.7320508075688772j) if z \
        else (3.141592653589793, 2.718281828459045)

    def run(lib, target):
        A, B, Cm = A0.copy(order="F"), B0.copy(order="F"), C0.copy(order="F")
        if z:
            st = lib.rb_zgemm(ord(target), ord(origin), m, n, k, nb, ol.Z(alpha), ol.P(A), m, ol.P(B), k,
                              ol.Z(beta), ol.P(Cm), m, 1, None)
        else:
            st = lib.rb_dgemm(ord(target), ord(origin), m, n, k, nb, alpha, ol.P(A), m, ol.P(B), k, beta,
                              ol.P(Cm), m, 1, None)
        assert st == 0
        return Cm

    host = run(L, "T")
    l0 = launches()
    dev = run(L, "D")
    assert launches() > l0, "Target::Devices did not launch the engine's kernels"
    err = np.abs(dev - host).max() / np.abs(host).max()
    assert err <= 50 * EPS, err
    # the reference's stock GPU path (cuBLAS under the same driver) must say the same, without
    # touching the engine
    l1 = launches()
    stock = run(binding("cublas"), "D")
    assert launches() == l1
    assert np.abs(stock - host).max() / np.abs(host).max() <= 50 * EPS


@pytest.mark.parametrize("origin", "HD")
@pytest.mark.parametrize("uplo", "LU")
@pytest.mark.parametrize("dtype", [np.float64, np.complex128])
def test_reference_potrf_devices_on_the_engine(origin, uplo, dtype):
    L = binding("te")
    n, nb = 1100, 256
    z = np.dtype(dtype) == np.complex128
    A0 = ol.matgen(7, n, n, dtype)
    A0 = (A0 + A0.conj().T) / 2
    A0[np.arange(n), np.arange(n)] = A0.diagonal().real + n
    tri = np.tril if uplo == "L" else np.triu

    def run(lib, target, la):
        A = A0.copy(order="F")
        fn = lib.rb_zpotrf if z else lib.rb_dpotrf
        info = fn(ord(target), ord(origin), ord(uplo), n, nb, ol.P(A), n, la, None)
        return info, tri(A)

    info_h, host = run(L, "T", 1)
    assert info_h == 0
    for la in (0, 1, 2):
        l0 = launches()
        info_d, dev = run(L, "D", la)
        assert launches() > l0
        assert info_d == 0
        assert np.abs(dev - host).max() / np.abs(host).max() <= 50 * EPS, (la, uplo)
        F = dev
        R = (F @ F.conj().T if uplo == "L" else F.conj().T @ F) - A0
        assert np.linalg.norm(R) / (n * np.linalg.norm(A0)) <= 50 * EPS
    info_s, stock = run(binding("cublas"), "D", 1)
    assert info_s == 0 and np.abs(stock - host).max() / np.abs(host).max() <= 50 * EPS


def test_reference_potrf_devices_reports_info():
    """a matrix that stops being positive definite at column 700: the reference's driver must
    return LAPACK's info from the engine's device info word"""
    L = binding("te")
    n, nb = 1000, 256
    A0 = ol.matgen(7, n, n)
    A0 = (A0 + A0.T) / 2
    A0[np.arange(n), np.arange(n)] += n
    A0[699, 699] = -3.0
    got = {}
    for target in "TD":
        A = A0.copy(order="F")
        got[target] = L.rb_dpotrf(ord(target), ord("H"), ord("L"), n, nb, ol.P(A), n, 1, None)
    assert got["T"] == got["D"] == 700
