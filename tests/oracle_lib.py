"""ctypes access to the CPU oracle (oracle/liboracle.so) and, when built, to the
unmodified reference (oracle/_ref/libslate_ref.so).  TEST INFRASTRUCTURE ONLY."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "liboracle.so")
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libslate_ref.so")
i64 = C.c_int64
dp = C.POINTER(C.c_double)


def P(a):
    assert a.dtype in (np.float64, np.complex128) and a.flags.f_contiguous or a.ndim == 1
    return a.ctypes.data_as(dp)


def Z(x):
    """complex scalar -> pointer to (re, im)"""
    arr = (C.c_double * 2)(complex(x).real, complex(x).imag)
    return arr


_oracle = None
_ref = None


def oracle():
    global _oracle
    if _oracle is None:
        if not os.path.exists(ORACLE_SO):
            subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "oracle"])
        L = C.CDLL(ORACLE_SO)
        L.oracle_matgen_d.argtypes = [C.c_int, i64, i64, i64, i64, i64, dp, i64]
        L.oracle_matgen_z.argtypes = L.oracle_matgen_d.argtypes
        L.oracle_tile_dgemm.argtypes = [C.c_int, C.c_int, i64, i64, i64, C.c_double, dp, i64, dp, i64, C.c_double, dp, i64]
        L.oracle_tile_zgemm.argtypes = [C.c_int, C.c_int, i64, i64, i64, dp, dp, i64, dp, i64, dp, dp, i64]
        L.oracle_tile_dsyrk.argtypes = [C.c_int, C.c_int, i64, i64, C.c_double, dp, i64, C.c_double, dp, i64]
        L.oracle_tile_zherk.argtypes = L.oracle_tile_dsyrk.argtypes
        L.oracle_tile_dtrsm.argtypes = [C.c_int] * 4 + [i64, i64, C.c_double, dp, i64, dp, i64]
        L.oracle_tile_ztrsm.argtypes = [C.c_int] * 4 + [i64, i64, dp, dp, i64, dp, i64]
        for f in (L.oracle_tile_dpotrf, L.oracle_tile_zpotrf):
            f.argtypes = [C.c_int, i64, dp, i64]
            f.restype = i64
        L.oracle_dgemm.argtypes = [C.c_int, C.c_int, i64, i64, i64, i64, C.c_double, dp, i64, dp, i64, C.c_double, dp, i64]
        L.oracle_zgemm.argtypes = [C.c_int, C.c_int, i64, i64, i64, i64, dp, dp, i64, dp, i64, dp, dp, i64]
        for f in (L.oracle_dpotrf, L.oracle_zpotrf):
            f.argtypes = [C.c_int, i64, i64, dp, i64]
            f.restype = i64
        ci = C.c_int
        L.oracle_drank2k.argtypes = [ci, ci, ci, i64, i64, i64, C.c_double, dp, i64, dp, i64, C.c_double, dp, i64]
        L.oracle_zrank2k.argtypes = [ci, ci, ci, i64, i64, i64, dp, dp, i64, dp, i64, dp, dp, i64]
        L.oracle_dhemm.argtypes = [ci, ci, ci, i64, i64, i64, C.c_double, dp, i64, dp, i64, C.c_double, dp, i64]
        L.oracle_zhemm.argtypes = [ci, ci, ci, i64, i64, i64, dp, dp, i64, dp, i64, dp, dp, i64]
        for f in (L.oracle_drank2k, L.oracle_zrank2k, L.oracle_dhemm, L.oracle_zhemm):
            f.restype = None
        for f in (L.oracle_dgetrf_nopiv, L.oracle_zgetrf_nopiv):
            f.argtypes = [i64, i64, i64, dp, i64]
            f.restype = i64
        for f in (L.oracle_matgen_d, L.oracle_matgen_z, L.oracle_tile_dgemm, L.oracle_tile_zgemm,
                  L.oracle_tile_dsyrk, L.oracle_tile_zherk, L.oracle_tile_dtrsm, L.oracle_tile_ztrsm,
                  L.oracle_dgemm, L.oracle_zgemm):
            f.restype = None
        _oracle = L
    return _oracle


def have_ref():
    return os.path.exists(REF_SO)


def ref():
    """The unmodified reference, HostTask, single rank (None if not built)."""
    global _ref
    if _ref is None:
        if not have_ref():
            return None
        L = C.CDLL(REF_SO)
        L.ref_version.restype = C.c_char_p
        L.ref_dgemm.argtypes = [C.c_int, C.c_int, i64, i64, i64, i64, C.c_double, dp, i64, dp, i64, C.c_double, dp, i64, C.c_int]
        L.ref_zgemm.argtypes = [C.c_int, C.c_int, i64, i64, i64, i64, dp, dp, i64, dp, i64, dp, dp, i64, C.c_int]
        for f in (L.ref_dpotrf, L.ref_zpotrf):
            f.argtypes = [C.c_int, i64, i64, dp, i64, C.c_int]
            f.restype = i64
        L.ref_internal_dgemm.argtypes = [C.c_int, C.c_int, i64, i64, i64, i64, C.c_double, dp, i64, dp, i64, C.c_double, dp, i64]
        L.ref_internal_zgemm.argtypes = [C.c_int, C.c_int, i64, i64, i64, i64, dp, dp, i64, dp, i64, dp, dp, i64]
        L.ref_internal_dherk.argtypes = [C.c_int, C.c_int, i64, i64, i64, C.c_double, dp, i64, C.c_double, dp, i64]
        L.ref_internal_zherk.argtypes = L.ref_internal_dherk.argtypes
        L.ref_internal_dtrsm.argtypes = [C.c_int] * 4 + [i64, i64, i64, C.c_double, dp, i64, dp, i64]
        L.ref_internal_ztrsm.argtypes = [C.c_int] * 4 + [i64, i64, i64, dp, dp, i64, dp, i64]
        for f in (L.ref_internal_dpotrf, L.ref_internal_zpotrf):
            f.argtypes = [C.c_int, i64, dp, i64]
            f.restype = i64
        # widened drivers: dtype 0 / 1, scalars as pointers to 2 doubles
        ci, vp = C.c_int, C.c_void_p
        if hasattr(L, "ref_rank_k_any"):
            L.ref_rank_k_any.argtypes = [ci, ci, ci, ci, i64, i64, i64, dp, vp, i64, dp, vp, i64, ci]
            L.ref_rank_2k_any.argtypes = [ci, ci, ci, ci, i64, i64, i64, dp, vp, i64, vp, i64, dp, vp, i64, ci]
            L.ref_hemm_any.argtypes = [ci, ci, ci, ci, i64, i64, i64, dp, vp, i64, vp, i64, dp, vp, i64, ci]
            L.ref_trsm_any.argtypes = [ci, ci, ci, ci, ci, i64, i64, i64, dp, vp, i64, vp, i64, ci]
            L.ref_posv_any.argtypes = [ci, ci, i64, i64, i64, vp, i64, vp, i64, ci]
            L.ref_gemmA_any.argtypes = [ci, ci, ci, i64, i64, i64, i64, dp, vp, i64, vp, i64, dp, vp, i64, ci]
            L.ref_getrf_nopiv_any.argtypes = [ci, i64, i64, i64, vp, i64, ci]
            for f in (L.ref_rank_k_any, L.ref_rank_2k_any, L.ref_hemm_any, L.ref_trsm_any,
                      L.ref_posv_any, L.ref_gemmA_any, L.ref_getrf_nopiv_any):
                f.restype = i64
        L.ref_matgen_d.argtypes = [C.c_int, i64, i64, i64, i64, i64, dp, i64]
        L.ref_matgen_z.argtypes = L.ref_matgen_d.argtypes
        L.ref_matgen_d.restype = None
        L.ref_matgen_z.restype = None
        _ref = L
    return _ref


def fmat(m, n, dtype=np.float64, ld=None):
    """column-major m x n view of an ld x n allocation (ld >= m)"""
    ld = ld or max(m, 1)
    buf = np.zeros((ld, n), dtype=dtype, order="F")
    return buf, buf[:m, :]


def matgen(seed, m, n, dtype=np.float64, dist=1, ioff=0, joff=0, lib=None):
    """matgen 'rand'-family matrix by the oracle (or the given library)."""
    A = np.zeros((m, n), dtype=dtype, order="F")
    L = lib or oracle()
    pre = "oracle_matgen_" if L is oracle() else "ref_matgen_"
    fn = getattr(L, pre + ("z" if dtype == np.complex128 else "d"))
    fn(dist, seed, m, n, ioff, joff, P(A), max(m, 1))
    return A
