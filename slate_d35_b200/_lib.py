"""ctypes binding of libtile_engine.so (the C ABI declared in include/tile_engine.h).

The library is the product; there is no Python or CPU fallback: if the shared
object is missing, loading raises.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libtile_engine.so")

TE_OK = 0
TE_F64, TE_C64 = 0, 1
TE_TILE_GENERAL, TE_TILE_LOWER, TE_TILE_UPPER, TE_TILE_HERMITIAN = 0, 1, 2, 4


class TileDesc(C.Structure):
    _fields_ = [
        ("a", C.c_void_p), ("b", C.c_void_p), ("c", C.c_void_p),
        ("m", C.c_int32), ("n", C.c_int32), ("k", C.c_int32),
        ("lda", C.c_int32), ("ldb", C.c_int32), ("ldc", C.c_int32),
        ("flags", C.c_int32), ("reserved", C.c_int32),
    ]


assert C.sizeof(TileDesc) == 56


class TrsmDesc(C.Structure):
    _fields_ = [("b", C.c_void_p), ("m", C.c_int32), ("n", C.c_int32),
                ("ldb", C.c_int32), ("reserved", C.c_int32)]


assert C.sizeof(TrsmDesc) == 24

_lib = None


def lib():
    """Load (once) and return the tile-engine library with typed prototypes."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'`"
            " (there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    L.te_last_error.restype = C.c_char_p
    L.te_version.restype = C.c_char_p
    L.te_device_count.restype = C.c_int
    L.te_gemm_batched.restype = C.c_int
    L.te_gemm_batched.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                  C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                  C.c_int32, C.c_int32]
    L.te_gemm_inline.restype = C.c_int
    L.te_gemm_inline.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                 C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
    vp, i32, i64, sz = C.c_void_p, C.c_int32, C.c_int64, C.c_size_t
    L.te_potrf_scratch_bytes.restype = sz
    L.te_potrf_scratch_bytes.argtypes = [C.c_int]
    L.te_potrf_tile.argtypes = [C.c_int, vp, C.c_int, C.c_int, C.c_int, vp, C.c_int, vp, vp, sz]
    L.te_getrf_nopiv_scratch_bytes.restype = sz
    L.te_getrf_nopiv_scratch_bytes.argtypes = [C.c_int]
    L.te_getrf_nopiv_tile.argtypes = [C.c_int, vp, C.c_int, C.c_int, C.c_int, vp, C.c_int, vp, vp, sz]
    L.te_trsm_scratch_bytes.restype = sz
    L.te_trsm_scratch_bytes.argtypes = [C.c_int, C.c_int, i64]
    L.te_trsm_batched.argtypes = [C.c_int, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp, vp,
                                  C.c_int, C.c_int, vp, i64, i32, i32, vp, sz]
    L.te_trsm_batched_static.argtypes = [C.c_int, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                         vp, vp, C.c_int, C.c_int, vp, i64, i32, i32]
    L.te_matgen_tile.argtypes = [C.c_int, vp, C.c_int, C.c_int, i64, C.c_int, C.c_int, i64, i64,
                                 vp, i64, C.c_double]
    L.te_tile_add.argtypes = [C.c_int, vp, C.c_int, C.c_int, C.c_int, C.c_double, vp, i64,
                              C.c_double, vp, i64]
    L.te_tile_sumsq.argtypes = [C.c_int, vp, C.c_int, C.c_int, C.c_int, C.c_int, vp, i64, vp]
    L.te_tile_keep_triangle.argtypes = [C.c_int, vp, C.c_int, C.c_int, C.c_int, C.c_int, vp, i64]
    L.te_tile_scale.argtypes = [C.c_int, vp, C.c_int, C.c_int, C.c_int, vp, vp, i64]
    L.te_tile_transpose.argtypes = [C.c_int, vp, C.c_int, C.c_int, C.c_int, C.c_int, vp, i64, vp, i64]
    L.te_tile_transpose_batch.argtypes = [C.c_int, vp, C.c_int, C.c_int, C.c_int, C.c_int, vp, i64,
                                          vp, i64, C.c_int]
    L.te_tile_symmetrize.argtypes = [C.c_int, vp, C.c_int, C.c_int, C.c_int, C.c_int, vp, i64]
    L.te_gemm_set_items_per_cta.argtypes = [C.c_int]
    L.te_device_mem_info.argtypes = [C.c_int, C.POINTER(sz), C.POINTER(sz)]
    L.te_pool_create.argtypes = [C.c_int, sz, C.POINTER(vp)]
    L.te_pool_reserve.argtypes = [vp, i64]
    L.te_pool_alloc.argtypes = [vp, C.POINTER(vp)]
    L.te_pool_free.argtypes = [vp, vp]
    L.te_pool_stats.argtypes = [vp, C.POINTER(sz), C.POINTER(sz), C.POINTER(sz)]
    L.te_pool_clear.argtypes = [vp]
    L.te_pool_destroy.argtypes = [vp]
    L.te_peak_fp64.restype = C.c_int
    L.te_peak_fp64.argtypes = [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double)]
    _lib = L
    return L


def check(status, what=""):
    if status != TE_OK:
        raise RuntimeError(f"tile_engine {what} failed: status {status}: "
                           f"{lib().te_last_error().decode()}")
