"""Python face of libslate_b200.so, the host C++ mirror of SLATE's interface
(slate::Matrix / HermitianMatrix, slate::gemm, slate::potrf, internal::* on
Target::Devices).  Used by the tests, bench.py and smoke(); the product is the
C++/CUDA underneath, this file only marshals arguments.
"""
import ctypes as C
import os

import numpy as np

from ._lib import lib as _te_lib

_HERE = os.path.dirname(os.path.abspath(__file__))
HOST_LIB_PATH = os.path.join(_HERE, "lib", "libslate_b200.so")

F64, C64 = 0, 1
GENERAL, HERMITIAN, SYMMETRIC, TRIANGULAR = 0, 1, 2, 3
TARGET_DEVICES, TARGET_HOSTTASK = ord("D"), ord("T")

_lib = None
i64, vp, dp = C.c_int64, C.c_void_p, C.POINTER(C.c_double)


def hostlib():
    global _lib
    if _lib is not None:
        return _lib
    _te_lib()     # loads libtile_engine.so first (and fails loudly if it is missing)
    if not os.path.exists(HOST_LIB_PATH):
        raise RuntimeError(f"{HOST_LIB_PATH} is missing: run __graft_entry__.build()")
    L = C.CDLL(HOST_LIB_PATH)
    L.sb_last_error.restype = C.c_char_p
    L.sb_initialize.argtypes = [C.c_int, C.c_int, C.c_int, vp]
    L.sb_matrix_create.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, i64, i64, i64, C.c_int,
                                   C.c_int, C.c_int, vp, i64, C.POINTER(vp)]
    L.sb_matrix_create_tiled.argtypes = [C.c_int, C.c_int, C.c_int, i64, i64, C.POINTER(i64), i64,
                                         C.POINTER(i64), i64, C.c_int, C.c_int, C.c_int, C.POINTER(vp)]
    L.sb_matrix_from_devices.argtypes = [C.c_int, C.c_int, C.c_int, i64, i64, vp, i64, i64, C.c_int,
                                         C.c_int, C.POINTER(vp)]
    L.sb_matrix_destroy.argtypes = [vp]
    L.sb_matrix_generate.argtypes = [vp, i64, C.c_int, C.c_int]
    L.sb_matrix_to_host.argtypes = [vp, vp, i64]
    L.sb_matrix_from_host.argtypes = [vp, vp, i64]
    L.sb_matrix_sumsq.argtypes = [vp, dp]
    L.sb_pool_stats.argtypes = [vp, C.POINTER(i64), C.POINTER(i64)]
    L.sb_gemm.argtypes = [C.c_int, C.c_int, dp, vp, vp, dp, vp, C.c_int, C.c_int,
                          C.POINTER(C.c_float), C.c_int]
    L.sb_potrf.argtypes = [vp, C.c_int, C.c_int, C.POINTER(i64), C.POINTER(C.c_float)]
    L.sb_rank_k.argtypes = [C.c_int, C.c_int, dp, vp, dp, vp, C.c_int, C.c_int]
    L.sb_rank_2k.argtypes = [C.c_int, C.c_int, dp, vp, vp, dp, vp, C.c_int, C.c_int]
    L.sb_getrf_nopiv.argtypes = [vp, C.c_int, C.c_int, C.POINTER(i64)]
    L.sb_hemm.argtypes = [C.c_int, dp, vp, vp, dp, vp, C.c_int, C.c_int]
    L.sb_trsm.argtypes = [C.c_int, C.c_int, dp, vp, vp, C.c_int, C.c_int]
    L.sb_potrs.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, C.POINTER(i64)]
    L.sb_potrf_residual.argtypes = [vp, i64, C.c_int, dp, dp]
    ip = C.POINTER(i64)
    L.sb_plan_gemm_step.argtypes = [vp, vp, vp, i64, ip, i64, ip]
    L.sb_plan_potrf_step.argtypes = [vp, i64, ip, i64, ip]
    L.sb_plan_gemmA_step.argtypes = [vp, vp, vp, i64, ip, i64, ip]
    L.sb_plan_getrf_nopiv_step.argtypes = [vp, i64, ip, i64, ip]
    L.sb_plan_hemm_step.argtypes = [vp, vp, vp, i64, ip, i64, ip]
    L.sb_plan_pairs.argtypes = [C.c_int, vp, vp, vp, i64, C.c_int, C.c_int, ip, i64, ip]
    L.sb_tile_rank.argtypes = [vp, i64, i64]
    L.sb_tile_layout.argtypes = [vp, i64, i64]
    L.sb_tile_layout_convert.argtypes = [vp, i64, i64, C.c_int]
    L.sb_internal_gemm.argtypes = [C.c_int, C.c_int, C.c_int, dp, vp, vp, dp, vp, C.c_int, C.c_int]
    L.sb_internal_herk.argtypes = [C.c_int, C.c_int, C.c_double, vp, C.c_double, vp, C.c_int]
    L.sb_internal_syrk.argtypes = [C.c_int, C.c_int, dp, vp, dp, vp, C.c_int]
    L.sb_internal_trsm.argtypes = [C.c_int, C.c_int, C.c_int, dp, vp, vp, C.c_int]
    L.sb_internal_potrf.argtypes = [vp, C.c_int, C.POINTER(i64)]
    _lib = L
    return L


class SlateError(RuntimeError):
    pass


def _check(status):
    if status != 0:
        raise SlateError(hostlib().sb_last_error().decode())


def _scalar(x):
    z = complex(x)
    return (C.c_double * 2)(z.real, z.imag)


_initialized = False


def initialize(device=None, rank=0, size=1, nccl_id=None):
    """Bind this process to a GPU (default $LOCAL_RANK) and, for size > 1, join the NCCL
    communicator whose unique id rank 0 created with nccl_unique_id()."""
    global _initialized
    if device is None:
        device = int(os.environ.get("LOCAL_RANK", "0"))
    buf = None
    if size > 1 and _te_lib().te_device_count() > 0:
        assert nccl_id is not None and len(nccl_id) == 128
        buf = C.create_string_buffer(bytes(nccl_id), 128)
    # (without a GPU a multi-rank process can still build index maps and broadcast plans)
    _check(hostlib().sb_initialize(device, rank, size, buf))
    _initialized = True


def finalize():
    global _initialized
    if _initialized:
        _check(hostlib().sb_finalize())
    _initialized = False


def nccl_unique_id():
    buf = C.create_string_buffer(128)
    st = _te_lib().te_comm_unique_id(buf, 128)
    if st != 0:
        raise SlateError(_te_lib().te_last_error().decode())
    return buf.raw


def sync():
    _check(hostlib().sb_sync())


def _dtype_code(dtype):
    return C64 if np.dtype(dtype) == np.complex128 else F64


class _Base:
    kind = GENERAL

    def __init__(self, m, n, nb, p=1, q=1, dtype=np.float64, origin="D", uplo="L", diag="N",
                 host_array=None):
        """host_array: column-major (ld x n) buffer whose first m rows are the matrix."""
        if not _initialized:
            initialize()
        self.m, self.n, self.nb, self.p, self.q = m, n, nb, p, q
        self.dtype = np.dtype(dtype)
        self.uplo = uplo
        self._host_array = host_array          # keep user memory alive (fromLAPACK)
        h = vp()
        ptr, ld = None, 0
        if host_array is not None:
            assert host_array.flags.f_contiguous and host_array.dtype == self.dtype
            if origin != "S":
                assert host_array.shape[0] >= m and host_array.shape[1] == n
            ptr, ld = host_array.ctypes.data, host_array.shape[0]
        _check(hostlib().sb_matrix_create(_dtype_code(dtype), self.kind, ord(uplo), ord(diag), m, n,
                                          nb, p, q, ord(origin), ptr, ld, C.byref(h)))
        self.h = h

    def __del__(self):
        try:
            if getattr(self, "h", None):
                hostlib().sb_matrix_destroy(self.h)
                self.h = None
        except Exception:
            pass

    def free(self):
        if self.h:
            _check(hostlib().sb_matrix_destroy(self.h))
            self.h = None

    def generate(self, seed, dominant=False, dist=1):
        """matgen 'rand' (dist 1) / 'rands' (dist 2); dominant adds n to the diagonal."""
        _check(hostlib().sb_matrix_generate(self.h, seed, int(dominant), dist))
        return self

    def to_numpy(self):
        """Gather this rank's tiles into a full-size column-major array (zeros elsewhere)."""
        out = np.zeros((self.m, self.n), dtype=self.dtype, order="F")
        _check(hostlib().sb_matrix_to_host(self.h, out.ctypes.data, max(self.m, 1)))
        return out

    def from_numpy(self, a):
        a = np.asfortranarray(a, dtype=self.dtype)
        assert a.shape == (self.m, self.n)
        _check(hostlib().sb_matrix_from_host(self.h, a.ctypes.data, max(self.m, 1)))
        return self

    def tile_layout(self, i, j):
        """'C' / 'R' of the device instance of tile (i, j); None if there is none"""
        r = hostlib().sb_tile_layout(self.h, i, j)
        return chr(r) if r else None

    def tile_layout_convert(self, i, j, layout):
        """BaseMatrix::tileLayoutConvert of the device instance (enqueued on compute queue 0)"""
        _check(hostlib().sb_tile_layout_convert(self.h, i, j, ord(layout)))
        return self

    def sumsq(self):
        out = C.c_double()
        _check(hostlib().sb_matrix_sumsq(self.h, C.byref(out)))
        return out.value

    def pool_stats(self):
        a, c = i64(), i64()
        _check(hostlib().sb_pool_stats(self.h, C.byref(a), C.byref(c)))
        return a.value, c.value


def _adopt(cls, h, m, n, nb, p, q, dtype, uplo="L", keep=None):
    """wrap a handle made by one of the alternative constructors"""
    self = cls.__new__(cls)
    self.m, self.n, self.nb, self.p, self.q = m, n, nb, p, q
    self.dtype = np.dtype(dtype)
    self.uplo = uplo
    self._host_array = keep            # keeps the caller's (device) memory alive
    self.h = h
    return self


def _sizes(v):
    return (i64 * len(v))(*[int(x) for x in v])


class Matrix(_Base):
    kind = GENERAL

    @classmethod
    def tiled(cls, row_sizes, col_sizes, p=1, q=1, dtype=np.float64, origin="D"):
        """Lambda constructor (include/slate/Matrix.hh:33-38): tile (i, j) is row_sizes[i] x
        col_sizes[j]; sizes need not be uniform.  Owners 2D block-cyclic over p x q."""
        if not _initialized:
            initialize()
        m, n = int(sum(row_sizes)), int(sum(col_sizes))
        h = vp()
        _check(hostlib().sb_matrix_create_tiled(_dtype_code(dtype), GENERAL, ord("L"), m, n, _sizes(row_sizes),
                                                len(row_sizes), _sizes(col_sizes), len(col_sizes), p, q,
                                                ord(origin), C.byref(h)))
        return _adopt(cls, h, m, n, max(max(row_sizes), max(col_sizes)), p, q, dtype)

    @classmethod
    def fromDevices(cls, m, n, dev_tensor, lda, nb, p=1, q=1, dtype=np.float64):
        """Matrix::fromDevices (include/slate/Matrix.hh:102-112): tiles are views into the caller's
        device array (a torch CUDA tensor holding this rank's local matrix, column major, leading
        dimension lda); results land in that memory."""
        if not _initialized:
            initialize()
        h = vp()
        _check(hostlib().sb_matrix_from_devices(_dtype_code(dtype), GENERAL, ord("L"), m, n,
                                                dev_tensor.data_ptr(), lda, nb, p, q, C.byref(h)))
        return _adopt(cls, h, m, n, nb, p, q, dtype, keep=dev_tensor)

    @classmethod
    def fromLAPACK(cls, a, nb, p=1, q=1, m=None):
        """View of the first m rows (default all) of the column-major array a (ld = a.shape[0])."""
        return cls(a.shape[0] if m is None else m, a.shape[1], nb, p, q, a.dtype, origin="H",
                   host_array=a)


    @classmethod
    def fromScaLAPACK(cls, m, n, local, nb, p, q):
        """`local` is this rank's 2D block-cyclic local array (column-major, ld = local.shape[0]),
        as ScaLAPACK lays it out (include/slate/Matrix.hh:486-520)."""
        return cls(m, n, nb, p, q, local.dtype, origin="S", host_array=local)


class HermitianMatrix(_Base):
    kind = HERMITIAN

    def __init__(self, uplo, n, nb, p=1, q=1, dtype=np.float64, origin="D", host_array=None):
        super().__init__(n, n, nb, p, q, dtype, origin, uplo=uplo, host_array=host_array)

    @classmethod
    def fromLAPACK(cls, uplo, a, nb, p=1, q=1):
        return cls(uplo, a.shape[0], nb, p, q, a.dtype, origin="H", host_array=a)

    @classmethod
    def tiled(cls, uplo, sizes, p=1, q=1, dtype=np.float64, origin="D"):
        """Lambda constructor: square, tile (i, j) is sizes[i] x sizes[j]."""
        if not _initialized:
            initialize()
        n = int(sum(sizes))
        h = vp()
        _check(hostlib().sb_matrix_create_tiled(_dtype_code(dtype), HERMITIAN, ord(uplo), n, n, None, 0,
                                                _sizes(sizes), len(sizes), p, q, ord(origin), C.byref(h)))
        return _adopt(cls, h, n, n, max(sizes), p, q, dtype, uplo=uplo)

    @classmethod
    def fromDevices(cls, uplo, n, dev_tensor, lda, nb, p=1, q=1, dtype=np.float64):
        if not _initialized:
            initialize()
        h = vp()
        _check(hostlib().sb_matrix_from_devices(_dtype_code(dtype), HERMITIAN, ord(uplo), n, n,
                                                dev_tensor.data_ptr(), lda, nb, p, q, C.byref(h)))
        return _adopt(cls, h, n, n, nb, p, q, dtype, uplo=uplo, keep=dev_tensor)


class SymmetricMatrix(HermitianMatrix):
    kind = SYMMETRIC


class TriangularMatrix(_Base):
    kind = TRIANGULAR

    def __init__(self, uplo, diag, n, nb, p=1, q=1, dtype=np.float64, origin="D", host_array=None):
        super().__init__(n, n, nb, p, q, dtype, origin, uplo=uplo, diag=diag, host_array=host_array)

    @classmethod
    def fromLAPACK(cls, uplo, diag, a, nb, p=1, q=1):
        return cls(uplo, diag, a.shape[0], nb, p, q, a.dtype, origin="H", host_array=a)


def gemm(alpha, A, B, beta, Cm, opa="N", opb="N", lookahead=1, target="D", timed=False,
         method="auto"):
    """slate::gemm: C = alpha op(A) op(B) + beta C.  Returns device milliseconds if timed.
    method: "auto" (gemmA when B has one block column, else gemmC), "A" or "C"."""
    ms = C.c_float()
    _check(hostlib().sb_gemm(ord(opa), ord(opb), _scalar(alpha), A.h, B.h, _scalar(beta), Cm.h,
                             lookahead, ord(target), C.byref(ms) if timed else None,
                             {"auto": 0, "A": 1, "C": 2}[method]))
    return ms.value if timed else None


def potrf(A, lookahead=1, target="D", timed=False):
    """slate::potrf.  Returns info, or (info, device milliseconds) if timed."""
    info, ms = i64(), C.c_float()
    _check(hostlib().sb_potrf(A.h, lookahead, ord(target), C.byref(info),
                              C.byref(ms) if timed else None))
    return (info.value, ms.value) if timed else info.value


def herk(alpha, A, beta, Cm, opa="N", lookahead=1, target="D"):
    """slate::herk: C = alpha op(A) op(A)^H + beta C on the stored triangle of C."""
    _check(hostlib().sb_rank_k(0, ord(opa), _scalar(alpha), A.h, _scalar(beta), Cm.h, lookahead,
                               ord(target)))


def syrk(alpha, A, beta, Cm, opa="N", lookahead=1, target="D"):
    """slate::syrk: C = alpha op(A) op(A)^T + beta C on the stored triangle of C."""
    _check(hostlib().sb_rank_k(1, ord(opa), _scalar(alpha), A.h, _scalar(beta), Cm.h, lookahead,
                               ord(target)))


def her2k(alpha, A, B, beta, Cm, opa="N", lookahead=1, target="D"):
    """slate::her2k: C = alpha op(A) op(B)^H + conj(alpha) op(B) op(A)^H + beta C."""
    _check(hostlib().sb_rank_2k(0, ord(opa), _scalar(alpha), A.h, B.h, _scalar(beta), Cm.h, lookahead,
                                ord(target)))


def syr2k(alpha, A, B, beta, Cm, opa="N", lookahead=1, target="D"):
    """slate::syr2k: C = alpha op(A) op(B)^T + alpha op(B) op(A)^T + beta C."""
    _check(hostlib().sb_rank_2k(1, ord(opa), _scalar(alpha), A.h, B.h, _scalar(beta), Cm.h, lookahead,
                                ord(target)))


def getrf_nopiv(A, lookahead=1, target="D"):
    """slate::getrf_nopiv: A = L U in place without pivoting.  Returns info."""
    info = i64()
    _check(hostlib().sb_getrf_nopiv(A.h, lookahead, ord(target), C.byref(info)))
    return info.value


def hemm(side, alpha, A, B, beta, Cm, lookahead=1, target="D"):
    """slate::hemm (A a HermitianMatrix) or slate::symm (A a SymmetricMatrix):
    C = alpha A B + beta C (side 'L') or alpha B A + beta C (side 'R')."""
    _check(hostlib().sb_hemm(ord(side), _scalar(alpha), A.h, B.h, _scalar(beta), Cm.h, lookahead,
                             ord(target)))


symm = hemm


def trsm(side, alpha, A, B, opa="N", lookahead=1, target="D"):
    """slate::trsm: op(A) X = alpha B (side 'L') or X op(A) = alpha B (side 'R'); B <- X."""
    _check(hostlib().sb_trsm(ord(side), ord(opa), _scalar(alpha), A.h, B.h, lookahead, ord(target)))


def potrs(A, B, lookahead=1, target="D"):
    """slate::potrs: solve A X = B with the Cholesky factor held in A; B <- X."""
    info = i64()
    _check(hostlib().sb_potrs(A.h, B.h, 0, lookahead, ord(target), C.byref(info)))


def posv(A, B, lookahead=1, target="D"):
    """slate::posv: factor A and solve A X = B; returns potrf's info."""
    info = i64()
    _check(hostlib().sb_potrs(A.h, B.h, 1, lookahead, ord(target), C.byref(info)))
    return info.value


def potrf_residual(A, seed, dominant=True):
    """(||A0 - L L^H||_F^2, ||A0||_F^2) over this rank's tiles; A holds the lower factor."""
    r, a0 = C.c_double(), C.c_double()
    _check(hostlib().sb_potrf_residual(A.h, seed, int(dominant), C.byref(r), C.byref(a0)))
    return r.value, a0.value


def _unflatten(buf, used):
    ops, pos = [], 0
    while pos < used:
        which, i, j, root, nd = (int(x) for x in buf[pos:pos + 5])
        ops.append((which, i, j, root, [int(x) for x in buf[pos + 5:pos + 5 + nd]]))
        pos += 5 + nd
    return ops


def plan_gemm_step(A, B, Cm, k):
    """Broadcasts slate::gemm issues for SUMMA step k: [(which, i, j, root, [dest ranks])],
    which = 0 for A(i, k), 1 for B(k, j).  Host logic only (no GPU needed)."""
    cap = 8 * (A.m // A.nb + Cm.n // Cm.nb + 8) * (Cm.p * Cm.q + 6)
    buf = (i64 * cap)()
    used = i64()
    _check(hostlib().sb_plan_gemm_step(A.h, B.h, Cm.h, k, buf, cap, C.byref(used)))
    return _unflatten(buf, used.value)


def plan_gemmA_step(A, B, Cm, k):
    """column k of slate::gemmA: [(0, i, j, root, dests)] for B's broadcast, then
    [(2, i, j, root, sources)] for the reduction of C(:, k)"""
    cap = 64 + 2 * (A.m // A.nb + A.n // A.nb + 8) * (A.p * A.q + 6)
    buf = (i64 * cap)()
    used = i64()
    _check(hostlib().sb_plan_gemmA_step(A.h, B.h, Cm.h, k, buf, cap, C.byref(used)))
    return _unflatten(buf, used.value)


def plan_pairs(kind, mats, k, which, rank):
    """Messages of one broadcast list grouped by (root, destination) pair as `rank` sees them
    (BaseMatrix::bcastPairs): kind "gemm" (mats = (A, B, C)) or "potrf" (mats = (A,)).
    Returns [(direction, peer, bytes, [(i, j), ...])], direction 0 = send, 1 = receive."""
    cap = 1 << 16
    buf = (i64 * cap)()
    used = i64()
    hs = [m.h for m in mats] + [None] * (3 - len(mats))
    _check(hostlib().sb_plan_pairs(0 if kind == "gemm" else 1, hs[0], hs[1], hs[2], k, which, rank, buf,
                                   cap, C.byref(used)))
    out, pos = [], 0
    while pos < used.value:
        d, peer, nt, nbytes = (int(x) for x in buf[pos:pos + 4])
        tiles = [(int(buf[pos + 4 + 2 * t]), int(buf[pos + 5 + 2 * t])) for t in range(nt)]
        out.append((d, peer, nbytes, tiles))
        pos += 4 + 2 * nt
    return out


def plan_hemm_step(A, B, Cm, k):
    """Broadcasts of step k of slate::hemm, side Left: which = 0 the tiles standing for block
    column k of the Hermitian A, 1 block row k of B."""
    cap = 8 * (Cm.m // Cm.nb + Cm.n // Cm.nb + 8) * (Cm.p * Cm.q + 6)
    buf = (i64 * cap)()
    used = i64()
    _check(hostlib().sb_plan_hemm_step(A.h, B.h, Cm.h, k, buf, cap, C.byref(used)))
    return _unflatten(buf, used.value)


def plan_getrf_nopiv_step(A, k):
    """Broadcasts of LU step k (slate::getrf_nopiv): which = 0 the diagonal tile, 1 the tiles
    of block column k along their rows, 2 the tiles of block row k down their columns."""
    cap = 8 * (A.m // A.nb + A.n // A.nb + 8) * (A.p * A.q + 6)
    buf = (i64 * cap)()
    used = i64()
    _check(hostlib().sb_plan_getrf_nopiv_step(A.h, k, buf, cap, C.byref(used)))
    return _unflatten(buf, used.value)


def plan_potrf_step(A, k):
    """Broadcasts slate::potrf issues in step k: which = 0 the diagonal tile (before the
    panel solve), which = 1 the panel tiles (after it)."""
    cap = 8 * (A.n // A.nb + 8) * (A.p * A.q + 6)
    buf = (i64 * cap)()
    used = i64()
    _check(hostlib().sb_plan_potrf_step(A.h, k, buf, cap, C.byref(used)))
    return _unflatten(buf, used.value)


def tile_rank(A, i, j):
    return hostlib().sb_tile_rank(A.h, i, j)


def internal_gemm(alpha, A, B, beta, Cm, opa="N", opb="N", opc="N", target="D", queue=0):
    _check(hostlib().sb_internal_gemm(ord(opa), ord(opb), ord(opc), _scalar(alpha), A.h, B.h,
                                      _scalar(beta), Cm.h, ord(target), queue))


def internal_herk(alpha, A, beta, Cm, opa="N", opc="N", queue=0):
    _check(hostlib().sb_internal_herk(ord(opa), ord(opc), float(alpha), A.h, float(beta), Cm.h, queue))


def internal_syrk(alpha, A, beta, Cm, opa="N", opc="N", queue=0):
    _check(hostlib().sb_internal_syrk(ord(opa), ord(opc), _scalar(alpha), A.h, _scalar(beta), Cm.h,
                                      queue))


def internal_trsm(side, alpha, A, B, opa="N", opb="N", queue=0):
    _check(hostlib().sb_internal_trsm(ord(side), ord(opa), ord(opb), _scalar(alpha), A.h, B.h, queue))


def internal_potrf(A, queue=0):
    info = i64()
    _check(hostlib().sb_internal_potrf(A.h, queue, C.byref(info)))
    return info.value
