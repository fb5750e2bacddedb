"""ctypes face of the ScaLAPACK-convention entry points (include/slate_b200_scalapack.h):
slate_pdgemm, slate_pdpotrf, ... exactly as a Fortran / C ScaLAPACK caller passes them
(everything by reference, 1-based offsets, 9-integer descriptors).  Helpers numroc /
descinit / local_part follow ScaLAPACK's TOOLS routines of the same names.
"""
import ctypes as C

import numpy as np

from . import host

_bound = False
ip = C.POINTER(C.c_int)


def _lib():
    global _bound
    L = host.hostlib()
    if not _bound:
        L.slate_scalapack_grid_set.argtypes = [C.c_int, C.c_char, C.c_int, C.c_int]
        L.slate_scalapack_last_error.restype = C.c_char_p
        _bound = True
    return L


class ScalapackError(RuntimeError):
    pass


def grid_set(ctxt, order, nprow, npcol):
    if _lib().slate_scalapack_grid_set(ctxt, order.encode()[:1], nprow, npcol) != 0:
        raise ValueError("slate_scalapack_grid_set: bad argument")


def numroc(n, nb, iproc, nprocs, isrc=0):
    """rows (columns) of a dimension of n in blocks of nb that land on process iproc"""
    dist = (nprocs + iproc - isrc) % nprocs
    nblocks = n // nb
    out = (nblocks // nprocs) * nb
    extra = nblocks % nprocs
    if dist < extra:
        out += nb
    elif dist == extra:
        out += n % nb
    return out


def descinit(m, n, mb, nb, ctxt, lld, rsrc=0, csrc=0):
    return (C.c_int * 9)(1, ctxt, m, n, mb, nb, rsrc, csrc, max(1, lld))


def local_part(full, mb, nb, myrow, mycol, p, q):
    """this process's 2D block-cyclic local array (column-major) and its global indices"""
    m, n = full.shape
    ri = [i for t in range(myrow, -(-m // mb), p) for i in range(t * mb, min((t + 1) * mb, m))]
    ci = [j for t in range(mycol, -(-n // nb), q) for j in range(t * nb, min((t + 1) * nb, n))]
    return np.asfortranarray(full[np.ix_(ri, ci)]), ri, ci


def _i(x):
    return C.byref(C.c_int(int(x)))


def _s(x, dtype):
    """scalar by reference in the routine's precision"""
    if np.dtype(dtype) == np.complex128:
        return (C.c_double * 2)(complex(x).real, complex(x).imag)
    return C.byref(C.c_double(float(x)))


def _p(a):
    assert a.flags.f_contiguous or a.size == 0
    return a.ctypes.data_as(C.c_void_p)


def _finish(L):
    if L.slate_scalapack_status() != 0:
        raise ScalapackError(L.slate_scalapack_last_error().decode())


def _z(a):
    return "z" if a.dtype == np.complex128 else "d"


def pgemm(transa, transb, m, n, k, alpha, a, ia, ja, desca, b, ib, jb, descb, beta, c, ic, jc, descc):
    L = _lib()
    f = getattr(L, f"slate_p{_z(c)}gemm")
    f(transa.encode(), transb.encode(), _i(m), _i(n), _i(k), _s(alpha, c.dtype), _p(a), _i(ia), _i(ja),
      desca, _p(b), _i(ib), _i(jb), descb, _s(beta, c.dtype), _p(c), _i(ic), _i(jc), descc)
    _finish(L)


def ppotrf(uplo, n, a, ia, ja, desca):
    L = _lib()
    info = C.c_int(-999)
    getattr(L, f"slate_p{_z(a)}potrf")(uplo.encode(), _i(n), _p(a), _i(ia), _i(ja), desca, C.byref(info))
    _finish(L)
    return info.value


def _solve(name, uplo, n, nrhs, a, ia, ja, desca, b, ib, jb, descb):
    L = _lib()
    info = C.c_int(-999)
    getattr(L, f"slate_p{_z(a)}{name}")(uplo.encode(), _i(n), _i(nrhs), _p(a), _i(ia), _i(ja), desca,
                                        _p(b), _i(ib), _i(jb), descb, C.byref(info))
    _finish(L)
    return info.value


def ppotrs(*args):
    return _solve("potrs", *args)


def pposv(*args):
    return _solve("posv", *args)


def ptrsm(side, uplo, transa, diag, m, n, alpha, a, ia, ja, desca, b, ib, jb, descb):
    L = _lib()
    getattr(L, f"slate_p{_z(b)}trsm")(side.encode(), uplo.encode(), transa.encode(), diag.encode(),
                                      _i(m), _i(n), _s(alpha, b.dtype), _p(a), _i(ia), _i(ja), desca,
                                      _p(b), _i(ib), _i(jb), descb)
    _finish(L)


def psyrk(uplo, trans, n, k, alpha, a, ia, ja, desca, beta, c, ic, jc, descc):
    L = _lib()
    getattr(L, f"slate_p{_z(c)}syrk")(uplo.encode(), trans.encode(), _i(n), _i(k), _s(alpha, c.dtype),
                                      _p(a), _i(ia), _i(ja), desca, _s(beta, c.dtype), _p(c), _i(ic),
                                      _i(jc), descc)
    _finish(L)


def pherk(uplo, trans, n, k, alpha, a, ia, ja, desca, beta, c, ic, jc, descc):
    L = _lib()
    assert c.dtype == np.complex128
    L.slate_pzherk(uplo.encode(), trans.encode(), _i(n), _i(k), _s(alpha, np.float64), _p(a), _i(ia),
                   _i(ja), desca, _s(beta, np.float64), _p(c), _i(ic), _i(jc), descc)
    _finish(L)


def phemm(side, uplo, m, n, alpha, a, ia, ja, desca, b, ib, jb, descb, beta, c, ic, jc, descc,
          hermitian=True):
    """pzhemm (hermitian, complex only) or p?symm"""
    L = _lib()
    name = "slate_pzhemm" if hermitian else f"slate_p{_z(c)}symm"
    assert not (hermitian and c.dtype != np.complex128)
    getattr(L, name)(side.encode(), uplo.encode(), _i(m), _i(n), _s(alpha, c.dtype), _p(a), _i(ia),
                     _i(ja), desca, _p(b), _i(ib), _i(jb), descb, _s(beta, c.dtype), _p(c), _i(ic),
                     _i(jc), descc)
    _finish(L)


def psymm(*args):
    return phemm(*args, hermitian=False)


def psyr2k(uplo, trans, n, k, alpha, a, ia, ja, desca, b, ib, jb, descb, beta, c, ic, jc, descc):
    L = _lib()
    getattr(L, f"slate_p{_z(c)}syr2k")(uplo.encode(), trans.encode(), _i(n), _i(k), _s(alpha, c.dtype),
                                       _p(a), _i(ia), _i(ja), desca, _p(b), _i(ib), _i(jb), descb,
                                       _s(beta, c.dtype), _p(c), _i(ic), _i(jc), descc)
    _finish(L)


def pher2k(uplo, trans, n, k, alpha, a, ia, ja, desca, b, ib, jb, descb, beta, c, ic, jc, descc):
    L = _lib()
    assert c.dtype == np.complex128
    L.slate_pzher2k(uplo.encode(), trans.encode(), _i(n), _i(k), _s(alpha, c.dtype), _p(a), _i(ia),
                    _i(ja), desca, _p(b), _i(ib), _i(jb), descb, _s(beta, np.float64), _p(c), _i(ic),
                    _i(jc), descc)
    _finish(L)
