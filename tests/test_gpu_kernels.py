"""Parity of the CUDA kernels, called through the C ABI (include/tile_engine.h),
against the CPU oracle on the same seeded (matgen) inputs.  Needs a GPU.

Tolerances follow the reference's own tests: 3*sqrt(k)*eps elementwise for one block
update (unit_test/test_internal_blas.cc:446-447), 50*eps*n-style bounds for the
factorisation / solve (test/test_posv.cc:302-343 uses tol = 50).  Complex uses 4k
real products per entry.
"""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as ol

pytestmark = pytest.mark.gpu
EPS = np.finfo(np.float64).eps


@pytest.fixture(scope="module")
def te():
    import torch
    from slate_d35_b200 import lib
    assert torch.cuda.is_available()
    torch.cuda.set_device(0)
    L = lib()
    assert L.te_device_count() >= 1
    return L


def dev(a):
    """numpy (F-order) -> torch cuda tensor holding the same bytes (col-major)."""
    import torch
    flat = np.ascontiguousarray(a.T).reshape(-1)           # column-major order
    if np.iscomplexobj(a):
        flat = flat.view(np.float64)
    return torch.from_numpy(flat.copy()).cuda()


def host(t, shape, dtype):
    flat = t.cpu().numpy()
    if dtype == np.complex128:
        flat = flat.view(np.complex128)
    m, n = shape
    return np.asfortranarray(flat.reshape(n, m).T)


def stream():
    import torch
    return torch.cuda.current_stream().cuda_stream


def scalar(x, dtype):
    if dtype == np.complex128:
        return (C.c_double * 2)(complex(x).real, complex(x).imag)
    return (C.c_double * 1)(float(x))


DT = {np.float64: 0, np.complex128: 1}


def run_gemm(te, dtype, opa, opb, m, n, k, alpha, beta, flags=0, count=2, inline=False, ld_pad=0):
    import torch
    from slate_d35_b200 import TileDesc, check
    descs = (TileDesc * count)()
    keep, cases = [], []
    for t in range(count):
        ash = (m, k) if opa == "N" else (k, m)
        bsh = (k, n) if opb == "N" else (n, k)
        bufs = []
        for seed, sh in ((10 + t, ash), (20 + t, bsh), (30 + t, (m, n))):
            full = ol.matgen(seed, sh[0] + ld_pad, sh[1], dtype, dist=2)
            bufs.append(full)
        A, B, Cm = (b[: sh[0], :] for b, sh in zip(bufs, (ash, bsh, (m, n))))
        dA, dB, dC = dev(bufs[0]), dev(bufs[1]), dev(bufs[2])
        keep += [dA, dB, dC]
        d = descs[t]
        d.a, d.b, d.c = dA.data_ptr(), dB.data_ptr(), dC.data_ptr()
        d.m, d.n, d.k = m, n, k
        d.lda, d.ldb, d.ldc = bufs[0].shape[0], bufs[1].shape[0], bufs[2].shape[0]
        d.flags = flags
        cases.append((A, B, bufs[2], dC))
    al, be = scalar(alpha, dtype), scalar(beta, dtype)
    if inline:
        check(te.te_gemm_inline(0, stream(), DT[dtype], ord(opa), ord(opb), al, be, descs, count))
    else:
        dd = torch.frombuffer(bytearray(bytes(descs)), dtype=torch.uint8).cuda()
        check(te.te_gemm_batched(0, stream(), DT[dtype], ord(opa), ord(opb), al, be,
                                 dd.data_ptr(), count, m, n))
    torch.cuda.synchronize()
    O = ol.oracle()
    for A, B, Cfull, dC in cases:
        want = Cfull.copy(order="F")
        got = host(dC, Cfull.shape, dtype)
        lda, ldb, ldc = A.base.shape[0] if A.base is not None else A.shape[0], \
            B.base.shape[0] if B.base is not None else B.shape[0], Cfull.shape[0]
        if dtype == np.float64:
            O.oracle_tile_dgemm(ord(opa), ord(opb), m, n, k, float(alpha), ol.P(np.asfortranarray(A)),
                                A.shape[0], ol.P(np.asfortranarray(B)), B.shape[0], float(beta),
                                ol.P(want), ldc)
        else:
            O.oracle_tile_zgemm(ord(opa), ord(opb), m, n, k, ol.Z(alpha), ol.P(np.asfortranarray(A)),
                                A.shape[0], ol.P(np.asfortranarray(B)), B.shape[0], ol.Z(beta),
                                ol.P(want), ldc)
        if flags & 1:
            want = np.where(np.tril(np.ones_like(want, dtype=bool)), want, Cfull)
        if flags & 2:
            want = np.where(np.triu(np.ones_like(want, dtype=bool)), want, Cfull)
        if (flags & 4) and dtype == np.complex128:
            idx = np.arange(min(m, n))
            want[idx, idx] = want[idx, idx].real
        # padding rows (ld > m) must be untouched
        assert np.array_equal(got[m:, :], Cfull[m:, :])
        scale = max(1.0, np.abs(want[:m]).max())
        kk = k * (4 if dtype == np.complex128 else 1)
        err = np.abs(got[:m] - want[:m]).max() / scale
        assert err <= 3 * np.sqrt(max(kk, 1)) * EPS, (err, opa, opb, m, n, k)


@pytest.mark.parametrize("opa", "NT")
@pytest.mark.parametrize("opb", "NT")
@pytest.mark.parametrize("shape", [(80, 64, 16), (512, 512, 512), (130, 257, 37), (1, 1, 1),
                                   (128, 128, 0), (129, 127, 513)])
def test_dgemm_tiles_vs_oracle(te, opa, opb, shape):
    m, n, k = shape
    run_gemm(te, np.float64, opa, opb, m, n, k, 3.141592653589793, 2.718281828459045)


@pytest.mark.parametrize("inline", [False, True])
@pytest.mark.parametrize("ld_pad", [0, 1, 2])
def test_dgemm_ragged_ld_and_inline(te, inline, ld_pad):
    """odd leading dimensions take the 8-byte copy path (fromLAPACK views with ld = m+1,
    unit_test/test_internal_blas.cc:300-312)"""
    run_gemm(te, np.float64, "N", "T", 80, 64, 16, -1.0, 1.0, inline=inline, ld_pad=ld_pad)
    run_gemm(te, np.float64, "T", "N", 200, 131, 77, 0.5, 0.0, inline=inline, ld_pad=ld_pad)


@pytest.mark.parametrize("flags", [1, 2])
@pytest.mark.parametrize("beta", [0.0, 1.0])
def test_dsyrk_diagonal_tile_mask(te, flags, beta):
    run_gemm(te, np.float64, "N", "T", 300, 300, 64, -1.0, beta, flags=flags)


def test_beta_zero_ignores_nan_in_c(te):
    import torch
    from slate_d35_b200 import TileDesc, check
    m = n = k = 64
    A, B = ol.matgen(1, m, k), ol.matgen(2, k, n)
    dA, dB = dev(A), dev(B)
    dC = torch.full((m * n,), float("nan"), dtype=torch.float64, device="cuda")
    d = (TileDesc * 1)()
    d[0].a, d[0].b, d[0].c = dA.data_ptr(), dB.data_ptr(), dC.data_ptr()
    d[0].m, d[0].n, d[0].k, d[0].lda, d[0].ldb, d[0].ldc = m, n, k, m, k, m
    check(te.te_gemm_inline(0, stream(), 0, ord("N"), ord("N"), scalar(1.0, np.float64),
                            scalar(0.0, np.float64), d, 1))
    got = host(dC, (m, n), np.float64)
    assert np.abs(got - A @ B).max() <= 3 * np.sqrt(k) * EPS * np.abs(A @ B).max()


@pytest.mark.parametrize("uplo", "LU")
@pytest.mark.parametrize("n", [1, 37, 128, 200, 512])
def test_dpotrf_tile_vs_oracle(te, uplo, n):
    import torch
    from slate_d35_b200 import check
    lda = n + 3
    full = ol.matgen(5, lda, n)
    A = full[:n, :]
    A[np.arange(n), np.arange(n)] += n
    want = full.copy(order="F")
    assert ol.oracle().oracle_tile_dpotrf(ord(uplo), n, ol.P(want), lda) == 0
    dA = dev(full)
    info = torch.full((1,), 7, dtype=torch.int32, device="cuda")
    sb = te.te_potrf_scratch_bytes(0)
    scratch = torch.empty(sb, dtype=torch.uint8, device="cuda")
    check(te.te_potrf_tile(0, stream(), 0, ord(uplo), n, dA.data_ptr(), lda, info.data_ptr(),
                           scratch.data_ptr(), sb))
    got = host(dA, (lda, n), np.float64)
    assert int(info.item()) == 0
    # the other triangle and the padding rows are untouched; the factor matches
    assert np.abs(got - want).max() / np.abs(want).max() <= 50 * EPS
    tri = (lambda x: np.triu(x[:n], 1)) if uplo == "L" else (lambda x: np.tril(x[:n], -1))
    assert np.array_equal(tri(got), tri(full))
    assert np.array_equal(got[n:], full[n:])


def test_dpotrf_tile_info(te):
    import torch
    from slate_d35_b200 import check
    n = 300
    A = ol.matgen(5, n, n)
    A[np.arange(n), np.arange(n)] += n
    A[170, 170] = -1.0
    dA = dev(A)
    info = torch.zeros(1, dtype=torch.int32, device="cuda")
    sb = te.te_potrf_scratch_bytes(0)
    scratch = torch.empty(sb, dtype=torch.uint8, device="cuda")
    check(te.te_potrf_tile(0, stream(), 0, ord("L"), n, dA.data_ptr(), n, info.data_ptr(),
                           scratch.data_ptr(), sb))
    want = A.copy(order="F")
    assert int(info.item()) == ol.oracle().oracle_tile_dpotrf(ord("L"), n, ol.P(want), n) == 171


@pytest.mark.parametrize("side,uplo,op,diag", [
    ("R", "L", "C", "N"),      # the potrf panel solve (src/potrf.cc:113-119)
    ("R", "L", "N", "N"), ("R", "U", "N", "U"), ("R", "U", "T", "N"),
    ("L", "L", "N", "N"), ("L", "L", "T", "U"), ("L", "U", "N", "N"), ("L", "U", "C", "N")])
@pytest.mark.parametrize("na", [100, 512])
def test_dtrsm_batched_vs_oracle(te, side, uplo, op, diag, na):
    import torch
    from slate_d35_b200 import TrsmDesc, check
    count = 3
    lda = na + 1
    Afull = ol.matgen(16, lda, na, dist=2)
    Afull[np.arange(na), np.arange(na)] += na ** 0.5 * 2
    dA = dev(Afull)
    descs = (TrsmDesc * count)()
    keep, cases = [], []
    sizes = [na, 70, 300]
    for t in range(count):
        m, n = (sizes[t], na) if side == "R" else (na, sizes[t])
        B = ol.matgen(40 + t, m, n, dist=2)
        dB = dev(B)
        keep.append(dB)
        descs[t].b, descs[t].m, descs[t].n, descs[t].ldb = dB.data_ptr(), m, n, m
        cases.append((B, dB, m, n))
    dd = torch.frombuffer(bytearray(bytes(descs)), dtype=torch.uint8).cuda()
    sb = te.te_trsm_scratch_bytes(0, na, count)
    scratch = torch.empty(sb, dtype=torch.uint8, device="cuda")
    alpha = 0.75
    check(te.te_trsm_batched(0, stream(), 0, ord(side), ord(uplo), ord(op), ord(diag),
                             scalar(alpha, np.float64), dA.data_ptr(), lda, na, dd.data_ptr(), count,
                             max(c[2] for c in cases), max(c[3] for c in cases),
                             scratch.data_ptr(), sb))
    torch.cuda.synchronize()
    for B, dB, m, n in cases:
        want = B.copy(order="F")
        ol.oracle().oracle_tile_dtrsm(ord(side), ord(uplo), ord(op), ord(diag), m, n, alpha,
                                      ol.P(Afull), lda, ol.P(want), m)
        got = host(dB, (m, n), np.float64)
        err = np.abs(got - want).max() / np.abs(want).max()
        assert err <= 50 * EPS * np.sqrt(na), (err, side, uplo, op, diag)


def test_matgen_device_bit_exact(te):
    import torch
    from slate_d35_b200 import check
    for dtype in (np.float64, np.complex128):
        m, n, ld = 77, 45, 80
        buf = torch.zeros(ld * n * (2 if dtype == np.complex128 else 1), dtype=torch.float64,
                          device="cuda")
        check(te.te_matgen_tile(0, stream(), DT[dtype], 1, 12345, m, n, 98300, 65530,
                                buf.data_ptr(), ld, 0.0))
        got = host(buf, (ld, n), dtype)[:m]
        assert np.array_equal(got, ol.matgen(12345, m, n, dtype, 1, 98300, 65530))
    buf = torch.zeros(64 * 64, dtype=torch.float64, device="cuda")
    check(te.te_matgen_tile(0, stream(), 0, 1, 3, 64, 64, 128, 128, buf.data_ptr(), 64, 1000.0))
    want = ol.matgen(3, 64, 64, ioff=128, joff=128)
    want[np.arange(64), np.arange(64)] += 1000.0
    assert np.array_equal(host(buf, (64, 64), np.float64), want)


def test_pool_semantics(te):
    """unit_test/test_Memory.cc:58-96,147-211: reserve, pop, grow by one, push back, clear."""
    from slate_d35_b200 import check
    vp, sz = C.c_void_p, C.c_size_t
    pool = vp()
    check(te.te_pool_create(0, 8 * 100 * 100, C.byref(pool)))
    av, cap, bb = sz(), sz(), sz()
    check(te.te_pool_stats(pool, C.byref(av), C.byref(cap), C.byref(bb)))
    assert (av.value, cap.value) == (0, 0) and bb.value >= 80000 and bb.value % 256 == 0
    check(te.te_pool_reserve(pool, 4))
    check(te.te_pool_stats(pool, C.byref(av), C.byref(cap), None))
    assert (av.value, cap.value) == (4, 4)
    blocks = []
    for i in range(6):                                   # 2 more than reserved: grows one by one
        b = vp()
        check(te.te_pool_alloc(pool, C.byref(b)))
        assert b.value and b.value % 256 == 0 and b.value not in blocks
        blocks.append(b.value)
        check(te.te_pool_stats(pool, C.byref(av), C.byref(cap), None))
        assert cap.value == max(4, i + 1) and av.value == cap.value - (i + 1)
    assert blocks[1] - blocks[0] == bb.value              # slab blocks are contiguous, ascending
    for b in blocks:
        check(te.te_pool_free(pool, b))
    check(te.te_pool_stats(pool, C.byref(av), C.byref(cap), None))
    assert (av.value, cap.value) == (6, 6)
    check(te.te_pool_clear(pool))
    check(te.te_pool_stats(pool, C.byref(av), C.byref(cap), None))
    assert (av.value, cap.value) == (0, 0)
    check(te.te_pool_destroy(pool))


# ------------------------------ complex-FP64 (TE_C64) ------------------------------
@pytest.mark.parametrize("opa", "NTC")
@pytest.mark.parametrize("opb", "NTC")
@pytest.mark.parametrize("shape", [(80, 64, 16), (384, 384, 384), (130, 257, 37), (128, 384, 128), (1, 1, 1)])
def test_zgemm_tiles_vs_oracle(te, opa, opb, shape):
    m, n, k = shape
    run_gemm(te, np.complex128, opa, opb, m, n, k, 3.141592653589793 + 1.4142135623730951j,
             2.718281828459045 + 1.7320508075688772j)


@pytest.mark.parametrize("inline", [False, True])
def test_zgemm_padding_beta0_and_inline(te, inline):
    run_gemm(te, np.complex128, "N", "C", 200, 131, 77, -1.0, 0.0, inline=inline, ld_pad=3)
    run_gemm(te, np.complex128, "C", "N", 65, 129, 17, 0.5 - 0.25j, 1.0, inline=inline, ld_pad=1)


@pytest.mark.parametrize("flags", [1 | 4, 2 | 4])
def test_zherk_diagonal_tile_mask(te, flags):
    """lower / upper mask and the herk rule: imaginary part of the diagonal set to zero"""
    run_gemm(te, np.complex128, "N", "C", 300, 300, 64, -1.0, 1.0, flags=flags)


@pytest.mark.parametrize("uplo", "LU")
@pytest.mark.parametrize("n", [1, 37, 64, 200, 384])
def test_zpotrf_tile_vs_oracle(te, uplo, n):
    import torch
    from slate_d35_b200 import check
    lda = n + 1
    full = ol.matgen(5, lda, n, np.complex128)
    A = full[:n, :]
    A[np.arange(n), np.arange(n)] += n
    want = full.copy(order="F")
    assert ol.oracle().oracle_tile_zpotrf(ord(uplo), n, ol.P(want), lda) == 0
    dA = dev(full)
    info = torch.full((1,), 7, dtype=torch.int32, device="cuda")
    sb = te.te_potrf_scratch_bytes(1)
    scratch = torch.empty(sb, dtype=torch.uint8, device="cuda")
    check(te.te_potrf_tile(0, stream(), 1, ord(uplo), n, dA.data_ptr(), lda, info.data_ptr(),
                           scratch.data_ptr(), sb))
    got = host(dA, (lda, n), np.complex128)
    assert int(info.item()) == 0
    tri = np.tril if uplo == "L" else np.triu
    # LAPACK leaves the imaginary part of the input diagonal unreferenced; compare the factor
    assert np.abs(tri(got[:n]) - tri(want[:n])).max() / np.abs(tri(want[:n])).max() <= 50 * EPS
    other = (lambda x: np.triu(x[:n], 1)) if uplo == "L" else (lambda x: np.tril(x[:n], -1))
    assert np.array_equal(other(got), other(full))


@pytest.mark.parametrize("side,uplo,op,diag", [
    ("R", "L", "C", "N"), ("R", "U", "T", "N"), ("L", "L", "N", "U"), ("L", "U", "C", "N")])
def test_ztrsm_batched_vs_oracle(te, side, uplo, op, diag):
    import torch
    from slate_d35_b200 import TrsmDesc, check
    na, count = 384, 2
    lda = na + 1
    Afull = ol.matgen(16, lda, na, np.complex128, dist=2)
    Afull[np.arange(na), np.arange(na)] += na ** 0.5 * 2
    dA = dev(Afull)
    descs = (TrsmDesc * count)()
    keep, cases = [], []
    for t, sz in enumerate((na, 70)):
        m, n = (sz, na) if side == "R" else (na, sz)
        B = ol.matgen(40 + t, m, n, np.complex128, dist=2)
        dB = dev(B)
        keep.append(dB)
        descs[t].b, descs[t].m, descs[t].n, descs[t].ldb = dB.data_ptr(), m, n, m
        cases.append((B, dB, m, n))
    dd = torch.frombuffer(bytearray(bytes(descs)), dtype=torch.uint8).cuda()
    sb = te.te_trsm_scratch_bytes(1, na, count)
    scratch = torch.empty(sb, dtype=torch.uint8, device="cuda")
    alpha = 0.75 - 0.5j
    check(te.te_trsm_batched(0, stream(), 1, ord(side), ord(uplo), ord(op), ord(diag),
                             scalar(alpha, np.complex128), dA.data_ptr(), lda, na, dd.data_ptr(),
                             count, max(c[2] for c in cases), max(c[3] for c in cases),
                             scratch.data_ptr(), sb))
    torch.cuda.synchronize()
    for B, dB, m, n in cases:
        want = B.copy(order="F")
        ol.oracle().oracle_tile_ztrsm(ord(side), ord(uplo), ord(op), ord(diag), m, n, ol.Z(alpha),
                                      ol.P(Afull), lda, ol.P(want), m)
        got = host(dB, (m, n), np.complex128)
        err = np.abs(got - want).max() / np.abs(want).max()
        assert err <= 50 * EPS * np.sqrt(na), (err, side, uplo, op, diag)


# ------------------------------ transposes, symmetric fill, scaling ------------------------------
@pytest.mark.parametrize("dt", [np.float64, np.complex128])
@pytest.mark.parametrize("m,n,pad", [(512, 512, 0), (384, 130, 5), (33, 97, 1), (1, 7, 0)])
@pytest.mark.parametrize("conj", [0, 1])
def test_tile_transpose_out_of_place(te, dt, m, n, pad, conj):
    """src/cuda/device_transpose.cu, rectangular form: AT = A^T (A^H), bit exact, padding rows
    of both arrays untouched."""
    from slate_d35_b200 import check
    A0 = ol.matgen(31, m + pad, n, dt, dist=2)
    T0 = ol.matgen(32, n + pad, m, dt, dist=2)
    a, t = dev(A0), dev(T0)
    check(te.te_tile_transpose(0, stream(), DT[dt], conj, m, n, a.data_ptr(), m + pad, t.data_ptr(), n + pad))
    got = host(t, (n + pad, m), dt)
    want = A0[:m].conj().T if (conj and dt == np.complex128) else A0[:m].T
    assert np.array_equal(got[:n], want)
    assert np.array_equal(got[n:], T0[n:])
    assert np.array_equal(host(a, (m + pad, n), dt), A0)


@pytest.mark.parametrize("dt", [np.float64, np.complex128])
@pytest.mark.parametrize("n,pad", [(512, 0), (200, 3), (31, 0), (65, 2)])
def test_tile_transpose_in_place_and_batched(te, dt, n, pad):
    """square in-place form and the batched forms (device arrays of tile pointers)."""
    import torch
    from slate_d35_b200 import check
    A0 = ol.matgen(33, n + pad, n, dt, dist=2)
    a = dev(A0)
    check(te.te_tile_transpose(0, stream(), DT[dt], 1, n, n, a.data_ptr(), n + pad, a.data_ptr(), n + pad))
    got = host(a, (n + pad, n), dt)
    assert np.array_equal(got[:n], A0[:n].conj().T) and np.array_equal(got[n:], A0[n:])
    # a mismatched in-place call is refused
    assert te.te_tile_transpose(0, stream(), DT[dt], 0, n, n - 1, a.data_ptr(), n + pad, a.data_ptr(), n + pad) != 0

    batch = 5
    tiles = [ol.matgen(40 + b, n + pad, n, dt, dist=2) for b in range(batch)]
    d_in = [dev(x) for x in tiles]
    d_out = [dev(np.zeros((n, n), dtype=dt, order="F")) for _ in range(batch)]
    p_in = torch.tensor([x.data_ptr() for x in d_in], dtype=torch.int64, device="cuda")
    p_out = torch.tensor([x.data_ptr() for x in d_out], dtype=torch.int64, device="cuda")
    check(te.te_tile_transpose_batch(0, stream(), DT[dt], 0, n, n, p_in.data_ptr(), n + pad, p_out.data_ptr(), n, batch))
    for b in range(batch):
        assert np.array_equal(host(d_out[b], (n, n), dt), tiles[b][:n].T)
    check(te.te_tile_transpose_batch(0, stream(), DT[dt], 0, n, n, p_in.data_ptr(), n + pad, p_in.data_ptr(), n + pad, batch))
    for b in range(batch):
        got = host(d_in[b], (n + pad, n), dt)
        assert np.array_equal(got[:n], tiles[b][:n].T) and np.array_equal(got[n:], tiles[b][n:])


@pytest.mark.parametrize("dt", [np.float64, np.complex128])
@pytest.mark.parametrize("uplo", "LU")
@pytest.mark.parametrize("herm", [0, 1])
@pytest.mark.parametrize("n,pad", [(512, 0), (100, 4), (33, 0)])
def test_tile_symmetrize(te, dt, uplo, herm, n, pad):
    from slate_d35_b200 import check
    A0 = ol.matgen(34, n + pad, n, dt, dist=2)
    a = dev(A0)
    check(te.te_tile_symmetrize(0, stream(), DT[dt], ord(uplo), herm, n, a.data_ptr(), n + pad))
    got = host(a, (n + pad, n), dt)
    S = A0[:n]
    strict = np.tril(S, -1) if uplo == "L" else np.triu(S, 1)
    d = np.diag(np.diag(S))
    if herm and dt == np.complex128:
        want = strict + strict.conj().T + d.real
    else:
        want = strict + strict.T + d
    assert np.array_equal(got[:n], want) and np.array_equal(got[n:], A0[n:])


@pytest.mark.parametrize("dt", [np.float64, np.complex128])
@pytest.mark.parametrize("beta", [0.0, 1.0, -0.75])
def test_tile_scale(te, dt, beta):
    """device_gescale / geset: A = beta A; beta = 0 stores zeros over NaN."""
    from slate_d35_b200 import check
    m, n, pad = 130, 70, 3
    A0 = ol.matgen(35, m + pad, n, dt, dist=2)
    if beta == 0.0:
        A0[:m][::7] = np.nan
    b = beta * (1 - 0.5j) if dt == np.complex128 else beta
    a = dev(A0)
    check(te.te_tile_scale(0, stream(), DT[dt], m, n, scalar(b, dt), a.data_ptr(), m + pad))
    got = host(a, (m + pad, n), dt)
    want = np.zeros_like(A0[:m]) if beta == 0.0 else b * A0[:m]
    assert np.allclose(got[:m], want, rtol=4 * EPS, atol=0)
    assert np.array_equal(got[m:], A0[m:])


# ------------------------------ tile LU without pivoting ------------------------------
@pytest.mark.parametrize("dt", [np.float64, np.complex128])
@pytest.mark.parametrize("m,n,pad", [(512, 512, 0), (300, 300, 5), (130, 130, 0), (64, 64, 1), (40, 40, 0),
                                     (384, 256, 0), (256, 384, 3), (1, 1, 0)])
def test_getrf_nopiv_tile(te, dt, m, n, pad):
    """te_getrf_nopiv_tile (replaces the host tile routine of internal::getrf_nopiv):
    L U = A on a diagonally dominant tile, square and rectangular, padding untouched."""
    import torch
    from slate_d35_b200 import check
    A0 = ol.matgen(51, m + pad, n, dt, dist=2)
    mn = min(m, n)
    A0[np.arange(mn), np.arange(mn)] += max(m, n)
    a = dev(A0)
    info = torch.full((1,), -7, dtype=torch.int32, device="cuda")
    sb = te.te_getrf_nopiv_scratch_bytes(DT[dt])
    scratch = torch.empty(sb, dtype=torch.uint8, device="cuda")
    check(te.te_getrf_nopiv_tile(0, stream(), DT[dt], m, n, a.data_ptr(), m + pad, info.data_ptr(),
                                 scratch.data_ptr(), sb))
    got = host(a, (m + pad, n), dt)
    assert int(info.item()) == 0
    Lf = np.tril(got[:m, :mn], -1) + np.eye(m, mn)
    Uf = np.triu(got[:mn, :])
    assert np.abs(Lf @ Uf - A0[:m]).max() <= 30 * EPS * max(m, n) * np.abs(A0[:m]).max()
    assert np.array_equal(got[m:], A0[m:])


def test_getrf_nopiv_tile_zero_pivot(te):
    """an exactly zero pivot is reported (first one), as LAPACK's unblocked LU does"""
    import torch
    from slate_d35_b200 import check
    n = 200
    A0 = np.asfortranarray(np.eye(n) * 3.0 + np.triu(np.ones((n, n)), 1))
    A0[141, 141] = 0.0
    a = dev(A0)
    info = torch.zeros(1, dtype=torch.int32, device="cuda")
    sb = te.te_getrf_nopiv_scratch_bytes(0)
    scratch = torch.empty(sb, dtype=torch.uint8, device="cuda")
    check(te.te_getrf_nopiv_tile(0, stream(), 0, n, n, a.data_ptr(), n, info.data_ptr(), scratch.data_ptr(), sb))
    torch.cuda.synchronize()
    assert int(info.item()) == 142
