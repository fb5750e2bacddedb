"""te_trsm_batched_static: the batched triangular solve for callers that cannot allocate.

Some users of the engine have no device scratch to pass and no way to get any when the
solve is due: their tile pool already owns all the HBM they are allowed, or they sit on a
panel's critical path where a cudaMalloc (which may stall the whole device) is not
acceptable.  For them the library carries a fixed scratch block of its own.  This test
calls that entry point through the C ABI twice in a row -- different shapes, so the second
call runs over what the first one left in the block -- and checks each result against
te_trsm_batched (caller-provided scratch) on the same inputs and against the CPU oracle,
to the tolerance the other trsm tests use.  A second test goes to the documented limits
(na = 512, count = 256, real and complex) and checks that the call leaves the device's
free memory unchanged.  It also checks that sizes beyond the documented limits are refused with
an explanatory message instead of overrunning the block.

Collected last on purpose (file name), so a run that stops at the first failure has
already been through every other test; it also passes on its own in a fresh process.
"""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as ol

pytestmark = pytest.mark.gpu
EPS = np.finfo(np.float64).eps


def _solve_both(te, side, uplo, op, diag, na, sizes, seed, dtype=np.float64, mem_probe=None):
    """Runs the static and the caller-scratch entry points on identical inputs and also returns
    what the CPU oracle (oracle_tile_{d,z}trsm) says the answer is.  mem_probe: a list that
    receives (free bytes before, free bytes after) around the static call."""
    import torch
    from slate_d35_b200 import TrsmDesc, check
    stream = torch.cuda.current_stream().cuda_stream
    cplx = np.dtype(dtype) == np.complex128
    code = 1 if cplx else 0
    lda = na + 2
    A = ol.matgen(seed, lda, na, dtype, dist=2)
    A[np.arange(na), np.arange(na)] += 2 * np.sqrt(na)
    dA = _dev(A)
    a = (0.75 - 0.5j) if cplx else 0.75
    alpha = (C.c_double * 2)(a.real, a.imag) if cplx else (C.c_double * 1)(a)
    inputs = []
    for t, s_ in enumerate(sizes):
        m, n = (s_, na) if side == "R" else (na, s_)
        inputs.append((ol.matgen(seed + 10 + t, m, n, dtype, dist=2), m, n))
    results = []
    for static in (True, False):
        count = len(sizes)
        descs = (TrsmDesc * count)()
        bufs = []
        for t, (B, m, n) in enumerate(inputs):
            dB = _dev(B)
            bufs.append((dB, m, n))
            descs[t].b, descs[t].m, descs[t].n, descs[t].ldb = dB.data_ptr(), m, n, m
        dd = torch.frombuffer(bytearray(bytes(descs)), dtype=torch.uint8).cuda()
        max_m, max_n = max(b[1] for b in bufs), max(b[2] for b in bufs)
        if static:
            torch.cuda.synchronize()
            free0, free1, total = C.c_size_t(), C.c_size_t(), C.c_size_t()
            te.te_device_mem_info(0, C.byref(free0), C.byref(total))
            check(te.te_trsm_batched_static(0, stream, code, ord(side), ord(uplo), ord(op), ord(diag),
                                            alpha, dA.data_ptr(), lda, na, dd.data_ptr(), count,
                                            max_m, max_n), "te_trsm_batched_static")
            torch.cuda.synchronize()
            te.te_device_mem_info(0, C.byref(free1), C.byref(total))
            if mem_probe is not None:
                mem_probe.append((free0.value, free1.value))
        else:
            sb = te.te_trsm_scratch_bytes(code, na, count)
            scratch = torch.empty(sb, dtype=torch.uint8, device="cuda")
            check(te.te_trsm_batched(0, stream, code, ord(side), ord(uplo), ord(op), ord(diag), alpha,
                                     dA.data_ptr(), lda, na, dd.data_ptr(), count, max_m, max_n,
                                     scratch.data_ptr(), sb), "te_trsm_batched")
        torch.cuda.synchronize()
        results.append([_host(dB, (m, n), dtype) for dB, m, n in bufs])
    oracle = []
    for B, m, n in inputs:
        want = B.copy(order="F")
        if cplx:
            ol.oracle().oracle_tile_ztrsm(ord(side), ord(uplo), ord(op), ord(diag), m, n, ol.Z(a),
                                          ol.P(A), lda, ol.P(want), m)
        else:
            ol.oracle().oracle_tile_dtrsm(ord(side), ord(uplo), ord(op), ord(diag), m, n, a,
                                          ol.P(A), lda, ol.P(want), m)
        oracle.append(want)
    return results[0], results[1], oracle


def _agree(got, want, oracle, na, what):
    for g, w, o in zip(got, want, oracle):
        assert np.isfinite(g).all(), what
        # same tolerance as the other trsm tests (tests/test_gpu_kernels.py)
        assert np.abs(g - w).max() / np.abs(w).max() <= 50 * EPS * np.sqrt(na), ("vs caller scratch", what)
        assert np.abs(g - o).max() / np.abs(o).max() <= 50 * EPS * np.sqrt(na), ("vs CPU oracle", what)


def test_static_scratch_trsm_matches_caller_scratch_twice_in_a_row():
    import torch
    from slate_d35_b200 import lib
    assert torch.cuda.is_available()
    torch.cuda.set_device(0)
    te = lib()
    # first call: the potrf panel solve shape (Right, Lower, ConjTrans), full-size tile
    # second call: smaller ragged tile, other side / triangle, so it reuses a dirty block
    for (side, uplo, op, diag, na, sizes, seed) in (
            ("R", "L", "C", "N", 512, [512, 70, 300], 16),
            ("L", "U", "N", "N", 200, [33, 257], 31)):
        got, want, oracle = _solve_both(te, side, uplo, op, diag, na, sizes, seed)
        _agree(got, want, oracle, na, (side, uplo, op, na))


def test_static_scratch_trsm_at_its_limits_real_and_complex():
    """na = 512 and count = 256 are the documented maxima; the complex case needs 753 920 of the
    block's 786 432 bytes, the closest any legal call comes to its end.  Each result is checked
    against the caller-scratch entry point and against the CPU oracle, and a warmed-up call
    must leave the device's free memory exactly as it found it (it allocates nothing)."""
    import torch
    from slate_d35_b200 import lib
    torch.cuda.set_device(0)
    te = lib()
    sizes = [8 + (t % 5) * 6 for t in range(256)]          # 256 small ragged right-hand sides
    for dtype in (np.complex128, np.float64):
        probe = []
        # the first call on a device also resolves the block's address: measure the second
        for rep in range(2):
            got, want, oracle = _solve_both(te, "R", "L", "C", "N", 512, sizes, 50 + rep, dtype, probe)
            _agree(got, want, oracle, 512, (str(np.dtype(dtype)), "count 256", rep))
        before, after = probe[-1]
        assert after == before, f"free device memory changed across the call: {before} -> {after}"


def _host(t, shape, dtype=np.float64):
    m, n = shape
    a = t.cpu().numpy()
    if np.dtype(dtype) == np.complex128:
        a = a.view(np.complex128)
    return np.asfortranarray(a.reshape(n, m).T)


def _dev(a):
    import torch
    flat = np.ascontiguousarray(a.T).reshape(-1)
    if flat.dtype == np.complex128:
        flat = flat.view(np.float64)
    return torch.from_numpy(flat.copy()).cuda()


def test_static_scratch_trsm_refuses_sizes_beyond_its_block():
    from slate_d35_b200 import lib
    te = lib()
    alpha = (C.c_double * 1)(1.0)
    dummy = C.c_void_p(256)          # never dereferenced: the size check comes first
    st = te.te_trsm_batched_static(0, None, 0, ord("R"), ord("L"), ord("C"), ord("N"), alpha, dummy,
                                   1024, 1024, dummy, 4, 1024, 1024)
    assert st == 6                    # TE_ERR_LIMIT
    msg = te.te_last_error().decode()
    assert "512" in msg and "te_trsm_batched" in msg
    st = te.te_trsm_batched_static(0, None, 0, ord("R"), ord("L"), ord("C"), ord("N"), alpha, dummy,
                                   512, 512, dummy, 257, 512, 512)
    assert st == 6
    assert "256" in te.te_last_error().decode()
