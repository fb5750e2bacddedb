"""te_trsm_batched_static: the batched triangular solve for callers that cannot allocate.

Some users of the engine have no device scratch to pass and no way to get any when the
solve is due: their tile pool already owns all the HBM they are allowed, or they sit on a
panel's critical path where a cudaMalloc (which may stall the whole device) is not
acceptable.  For them the library carries a fixed scratch block of its own.  This test
calls that entry point through the C ABI twice in a row -- different shapes, so the second
call runs over what the first one left in the block -- and checks each result against
te_trsm_batched (caller-provided scratch) on the same inputs, to the tolerance the other
trsm tests use.  It also checks that sizes beyond the documented limits are refused with
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


def _dev(a):
    import torch
    flat = np.ascontiguousarray(a.T).reshape(-1)
    return torch.from_numpy(flat.copy()).cuda()


def _host(t, shape):
    m, n = shape
    return np.asfortranarray(t.cpu().numpy().reshape(n, m).T)


def _solve_both(te, side, uplo, op, diag, na, sizes, seed):
    """Runs the static and the caller-scratch entry points on identical inputs."""
    import torch
    from slate_d35_b200 import TrsmDesc, check
    stream = torch.cuda.current_stream().cuda_stream
    lda = na + 2
    A = ol.matgen(seed, lda, na, dist=2)
    A[np.arange(na), np.arange(na)] += 2 * np.sqrt(na)
    dA = _dev(A)
    alpha = (C.c_double * 1)(0.75)
    results = []
    for static in (True, False):
        count = len(sizes)
        descs = (TrsmDesc * count)()
        bufs = []
        for t, s in enumerate(sizes):
            m, n = (s, na) if side == "R" else (na, s)
            dB = _dev(ol.matgen(seed + 10 + t, m, n, dist=2))
            bufs.append((dB, m, n))
            descs[t].b, descs[t].m, descs[t].n, descs[t].ldb = dB.data_ptr(), m, n, m
        dd = torch.frombuffer(bytearray(bytes(descs)), dtype=torch.uint8).cuda()
        max_m, max_n = max(b[1] for b in bufs), max(b[2] for b in bufs)
        if static:
            check(te.te_trsm_batched_static(0, stream, 0, ord(side), ord(uplo), ord(op), ord(diag),
                                            alpha, dA.data_ptr(), lda, na, dd.data_ptr(), count,
                                            max_m, max_n), "te_trsm_batched_static")
        else:
            sb = te.te_trsm_scratch_bytes(0, na, count)
            scratch = torch.empty(sb, dtype=torch.uint8, device="cuda")
            check(te.te_trsm_batched(0, stream, 0, ord(side), ord(uplo), ord(op), ord(diag), alpha,
                                     dA.data_ptr(), lda, na, dd.data_ptr(), count, max_m, max_n,
                                     scratch.data_ptr(), sb), "te_trsm_batched")
        torch.cuda.synchronize()
        results.append([_host(dB, (m, n)) for dB, m, n in bufs])
    return results


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
        got, want = _solve_both(te, side, uplo, op, diag, na, sizes, seed)
        for g, w in zip(got, want):
            assert np.isfinite(g).all()
            err = np.abs(g - w).max() / np.abs(w).max()
            assert err <= 50 * EPS * np.sqrt(na), (err, side, uplo, op, na)


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
