"""N > 1 on CPU: two gloo ranks execute the broadcast plans of slate::gemm and slate::potrf
(the product's own host C++: BaseMatrix::bcastPlan over the 2D block-cyclic map, fed by the
same list builders the drivers use) and must reproduce the oracle's result bit for bit.
No GPU, no NCCL: see tests/mp_worker.py."""
import os
import socket
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("grid", [(1, 2), (2, 1)])
def test_two_ranks_gloo(grid):
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", OMP_NUM_THREADS="2")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", str(free_port()),
           os.path.join(ROOT, "tests", "mp_worker.py"), str(grid[0]), str(grid[1])]
    out = subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=600)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-3000:]
    assert f"MP_CPU_OK {grid[0]}x{grid[1]}" in out.stdout


def test_plan_matches_block_cyclic_ownership():
    """single process, pretending to be each rank of a 2 x 4 grid: every SUMMA broadcast goes
    from the owner (i mod p) + (k mod q) p to exactly the other ranks of its process row."""
    from slate_d35_b200 import host as sl
    p, q, nb = 2, 4, 8
    m = n = k = 8 * 9
    sl.initialize(0, 0, p * q, None)
    A = sl.Matrix(m, k, nb, p, q, origin="N")
    B = sl.Matrix(k, n, nb, p, q, origin="N")
    Cm = sl.Matrix(m, n, nb, p, q, origin="N")
    for kk in (0, 3, 8):
        for which, i, j, root, dests in sl.plan_gemm_step(A, B, Cm, kk):
            if which == 0:
                assert (i, j) == (i, kk) and root == i % p + (kk % q) * p
                assert sorted(dests + [root]) == [i % p + c * p for c in range(q)]
            else:
                assert (i, j) == (kk, j) and root == kk % p + (j % q) * p
                assert sorted(dests + [root]) == [r + (j % q) * p for r in range(p)]
    H = sl.HermitianMatrix("L", n, nb, p, q, origin="N")
    ops = sl.plan_potrf_step(H, 2)
    diag = [o for o in ops if o[0] == 0]
    assert len(diag) == 1 and diag[0][1:3] == (2, 2)
    # A(i, 2) must reach every rank owning a tile of row i right of column 2 or of column i below
    for which, i, j, root, dests in ops:
        if which != 1:
            continue
        need = {sl.tile_rank(H, i, c) for c in range(3, i + 1)} | {sl.tile_rank(H, r, i) for r in range(i, 9)}
        assert set(dests) == need - {root}
    sl.finalize()
