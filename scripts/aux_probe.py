"""Runs the auxiliary and panel kernels once or twice on one 512 x 512 tile each (real), for ncu
captures: transposes (in place / out of place), symmetric fill, tile scale / add, the no-pivot LU
block kernel, the Cholesky + inverse block kernel, the batched trsm (descriptor builder)."""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from slate_d35_b200 import lib, check, TrsmDesc
te = lib(); torch.cuda.set_device(0)
st = torch.cuda.current_stream().cuda_stream
n = 512
rng = np.random.default_rng(0)
A = torch.from_numpy(rng.random((n, n)) + n * np.eye(n)).cuda()
B = torch.from_numpy(rng.random((n, n))).cuda()
info = torch.zeros(1, dtype=torch.int32, device="cuda")
one = (C.c_double * 1)(0.5)
for rep in range(2):
    check(te.te_tile_transpose(0, st, 0, 0, n, n, A.data_ptr(), n, A.data_ptr(), n))          # in place
    check(te.te_tile_transpose(0, st, 0, 0, n, n, A.data_ptr(), n, B.data_ptr(), n))          # out of place
    check(te.te_tile_symmetrize(0, st, 0, ord("L"), 0, n, B.data_ptr(), n))
    check(te.te_tile_scale(0, st, 0, n, n, one, B.data_ptr(), n))
    check(te.te_tile_add(0, st, 0, n, n, 1.0, A.data_ptr(), n, 1.0, B.data_ptr(), n))
    S = (A + A.T).contiguous()
    sb = te.te_getrf_nopiv_scratch_bytes(0); scratch = torch.empty(sb, dtype=torch.uint8, device="cuda")
    check(te.te_getrf_nopiv_tile(0, st, 0, n, n, S.data_ptr(), n, info.data_ptr(), scratch.data_ptr(), sb))
    S = (A + A.T).contiguous()
    sb = te.te_potrf_scratch_bytes(0); scratch = torch.empty(sb, dtype=torch.uint8, device="cuda")
    check(te.te_potrf_tile(0, st, 0, ord("L"), n, S.data_ptr(), n, info.data_ptr(), scratch.data_ptr(), sb))
    count = 16
    Bs = [torch.from_numpy(rng.random((n, n))).cuda() for _ in range(count)]
    d = (TrsmDesc * count)()
    for t in range(count):
        d[t].b, d[t].m, d[t].n, d[t].ldb = Bs[t].data_ptr(), n, n, n
    dd = torch.frombuffer(bytearray(bytes(d)), dtype=torch.uint8).cuda()
    sb = te.te_trsm_scratch_bytes(0, n, count); scratch = torch.empty(sb, dtype=torch.uint8, device="cuda")
    check(te.te_trsm_batched(0, st, 0, ord("R"), ord("L"), ord("C"), ord("N"), one, S.data_ptr(), n, n,
                             dd.data_ptr(), count, n, n, scratch.data_ptr(), sb))
    torch.cuda.synchronize()
print("aux probe done")
