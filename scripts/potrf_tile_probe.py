import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, torch
import oracle_lib as ol
from slate_d35_b200 import lib, check
te = lib(); torch.cuda.set_device(0)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 128
A = ol.matgen(5, n, n); A[np.arange(n), np.arange(n)] += n
dA = torch.from_numpy(np.ascontiguousarray(A.T).reshape(-1).copy()).cuda()
info = torch.zeros(1, dtype=torch.int32, device="cuda")
sb = te.te_potrf_scratch_bytes(0); scratch = torch.empty(sb, dtype=torch.uint8, device="cuda")
st = torch.cuda.current_stream().cuda_stream
for rep in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    check(te.te_potrf_tile(0, st, 0, ord("L"), n, dA.data_ptr(), n, info.data_ptr(), scratch.data_ptr(), sb))
    e1.record(); torch.cuda.synchronize()
    print(f"potrf_tile n={n}: {e0.elapsed_time(e1)*1e3:.1f} us")
