"""PCIe probe: tile-wise 2D copies (512 x 512 doubles out of a strided host array) vs one flat copy."""
import ctypes as C, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from slate_d35_b200 import lib
te = lib(); torch.cuda.set_device(0)
vp, sz = C.c_void_p, C.c_size_t
te.te_host_alloc_pinned.argtypes = [sz, C.POINTER(vp)]
te.te_memcpy2d_async.argtypes = [vp, sz, vp, sz, sz, sz, C.c_int, vp]
te.te_memcpy_async.argtypes = [vp, vp, sz, C.c_int, vp]
n, nb = 16384, 512                    # 2 GiB
h = vp(); assert te.te_host_alloc_pinned(n * n * 8, C.byref(h)) == 0
d = torch.empty(n * n, dtype=torch.float64, device="cuda")
st = torch.cuda.Stream()
def run(kind, tiled):
    torch.cuda.synchronize(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    if tiled:
        t = 0
        for j in range(n // nb):
            for i in range(n // nb):
                hp = h.value + (i * nb + j * nb * n) * 8
                dp = d.data_ptr() + t * nb * nb * 8; t += 1
                if kind == 1: te.te_memcpy2d_async(dp, nb * 8, hp, n * 8, nb * 8, nb, 1, st.cuda_stream)
                else: te.te_memcpy2d_async(hp, n * 8, dp, nb * 8, nb * 8, nb, 2, st.cuda_stream)
    else:
        if kind == 1: te.te_memcpy_async(d.data_ptr(), h.value, n * n * 8, 1, st.cuda_stream)
        else: te.te_memcpy_async(h.value, d.data_ptr(), n * n * 8, 2, st.cuda_stream)
    e1.record(st); torch.cuda.synchronize()
    return n * n * 8 / e0.elapsed_time(e1) / 1e6
for kind, name in ((1, "H2D"), (2, "D2H")):
    for tiled in (False, True):
        run(kind, tiled)
        print(name, "tiled 2D (4 KB rows)" if tiled else "flat", f"{run(kind, tiled):.1f} GB/s", flush=True)
