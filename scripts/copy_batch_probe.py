"""Host cost of enqueueing tile copies one by one (te_memcpy2d_async): 2048 tiles of 512 x 512
doubles between a pinned host array (ScaLAPACK layout, pitch = local rows) and contiguous device
blocks, and the cost of a stream-wait-event in front of each.  Prints enqueue time (host) and
transfer rate (device events)."""
import ctypes as C, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from slate_d35_b200 import lib
te = lib(); torch.cuda.set_device(0)
vp, sz = C.c_void_p, C.c_size_t
te.te_host_alloc_pinned.argtypes = [sz, C.POINTER(vp)]
te.te_memcpy2d_async.argtypes = [vp, sz, vp, sz, sz, sz, C.c_int, vp]
te.te_memcpy_async.argtypes = [vp, vp, sz, C.c_int, vp]
te.te_event_create.argtypes = [C.c_int, C.c_int, C.POINTER(vp)]
te.te_event_record.argtypes = [vp, vp]
te.te_stream_wait_event.argtypes = [vp, vp]
rows, cols, nb = 32768, 16384, 512                 # 4 GiB local array = 2048 tiles
h = vp(); assert te.te_host_alloc_pinned(rows * cols * 8, C.byref(h)) == 0
d = torch.empty(rows * cols, dtype=torch.float64, device="cuda")
st, st2 = torch.cuda.Stream(), torch.cuda.Stream()
ev = vp(); assert te.te_event_create(0, 0, C.byref(ev)) == 0
te.te_event_record(ev, st2.cuda_stream)
ntile = (rows // nb) * (cols // nb)
hp = [h.value + (i * nb + j * nb * rows) * 8 for j in range(cols // nb) for i in range(rows // nb)]
dp = [d.data_ptr() + t * nb * nb * 8 for t in range(ntile)]
for name, kind in (("H2D", 1), ("D2H", 2)):
    for mode in ("2d", "2d+wait", "1d-contiguous", "2d"):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        t0 = time.perf_counter()
        for t in range(ntile):
            if mode == "2d+wait":
                te.te_stream_wait_event(st.cuda_stream, ev)
            if mode == "1d-contiguous":
                a, b = (dp[t], h.value + t * nb * nb * 8) if kind == 1 else (h.value + t * nb * nb * 8, dp[t])
                rc = te.te_memcpy_async(a, b, nb * nb * 8, kind, st.cuda_stream)
            elif kind == 1: rc = te.te_memcpy2d_async(dp[t], nb * 8, hp[t], rows * 8, nb * 8, nb, 1, st.cuda_stream)
            else: rc = te.te_memcpy2d_async(hp[t], rows * 8, dp[t], nb * 8, nb * 8, nb, 2, st.cuda_stream)
            assert rc == 0
        t1 = time.perf_counter()
        e1.record(st); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        print(f"{name} {mode:14s}: host enqueue {1e3 * (t1 - t0):7.2f} ms for {ntile} tiles ({1e6 * (t1 - t0) / ntile:5.1f} us each, incl. ctypes), "
              f"device {ms:7.1f} ms = {rows * cols * 8 / ms / 1e6:5.1f} GB/s", flush=True)
