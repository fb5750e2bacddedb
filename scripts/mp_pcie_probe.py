"""All ranks stream between pinned host memory and their GPU at the same time: what the box's
PCIe / host-memory fabric gives each of N GPUs when the e2e arm's host buffers of all ranks move
at once.  H2D alone, D2H alone, both directions together; 2 GiB per direction per rank, flat
copies of 64 MiB.  Run under torchrun."""
import ctypes as C, os, sys
import torch, torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from slate_d35_b200 import lib
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
te = lib()
vp, sz = C.c_void_p, C.c_size_t
te.te_host_alloc_pinned.argtypes = [sz, C.POINTER(vp)]
te.te_memcpy_async.argtypes = [vp, vp, sz, C.c_int, vp]
nbytes, chunk = 2 << 30, 64 << 20
hin, hout = vp(), vp()
assert te.te_host_alloc_pinned(nbytes, C.byref(hin)) == 0 and te.te_host_alloc_pinned(nbytes, C.byref(hout)) == 0
din = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
dout = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()
for mode in ("H2D", "D2H", "both"):
    for rep in range(2):
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        s_in.wait_stream(torch.cuda.current_stream()); s_out.wait_stream(torch.cuda.current_stream())
        for off in range(0, nbytes, chunk):
            if mode in ("H2D", "both"):
                te.te_memcpy_async(din.data_ptr() + off, hin.value + off, chunk, 1, s_in.cuda_stream)
            if mode in ("D2H", "both"):
                te.te_memcpy_async(hout.value + off, dout.data_ptr() + off, chunk, 2, s_out.cuda_stream)
        torch.cuda.current_stream().wait_stream(s_in); torch.cuda.current_stream().wait_stream(s_out)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
    moved = nbytes * (2 if mode == "both" else 1)
    t = torch.tensor([moved / ms / 1e6], dtype=torch.float64, device="cuda")
    lo, hi, sm = t.clone(), t.clone(), t.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN); dist.all_reduce(hi, op=dist.ReduceOp.MAX); dist.all_reduce(sm)
    if rank == 0:
        print(f"{world} ranks, {mode:4s}: per rank {lo.item():.1f} .. {hi.item():.1f} GB/s, aggregate {sm.item():.1f} GB/s", flush=True)
dist.barrier(); dist.destroy_process_group()
