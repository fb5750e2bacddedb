"""NCCL send/recv bandwidth between ranks through te_comm_* (the same calls listBcast makes)."""
import ctypes as C, os, sys, time
import torch, torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from slate_d35_b200 import lib, host as sl
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
te = lib()
vp = C.c_void_p
te.te_comm_init_rank.argtypes = [vp, C.c_size_t, C.c_int, C.c_int, C.c_int, C.POINTER(vp)]
te.te_comm_send.argtypes = [vp, vp, C.c_size_t, C.c_int, vp]
te.te_comm_recv.argtypes = [vp, vp, C.c_size_t, C.c_int, vp]
obj = [sl.nccl_unique_id() if rank == 0 else None]
dist.broadcast_object_list(obj, src=0)
comm = vp()
assert te.te_comm_init_rank(C.create_string_buffer(obj[0], 128), 128, world, rank, lr, C.byref(comm)) == 0
st = torch.cuda.Stream()
peer = rank ^ 1
for ntile, tile_mb in ((64, 2), (1, 128), (256, 2)):
    nbytes = tile_mb << 20
    sbuf = torch.empty(ntile * nbytes, dtype=torch.uint8, device="cuda")
    rbuf = torch.empty(ntile * nbytes, dtype=torch.uint8, device="cuda")
    for rep in range(4):
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(st):
            e0.record(st)
            te.te_comm_group_start()
            for t in range(ntile):
                te.te_comm_send(comm, sbuf.data_ptr() + t * nbytes, nbytes, peer, st.cuda_stream)
                te.te_comm_recv(comm, rbuf.data_ptr() + t * nbytes, nbytes, peer, st.cuda_stream)
            te.te_comm_group_end()
            e1.record(st)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
    if rank == 0:
        print(f"bidirectional exchange {ntile} x {tile_mb} MiB: {ms:.3f} ms -> {ntile*nbytes/ms/1e6:.1f} GB/s per direction", flush=True)
dist.barrier(); dist.destroy_process_group()
