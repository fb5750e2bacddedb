"""Copy-engine tile transport (te_peer_*) between ranks: bandwidth of per-tile 2 MiB puts and of
merged runs, next to the NCCL send/recv numbers of scripts/mp_commbench.py.  Run under torchrun
with an even number of ranks; rank r exchanges with rank r ^ 1 in both directions at once."""
import ctypes as C, os, sys
import torch, torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from slate_d35_b200 import lib, host as sl
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
te = lib()
vp, u64 = C.c_void_p, C.c_uint64
te.te_comm_init_rank.argtypes = [vp, C.c_size_t, C.c_int, C.c_int, C.c_int, C.POINTER(vp)]
te.te_peer_open.argtypes = [vp, C.c_size_t]
te.te_peer_max_message.argtypes = [vp]
te.te_peer_max_message.restype = C.c_size_t
te.te_peer_post.argtypes = [vp, C.c_int, C.c_size_t, C.POINTER(u64), C.POINTER(vp)]
te.te_peer_fetch.argtypes = [vp, C.c_int, C.c_size_t, C.POINTER(u64), C.POINTER(u64)]
te.te_peer_put.argtypes = [vp, C.c_int, u64, vp, C.c_size_t, vp]
te.te_peer_signal.argtypes = [vp, C.c_int, u64, vp]
te.te_peer_wait.argtypes = [vp, C.c_int, u64, vp]
te.te_peer_consumed.argtypes = [vp, C.c_int, u64, vp]
te.te_peer_stats.argtypes = [vp, C.POINTER(u64), C.POINTER(u64), C.POINTER(u64), C.POINTER(u64), C.POINTER(C.c_int)]
te.te_last_error.restype = C.c_char_p


def ck(st):
    assert st == 0, te.te_last_error().decode()


obj = [sl.nccl_unique_id() if rank == 0 else None]
dist.broadcast_object_list(obj, src=0)
comm = vp()
ck(te.te_comm_init_rank(C.create_string_buffer(obj[0], 128), 128, world, rank, lr, C.byref(comm)))
ck(te.te_peer_open(comm, 1024 << 20))
maxmsg = te.te_peer_max_message(comm)
st = torch.cuda.Stream()
peer = rank ^ 1
tile = 2 << 20
for ntile, merged in ((1, False), (16, False), (16, True), (64, False), (64, True), (128, False), (128, True)):
    nbytes = ntile * tile
    assert nbytes <= maxmsg, (nbytes, maxmsg)
    src = torch.full((nbytes // 8,), float(rank + 1), dtype=torch.float64, device="cuda")
    dst = torch.zeros(nbytes // 8, dtype=torch.float64, device="cuda")
    for rep in range(5):
        src.fill_(float(rank + 1 + 10 * rep))
        torch.cuda.synchronize(); dist.barrier()
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        seq, zone, rseq, off = u64(), vp(), u64(), u64()
        ck(te.te_peer_post(comm, peer, nbytes, C.byref(seq), C.byref(zone)))
        ck(te.te_peer_fetch(comm, peer, nbytes, C.byref(rseq), C.byref(off)))
        s = st.cuda_stream
        e0.record(st)
        if merged:
            ck(te.te_peer_put(comm, peer, off.value, src.data_ptr(), nbytes, s))
        else:
            for t in range(ntile):
                ck(te.te_peer_put(comm, peer, off.value + t * tile, src.data_ptr() + t * tile, tile, s))
        ck(te.te_peer_signal(comm, peer, rseq.value, s))
        e1.record(st)
        ck(te.te_peer_wait(comm, peer, seq.value, s))
        for t in range(ntile):
            ck(te.te_memcpy_async(vp(dst.data_ptr() + t * tile), vp(zone.value + t * tile), C.c_size_t(tile), 3, vp(s)))
        ck(te.te_peer_consumed(comm, peer, seq.value, s))
        e2.record(st)
        torch.cuda.synchronize()
        ms_put, ms_all = e0.elapsed_time(e1), e0.elapsed_time(e2)
        want = float(peer + 1 + 10 * rep)
        ok = bool((dst == want).all().item())
        assert ok, f"rank {rank}: wrong data after peer transfer (want {want}, got {dst[:4].tolist()} .. {dst[-4:].tolist()})"
    if rank == 0:
        print(f"peer exchange {ntile:4d} x 2 MiB {'one copy ' if merged else 'per tile '}: push {ms_put:.3f} ms "
              f"-> {nbytes / ms_put / 1e6:7.1f} GB/s per direction; with arrival + copy-out {ms_all:.3f} ms", flush=True)
p_, b_, s_, w_, m_ = u64(), u64(), u64(), u64(), C.c_int()
ck(te.te_peer_stats(comm, C.byref(p_), C.byref(b_), C.byref(s_), C.byref(w_), C.byref(m_)))
if rank == 0:
    print(f"stats: {p_.value} puts, {b_.value >> 20} MiB, {s_.value} signals, {w_.value} waits, memops mask {m_.value} "
          f"(1 = wait, 2 = signal by stream memory operation)", flush=True)
dist.barrier()
te.te_comm_destroy.argtypes = [vp]
te.te_comm_destroy(comm)
dist.destroy_process_group()
