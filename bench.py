#!/usr/bin/env python
"""bench.py -- dgemm FP64 TFLOP/s of slate::gemm (gemmC SUMMA) on the B200 tile engine.

Workload (BASELINE.json configs[1]): slate::gemm<double>, m = n = 65536, nb = 512,
Target::Devices, device-origin tiles filled by the matgen generator (seeds 1, 2, 3), on a
p x q grid of 1x1 / 1x2 / 2x2 / 2x4 for 1 / 2 / 4 / 8 GPUs (one rank per GPU, NCCL).

A "step" is one pass of the hot path: one SUMMA block step, i.e. one rank-512 update of
the whole 65536 x 65536 C (internal::gemm<Devices> = one batched DMMA launch per GPU,
behind the panel broadcast).  `--steps K` runs slate::gemm with inner dimension k = 512*K,
so K = 128 (the default) is the complete named configuration; `--warmup W` runs W such
block steps first in a separate call.  Timing is on the device (CUDA events spanning all
streams), bracketed by barriers, MAX over ranks.

Extra objects on the JSON line:
  roofline      dominant kernel (dgemm_tiles_kernel) against the DMMA peak measured in the
                same process by te_peak_fp64 (MEASURED_PEAKS.json has no FP64 entry)
  cpu_baseline  the reference's own Target::HostTask path (oracle/_ref) on the host cores,
                bounded sample, rank 0 at N = 1 only
  e2e           the same metric through host buffers (Matrix::fromLAPACK / fromScaLAPACK on
                pinned host arrays): H2D of A, B, C and D2H of C inside the timed region
  potrf         (unless --no-potrf) dpotrf n = 98304 nb = 512 lookahead 1, info and
                backward error; secondary metric of BASELINE.json

--impl reference times the reference CPU implementation (oracle/_ref, else the oracle
port) on the host cores; under torchrun only rank 0 works.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
os.environ.setdefault("OPENBLAS_NUM_THREADS", "1")

N_GEMM = 65536
NB = 512
N_POTRF = 98304
ALPHA, BETA = 3.141592653589793, 2.718281828459045      # test/test.cc:392-393
GRIDS = {1: (1, 1), 2: (1, 2), 4: (2, 2), 8: (2, 4)}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=128)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--n", "--gemm-n", dest="n", type=int, default=N_GEMM, help="override m = n (debug only)")
    ap.add_argument("--no-potrf", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--potrf-n", type=int, default=N_POTRF)
    ap.add_argument("--lookahead", type=int, default=1)
    ap.add_argument("--zgemm", action="store_true", help="(default now; kept for old command lines)")
    ap.add_argument("--no-zgemm", action="store_true",
                    help="skip config 5: slate::gemm complex<double> n=32768 nb=384 (ragged tiles)")
    ap.add_argument("--zgemm-n", type=int, default=32768)
    ap.add_argument("--herk", action="store_true", help="(default now; kept for old command lines)")
    ap.add_argument("--no-herk", action="store_true",
                    help="skip config 4: one internal::herk / syrk tile update, k = nb = 512 (1 GPU)")
    ap.add_argument("--no-parity", action="store_true",
                    help="skip the small multi-rank parity block against the CPU reference (N > 1)")
    ap.add_argument("--herk-n", type=int, default=32768)
    return ap.parse_args()


# ------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_indices):
        # rank 0 samples every GPU of the job: one throttled GPU slows all ranks
        self.rows, self.proc = [], None
        self.gpu = ",".join(str(g) for g in gpu_indices)

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.gpu}", f"--query-gpu={self.Q}",
                 "--format=csv,noheader,nounits", "-lms", "200"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()
        sm, mx, reasons, per_gpu = [], 0.0, set(), {}
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                per_gpu.setdefault(int(r[0]), []).append(float(r[1]))
                mx = max(mx, float(r[2]))
                for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown",
                                      "sw_power_cap"), r[5:9]):
                    if val.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                pass
        sm.sort()
        # under load = the upper half of the samples (idle gaps between phases are excluded)
        load = sm[len(sm) // 2:] if sm else []
        med = load[len(load) // 2] if load else None
        out = {"sm_mhz": med, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
               "samples": len(sm)}
        if len(per_gpu) > 1:
            # median of the upper half of each GPU's samples
            out["sm_mhz_per_gpu"] = [sorted(v)[(len(v) // 2 + len(v)) // 2] for _, v in sorted(per_gpu.items())]
        return out


def host_cores():
    """cores this process may run on (torchrun exports OMP_NUM_THREADS=1: not a core count)"""
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def reference_lib():
    """(ctypes lib, kind): oracle/_ref when built, else the oracle port; its OpenMP team is
    set to every host core explicitly, whatever OMP_NUM_THREADS the launcher exported."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as ol
    if ol.have_ref():
        L, kind = ol.ref(), "reference"
        if hasattr(L, "ref_set_num_threads"):
            L.ref_set_num_threads(host_cores())
    else:
        L, kind = ol.oracle(), "port"
        if hasattr(L, "oracle_set_num_threads"):
            L.oracle_set_num_threads(host_cores())
    return ol, L, kind


def time_reference_gemm(n, k, nb, reps):
    """reference Target::HostTask slate::gemm (or the oracle port) on host cores -> TFLOP/s"""
    import numpy as np
    ol, L, kind = reference_lib()
    A, B, Cm = ol.matgen(1, n, k), ol.matgen(2, k, n), ol.matgen(3, n, n)
    best = None
    for _ in range(reps):
        t0 = time.perf_counter()
        if kind == "reference":
            L.ref_dgemm(ord("N"), ord("N"), n, n, k, nb, ALPHA, ol.P(A), n, ol.P(B), k, BETA,
                        ol.P(Cm), n, 1)
        else:
            L.oracle_dgemm(ord("N"), ord("N"), n, n, k, nb, ALPHA, ol.P(A), n, ol.P(B), k, BETA,
                           ol.P(Cm), n)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    cores = L.ref_num_threads() if kind == "reference" else L.oracle_num_threads()
    return 2.0 * n * n * k / best * 1e-12, best, cores, kind


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    os.environ["OMP_NUM_THREADS"] = str(host_cores())      # override torchrun's OMP_NUM_THREADS=1
    # each step = one block outer product (k = 512) of a bounded m = n sample
    n_s = 16384 if args.n == N_GEMM else min(args.n, 16384)
    k = NB * args.steps
    if args.warmup > 0:
        time_reference_gemm(n_s, NB * min(args.warmup, 2), NB, 1)
    tf, secs, cores, kind = time_reference_gemm(n_s, k, NB, 1)
    line = {
        "impl": "reference", "metric": "dgemm FP64 TFLOP/s (slate::gemm, gemmC SUMMA)",
        "value": tf, "unit": "TFLOP/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": secs * 1e3 / args.steps, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"slate::gemm double m=n={N_GEMM} nb={NB} k=nb*steps",
                   "reference_sample": f"Target::HostTask m=n={n_s} k={k} nb={NB}, 1 rank"},
        "cpu_baseline": {"value": tf, "unit": "TFLOP/s", "cores": cores, "kind": kind,
                         "sample": f"m=n={n_s}, k={k}, nb={NB}"},
        "e2e": {"value": tf, "unit": "TFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------
def multirank_parity(sl, world, rank, p, q, lookahead):
    """Small ragged problems on this job's p x q grid, compared on rank 0 with the reference's
    own Target::HostTask results (oracle/_ref; the oracle port if that is not built) on the same
    matgen inputs.  The CPU side is the CHECKER, outside every timed region.  Errors in units of
    eps = 2^-52; bounds: the reference's 3 sqrt(k) eps for a block update
    (unit_test/test_internal_blas.cc:446-447), 50 eps for factors, solutions and backward errors
    (north star), 25 eps for the tester's solve residual (test/test_posv.cc:302-343)."""
    import numpy as np
    import torch
    import torch.distributed as dist
    eps = 2.0 ** -52

    def assemble(M):
        a = M.to_numpy()
        if world > 1:
            t = torch.from_numpy(np.ascontiguousarray(a)).cuda()
            dist.all_reduce(t)
            a = t.cpu().numpy()
        return np.asfortranarray(a)

    ol = L = kind = None
    if rank == 0:
        ol, L, kind = reference_lib()
    out = {"checker": None, "grid": f"{p}x{q}", "cases": {}}
    worst = 0.0
    ok_all = True

    def record(name, err, bound):
        nonlocal worst, ok_all
        out["cases"][name] = {"err_eps": err / eps, "bound_eps": bound / eps}
        worst = max(worst, err / eps)
        ok_all = ok_all and (err <= bound)

    # ---- gemm: ragged in every dimension, tester alpha / beta ----
    m, n, k, nb = 2000, 1900, 1100, 256
    A = sl.Matrix(m, k, nb, p, q).generate(1)
    B = sl.Matrix(k, n, nb, p, q).generate(2)
    Cm = sl.Matrix(m, n, nb, p, q).generate(3)
    sl.gemm(ALPHA, A, B, BETA, Cm, lookahead=lookahead, method="C")
    got = assemble(Cm)
    for M in (A, B, Cm):
        M.free()
    if rank == 0:
        A0, B0, want = ol.matgen(1, m, k), ol.matgen(2, k, n), ol.matgen(3, m, n)
        if kind == "reference":
            L.ref_dgemm(ord("N"), ord("N"), m, n, k, nb, ALPHA, ol.P(A0), m, ol.P(B0), k, BETA, ol.P(want), m, 1)
        else:
            L.oracle_dgemm(ord("N"), ord("N"), m, n, k, nb, ALPHA, ol.P(A0), m, ol.P(B0), k, BETA, ol.P(want), m)
        record("gemm m=2000 n=1900 k=1100 nb=256", float(np.abs(got - want).max() / np.abs(want).max()),
               3 * k ** 0.5 * eps)

    # ---- potrf, both triangles: factor against the reference's, and the backward error ----
    n, nb = 2000, 256
    for uplo in "LU":
        H = sl.HermitianMatrix(uplo, n, nb, p, q).generate(7, dominant=True)
        info = sl.potrf(H, lookahead=lookahead)
        got = assemble(H)
        H.free()
        if rank == 0:
            A0 = ol.matgen(7, n, n)
            A0[np.arange(n), np.arange(n)] += n
            want = A0.copy(order="F")
            fn = L.ref_dpotrf if kind == "reference" else L.oracle_dpotrf
            args = (ord(uplo), n, nb, ol.P(want), n) + ((1,) if kind == "reference" else ())
            info_ref = fn(*args)
            tri = np.tril if uplo == "L" else np.triu
            F = tri(got)
            S = tri(A0) + tri(A0, -1 if uplo == "L" else 1).T          # the matrix potrf sees
            R = (F @ F.T if uplo == "L" else F.T @ F) - S
            record(f"potrf {uplo} n=2000 nb=256 factor", float(np.abs(F - tri(want)).max() / np.abs(tri(want)).max()),
                   50 * eps)
            record(f"potrf {uplo} n=2000 nb=256 backward", float(np.linalg.norm(R) / (n * np.linalg.norm(S))), 50 * eps)
            ok_all = ok_all and info == 0 and info_ref == 0

    # ---- posv: solution against the reference's, and the tester's residual ----
    n, nrhs, nb = 2000, 10, 256
    H = sl.HermitianMatrix("L", n, nb, p, q).generate(7, dominant=True)
    Bm = sl.Matrix(n, nrhs, nb, p, q).generate(8)
    info = sl.posv(H, Bm, lookahead=lookahead)
    X = assemble(Bm)
    H.free()
    Bm.free()
    if rank == 0:
        A0 = ol.matgen(7, n, n)
        A0[np.arange(n), np.arange(n)] += n
        S = np.tril(A0) + np.tril(A0, -1).T
        B0 = ol.matgen(8, n, nrhs)
        if kind == "reference" and hasattr(L, "ref_posv_any"):
            Af, Xr = A0.copy(order="F"), B0.copy(order="F")
            L.ref_posv_any(0, ord("L"), n, nrhs, nb, Af.ctypes.data, n, Xr.ctypes.data, n, 1)
        else:
            Xr = np.linalg.solve(S, B0)
        record("posv n=2000 nrhs=10 solution", float(np.abs(X - Xr).max() / np.abs(Xr).max()), 50 * eps)
        record("posv n=2000 nrhs=10 residual",
               float(np.linalg.norm(B0 - S @ X, 1) / (n * np.linalg.norm(S, 1) * np.linalg.norm(X, 1))), 25 * eps)
        ok_all = ok_all and info == 0
        out["checker"] = ("oracle/_ref: the unmodified reference, Target::HostTask" if kind == "reference"
                          else "oracle port (oracle/_ref not built)")
        out["max_err_eps"] = worst
        out["ok"] = bool(ok_all)
    return out if rank == 0 else None


# ------------------------------------------------------------------------------------------
def pinned_array(shape, te):
    """column-major float64 array in pinned host memory (te_host_alloc_pinned)"""
    import numpy as np
    nbytes = int(shape[0]) * int(shape[1]) * 8
    ptr = C.c_void_p()
    st = te.te_host_alloc_pinned(C.c_size_t(max(nbytes, 8)), C.byref(ptr))
    if st != 0:
        raise RuntimeError("te_host_alloc_pinned: " + te.te_last_error().decode())
    buf = (C.c_double * (nbytes // 8)).from_address(ptr.value)
    arr = np.frombuffer(buf, dtype=np.float64).reshape(shape[1], shape[0]).T
    return arr, ptr


def local_extent(n, nb, p, r):
    """rows (or cols) of the 2D block-cyclic local array of process coordinate r of p"""
    nt = (n + nb - 1) // nb
    total = 0
    for t in range(r, nt, p):
        total += min(nb, n - t * nb)
    return total


def bind_near_gpu(local_rank, world):
    """Multi-rank runs: keep this rank's threads (and therefore the first-touch placement of its
    pinned host buffers) on the CPU cores of the GPU's own NUMA node, so that the e2e arm's host
    tiles do not cross the socket interconnect on their way to PCIe.  Returns a description."""
    if world == 1:
        return None
    try:
        import torch
        pr = torch.cuda.get_device_properties(local_rank)
        bdf = f"{pr.pci_domain_id:04x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
        base = f"/sys/bus/pci/devices/{bdf}"
        cpulist = open(f"{base}/local_cpulist").read().strip()
        node = open(f"{base}/numa_node").read().strip()
        cpus = set()
        for part in cpulist.split(","):
            if "-" in part:
                a, b = part.split("-")
                cpus.update(range(int(a), int(b) + 1))
            elif part:
                cpus.add(int(part))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
        return {"pci": bdf, "numa_node": int(node), "cpus": len(cpus)}
    except Exception as e:          # not fatal: the run goes on unbound
        return {"error": str(e)[:80]}


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
        return

    import numpy as np
    import torch
    import torch.distributed as dist
    from slate_d35_b200 import host as sl, lib

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world}"
    torch.cuda.set_device(local_rank)
    binding = bind_near_gpu(local_rank, world)
    te = lib()
    te.te_host_alloc_pinned.argtypes = [C.c_size_t, C.POINTER(C.c_void_p)]
    te.te_host_free_pinned.argtypes = [C.c_void_p]
    te.te_launch_count.restype = C.c_longlong

    nccl_id = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        obj = [sl.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(obj, src=0)
        nccl_id = obj[0]
    sl.initialize(local_rank, rank, world, nccl_id)
    p, q = GRIDS[world]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    n, nb, K, W = args.n, NB, args.steps, args.warmup
    k = nb * K

    # ---- FP64 tensor (DMMA) peak of this GPU: the roofline denominator -------------------
    # (first call: clocks still ramping and the kernel image loading -- one run here measured
    #  29.7 instead of 37.0 TFLOP/s on a single cold call; so warm up, then best of three)
    peak = C.c_double()
    peak_tf = 0.0
    for rep in range(4):
        st = te.te_peak_fp64(local_rank, 0, 40000, C.byref(peak))
        assert st == 0, te.te_last_error()
        if rep > 0:
            peak_tf = max(peak_tf, peak.value)
    peaks_all = [peak_tf]
    if world > 1:
        # aggregate peak = sum over the GPUs of the job; peak_tf becomes their mean
        t = torch.zeros(world, dtype=torch.float64, device="cuda")
        t[rank] = peak_tf
        dist.all_reduce(t)
        peaks_all = [float(x) for x in t.tolist()]
        peak_tf = sum(peaks_all) / world

    # ---- device-resident run: `value` -----------------------------------------------------
    A = sl.Matrix(n, k, nb, p, q).generate(1)
    B = sl.Matrix(k, n, nb, p, q).generate(2)
    Cm = sl.Matrix(n, n, nb, p, q).generate(3)
    if W > 0:
        Aw = sl.Matrix(n, nb * W, nb, p, q).generate(4)
        Bw = sl.Matrix(nb * W, n, nb, p, q).generate(5)
        sl.gemm(ALPHA, Aw, Bw, BETA, Cm, lookahead=args.lookahead, method="C")
        Aw.free()
        Bw.free()
        Cm.generate(3)
    sampler = ClockSampler(list(range(world)) if world > 1 else [local_rank])
    if rank == 0:
        sampler.start()
    barrier()
    launches0 = te.te_launch_count()
    ms = sl.gemm(ALPHA, A, B, BETA, Cm, lookahead=args.lookahead, timed=True, method="C")
    launches = te.te_launch_count() - launches0
    barrier()
    ms = max_over_ranks(ms)
    clocks = sampler.stop() if rank == 0 else None
    flops = 2.0 * n * n * k
    value = flops / (ms * 1e-3) * 1e-12

    # the tester's own check on the result (test/test_gemm.cc:136-158, 196-212):
    # || C X - (alpha A (B X) + beta C0 X) ||_1 / || Y ||_1, X = nrhs columns of kind rand
    # (the check products run through gemmC, the route these logs were verified with on 1, 4
    # and 8 GPUs; slate::gemm would pick gemmA for a single block column by itself)
    nrhs = 10
    X = sl.Matrix(n, nrhs, nb, p, q).generate(6)
    Wm = sl.Matrix(k, nrhs, nb, p, q)
    Y = sl.Matrix(n, nrhs, nb, p, q)
    Z = sl.Matrix(n, nrhs, nb, p, q)
    sl.gemm(1.0, B, X, 0.0, Wm, method="C")               # W = B X
    sl.gemm(ALPHA, A, Wm, 0.0, Y, method="C")             # Y = alpha A W
    sl.gemm(1.0, Cm, X, 0.0, Z, method="C")               # Z = C X
    A.free()
    B.free()
    Wm.free()
    Cm.generate(3)                            # C0 again
    sl.gemm(BETA, Cm, X, 1.0, Y, method="C")              # Y += beta C0 X
    yh, zh = Y.to_numpy(), Z.to_numpy()       # n x nrhs, this rank's tiles (zeros elsewhere)
    if world > 1:
        t = torch.from_numpy(np.ascontiguousarray(np.stack([yh, zh]))).cuda()
        dist.all_reduce(t)
        yh, zh = t[0].cpu().numpy(), t[1].cpu().numpy()
    gemm_error = float(np.abs(zh - yh).sum(axis=0).max() / np.abs(yh).sum(axis=0).max())
    for M in (X, Y, Z, Cm):
        M.free()

    # ---- e2e: host buffers, copies inside the timed region -----------------------------
    e2e = None
    if not args.no_e2e:
        myrow, mycol = rank % p, rank // p
        la_rows, la_k_rows = local_extent(n, nb, p, myrow), local_extent(k, nb, p, myrow)
        la_cols, la_k_cols = local_extent(n, nb, q, mycol), local_extent(k, nb, q, mycol)
        Ah, pa = pinned_array((la_rows, la_k_cols), te)
        Bh, pb = pinned_array((la_k_rows, la_cols), te)
        Ch, pc = pinned_array((la_rows, la_cols), te)
        mk = (lambda m_, n_, arr: sl.Matrix.fromScaLAPACK(m_, n_, arr, nb, p, q))
        Ae, Be, Ce = mk(n, k, Ah), mk(k, n, Bh), mk(n, n, Ch)
        Ae.generate(1)
        Be.generate(2)
        Ce.generate(3)
        barrier()
        ms_e = sl.gemm(ALPHA, Ae, Be, BETA, Ce, lookahead=args.lookahead, timed=True, method="C")
        barrier()
        ms_e = max_over_ranks(ms_e)
        h2d = 8.0 * (n * k + k * n + n * n)
        d2h = 8.0 * n * n
        e2e = {"value": flops / (ms_e * 1e-3) * 1e-12, "unit": "TFLOP/s",
               "h2d_bytes_per_step": h2d / K, "d2h_bytes_per_step": d2h / K,
               "ms_per_step": ms_e / K, "host_buffers": "pinned, ScaLAPACK block-cyclic local arrays",
               "steps": K,
               "pcie_gbs_per_rank": (h2d + d2h) / world / (ms_e * 1e-3) * 1e-9,
               "cpu_binding_rank0": binding}
        for M in (Ae, Be, Ce):
            M.free()
        for ptr in (pa, pb, pc):
            te.te_host_free_pinned(ptr)

    # ---- dpotrf (secondary metric), N = 1 ------------------------------------------------
    potrf = None
    if not args.no_potrf:
        npo = args.potrf_n
        H = sl.HermitianMatrix("L", npo, nb, p, q).generate(7, dominant=True)
        Hw = sl.HermitianMatrix("L", 4 * nb, nb, p, q).generate(8, dominant=True)
        sl.potrf(Hw, lookahead=args.lookahead)
        Hw.free()
        l0 = te.te_launch_count()
        barrier()
        info, ms_p = sl.potrf(H, lookahead=args.lookahead, timed=True)
        barrier()
        ms_p = max_over_ranks(ms_p)
        lp = te.te_launch_count() - l0
        pf = (npo ** 3 / 3.0 + npo ** 2 / 2.0 + npo / 6.0)      # lapack::Gflop<double>::potrf
        potrf = {"metric": "dpotrf FP64 TFLOP/s (slate::potrf)", "n": npo, "nb": nb,
                 "lookahead": args.lookahead, "info": int(info), "ms": ms_p,
                 "value": pf / (ms_p * 1e-3) * 1e-12, "unit": "TFLOP/s",
                 "frac_of_fp64_tensor_peak": pf / (ms_p * 1e-3) * 1e-12 / (peak_tf * world),
                 "gpu_launches": int(lp),
                 # scheduling of the trailing update during potrf
                 "update_blocks_per_cta_cap": int(os.environ.get("TE_ITEMS_PER_CTA", 0)) or (4 if world > 1 else 16)}
        # north-star check at full size: ||A - L L^H||_F / (n ||A||_F) <= 50 eps, evaluated with
        # the engine's own update kernels on a regenerated A (n^3/3 flops, outside the timing)
        free_b, total_b = C.c_size_t(), C.c_size_t()
        te.te_device_mem_info(local_rank, C.byref(free_b), C.byref(total_b))
        need_b = 8.0 * npo * npo / 2 / world * 1.15 + (2 << 30)
        if free_b.value > need_b:
            r, a0 = sl.potrf_residual(H, 7, True)
            if world > 1:
                t = torch.tensor([r, a0], dtype=torch.float64, device="cuda")
                dist.all_reduce(t)
                r, a0 = float(t[0].item()), float(t[1].item())
            potrf["backward_error"] = (r ** 0.5) / (npo * a0 ** 0.5)
            potrf["backward_error_bound_50eps"] = 50 * 2.0 ** -52
        H.free()

    # ---- zgemm, BASELINE.json configs[4]: complex<double>, nb = 384, ragged last tiles -------
    zg = None
    if not args.no_zgemm:
        nz, nbz = args.zgemm_n, 384
        za, zb = ALPHA + 1.4142135623730951j, BETA + 1.7320508075688772j
        Az = sl.Matrix(nz, nz, nbz, p, q, dtype=np.complex128).generate(1)
        Bz = sl.Matrix(nz, nz, nbz, p, q, dtype=np.complex128).generate(2)
        Cz = sl.Matrix(nz, nz, nbz, p, q, dtype=np.complex128).generate(3)
        Aw = sl.Matrix(nz, 2 * nbz, nbz, p, q, dtype=np.complex128).generate(4)
        Bw = sl.Matrix(2 * nbz, nz, nbz, p, q, dtype=np.complex128).generate(5)
        sl.gemm(za, Aw, Bw, zb, Cz, lookahead=args.lookahead, method="C")
        Aw.free()
        Bw.free()
        Cz.generate(3)
        barrier()
        ms_z = sl.gemm(za, Az, Bz, zb, Cz, lookahead=args.lookahead, timed=True, method="C")
        barrier()
        ms_z = max_over_ranks(ms_z)
        # tester check with X of kind rand (complex): ||C X - (alpha A (B X) + beta C0 X)||_1 / ||Y||_1
        Xz = sl.Matrix(nz, 10, nbz, p, q, dtype=np.complex128).generate(6)
        Wz = sl.Matrix(nz, 10, nbz, p, q, dtype=np.complex128)
        Yz = sl.Matrix(nz, 10, nbz, p, q, dtype=np.complex128)
        Zz = sl.Matrix(nz, 10, nbz, p, q, dtype=np.complex128)
        sl.gemm(1.0, Bz, Xz, 0.0, Wz, method="C")
        sl.gemm(za, Az, Wz, 0.0, Yz, method="C")
        sl.gemm(1.0, Cz, Xz, 0.0, Zz, method="C")
        Cz.generate(3)
        sl.gemm(zb, Cz, Xz, 1.0, Yz, method="C")
        yh, zh = Yz.to_numpy(), Zz.to_numpy()
        if world > 1:
            t = torch.from_numpy(np.ascontiguousarray(np.stack([yh, zh]))).cuda()
            t = torch.view_as_real(t).contiguous()
            dist.all_reduce(t)
            t = torch.view_as_complex(t)
            yh, zh = t[0].cpu().numpy(), t[1].cpu().numpy()
        zerr = float(np.abs(zh - yh).sum(axis=0).max() / np.abs(yh).sum(axis=0).max())
        fz = 8.0 * nz ** 3
        zg = {"metric": "zgemm FP64 TFLOP/s (slate::gemm complex<double>)", "n": nz, "nb": nbz,
              "ms": ms_z, "value": fz / (ms_z * 1e-3) * 1e-12, "unit": "TFLOP/s",
              "frac_of_fp64_tensor_peak": fz / (ms_z * 1e-3) * 1e-12 / (peak_tf * world),
              "error": zerr, "error_bound": 3 * 2.0 ** -52 * 2 * 2 ** 0.5}
        for M in (Az, Bz, Cz, Xz, Wz, Yz, Zz):
            M.free()

    # ---- herk / syrk tile-update microbench, BASELINE.json configs[3] (1 GPU) ---------------
    hk = None
    if not args.no_herk and world == 1:
        nh = args.herk_n
        Ah = sl.Matrix(nh, nb, nb).generate(1, dist=2)
        hk = {"n": nh, "k": nb, "nb": nb, "flops": "k*n*(n+1) (blaspp flops.hh:128-132)"}
        for name, cls, fn in (("herk", sl.HermitianMatrix, lambda Cx: sl.internal_herk(-1.0, Ah, 1.0, Cx)),
                              ("syrk", sl.SymmetricMatrix, lambda Cx: sl.internal_syrk(-1.0, Ah, 1.0, Cx))):
            Ch = cls("L", nh, nb).generate(2)
            best = None
            for rep in range(5):
                sl.sync()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                # the engine's queue 0 is its own stream: bracket with full syncs
                t0 = time.perf_counter()
                fn(Ch)
                sl.sync()
                dt = time.perf_counter() - t0
                if rep >= 2:
                    best = dt if best is None else min(best, dt)
            hk[name] = {"ms": best * 1e3, "value": nb * nh * (nh + 1.0) / best * 1e-12, "unit": "TFLOP/s",
                        "frac_of_fp64_tensor_peak": nb * nh * (nh + 1.0) / best * 1e-12 / peak_tf,
                        "timing": "host wall clock around one call + sync (one launch of ~2080 tiles)"}
            Ch.free()
        Ah.free()

    # ---- parity on this job's grid against the CPU reference (checker, untimed) ------------
    parity = None
    if not args.no_parity:
        parity = multirank_parity(sl, world, rank, p, q, args.lookahead)

    # ---- CPU baseline: the reference's HostTask path on the host cores --------------------
    cpu = None
    if not args.no_cpu and rank == 0 and world == 1:
        n_s, k_s = (16384, 8192) if n >= 16384 else (n, min(k, 4 * nb))
        tf_c, secs, cores, kind = time_reference_gemm(n_s, k_s, nb, 1)
        cpu = {"value": tf_c, "unit": "TFLOP/s", "cores": cores, "kind": kind,
               "sample": f"Target::HostTask slate::gemm m=n={n_s}, k={k_s}, nb={nb}, "
                         f"{secs:.1f} s wall, OpenBLAS sequential under OpenMP tasks"}

    if rank == 0:
        launch_flops = flops / K / world      # one dgemm_tiles_kernel launch per step per GPU
        # dram__bytes of ONE launch, from ncu captures at this launch shape (tiles per launch on
        # one GPU); null when that shape has not been captured
        traffic = None
        tiles_per_launch = ((n + nb - 1) // nb) ** 2 // world
        tfile = os.path.join(ROOT, "profiles", "dgemm_traffic_per_launch.json")
        if os.path.exists(tfile):
            try:
                traffic = json.load(open(tfile)).get("dram_bytes_per_launch_by_tiles", {}).get(str(tiles_per_launch))
            except Exception:
                traffic = None
        line = {
            "metric": "dgemm FP64 TFLOP/s (slate::gemm, gemmC SUMMA)",
            "value": value, "unit": "TFLOP/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"slate::gemm double m=n={n} nb={nb} k=nb*steps={k} "
                                   f"Target::Devices origin=Devices lookahead={args.lookahead}",
                       "grid": f"{p}x{q}", "matgen": "rand seeds 1,2,3",
                       "l2": "inputs_exceed_l2 (C alone is %.0f GiB)" % (8.0 * n * n / 2 ** 30),
                       "step": "one SUMMA block step (rank-512 update of C); steps=128 is the full gemm"},
            "frac_of_fp64_tensor_peak": value / (peak_tf * world),
            "fp64_dmma_peak_tflops_per_gpu": peak_tf,
            "fp64_dmma_peak_each_gpu": [round(x, 2) for x in peaks_all],
            "gemm_error": gemm_error, "gemm_error_bound_3eps": 3 * 2.0 ** -52,
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"bound": "tensor", "achieved": launch_flops / (ms / K * 1e-3) * 1e-12,
                         "peak": peak_tf, "unit": "TFLOP/s",
                         "frac": launch_flops / (ms / K * 1e-3) * 1e-12 / peak_tf,
                         "traffic": traffic, "tiles_per_launch": tiles_per_launch,
                         "kernel": "dgemm_tiles_kernel",
                         "peak_source": "measured in this run by te_peak_fp64 (register-resident "
                                        "DMMA.8x8x4); MEASURED_PEAKS.json has no FP64 entry"},
            "e2e": e2e, "cpu_baseline": cpu, "potrf": potrf, "zgemm": zg, "herk_microbench": hk,
            "parity": parity,
        }
        print(json.dumps(line), flush=True)
    sl.finalize()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
