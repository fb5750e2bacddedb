"""First GPU contact: box facts, FP64 peaks, dgemm tile kernel correctness + speed."""
import ctypes as C, os, sys, time, json, subprocess
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from slate_d35_b200 import lib, check, TileDesc
L = lib()
print("nproc", os.cpu_count()); print(subprocess.run("free -g | head -2; nvidia-smi -L", shell=True, capture_output=True, text=True).stdout)
dev = 0
torch.cuda.set_device(dev)
res = {}
for kind, name in ((0, "dmma"), (1, "dfma")):
    tf = C.c_double()
    check(L.te_peak_fp64(dev, kind, 20000, C.byref(tf)), "peak")
    res[name] = tf.value
    print(f"peak {name}: {tf.value:.2f} TFLOP/s", flush=True)

def run_case(m, n, k, opa, opb, lda_pad=0, alpha=1.5, beta=-0.5, flags=0, count=3, inline=False, off=0):
    g = torch.Generator(device="cuda"); g.manual_seed(m * 7 + n * 3 + k)
    def mk(r, c):
        ld = r + lda_pad
        buf = torch.rand(ld * c + off + 8, dtype=torch.float64, device="cuda", generator=g) - 0.5
        view = buf[off:off + ld * c].view(c, ld).t()[:r, :]     # col-major r x c, ld
        return buf, view, ld
    descs = (TileDesc * count)(); keep = []; refs = []
    for i in range(count):
        ar, ac = (m, k) if opa == 'N' else (k, m)
        br, bc = (k, n) if opb == 'N' else (n, k)
        ab, av, lda = mk(ar, ac); bb, bv, ldb = mk(br, bc); cb, cv, ldc = mk(m, n)
        keep += [ab, bb, cb]
        opA = av if opa == 'N' else av.t(); opB = bv if opb == 'N' else bv.t()
        ref = alpha * (opA @ opB) + (beta * cv if beta != 0 else 0)
        full = cv.clone()
        if flags & 1: ref = torch.where(torch.ones_like(ref).tril().bool(), ref, full)
        if flags & 2: ref = torch.where(torch.ones_like(ref).triu().bool(), ref, full)
        refs.append((ref, cv, cb.clone(), cb))
        d = descs[i]
        d.a = ab.data_ptr() + off * 8; d.b = bb.data_ptr() + off * 8; d.c = cb.data_ptr() + off * 8
        d.m, d.n, d.k, d.lda, d.ldb, d.ldc, d.flags = m, n, k, lda, ldb, ldc, flags
    al = C.c_double(alpha); be = C.c_double(beta)
    st = torch.cuda.current_stream().cuda_stream
    if inline:
        check(L.te_gemm_inline(dev, st, 0, ord(opa), ord(opb), C.byref(al), C.byref(be), descs, count), "inline")
    else:
        dd = torch.frombuffer(bytearray(bytes(descs)), dtype=torch.uint8).cuda()
        check(L.te_gemm_batched(dev, st, 0, ord(opa), ord(opb), C.byref(al), C.byref(be), dd.data_ptr(), count, m, n), "batched")
    torch.cuda.synchronize()
    worst = 0.0
    for ref, cv, before, cb in refs:
        err = (cv - ref).abs().max().item() / max(1.0, ref.abs().max().item())
        worst = max(worst, err)
        # padding rows (ld > m) and slack must be untouched
        ld = cv.stride(1)
        mask = torch.ones_like(cb, dtype=torch.bool)
        idx = torch.arange(cb.numel(), device="cuda") - off
        inside = (idx >= 0) & (idx < ld * n) & ((idx % ld) < m)
        assert torch.equal(cb[~inside], before[~inside]), "wrote outside the tile"
    return worst

eps = 2.0 ** -52
bad = 0
cases = []
for opa in "NT":
    for opb in "NT":
        for (m, n, k) in [(512, 512, 512), (16, 16, 16), (80, 64, 16), (130, 257, 37), (128, 128, 0), (1, 1, 1), (384, 128, 128), (129, 127, 513)]:
            for pad, off in ((0, 0), (1, 0), (2, 1)):
                for inline in (False, True):
                    e = run_case(m, n, k, opa, opb, pad, inline=inline, off=off)
                    tol = 3 * max(k, 1) ** 0.5 * eps * 4
                    ok = e <= tol
                    bad += (not ok)
                    if not ok: print("FAIL", opa, opb, m, n, k, pad, off, inline, e, tol, flush=True)
for flags in (1, 2):
    for beta in (0.0, 1.0):
        e = run_case(512, 512, 512, 'N', 'T', 0, alpha=-1.0, beta=beta, flags=flags)
        ok = e <= 3 * 512 ** 0.5 * eps * 4; bad += (not ok)
        print("tri flags", flags, "beta", beta, "err", e, "ok", ok, flush=True)
print("correctness failures:", bad, flush=True)

# ---- speed: 2048 independent 512^3 tiles (the G per-step per-GPU shape on 2x4) and the 1-GPU shape
def speed(nt_rows, nt_cols, nb, opb, reps=3):
    A = torch.rand(nt_rows, nb * nb, dtype=torch.float64, device="cuda")
    B = torch.rand(nt_cols, nb * nb, dtype=torch.float64, device="cuda")
    Cm = torch.rand(nt_rows * nt_cols, nb * nb, dtype=torch.float64, device="cuda")
    count = nt_rows * nt_cols
    descs = (TileDesc * count)()
    idx = 0
    # 8x8 super-blocks for L2 reuse
    SB = 8
    for jb in range(0, nt_cols, SB):
        for ib in range(0, nt_rows, SB):
            for j in range(jb, min(jb + SB, nt_cols)):
                for i in range(ib, min(ib + SB, nt_rows)):
                    d = descs[idx]; idx += 1
                    d.a = A[i].data_ptr(); d.b = B[j].data_ptr(); d.c = Cm[i + j * nt_rows].data_ptr()
                    d.m = d.n = d.k = nb; d.lda = d.ldb = d.ldc = nb; d.flags = 0
    dd = torch.frombuffer(bytearray(bytes(descs)), dtype=torch.uint8).cuda()
    al = C.c_double(1.0); be = C.c_double(1.0)
    st = torch.cuda.current_stream().cuda_stream
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    best = 1e9
    for r in range(reps + 2):
        e0.record()
        check(L.te_gemm_batched(dev, st, 0, ord('N'), ord(opb), C.byref(al), C.byref(be), dd.data_ptr(), count, nb, nb))
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        if r >= 2: best = min(best, ms)
    tf = 2.0 * nb ** 3 * count / (best * 1e-3) * 1e-12
    print(f"speed {nt_rows}x{nt_cols} tiles nb={nb} opb={opb}: {best:.3f} ms  {tf:.2f} TFLOP/s", flush=True)
    return tf
res["gemm_64x32_NN"] = speed(64, 32, 512, 'N')
res["gemm_64x32_NT"] = speed(64, 32, 512, 'T')
res["gemm_96x96_NN"] = speed(96, 96, 512, 'N')
os.makedirs("gpurun_out", exist_ok=True)
json.dump(res, open("gpurun_out/first.json", "w"), indent=1)
print(json.dumps(res))
sys.exit(1 if bad else 0)
