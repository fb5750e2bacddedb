"""FP64 issue-rate sweep behind the roofline denominator (te_peak_fp64_ex): DMMA (mma.sync.m8n8k4.f64)
and DFMA throughput against warps per CTA, CTAs per SM and independent accumulator chains per
thread.  Writes one JSON object (default profiles/r02_peak_sweep.json).  One GPU, ~20 s."""
import ctypes as C, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from slate_d35_b200 import lib
te = lib()
te.te_peak_fp64_ex.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double)]
out = {"what": "te_peak_fp64_ex(kind, iters=20000, warps_per_cta, ctas_per_sm, chains) -> TFLOP/s, register resident",
       "kinds": {"0": "mma.sync.aligned.m8n8k4.f64 (DMMA.8x8x4)", "1": "fma.rn.f64 (DFMA)"}, "rows": []}
v = C.c_double()
te.te_peak_fp64(0, 0, 20000, C.byref(v))      # warm-up / clocks
for kind in (0, 1):
    for warps, ctas in ((1, 1), (2, 1), (4, 1), (8, 1), (4, 2), (8, 2), (8, 4), (16, 2)):
        for chains in (1, 2, 4, 8, 16, 32):
            st = te.te_peak_fp64_ex(0, kind, 20000, warps, ctas, chains, C.byref(v))
            if st == 0:
                out["rows"].append({"kind": kind, "warps_per_cta": warps, "ctas_per_sm": ctas,
                                    "chains": chains, "tflops": round(v.value, 3)})
best = {k: max(r["tflops"] for r in out["rows"] if r["kind"] == k) for k in (0, 1)}
out["best_dmma_tflops"], out["best_dfma_tflops"] = best[0], best[1]
te.te_peak_fp64(0, 0, 20000, C.byref(v))
out["te_peak_fp64_default_shape_tflops"] = round(v.value, 3)
dst = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "profiles", "r02_peak_sweep.json")
json.dump(out, open(dst, "w"), indent=1)
print(json.dumps({k: out[k] for k in ("best_dmma_tflops", "best_dfma_tflops", "te_peak_fp64_default_shape_tflops")}))
