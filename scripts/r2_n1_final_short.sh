#!/bin/bash
# round 2, one GPU: last check of the final build (tests, smoke, default bench line)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_gpu_final.log 2>&1; echo "pytest rc=$?"; tail -n 2 gpurun_out/r2_pytest_gpu_final.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()"; echo "smoke rc=$?"
timeout 600 python bench.py --no-cpu > gpurun_out/r2_bench_n1_default.json 2> gpurun_out/r2_bench_n1_default.err; echo "bench(default) rc=$?"
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r2_bench_n1_default.json").read().strip().splitlines()[-1])
print("default: value", round(d["value"],2), "frac", round(d["roofline"]["frac"],4), "e2e", round(d["e2e"]["value"],2), "potrf", round(d["potrf"]["value"],2), round(d["potrf"]["frac_of_fp64_tensor_peak"],4), d["potrf"]["backward_error"], "zgemm", round(d["zgemm"]["value"],2), "herk", round(d["herk_microbench"]["herk"]["value"],2), "parity", d["parity"]["ok"], d["parity"]["max_err_eps"], "gemm_error", d["gemm_error"])
PY
