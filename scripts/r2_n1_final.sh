#!/bin/bash
# round 2, one GPU: what the driver runs at round end (tests, smoke, bench at its step count) + the default bench
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_gpu_final.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/r2_pytest_gpu_final.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()"; echo "smoke rc=$?"
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2_bench_n1_driver.json 2> gpurun_out/r2_bench_n1_driver.err; echo "bench(driver args) rc=$?"
timeout 900 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2_bench_n1_reference.json 2> gpurun_out/r2_bench_n1_reference.err; echo "reference arm rc=$?"; cat gpurun_out/r2_bench_n1_reference.json | cut -c1-400
timeout 900 python bench.py > gpurun_out/r2_bench_n1_default.json 2> gpurun_out/r2_bench_n1_default.err; echo "bench(default) rc=$?"
python - <<'PY'
import json
for f in ("driver","default"):
    d=json.loads(open(f"gpurun_out/r2_bench_n1_{f}.json").read().strip().splitlines()[-1])
    print(f, "value", round(d["value"],2), "frac", round(d["roofline"]["frac"],4), "e2e", round(d["e2e"]["value"],2), "potrf", round(d["potrf"]["value"],2), "zgemm", round(d["zgemm"]["value"],2), "herk", round(d["herk_microbench"]["herk"]["value"],2), "cpu", d["cpu_baseline"] and round(d["cpu_baseline"]["value"],3), "parity", d["parity"]["ok"], "clocks", d["clocks"]["sm_mhz"], d["clocks"]["reasons"])
PY
