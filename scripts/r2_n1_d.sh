#!/bin/bash
# round 2, one GPU: full GPU test suite, then the e2e arm at the driver's 20 steps with a launch trace
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_gpu_d.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/r2_pytest_gpu_d.log
timeout 600 python bench.py --steps 20 --warmup 3 --no-potrf --no-zgemm --no-herk --no-cpu --no-parity > gpurun_out/r2_bench_n1_k20.json 2> gpurun_out/r2_bench_n1_k20.err; echo "bench k20 rc=$?"; python -c "
import json
d=json.loads(open('gpurun_out/r2_bench_n1_k20.json').read().strip().splitlines()[-1]); print('value', d['value'], 'e2e', d['e2e'])"
SLATE_B200_TIMING=1 SLATE_B200_TRACE_FILE=gpurun_out/r2_trace_e2e_k20 timeout 600 python bench.py --steps 20 --warmup 3 --no-potrf --no-zgemm --no-herk --no-cpu --no-parity > gpurun_out/r2_bench_n1_k20_traced.json 2> gpurun_out/r2_bench_n1_k20_traced.err; echo "traced rc=$?"
timeout 300 python scripts/h2d_probe.py > gpurun_out/r2_h2d_probe.log 2>&1; tail -n 12 gpurun_out/r2_h2d_probe.log
ls -la gpurun_out | grep trace
