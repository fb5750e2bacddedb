#!/bin/bash
# round 2, one GPU: e2e arm at the driver's step count after the host-origin schedule changes
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_drivers.py -x -q > gpurun_out/r2_pytest_drivers_e.log 2>&1; echo "pytest rc=$?"; tail -n 2 gpurun_out/r2_pytest_drivers_e.log
SLATE_B200_TIMING=1 SLATE_B200_TRACE_FILE=gpurun_out/r2_trace_e2e_k20b timeout 600 python bench.py --steps 20 --warmup 5 --no-potrf --no-zgemm --no-herk --no-cpu --no-parity > gpurun_out/r2_bench_n1_k20b_traced.json 2> gpurun_out/r2_bench_n1_k20b_traced.err; echo "traced rc=$?"
timeout 600 python bench.py --steps 20 --warmup 5 --no-potrf --no-zgemm --no-herk --no-cpu --no-parity > gpurun_out/r2_bench_n1_k20b.json 2> gpurun_out/r2_bench_n1_k20b.err; echo "bench k20 rc=$?"; python -c "
import json
d=json.loads(open('gpurun_out/r2_bench_n1_k20b.json').read().strip().splitlines()[-1]); print('value', d['value'], 'e2e', d['e2e'])"
