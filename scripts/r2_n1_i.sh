#!/bin/bash
# round 2, one GPU: run-based host<->device tile transfers: tests, then e2e at the driver's step count
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_gpu_i.log 2>&1; echo "pytest rc=$?"; tail -n 6 gpurun_out/r2_pytest_gpu_i.log
for k in 20 128; do
timeout 900 python bench.py --steps $k --warmup 5 --no-potrf --no-zgemm --no-herk --no-cpu --no-parity > gpurun_out/r2_bench_n1_k${k}_runs.json 2> gpurun_out/r2_bench_n1_k${k}_runs.err; echo "bench k=$k rc=$?"; python -c "
import json
d=json.loads(open('gpurun_out/r2_bench_n1_k${k}_runs.json').read().strip().splitlines()[-1]); print('steps $k value', round(d['value'],2), 'e2e', round(d['e2e']['value'],2), 'pcie', round(d['e2e']['pcie_gbs_per_rank'],1), 'gemm_error', d['gemm_error'])"
done
