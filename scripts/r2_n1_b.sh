#!/bin/bash
# round 2: the reference's own sources over the engine (ref_binding) on the box's GPUs
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ref_binding.py -x -q > gpurun_out/r2_pytest_ref_binding.log 2>&1; echo "pytest rc=$?"; tail -n 30 gpurun_out/r2_pytest_ref_binding.log
timeout 900 python scripts/ref_binding_bench.py 16384 24576 512 > gpurun_out/r2_ref_binding_bench.json 2> gpurun_out/r2_ref_binding_bench.err; echo "bench rc=$?"; cat gpurun_out/r2_ref_binding_bench.json; tail -n 5 gpurun_out/r2_ref_binding_bench.err
