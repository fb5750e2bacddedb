#!/bin/bash
# round 2, one GPU: full suite, then one --set full capture (with source) of the dgemm kernel
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_gpu_g.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/r2_pytest_gpu_g.log
timeout 600 python scripts/launch_shape.py 2 4 3 | tail -n 1
timeout 900 ncu --set full --import-source on --clock-control none -k regex:dgemm_tiles -s 1 -c 1 -f -o gpurun_out/r2_dgemm_full python scripts/launch_shape.py 2 4 3 > gpurun_out/r2_ncu_full.log 2>&1; echo "ncu rc=$?"; ls -la gpurun_out/r2_dgemm_full.ncu-rep
