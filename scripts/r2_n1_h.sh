#!/bin/bash
# round 2, one GPU: dgemm K-slice 32 (wide32) against the default shape
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TE_DGEMM_SHAPE=wide32 timeout 600 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_drivers.py -x -q > gpurun_out/r2_pytest_wide32.log 2>&1; echo "pytest wide32 rc=$?"; tail -n 2 gpurun_out/r2_pytest_wide32.log
for shape in wide32 wide wide32 wide; do
  echo "== $shape"
  TE_DGEMM_SHAPE=$shape timeout 300 python scripts/launch_shape.py 1 1 12 2>&1 | tail -n 1
  TE_DGEMM_SHAPE=$shape timeout 300 python scripts/launch_shape.py 2 4 24 2>&1 | tail -n 1
done
