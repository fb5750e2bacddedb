#!/bin/bash
# round 2, one GPU: beta == 1 epilogue by red.global.add.f64 against the load / fma / store epilogue
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TE_DGEMM_RED=1 timeout 600 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_drivers.py tests/test_constructors.py -x -q > gpurun_out/r2_pytest_red.log 2>&1; echo "pytest red rc=$?"; tail -n 2 gpurun_out/r2_pytest_red.log
for red in 1 0 1 0; do
  echo "== TE_DGEMM_RED=$red"
  TE_DGEMM_RED=$red timeout 300 python scripts/launch_shape.py 1 1 12 2>&1 | tail -n 1
  TE_DGEMM_RED=$red timeout 300 python scripts/launch_shape.py 2 4 24 2>&1 | tail -n 1
done
