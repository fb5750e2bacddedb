#!/bin/bash
# round 2: dgemm kernel shapes (wide / pair / pairfull) on one GPU: parity + launch-shape timing
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for shape in pair pairfull wide; do
  echo "== shape $shape"
  TE_DGEMM_SHAPE=$shape timeout 600 python -m pytest tests/test_gpu_kernels.py -x -q -k "dgemm or herk or syrk or trsm or potrf or getrf" > gpurun_out/r2_pytest_shape_$shape.log 2>&1; echo "pytest $shape rc=$?"; tail -n 2 gpurun_out/r2_pytest_shape_$shape.log
  TE_DGEMM_SHAPE=$shape timeout 300 python scripts/launch_shape.py 1 1 8 2>&1 | tail -n 1
  TE_DGEMM_SHAPE=$shape timeout 300 python scripts/launch_shape.py 2 4 16 2>&1 | tail -n 1
done
