#!/bin/bash
# round 2, one GPU: complex kernel with the branch-free interior epilogue against the default
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TE_ZGEMM_EPILOGUE=1 timeout 300 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_drivers.py -x -q -k "z or complex or C64 or c128 or gemm or herk or her2k or hemm" > gpurun_out/r2_pytest_zepi.log 2>&1; echo "pytest rc=$?"; tail -n 2 gpurun_out/r2_pytest_zepi.log
for z in 1 0; do
TE_ZGEMM_EPILOGUE=$z timeout 200 python bench.py --n 8192 --steps 2 --warmup 1 --no-e2e --no-potrf --no-herk --no-parity --no-cpu --zgemm-n 24576 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('TE_ZGEMM_EPILOGUE=$z zgemm', round(d['zgemm']['value'],2), 'err', d['zgemm']['error'])"
done
