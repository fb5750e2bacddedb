#!/bin/bash
# round 2, one GPU: constructor tests, full GPU suite, ncu of the aux / panel kernels
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_constructors.py -x -q > gpurun_out/r2_pytest_constructors.log 2>&1; echo "constructors rc=$?"; tail -n 15 gpurun_out/r2_pytest_constructors.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_gpu_f.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/r2_pytest_gpu_f.log
timeout 120 python scripts/aux_probe.py; echo "probe rc=$?"
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,sm__throughput.avg.pct_of_peak_sustained_elapsed,smsp__issue_active.avg.pct_of_peak_sustained_active,launch__grid_size,launch__block_size,launch__registers_per_thread
timeout 600 ncu --metrics $M --clock-control none -s 20 -c 80 --csv --log-file gpurun_out/r2_ncu_aux_kernels.csv python scripts/aux_probe.py > /dev/null 2>&1; echo "ncu aux rc=$?"
grep -c . gpurun_out/r2_ncu_aux_kernels.csv
