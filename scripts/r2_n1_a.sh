#!/bin/bash
# round 2, first 1-GPU call: GPU tests, default bench, peak sweep, DRAM traffic per launch shape
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_gpu_a.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/r2_pytest_gpu_a.log
timeout 900 python bench.py > gpurun_out/r2_bench_n1_a.json 2> gpurun_out/r2_bench_n1_a.err; echo "bench rc=$?"; tail -c 3000 gpurun_out/r2_bench_n1_a.json
timeout 300 python scripts/peak_sweep.py gpurun_out/r02_peak_sweep.json; echo "sweep rc=$?"
for pq in "1 2" "2 2" "2 4"; do
  set -- $pq
  timeout 300 python scripts/launch_shape.py $1 $2 3 > gpurun_out/r2_shape_$1x$2.log 2>&1
  timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none \
      -k regex:dgemm_tiles -s 1 -c 2 --csv --log-file gpurun_out/r2_traffic_$1x$2.csv python scripts/launch_shape.py $1 $2 3 > /dev/null 2>&1
  echo "traffic $1x$2 rc=$?"; tail -n 2 gpurun_out/r2_shape_$1x$2.log
done
timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_fp64.sum --clock-control none \
    -k regex:dgemm_tiles -s 1 -c 2 --csv --log-file gpurun_out/r2_herk_shape.csv python scripts/launch_shape.py 1 1 3 --herk > /dev/null 2>&1; echo "herk ncu rc=$?"
