#!/bin/bash
# round 2, N GPUs: the bench at the driver's step count (20), plain and with a launch trace
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29621 bench.py --gpus $N --steps 20 --warmup 5 --no-zgemm --no-parity > gpurun_out/r2_bench_n${N}_k20.json 2> gpurun_out/r2_bench_n${N}_k20.err; echo "bench rc=$?"
python -c "
import json
d=json.loads(open('gpurun_out/r2_bench_n${N}_k20.json').read().strip().splitlines()[-1]); print('value', d['value'], d['frac_of_fp64_tensor_peak'], 'e2e', d['e2e']['value'], d['e2e']['pcie_gbs_per_rank'], 'potrf', d['potrf']['value'], d['potrf']['frac_of_fp64_tensor_peak'])"
SLATE_B200_TIMING=1 SLATE_B200_TRACE_FILE=gpurun_out/r2_trace_n${N}_k20 timeout 600 $TR --master-port 29622 bench.py --gpus $N --steps 20 --warmup 5 --no-zgemm --no-parity --no-potrf > gpurun_out/r2_bench_n${N}_k20_traced.json 2> gpurun_out/r2_bench_n${N}_k20_traced.err; echo "traced rc=$?"
rm -f gpurun_out/r2_trace_n${N}_k20_r[1-9]*; ls gpurun_out | grep "trace_n${N}_k20"
