#!/bin/bash
# round 2, 4 GPUs: full bench on the copy-engine plane, potrf launch trace, multi-rank parity
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${1:-4}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 900 $TR --master-port 29611 bench.py --gpus $N > gpurun_out/r2_bench_n$N.json 2> gpurun_out/r2_bench_n$N.err; echo "bench rc=$?"; tail -c 2500 gpurun_out/r2_bench_n$N.json
SLATE_B200_TIMING=1 SLATE_B200_TRACE_FILE=gpurun_out/r2_trace_n$N timeout 600 $TR --master-port 29612 bench.py --gpus $N --steps 4 --warmup 1 --no-e2e --no-zgemm --no-parity --no-cpu > gpurun_out/r2_bench_trace_n$N.json 2> gpurun_out/r2_bench_trace_n$N.err; echo "trace rc=$?"
ls gpurun_out | grep trace | head; rm -f gpurun_out/r2_trace_n${N}_r[1-9]*
timeout 900 $TR --master-port 29613 scripts/mp_parity.py > gpurun_out/r2_parity_n$N.log 2>&1; echo "parity rc=$?"; grep -v "^\*\|OMP_NUM" gpurun_out/r2_parity_n$N.log | tail -n 12
