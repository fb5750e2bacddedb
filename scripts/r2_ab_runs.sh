#!/bin/bash
# round 2: A/B of the panel path for host-origin tiles at N GPUs, e2e at the driver's step count
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
for mode in ${MODES:-1 0 1 0}; do
  SLATE_B200_TILE_RUNS=$mode timeout 600 $TR --master-port 29651 bench.py --gpus $N --steps 20 --warmup 5 --no-potrf --no-zgemm --no-parity --no-cpu > gpurun_out/r2_ab_runs_n${N}_$mode.json 2> gpurun_out/r2_ab_runs_n${N}_$mode.err
  python -c "
import json
d=json.loads(open('gpurun_out/r2_ab_runs_n${N}_$mode.json').read().strip().splitlines()[-1]); print('N=$N tile runs $mode: value', round(d['value'],2), 'e2e', round(d['e2e']['value'],2), 'pcie/rank', round(d['e2e']['pcie_gbs_per_rank'],1))"
done
