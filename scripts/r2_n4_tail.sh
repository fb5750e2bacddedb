#!/bin/bash
# round 2, N GPUs: potrf with and without the short-CTA policy for the tail of the factorisation
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${1:-4}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
for mode in 1 0; do
  for pn in 98304 49152; do
    SLATE_B200_POTRF_TAIL_CAP=$mode timeout 600 $TR --master-port 29631 bench.py --gpus $N --steps 4 --warmup 1 --no-e2e --no-zgemm --no-parity --no-cpu --potrf-n $pn > gpurun_out/r2_tail_n${N}_${mode}_$pn.json 2> gpurun_out/r2_tail_n${N}_${mode}_$pn.err
    python -c "
import json
d=json.loads(open('gpurun_out/r2_tail_n${N}_${mode}_$pn.json').read().strip().splitlines()[-1]); print('tail policy $mode n=$pn', 'potrf', round(d['potrf']['value'],2), round(d['potrf']['frac_of_fp64_tensor_peak'],4), 'info', d['potrf']['info'])"
  done
done
