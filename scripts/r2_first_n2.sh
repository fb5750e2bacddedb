#!/bin/bash
# round 2, first 2-GPU call: copy-engine plane microbench + parity on all three data planes + short benches
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
nvidia-smi topo -m > gpurun_out/r2_topo.txt 2>&1
timeout 300 $TR --master-port 29601 scripts/mp_peerbench.py > gpurun_out/r2_peerbench_n2.log 2>&1; echo "peerbench rc=$?"
TE_PEER_MEMOPS=0 timeout 300 $TR --master-port 29602 scripts/mp_peerbench.py > gpurun_out/r2_peerbench_n2_nomemops.log 2>&1; echo "peerbench(no memops) rc=$?"
MP_PARITY_EXTRA=1 timeout 900 $TR --master-port 29603 scripts/mp_parity.py > gpurun_out/r2_parity_peer_n2.log 2>&1; echo "parity peer rc=$?"
SLATE_B200_BCAST=coalesce timeout 900 $TR --master-port 29604 scripts/mp_parity.py > gpurun_out/r2_parity_coalesce_n2.log 2>&1; echo "parity coalesce rc=$?"
for plane in peer nccl coalesce; do
  SLATE_B200_BCAST=$plane timeout 600 $TR --master-port 29605 bench.py --gpus 2 --steps 16 --warmup 3 --no-e2e --no-cpu --potrf-n 49152 > gpurun_out/r2_bench_n2_$plane.log 2>&1; echo "bench $plane rc=$?"
done
timeout 600 python -m pytest tests/test_gpu_ref_binding.py -x -q > gpurun_out/r2_pytest_ref_binding_2dev.log 2>&1; echo "ref_binding, one process x 2 devices rc=$?"; tail -n 3 gpurun_out/r2_pytest_ref_binding_2dev.log
timeout 600 python scripts/ref_binding_bench.py 16384 24576 512 > gpurun_out/r2_ref_binding_bench_2dev.json 2>/dev/null; cat gpurun_out/r2_ref_binding_bench_2dev.json
tail -n 3 gpurun_out/r2_peerbench_n2.log gpurun_out/r2_parity_peer_n2.log gpurun_out/r2_parity_coalesce_n2.log
