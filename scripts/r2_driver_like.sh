#!/bin/bash
# round 2: what the driver's scaling run does at N GPUs (bench.py --steps 20 --warmup 5, both arms)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29641 bench.py --impl reference --gpus $N --steps 20 --warmup 5 > gpurun_out/r2_driverlike_n${N}_reference.json 2> gpurun_out/r2_driverlike_n${N}_reference.err; echo "reference rc=$?"
timeout 900 $TR --master-port 29642 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r2_driverlike_n${N}.json 2> gpurun_out/r2_driverlike_n${N}.err; echo "bench rc=$?"
python - <<PY
import json
d=json.loads(open("gpurun_out/r2_driverlike_n${N}.json").read().strip().splitlines()[-1])
r=json.loads(open("gpurun_out/r2_driverlike_n${N}_reference.json").read().strip().splitlines()[-1])
print("N=${N}", "value", round(d["value"],2), "frac", round(d["frac_of_fp64_tensor_peak"],4), "e2e", round(d["e2e"]["value"],2), "pcie/rank", round(d["e2e"]["pcie_gbs_per_rank"],1), "potrf", round(d["potrf"]["value"],2), round(d["potrf"]["frac_of_fp64_tensor_peak"],4), "zgemm", round(d["zgemm"]["value"],2), round(d["zgemm"]["frac_of_fp64_tensor_peak"],4), "parity", d["parity"]["ok"], round(d["parity"]["max_err_eps"],2), "traffic", d["roofline"]["traffic"])
print("reference", round(r["value"],3), "cores", r["cpu_baseline"]["cores"], "ratio", round(d["value"]/r["value"],1), "e2e ratio", round(d["e2e"]["value"]/r["value"],1))
PY
