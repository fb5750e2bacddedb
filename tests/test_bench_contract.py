"""bench.py contract checks that need no GPU: the reference arm (--impl reference) as the driver
launches it under torch.distributed.run, where the environment carries OMP_NUM_THREADS=1 -- rank 0
must still time the reference's HostTask path on ALL host cores and say how many, the other ranks
exit 0 without output (round-1 finding: the N >= 2 reference lines ran on one core)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(rank):
    env = dict(os.environ, OMP_NUM_THREADS="1", RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE="2")
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2",
                           "--steps", "1", "--warmup", "0", "--n", "1024"],
                          env=env, capture_output=True, text=True, timeout=600)


def test_reference_arm_uses_all_cores_under_torchrun():
    out = _run(0)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["unit"] == "TFLOP/s" and line["higher_is_better"] is True
    assert line["n_gpus"] == 2 and line["steps"] == 1
    cores = len(os.sched_getaffinity(0))
    assert line["cpu_baseline"]["cores"] == cores, (line["cpu_baseline"], cores)
    assert line["cpu_baseline"]["kind"] in ("reference", "port")
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["e2e"]["value"] == line["value"]
    assert line["gpu_launches"] == 0


def test_reference_arm_other_ranks_stay_silent():
    out = _run(1)
    assert out.returncode == 0 and out.stdout.strip() == ""
