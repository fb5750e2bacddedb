"""One GPU, a few SUMMA block steps of slate::gemm with the LOCAL problem one rank of a p x q
grid sees at n = 65536, nb = 512 (local C = (128/p) x (128/q) tiles): the launch shape of
dgemm_tiles_kernel at N = p*q GPUs, for ncu captures (DRAM traffic per launch, tensor pipe).
usage: launch_shape.py P Q [steps] [--herk]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from slate_d35_b200 import host as sl
p, q = int(sys.argv[1]), int(sys.argv[2])
steps = int(sys.argv[3]) if len(sys.argv) > 3 and sys.argv[3].isdigit() else 3
nb = 512
m, n, k = 65536 // p, 65536 // q, nb * steps
sl.initialize(0)
if "--herk" in sys.argv:
    # config 4: one internal::herk tile update, n x n lower, k = nb (diagonal tiles masked in-kernel)
    nh = 32768
    A = sl.Matrix(nh, nb, nb).generate(1, dist=2)
    Cm = sl.HermitianMatrix("L", nh, nb).generate(2)
    for _ in range(steps):
        sl.internal_herk(-1.0, A, 1.0, Cm)
    sl.sync()
    print(f"herk n={nh} k={nb}: {steps} launches")
else:
    A = sl.Matrix(m, k, nb).generate(1)
    B = sl.Matrix(k, n, nb).generate(2)
    Cm = sl.Matrix(m, n, nb).generate(3)
    ms = sl.gemm(3.141592653589793, A, B, 2.718281828459045, Cm, timed=True, method="C")
    print(f"local shape of a {p}x{q} grid: C {m // nb} x {n // nb} tiles = {(m // nb) * (n // nb)} tiles per launch, "
          f"{steps} launches, {ms / steps:.3f} ms per launch, {2.0 * m * n * nb / (ms / steps * 1e-3) * 1e-12:.2f} TFLOP/s")
