"""The reference's own slate::gemm / slate::potrf, Target::Devices, device-origin tiles, one GPU:
the same unmodified driver code over (a) the B200 tile engine (libslate_ref_te.so) and (b) BLAS++'s
cuBLAS / cuSOLVER wrappers (libslate_ref_cublas.so, the reference's stock GPU path).  Wall time of
the driver call alone (the reference synchronises the host after every internal op, so the host
overhead is the same in both).  usage: ref_binding_bench.py [n_gemm] [n_potrf] [nb]"""
import ctypes as C, os, sys, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from slate_d35_b200 import lib
lib()
i64, dp = C.c_int64, C.POINTER(C.c_double)
n_g = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
n_p = int(sys.argv[2]) if len(sys.argv) > 2 else 24576
nb = int(sys.argv[3]) if len(sys.argv) > 3 else 512
rng = np.random.default_rng(1)
P = lambda a: a.ctypes.data_as(dp)
out = {"n_gemm": n_g, "n_potrf": n_p, "nb": nb, "origin": "Devices", "lookahead": 1}
A = np.asfortranarray(rng.random((n_g, n_g)))
B = np.asfortranarray(rng.random((n_g, n_g)))
S = rng.random((n_p, n_p))
S = np.asfortranarray((S + S.T) / 2 + n_p * np.eye(n_p))
res = {}
for fl in ("te", "cublas"):
    L = C.CDLL(os.path.join(ROOT, "slate_d35_b200", "host", "ref_binding", "_build", f"libslate_ref_{fl}.so"))
    L.rb_dgemm.argtypes = [C.c_int, C.c_int, i64, i64, i64, i64, C.c_double, dp, i64, dp, i64, C.c_double, dp, i64, C.c_int, dp]
    L.rb_dpotrf.argtypes = [C.c_int, C.c_int, C.c_int, i64, i64, dp, i64, C.c_int, dp]
    L.rb_dpotrf.restype = i64
    L.rb_num_devices.restype = C.c_int
    out["devices_in_this_process"] = L.rb_num_devices()
    s = C.c_double()
    best_g = best_p = None
    for rep in range(3):
        Cm = np.zeros((n_g, n_g), order="F")
        assert L.rb_dgemm(ord("D"), ord("D"), n_g, n_g, n_g, nb, 1.0, P(A), n_g, P(B), n_g, 0.0, P(Cm), n_g, 1, C.byref(s)) == 0
        best_g = s.value if best_g is None else min(best_g, s.value)
        F = S.copy(order="F")
        assert L.rb_dpotrf(ord("D"), ord("D"), ord("L"), n_p, nb, P(F), n_p, 1, C.byref(s)) == 0
        best_p = s.value if best_p is None else min(best_p, s.value)
    res[fl] = {"C": Cm, "F": np.tril(F)}
    out[fl] = {"dgemm_s": best_g, "dgemm_tflops": 2.0 * n_g ** 3 / best_g * 1e-12,
               "dpotrf_s": best_p, "dpotrf_tflops": (n_p ** 3 / 3.0) / best_p * 1e-12}
out["agreement"] = {"dgemm": float(np.abs(res["te"]["C"] - res["cublas"]["C"]).max() / np.abs(res["cublas"]["C"]).max()),
                    "dpotrf": float(np.abs(res["te"]["F"] - res["cublas"]["F"]).max() / np.abs(res["cublas"]["F"]).max())}
print(json.dumps(out))
