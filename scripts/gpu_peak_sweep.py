import ctypes as C, os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from slate_d35_b200 import lib, check
L = lib()
L.te_peak_fp64_ex.argtypes = [C.c_int]*6 + [C.POINTER(C.c_double)]
out = {}
for kind in (0, 1):
    for (w, c) in ((1, 1), (4, 1), (8, 1), (4, 2), (16, 1), (8, 4)):
        for ch in (1, 2, 4, 8, 16, 32):
            tf = C.c_double()
            check(L.te_peak_fp64_ex(0, kind, 4000, w, c, ch, C.byref(tf)))
            out[f"{'dmma' if kind == 0 else 'dfma'}_w{w}_c{c}_ch{ch}"] = round(tf.value, 2)
            print(kind, w, c, ch, round(tf.value, 2), flush=True)
json.dump(out, open("gpurun_out/peak_sweep.json", "w"), indent=1)
