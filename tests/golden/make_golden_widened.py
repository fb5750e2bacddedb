"""Generate tests/golden/ref_widened.npz from the UNMODIFIED reference (oracle/_ref/
libslate_ref.so, `make -C oracle ref`): the Target::HostTask drivers of SURVEY.md section 8f
on the cases of widened_cases.py.  Run in the authoring container:

    OPENBLAS_NUM_THREADS=1 python tests/golden/make_golden_widened.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)
import oracle_lib as ol  # noqa: E402
import widened_cases as wc  # noqa: E402

R = ol.ref()
assert R is not None and hasattr(R, "ref_rank_k_any"), "build oracle/_ref first: make -C oracle ref"
Z = ol.Z


def vp(a):
    return a.ctypes.data


def run_reference(spec, x):
    """runs the reference driver in place on copies; returns the result array(s)"""
    pre, k = spec["pre"], spec["kind"]
    dt = 1 if pre == "z" else 0
    a, b = Z(wc.ALPHA[pre]), Z(wc.BETA[pre])
    nb, la = wc.NB, 1
    y = {n: np.asfortranarray(v.copy()) for n, v in x.items()}
    if k in ("herk", "syrk"):
        st = R.ref_rank_k_any(dt, 0 if k == "herk" else 1, ord(spec["uplo"]), ord(spec["opa"]), spec["n"],
                              spec["k"], nb, a, vp(y["A"]), y["A"].shape[0], b, vp(y["C"]), spec["n"], la)
        return st, y["C"]
    if k in ("her2k", "syr2k"):
        st = R.ref_rank_2k_any(dt, 0 if k == "her2k" else 1, ord(spec["uplo"]), ord(spec["opa"]), spec["n"],
                               spec["k"], nb, a, vp(y["A"]), y["A"].shape[0], vp(y["B"]), y["B"].shape[0], b,
                               vp(y["C"]), spec["n"], la)
        return st, y["C"]
    if k in ("hemm", "symm"):
        st = R.ref_hemm_any(dt, 0 if k == "hemm" else 1, ord(spec["side"]), ord(spec["uplo"]), spec["m"],
                            spec["n"], nb, a, vp(y["A"]), y["A"].shape[0], vp(y["B"]), spec["m"], b,
                            vp(y["C"]), spec["m"], la)
        return st, y["C"]
    if k == "trsm":
        st = R.ref_trsm_any(dt, ord(spec["side"]), ord(spec["uplo"]), ord(spec["opa"]), ord("N"), spec["m"],
                            spec["n"], nb, a, vp(y["A"]), y["A"].shape[0], vp(y["B"]), spec["m"], la)
        return st, y["B"]
    if k == "posv":
        st = R.ref_posv_any(dt, ord(spec["uplo"]), spec["n"], spec["nrhs"], nb, vp(y["A"]), spec["n"],
                            vp(y["B"]), spec["n"], la)
        return st, np.concatenate([y["A"], y["B"]], axis=1)
    if k == "gemmA":
        st = R.ref_gemmA_any(dt, ord(spec["opa"]), ord(spec["opb"]), spec["m"], spec["n"], spec["k"], nb, a,
                             vp(y["A"]), y["A"].shape[0], vp(y["B"]), y["B"].shape[0], b, vp(y["C"]),
                             spec["m"], la)
        return st, y["C"]
    if k == "getrf_nopiv":
        st = R.ref_getrf_nopiv_any(dt, spec["m"], spec["n"], nb, vp(y["A"]), spec["m"], la)
        return st, y["A"]
    raise ValueError(k)


out = {}
for name, spec in wc.cases():
    x = wc.inputs(spec, lambda seed, m, n, dt: ol.matgen(seed, m, n, dt, lib=R))
    st, res = run_reference(spec, x)
    assert st == 0, (name, st)
    out[name] = res
np.savez_compressed(os.path.join(HERE, "ref_widened.npz"), **out)
print(len(out), "cases,", os.path.getsize(os.path.join(HERE, "ref_widened.npz")), "bytes")
