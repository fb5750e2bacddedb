"""Generate tests/golden/*.npz from the UNMODIFIED reference (oracle/_ref/libslate_ref.so,
built by `make -C oracle ref` from /root/reference).  Run in the authoring container:

    OPENBLAS_NUM_THREADS=1 python tests/golden/make_golden.py

Inputs are regenerated in the tests from the seeds (matgen is a pure function of
(seed, i, j)); only outputs and the parameters are stored.  Sizes are small on purpose.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_lib as ol  # noqa: E402

R = ol.ref()
assert R is not None, "build oracle/_ref first: make -C oracle ref"
P, Z = ol.P, ol.Z
out = {}

# 1. matgen raw values (bit-exact pin of the Philox restatement), incl. offsets
out["matgen_d_seed7_off_3_5"] = ol.matgen(7, 9, 11, np.float64, 1, 3, 5, lib=R)
out["matgen_z_seed7_off_3_5"] = ol.matgen(7, 9, 11, np.complex128, 1, 3, 5, lib=R)
out["matgen_d_signed_seed42"] = ol.matgen(42, 6, 4, np.float64, 2, 0, 0, lib=R)
out["matgen_d_seed1_big_off"] = ol.matgen(1, 4, 4, np.float64, 1, 98300, 65530, lib=R)

# 2. slate::gemm, ragged tiles, all op pairs (d) / selected (z)
m, n, k, nb = 70, 53, 45, 16
alpha, beta = 3.141592653589793, 2.718281828459045           # test/test.cc:392-393 (real part)
for opa in "NTC":
    for opb in "NTC":
        A = ol.matgen(1, *((m, k) if opa == "N" else (k, m)), lib=R)
        B = ol.matgen(2, *((k, n) if opb == "N" else (n, k)), lib=R)
        Cm = ol.matgen(3, m, n, lib=R)
        R.ref_dgemm(ord(opa), ord(opb), m, n, k, nb, alpha, P(A), A.shape[0], P(B), B.shape[0],
                    beta, P(Cm), m, 1)
        out[f"dgemm_{opa}{opb}"] = Cm
za, zb = alpha + 1.4142135623730951j, beta + 1.7320508075688772j
for opa in "NTC":
    for opb in "NTC":
        A = ol.matgen(1, *((m, k) if opa == "N" else (k, m)), np.complex128, lib=R)
        B = ol.matgen(2, *((k, n) if opb == "N" else (n, k)), np.complex128, lib=R)
        Cm = ol.matgen(3, m, n, np.complex128, lib=R)
        R.ref_zgemm(ord(opa), ord(opb), m, n, k, nb, Z(za), P(A), A.shape[0], P(B), B.shape[0],
                    Z(zb), P(Cm), m, 1)
        out[f"zgemm_{opa}{opb}"] = Cm

# 3. slate::potrf, rand_dominant (generate_type_rand.hh:70-78), ragged last tile
n, nb = 90, 32
for dt, pre in ((np.float64, "d"), (np.complex128, "z")):
    for uplo in "LU":
        A = ol.matgen(5, n, n, dt, lib=R)
        A[np.arange(n), np.arange(n)] += n
        info = getattr(R, f"ref_{pre}potrf")(ord(uplo), n, nb, P(A), n, 1)
        out[f"{pre}potrf_{uplo}"] = A
        out[f"{pre}potrf_{uplo}_info"] = np.array([info])
# not positive definite: info must name the first bad minor
A = ol.matgen(5, n, n, lib=R)
A[np.arange(n), np.arange(n)] += n
A[40, 40] = -1.0
out["dpotrf_L_bad_info"] = np.array([R.ref_dpotrf(ord("L"), n, nb, P(A), n, 1)])

# 4. internal ops (one block column), as unit_test/test_internal_blas.cc shapes
m, n, k, nb = 80, 64, 16, 16
for opa in "NT":
    for opb in "NT":
        A = ol.matgen(11, *((m, k) if opa == "N" else (k, m)), lib=R)
        B = ol.matgen(12, *((k, n) if opb == "N" else (n, k)), lib=R)
        Cm = ol.matgen(13, m, n, lib=R)
        R.ref_internal_dgemm(ord(opa), ord(opb), m, n, k, nb, alpha, P(A), A.shape[0], P(B),
                             B.shape[0], beta, P(Cm), m)
        out[f"internal_dgemm_{opa}{opb}"] = Cm
for dt, pre in ((np.float64, "d"), (np.complex128, "z")):
    for uplo in "LU":
        A = ol.matgen(14, n, k, dt, lib=R)
        Cm = ol.matgen(15, n, n, dt, lib=R)
        if pre == "z":
            Cm[np.arange(n), np.arange(n)] = Cm[np.arange(n), np.arange(n)].real
        getattr(R, f"ref_internal_{pre}herk")(ord(uplo), ord("N"), n, k, nb, -1.0, P(A), n, 1.0,
                                              P(Cm), n)
        out[f"internal_{pre}herk_{uplo}"] = Cm
    # trsm as potrf uses it: Right, Lower, ConjTrans, NonUnit, one triangular tile
    nbt = 32
    T = ol.matgen(16, nbt, nbt, dt, lib=R)
    T[np.arange(nbt), np.arange(nbt)] += nbt
    Bm = ol.matgen(17, 75, nbt, dt, lib=R)
    if pre == "d":
        R.ref_internal_dtrsm(ord("R"), ord("L"), ord("C"), ord("N"), 75, nbt, nbt, 1.0, P(T), nbt,
                             P(Bm), 75)
    else:
        R.ref_internal_ztrsm(ord("R"), ord("L"), ord("C"), ord("N"), 75, nbt, nbt, Z(1.0), P(T),
                             nbt, P(Bm), 75)
    out[f"internal_{pre}trsm_RLC"] = Bm
    T2 = ol.matgen(18, nbt, nbt, dt, lib=R)
    T2[np.arange(nbt), np.arange(nbt)] += nbt
    info = getattr(R, f"ref_internal_{pre}potrf")(ord("L"), nbt, P(T2), nbt)
    out[f"internal_{pre}potrf_L"] = T2
    out[f"internal_{pre}potrf_L_info"] = np.array([info])

np.savez_compressed(os.path.join(HERE, "ref_hosttask.npz"), **out)
print("wrote", len(out), "arrays,", os.path.getsize(os.path.join(HERE, "ref_hosttask.npz")), "bytes")
