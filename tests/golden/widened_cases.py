"""The cases of tests/golden/ref_widened.npz: drivers of SURVEY.md section 8f (herk, syrk, her2k,
syr2k, hemm, symm, trsm, posv, gemmA, getrf_nopiv) on small ragged problems.  One table, three
users: make_golden_widened.py runs the UNMODIFIED reference (oracle/_ref) on it, the CPU tests
check the stored outputs against plain numpy, the GPU tests run the B200 drivers on the same
inputs and compare with the stored outputs.

Inputs are regenerated from matgen seeds (a pure function of (seed, i, j)); only outputs are
stored.  Every case is (name, spec) with spec a dict of parameters.
"""
import numpy as np

NB = 8
DT = {"d": np.float64, "z": np.complex128}
ALPHA = {"d": 3.141592653589793, "z": 3.141592653589793 + 1.4142135623730951j}
BETA = {"d": 2.718281828459045, "z": 2.718281828459045 + 1.7320508075688772j}


def cases():
    out = []
    for pre in "dz":
        for uplo in "LU":
            for opa in ("N", "C"):
                out.append((f"{pre}herk_{uplo}{opa}", dict(kind="herk", pre=pre, uplo=uplo, opa=opa, n=38, k=21)))
                tr = "T" if opa == "C" else "N"
                out.append((f"{pre}syrk_{uplo}{tr}", dict(kind="syrk", pre=pre, uplo=uplo, opa=tr, n=38, k=21)))
                out.append((f"{pre}her2k_{uplo}{opa}", dict(kind="her2k", pre=pre, uplo=uplo, opa=opa, n=38, k=21)))
                out.append((f"{pre}syr2k_{uplo}{tr}", dict(kind="syr2k", pre=pre, uplo=uplo, opa=tr, n=38, k=21)))
            for side in "LR":
                out.append((f"{pre}hemm_{side}{uplo}", dict(kind="hemm", pre=pre, side=side, uplo=uplo, m=38, n=29)))
                out.append((f"{pre}symm_{side}{uplo}", dict(kind="symm", pre=pre, side=side, uplo=uplo, m=38, n=29)))
                for opa in "NTC":
                    out.append((f"{pre}trsm_{side}{uplo}{opa}", dict(kind="trsm", pre=pre, side=side, uplo=uplo,
                                                                      opa=opa, m=38, n=29)))
            out.append((f"{pre}posv_{uplo}", dict(kind="posv", pre=pre, uplo=uplo, n=42, nrhs=5)))
        for opa, opb in (("N", "N"), ("T", "N"), ("N", "C")):
            out.append((f"{pre}gemmA_{opa}{opb}", dict(kind="gemmA", pre=pre, opa=opa, opb=opb, m=38, n=29, k=21)))
        for m, n in ((42, 42), (38, 24), (24, 38)):   # the reference needs square diagonal tiles
            out.append((f"{pre}getrf_nopiv_{m}x{n}", dict(kind="getrf_nopiv", pre=pre, m=m, n=n)))
    return out


def op(a, t):
    return {"N": a, "T": a.T, "C": a.conj().T}[t]


def inputs(spec, matgen):
    """the input arrays of a case, from matgen(seed, m, n, dtype)"""
    dt, k = DT[spec["pre"]], spec["kind"]
    if k in ("herk", "syrk", "her2k", "syr2k"):
        n, kk = spec["n"], spec["k"]
        ash = (n, kk) if spec["opa"] == "N" else (kk, n)
        C0 = matgen(23, n, n, dt)
        if k in ("herk", "her2k"):
            C0[np.arange(n), np.arange(n)] = C0[np.arange(n), np.arange(n)].real
        return dict(A=matgen(21, *ash, dt), B=matgen(22, *ash, dt), C=C0)
    if k in ("hemm", "symm"):
        an = spec["m"] if spec["side"] == "L" else spec["n"]
        A0 = matgen(24, an, an, dt)
        if k == "hemm":
            A0[np.arange(an), np.arange(an)] = A0[np.arange(an), np.arange(an)].real
        return dict(A=A0, B=matgen(25, spec["m"], spec["n"], dt), C=matgen(26, spec["m"], spec["n"], dt))
    if k == "trsm":
        an = spec["m"] if spec["side"] == "L" else spec["n"]
        A0 = matgen(27, an, an, dt)
        A0[np.arange(an), np.arange(an)] += an
        return dict(A=A0, B=matgen(28, spec["m"], spec["n"], dt))
    if k == "posv":
        n = spec["n"]
        A0 = matgen(29, n, n, dt)
        A0[np.arange(n), np.arange(n)] = A0[np.arange(n), np.arange(n)].real + n
        return dict(A=A0, B=matgen(30, n, spec["nrhs"], dt))
    if k == "gemmA":
        m, n, kk = spec["m"], spec["n"], spec["k"]
        ash = (m, kk) if spec["opa"] == "N" else (kk, m)
        bsh = (kk, n) if spec["opb"] == "N" else (n, kk)
        return dict(A=matgen(31, *ash, dt), B=matgen(32, *bsh, dt), C=matgen(33, m, n, dt))
    if k == "getrf_nopiv":
        m, n = spec["m"], spec["n"]
        A0 = matgen(34, m, n, dt)
        mn = min(m, n)
        A0[np.arange(mn), np.arange(mn)] += max(m, n)
        return dict(A=A0)
    raise ValueError(k)


def full_from_triangle(A, uplo, hermitian):
    """the matrix a driver sees when only the `uplo` triangle of A is referenced"""
    s = np.tril(A, -1) if uplo == "L" else np.triu(A, 1)
    d = np.diag(np.diag(A).real if hermitian else np.diag(A))
    return s + (s.conj().T if hermitian else s.T) + d


def expected(spec, x):
    """plain numpy value of the result array of a case (the part that is defined)"""
    pre, k = spec["pre"], spec["kind"]
    a, b = ALPHA[pre], BETA[pre]
    if k == "herk":
        Ao = op(x["A"], spec["opa"])
        return a.real * Ao @ Ao.conj().T + b.real * full_from_triangle(x["C"], spec["uplo"], True)
    if k == "syrk":
        Ao = op(x["A"], spec["opa"])
        return a * Ao @ Ao.T + b * full_from_triangle(x["C"], spec["uplo"], False)
    if k == "her2k":
        Ao, Bo = op(x["A"], spec["opa"]), op(x["B"], spec["opa"])
        return a * Ao @ Bo.conj().T + np.conj(a) * Bo @ Ao.conj().T + b.real * full_from_triangle(x["C"], spec["uplo"], True)
    if k == "syr2k":
        Ao, Bo = op(x["A"], spec["opa"]), op(x["B"], spec["opa"])
        return a * Ao @ Bo.T + a * Bo @ Ao.T + b * full_from_triangle(x["C"], spec["uplo"], False)
    if k in ("hemm", "symm"):
        S = full_from_triangle(x["A"], spec["uplo"], k == "hemm")
        return a * (S @ x["B"] if spec["side"] == "L" else x["B"] @ S) + b * x["C"]
    if k == "gemmA":
        return a * op(x["A"], spec["opa"]) @ op(x["B"], spec["opb"]) + b * x["C"]
    raise ValueError(k)
