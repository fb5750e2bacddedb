"""tests/golden/ref_widened.npz holds what the UNMODIFIED reference (Target::HostTask, built by
oracle/Makefile) returns for the drivers of SURVEY.md section 8f.  CPU part: the stored outputs
agree with plain numpy on the regenerated inputs, which pins the numpy expressions the GPU tests
use as their checker to the reference's semantics (which triangle is read, where beta applies,
conjugation, what trsm / posv / getrf_nopiv leave in place).  GPU part: the B200 drivers on the
same inputs against the stored outputs.
"""
import os
import sys

import numpy as np
import pytest

import oracle_lib as ol

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import widened_cases as wc  # noqa: E402

EPS = np.finfo(np.float64).eps
GOLD = np.load(os.path.join(HERE, "golden", "ref_widened.npz"))
CASES = wc.cases()


def test_every_case_is_stored():
    assert sorted(GOLD.files) == sorted(name for name, _ in CASES) and len(CASES) >= 80


def tolerance(spec, scale):
    k = {"herk": spec.get("k"), "syrk": spec.get("k"), "her2k": 2 * spec.get("k", 0), "syr2k": 2 * spec.get("k", 0),
         "hemm": max(spec.get("m", 0), spec.get("n", 0)), "symm": max(spec.get("m", 0), spec.get("n", 0)),
         "gemmA": spec.get("k")}[spec["kind"]]
    return 3 * np.sqrt(k * (4 if spec["pre"] == "z" else 1)) * EPS * scale


def check_against(spec, x, res, ref_is_golden):
    """res: result array of a driver run on inputs x; numpy is the judge"""
    k = spec["kind"]
    tri = (np.tril if spec.get("uplo") == "L" else np.triu)
    if k in ("herk", "syrk", "her2k", "syr2k"):
        want = wc.expected(spec, x)
        assert np.abs(tri(res) - tri(want)).max() <= tolerance(spec, np.abs(want).max())
        other = (lambda a: np.triu(a, 1)) if spec["uplo"] == "L" else (lambda a: np.tril(a, -1))
        assert np.array_equal(other(res), other(x["C"]))           # the other triangle is not touched
        if k in ("herk", "her2k"):
            assert not np.diag(res).imag.any()
    elif k in ("hemm", "symm", "gemmA"):
        want = wc.expected(spec, x)
        assert np.abs(res - want).max() <= tolerance(spec, np.abs(want).max())
    elif k == "trsm":
        T = tri(x["A"])
        To = wc.op(T, spec["opa"])
        back = To @ res if spec["side"] == "L" else res @ To
        assert np.abs(back - wc.ALPHA[spec["pre"]] * x["B"]).max() <= 50 * EPS * np.abs(x["B"]).max() * np.linalg.cond(T, 1)
    elif k == "posv":
        n = spec["n"]
        F, X = tri(res[:, :n]), res[:, n:]
        S = wc.full_from_triangle(x["A"], spec["uplo"], True)
        back = F @ F.conj().T if spec["uplo"] == "L" else F.conj().T @ F
        assert np.linalg.norm(back - S) / (n * np.linalg.norm(S)) <= 50 * EPS
        assert np.linalg.norm(x["B"] - S @ X, 1) / (n * np.linalg.norm(S, 1) * np.linalg.norm(X, 1)) <= 25 * EPS
    elif k == "getrf_nopiv":
        m, n = spec["m"], spec["n"]
        mn = min(m, n)
        Lf = np.tril(res[:, :mn], -1) + np.eye(m, mn)
        Uf = np.triu(res[:mn, :])
        assert np.linalg.norm(Lf @ Uf - x["A"]) / (max(m, n) * np.linalg.norm(x["A"])) <= 50 * EPS
    else:
        raise AssertionError(k)


@pytest.mark.parametrize("name,spec", CASES, ids=[n for n, _ in CASES])
def test_reference_output_matches_numpy(name, spec):
    x = wc.inputs(spec, lambda seed, m, n, dt: ol.matgen(seed, m, n, dt))
    check_against(spec, x, GOLD[name], True)


ORACLE_KINDS = ("her2k", "syr2k", "hemm", "symm", "getrf_nopiv")


@pytest.mark.parametrize("name,spec", [c for c in CASES if c[1]["kind"] in ORACLE_KINDS],
                         ids=[n for n, s in CASES if s["kind"] in ORACLE_KINDS])
def test_c_oracle_matches_reference_output(name, spec):
    """the plain-C restatements of the widened drivers (oracle/oracle_tiles_body.inc) against
    what the unmodified reference returned for the same inputs"""
    O = ol.oracle()
    x = wc.inputs(spec, lambda seed, m, n, dt: ol.matgen(seed, m, n, dt))
    pre, k, nb = spec["pre"], spec["kind"], wc.NB
    a, b = wc.ALPHA[pre], wc.BETA[pre]
    z = pre == "z"
    sc = (lambda v: ol.Z(v)) if z else (lambda v: float(np.real(v)))
    y = {n_: np.asfortranarray(v.copy()) for n_, v in x.items()}
    gold = GOLD[name]
    if k in ("her2k", "syr2k"):
        f = O.oracle_zrank2k if z else O.oracle_drank2k
        beta = b.real if k == "her2k" else b
        f(1 if k == "her2k" else 0, ord(spec["uplo"]), ord(spec["opa"]), spec["n"], spec["k"], nb, sc(a),
          ol.P(y["A"]), y["A"].shape[0], ol.P(y["B"]), y["B"].shape[0], sc(beta), ol.P(y["C"]), spec["n"])
        tri = np.tril if spec["uplo"] == "L" else np.triu
        assert np.abs(tri(y["C"]) - tri(gold)).max() <= tolerance(spec, np.abs(gold).max())
        other = (lambda q: np.triu(q, 1)) if spec["uplo"] == "L" else (lambda q: np.tril(q, -1))
        assert np.array_equal(other(y["C"]), other(gold))
    elif k in ("hemm", "symm"):
        f = O.oracle_zhemm if z else O.oracle_dhemm
        f(1 if k == "hemm" else 0, ord(spec["side"]), ord(spec["uplo"]), spec["m"], spec["n"], nb, sc(a),
          ol.P(y["A"]), y["A"].shape[0], ol.P(y["B"]), spec["m"], sc(b), ol.P(y["C"]), spec["m"])
        assert np.abs(y["C"] - gold).max() <= tolerance(spec, np.abs(gold).max())
    else:
        f = O.oracle_zgetrf_nopiv if z else O.oracle_dgetrf_nopiv
        assert f(spec["m"], spec["n"], nb, ol.P(y["A"]), spec["m"]) == 0
        assert np.abs(y["A"] - gold).max() <= 50 * EPS * np.abs(gold).max()


# ------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def sl():
    from slate_d35_b200 import host
    host.initialize(0)
    yield host
    host.sync()


def run_b200(sl, spec, x):
    pre, k = spec["pre"], spec["kind"]
    dt, nb = wc.DT[pre], wc.NB
    a, b = wc.ALPHA[pre], wc.BETA[pre]
    if k in ("herk", "syrk", "her2k", "syr2k"):
        A = sl.Matrix(*x["A"].shape, nb, dtype=dt).from_numpy(x["A"])
        c_h = x["C"].copy(order="F")
        cls = sl.HermitianMatrix if k in ("herk", "her2k") else sl.SymmetricMatrix
        Cm = cls.fromLAPACK(spec["uplo"], c_h, nb)
        if k == "herk":
            sl.herk(a.real, A, b.real, Cm, opa=spec["opa"])
        elif k == "syrk":
            sl.syrk(a, A, b, Cm, opa=spec["opa"])
        else:
            B = sl.Matrix(*x["B"].shape, nb, dtype=dt).from_numpy(x["B"])
            (sl.her2k if k == "her2k" else sl.syr2k)(a, A, B, b.real if k == "her2k" else b, Cm, opa=spec["opa"])
        return c_h
    if k in ("hemm", "symm"):
        cls = sl.HermitianMatrix if k == "hemm" else sl.SymmetricMatrix
        A = cls.fromLAPACK(spec["uplo"], x["A"].copy(order="F"), nb)
        B = sl.Matrix(*x["B"].shape, nb, dtype=dt).from_numpy(x["B"])
        Cm = sl.Matrix(*x["C"].shape, nb, dtype=dt).from_numpy(x["C"])
        sl.hemm(spec["side"], a, A, B, b, Cm)
        return Cm.to_numpy()
    if k == "trsm":
        A = sl.TriangularMatrix.fromLAPACK(spec["uplo"], "N", x["A"].copy(order="F"), nb)
        B = sl.Matrix(*x["B"].shape, nb, dtype=dt).from_numpy(x["B"])
        sl.trsm(spec["side"], a, A, B, opa=spec["opa"])
        return B.to_numpy()
    if k == "posv":
        a_h = x["A"].copy(order="F")
        A = sl.HermitianMatrix.fromLAPACK(spec["uplo"], a_h, nb)
        B = sl.Matrix(*x["B"].shape, nb, dtype=dt).from_numpy(x["B"])
        assert sl.posv(A, B) == 0
        return np.concatenate([a_h, B.to_numpy()], axis=1)
    if k == "gemmA":
        A = sl.Matrix(*x["A"].shape, nb, dtype=dt).from_numpy(x["A"])
        B = sl.Matrix(*x["B"].shape, nb, dtype=dt).from_numpy(x["B"])
        Cm = sl.Matrix(*x["C"].shape, nb, dtype=dt).from_numpy(x["C"])
        sl.gemm(a, A, B, b, Cm, opa=spec["opa"], opb=spec["opb"], method="A")
        return Cm.to_numpy()
    if k == "getrf_nopiv":
        A = sl.Matrix(*x["A"].shape, nb, dtype=dt).from_numpy(x["A"])
        assert sl.getrf_nopiv(A) == 0
        return A.to_numpy()
    raise AssertionError(k)


@pytest.mark.gpu
@pytest.mark.parametrize("name,spec", CASES, ids=[n for n, _ in CASES])
def test_b200_driver_matches_reference_output(sl, name, spec):
    """same inputs, the B200 driver against what the reference returned: held to the reference's own
    test tolerance (3 sqrt(k) eps for the updates, 50 eps for factors and solves); the parts
    the reference leaves alone (other triangle, strictly other part of a factor) must be equal"""
    x = wc.inputs(spec, lambda seed, m, n, dt: ol.matgen(seed, m, n, dt))
    res = run_b200(sl, spec, x)
    check_against(spec, x, res, False)
    gold, k = GOLD[name], spec["kind"]
    if k in ("hemm", "symm", "gemmA"):
        assert np.abs(res - gold).max() <= tolerance(spec, np.abs(gold).max())
    elif k in ("herk", "syrk", "her2k", "syr2k"):
        tri = np.tril if spec["uplo"] == "L" else np.triu
        assert np.abs(tri(res) - tri(gold)).max() <= tolerance(spec, np.abs(gold).max())
    elif k == "trsm":
        T = (np.tril if spec["uplo"] == "L" else np.triu)(x["A"])
        assert np.abs(res - gold).max() <= 50 * EPS * np.abs(gold).max() * np.linalg.cond(T, 1)
    elif k in ("posv", "getrf_nopiv"):
        # factors of a diagonally dominant matrix: well conditioned, compare entry by entry
        mask = np.ones(gold.shape, dtype=bool)
        if k == "posv":
            n = spec["n"]
            stored = np.tril(np.ones((n, n), dtype=bool)) if spec["uplo"] == "L" else np.triu(np.ones((n, n), dtype=bool))
            mask[:, :n] = stored
            assert np.array_equal(res[:, :n][~stored], gold[:, :n][~stored])
        assert np.abs(res[mask] - gold[mask]).max() <= 50 * EPS * np.abs(gold[mask]).max()
