"""CPU checks of the drop-in boundary: the C-ABI library loads without a GPU and exports
every symbol include/tile_engine.h declares; the host mirror loads on top of it; host-side
logic that needs no device (grid maps, options, exceptions) behaves like the reference."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "tile_engine.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(te_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from slate_d35_b200 import lib
    L = lib()
    names = declared_symbols()
    assert len(names) >= 50
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing


def test_no_gpu_calls_fail_cleanly():
    from slate_d35_b200 import lib
    L = lib()
    assert L.te_version().startswith(b"tile_engine")
    if L.te_device_count() == 0:
        # argument errors are reported before any CUDA call
        assert L.te_gemm_batched(0, None, 0, ord("N"), ord("X"), None, None, None, 0, 0, 0) == 1


def test_host_mirror_loads_and_refuses_cpu_targets():
    from slate_d35_b200 import host, lib
    H = host.hostlib()
    assert H.sb_last_error() is not None
    if lib().te_device_count() == 0:
        # without a device the product path must fail loudly, never fall back to the CPU
        host.initialize(0)
        with pytest.raises(host.SlateError):
            A = host.Matrix(64, 64, 32)            # device-origin tiles need the pool
            host.gemm(1.0, A, A, 0.0, A)


def test_product_never_links_the_oracle():
    """the shipped libraries must not depend on oracle/ (checked on the ELF NEEDED list)."""
    import subprocess
    for so in ("libtile_engine.so", "libslate_b200.so", "libslate_b200_scalapack.so"):
        out = subprocess.run(["readelf", "-d", os.path.join(ROOT, "slate_d35_b200", "lib", so)],
                             capture_output=True, text=True).stdout
        assert "oracle" not in out and "slate_ref" not in out
    for dirpath, _, files in os.walk(os.path.join(ROOT, "slate_d35_b200")):
        for f in files:
            if f.endswith((".py", ".cc", ".cu", ".hh", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "oracle_lib" not in src and "liboracle" not in src, f


def scalapack_declared():
    text = open(os.path.join(ROOT, "include", "slate_b200_scalapack.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(slate_[a-z0-9_]+)\s*\(", text)))


def test_scalapack_entry_points_are_exported():
    """include/slate_b200_scalapack.h against libslate_b200.so, and the Fortran spellings
    (pdgemm_, pdgemm, PDGEMM, ...) against libslate_b200_scalapack.so."""
    from slate_d35_b200 import host
    H = host.hostlib()
    names = scalapack_declared()
    assert len(names) >= 16
    assert not [n for n in names if not hasattr(H, n)]
    F = C.CDLL(os.path.join(ROOT, "slate_d35_b200", "lib", "libslate_b200_scalapack.so"))
    for n in names:
        if n.startswith("slate_p"):
            base = n[len("slate_"):]
            for spelling in (base, base + "_", base.upper()):
                assert hasattr(F, spelling), spelling


def test_scalapack_argument_errors_need_no_gpu():
    import numpy as np
    from slate_d35_b200 import scalapack as sc
    with pytest.raises(ValueError):
        sc.grid_set(0, "X", 1, 1)
    with pytest.raises(ValueError):
        sc.grid_set(0, "C", 0, 1)
    a = np.zeros((64, 64), order="F")
    d = sc.descinit(64, 64, 32, 32, 4242, 64)         # context nobody registered
    with pytest.raises(sc.ScalapackError, match="no grid registered"):
        sc.pgemm("N", "N", 64, 64, 64, 1.0, a, 1, 1, d, a, 1, 1, d, 0.0, a, 1, 1, d)
    # 15 full blocks of 64 over 3 processes, the 40 left over follow the last full block
    assert [sc.numroc(1000, 64, r, 3) for r in range(3)] == [360, 320, 320]
    assert [sc.numroc(1100, 64, r, 3) for r in range(3)] == [384, 384, 332]


def test_process_grid_orders():
    """func::process_2d_grid (include/slate/func.hh:100-118): BLACS column-major and row-major
    rank layouts, as slate_scalapack_grid_set / fromScaLAPACK use them."""
    from slate_d35_b200 import host
    H = host.hostlib()
    H.sb_grid_tile_rank.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int64, C.c_int64]
    for p, q in ((1, 1), (2, 4), (4, 2), (3, 5)):
        for i in range(9):
            for j in range(11):
                assert H.sb_grid_tile_rank(ord("C"), p, q, i, j) == (i % p) + (j % q) * p
                assert H.sb_grid_tile_rank(ord("R"), p, q, i, j) == (i % p) * q + (j % q)


def test_scalapack_style_c_program_links_and_fails_loudly_without_gpu(tmp_path):
    """examples/example_pdgemm_pdpotrf.c: a plain C caller of pdgemm_ / pdpotrf_ compiles against
    include/slate_b200_scalapack.h (valid C, not only C++) and links against the three shared
    libraries.  Without a GPU the calls report an error and leave the output alone -- there is
    no CPU path to fall back to; with one the program factors its matrix."""
    import subprocess
    from slate_d35_b200 import lib
    libdir = os.path.join(ROOT, "slate_d35_b200", "lib")
    exe = str(tmp_path / "example_pd")
    subprocess.check_call(["gcc", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "examples", "example_pdgemm_pdpotrf.c"), "-L" + libdir,
                           "-lslate_b200_scalapack", "-lslate_b200", "-ltile_engine",
                           "-Wl,-rpath," + libdir, "-lm", "-o", exe])
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    if lib().te_device_count() == 0:
        assert r.returncode != 0 and "status = 1" in r.stdout
        assert "pdgemm" in r.stderr and "pdpotrf" in r.stderr
    else:
        assert r.returncode == 0 and "info = 0" in r.stdout and "status = 0" in r.stdout
