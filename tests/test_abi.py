"""CPU tests: the C-ABI library loads and exports exactly what include/cassmantle_b200.h declares."""
import ctypes
import subprocess

import pytest

from cassmantle_b200 import _lib, build


@pytest.fixture(scope="module")
def lib():
    build.build()
    return _lib.load()


def test_header_symbols_are_exported(lib):
    names = _lib.header_symbols()
    assert len(names) >= 15
    for n in names:
        assert hasattr(lib, n), f"{n} declared in the header but not exported"
    assert set(names) == set(_lib._SIGNATURES), "ctypes signatures and header disagree"


def test_exports_are_c_linkage():
    out = subprocess.check_output(["nm", "-D", "--defined-only", _lib.LIB_PATH], text=True)
    exported = {line.split()[-1] for line in out.splitlines() if " T " in line}
    for n in _lib.header_symbols():
        assert n in exported


def test_library_is_sm100a_only():
    out = subprocess.run(["cuobjdump", "-lelf", _lib.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in out
    assert all("sm_100a" in line for line in out.splitlines() if "sm_" in line)


def test_version_and_errors_without_compute(lib):
    assert lib.cm_version().decode().count(".") == 2
    assert lib.cm_engine_destroy(None) == 0
    h = ctypes.c_void_p()
    assert lib.cm_engine_create(ctypes.byref(h), 0, 10, 301) == -1          # dim % 4
    assert b"multiple of 4" in lib.cm_last_error()
    assert lib.cm_engine_create(ctypes.byref(h), 0, 0, 300) == -1           # n_rows
    assert lib.cm_engine_create(None, 0, 10, 300) == -1
    assert lib.cm_score_batch_host(None, None, None, 1, 1, None, None, 0.01, None, None) == -1
    assert lib.cm_kernel_launch_count() >= 0


def test_no_cpu_fallback_without_gpu(lib):
    """On a box without a GPU, engine creation must fail loudly (no silent CPU path)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from cassmantle_b200.engine import ScoringEngine
    with pytest.raises(_lib.CassMantleError):
        ScoringEngine(100, 300)


def test_product_package_never_touches_the_oracle():
    """oracle/ is test infrastructure: nothing under cassmantle_b200/ may import, load or run it,
    and there is no CPU fallback behind the CUDA extension."""
    import ast
    import glob
    import os
    pkg = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "cassmantle_b200")
    files = glob.glob(os.path.join(pkg, "**", "*.py"), recursive=True)
    assert len(files) >= 8
    for path in files:
        src = open(path).read()
        for node in ast.walk(ast.parse(src)):
            names = []
            if isinstance(node, ast.Import):
                names = [a.name for a in node.names]
            elif isinstance(node, ast.ImportFrom):
                names = [node.module or ""]
            assert not any(n == "oracle" or n.startswith("oracle.") for n in names), path
        assert "liboracle" not in src and "oracle/" not in src.replace("oracle/ is", ""), path
    for path in glob.glob(os.path.join(pkg, "csrc", "*")):
        text = open(path, errors="ignore").read()
        assert "oracle/" not in text and "liboracle" not in text and "dlopen" not in text, path
