"""GPU test: a plain-C program (no Python, no torch) drives the C ABI end to end."""
import os
import subprocess

import pytest

from cassmantle_b200 import _lib, build

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_plain_c_program_uses_the_library(tmp_path):
    build.build()
    exe = str(tmp_path / "abi_smoke")
    subprocess.check_call(["gcc", "-std=c11", "-Wall", "-I", os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "c", "abi_smoke.c"), "-o", exe, _lib.LIB_PATH, "-lm",
                           "-Wl,-rpath," + os.path.dirname(_lib.LIB_PATH)])
    out = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "abi_smoke ok" in out.stdout
