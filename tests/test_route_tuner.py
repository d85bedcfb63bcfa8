"""CPU test: the pipe's route tuner (cassmantle_b200/csrc/cm_route_tuner.h) is plain C++, so its
state machine is exercised here with a fake clock -- start-up, trial windows, the 3 % hysteresis,
pinned routes, changing conditions -- without a GPU.  tests/test_gpu_parity.py checks on the device
that results do not depend on the route it picks."""
import os
import shutil
import subprocess

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


@pytest.mark.skipif(shutil.which("g++") is None, reason="g++ not available")
def test_route_tuner_state_machine(tmp_path):
    exe = str(tmp_path / "route_tuner_test")
    subprocess.run(["g++", "-std=c++17", "-O1", "-Wall", "-Werror", "-I", os.path.join(ROOT, "cassmantle_b200", "csrc"),
                    os.path.join(HERE, "c", "route_tuner_test.cpp"), "-o", exe], check=True)
    out = subprocess.run([exe], capture_output=True, text=True)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "all checks passed" in out.stdout
