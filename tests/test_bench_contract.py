"""CPU test of bench.py's JSON contract on the arm that needs no GPU (--impl reference), and a GPU
test of the own arm's line.  The driver parses exactly one JSON line from stdout."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BASE_KEYS = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
             "vs_baseline", "dtype", "data", "config", "e2e", "cpu_baseline"}


def _run(args, timeout):
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py")] + args, cwd=ROOT, capture_output=True,
                         text=True, timeout=timeout)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [ln for ln in out.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1, lines                           # ONE JSON line, nothing else on stdout
    return json.loads(lines[0])


def test_reference_arm_line():
    d = _run(["--impl", "reference", "--steps", "2", "--warmup", "1", "--cpu-workers", "2", "--ref-budget-s", "4"], 300)   # W < 3 is raised to 3
    assert BASE_KEYS <= set(d) and d["impl"] == "reference"
    assert d["metric"] == "guesses_scored_per_sec" and d["unit"] == "guesses/s" and d["higher_is_better"] is True
    assert d["n_gpus"] == 1 and d["steps"] == 2 and d["warmup"] == 3 and d["value"] > 0
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] == 2 and cb["value"] == d["value"] and cb["sample"]
    assert "workload" in d["config"] and "model" not in d["config"] and d["vs_baseline"] is None


@pytest.mark.gpu
def test_own_arm_line():
    d = _run(["--steps", "60", "--warmup", "5", "--cpu-budget-s", "2", "--roofline-s", "0.2", "--no-other-shapes"], 900)
    assert BASE_KEYS | {"gpu_launches", "roofline", "clocks"} <= set(d) and "impl" not in d
    assert d["metric"] == "guesses_scored_per_sec" and d["steps"] == 60 and d["warmup"] == 5
    assert d["gpu_launches"] == 60 and d["dtype"] == "f32" and d["scaling"] == "weak" and d["vs_baseline"] is None
    r = d["roofline"]
    assert r["bound"] == "hbm" and r["unit"] == "GB/s" and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9
    assert r["traffic"] and r["traffic_source"] and 0.3 < r["frac"] < 1.3
    # the fraction is on the bytes the timed batches require (rows of live guesses only); SURVEY's
    # formula, which charges every guess all its rows, is reported beside it and is larger
    assert abs(r["achieved"] - r["algorithmic_bytes_per_launch"] / (r["kernel_us"] * 1e-6) / 1e9) < 1e-6 * r["achieved"]
    assert 0.92 < r["live_row_share"] < 0.96 and r["frac_survey_formula"] > r["frac"]
    assert r["algorithmic_bytes_per_launch"] < r["survey_formula_bytes_per_launch"]
    assert d["sustained"]["steps"] >= 60 and d["sustained"]["value"] > 0
    e = d["e2e"]
    assert e["h2d_bytes_per_step"] > 0 and e["d2h_bytes_per_step"] > 0 and e["value"] != d["value"] and e["value"] > 0
    assert e["sustained"]["steps"] >= 60 and e["one_step_per_call"]["value"] > 0
    cb = d["cpu_baseline"]
    assert {"value", "unit", "cores", "kind", "sample"} <= set(cb) and cb["kind"] == "port"
    assert {"sm_mhz", "sm_max_mhz", "reasons"} <= set(d["clocks"])
    assert d["parity"]["max_abs_diff_vs_oracle"] <= 2e-6
