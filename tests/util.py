"""Helpers shared by the parity tests (test infrastructure)."""
from __future__ import annotations

import numpy as np

from cassmantle_b200 import synth

# north_star allows |cos - reference| <= 1e-3; the kernel follows gensim's operation order
# (fp64 norm, fp32 unit vectors, fp32 dot) so only the dot's summation order differs and the
# tests hold it to this much tighter bound.
COS_TOL = 2e-6


def dense_kat_table(golden) -> np.ndarray:
    dim = golden["dim"]
    rows = []
    for k in golden["kat"]["keys"]:
        desc = golden["kat"]["rows"][k]
        v = np.zeros(dim, dtype=np.float32)
        if isinstance(desc, list):
            v[: desc[2]] = np.float32(desc[1])
        else:
            for i, x in desc.items():
                v[int(i)] = np.float32(x)
        rows.append(v)
    return np.stack(rows)


def compact_problem(idx: np.ndarray, answers: np.ndarray, seed: int = 0):
    """Materialise only the table rows a batch touches; remap indices onto that small table."""
    used = np.unique(np.concatenate([idx[idx >= 0].ravel(), answers[answers >= 0].ravel()]))
    table = synth.table_rows(used, seed)
    remap_idx = np.where(idx >= 0, np.searchsorted(used, np.maximum(idx, 0)), -1).astype(np.int32)
    remap_ans = np.where(answers >= 0, np.searchsorted(used, np.maximum(answers, 0)), -1).astype(np.int32)
    return table, remap_idx, remap_ans


def assert_scores_match(got: np.ndarray, want: np.ndarray, min_score: float, tol: float = COS_TOL):
    """Exact classes bit-exact (1.0 and the floor), cosines within tol."""
    got = np.asarray(got, dtype=np.float32)
    want = np.asarray(want, dtype=np.float32)
    floor = np.float32(min_score)
    # a cosine within tol of the floor may legitimately land on either side of it
    near_floor = np.abs(want - floor) <= tol
    is_floor_w = (want == floor) & ~near_floor
    is_floor_g = (got == floor) & ~near_floor
    assert np.array_equal(is_floor_w, is_floor_g), "floor (min_score) classification differs"
    diff = np.abs(got.astype(np.float64) - want.astype(np.float64))
    assert diff.max(initial=0.0) <= tol, f"max |cos diff| {diff.max()} > {tol}"
