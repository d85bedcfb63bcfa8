"""The error bound behind cm_most_similar_batch's candidate filter (cassmantle_b200/csrc/cm_similar_batch.cu: kBand).

The tensor-core pass multiplies tf32 operands (fp32 with the low 13 mantissa bits dropped) and accumulates in fp32;
its cosines only FILTER candidates, which are then re-scored exactly.  The filter is safe if a tf32 cosine is never
further than eps from the exact one, with 2 * eps <= kBand.  This test restates the arithmetic in numpy and checks the
bound on random, adversarial (all low mantissa bits set, all products of one sign) and synthetic-table vectors.
"""
import re
from pathlib import Path

import numpy as np

from cassmantle_b200 import synth


def tf32(x: np.ndarray) -> np.ndarray:
    """fp32 -> tf32 by truncation: what the tensor core sees of an fp32 operand."""
    return (x.astype(np.float32).view(np.uint32) & np.uint32(0xFFFFE000)).view(np.float32)


def band() -> float:
    src = Path(__file__).resolve().parents[1] / "cassmantle_b200" / "csrc" / "cm_similar_batch.cu"
    return float(re.search(r"constexpr float kBand = ([0-9.e+-]+)f;", src.read_text()).group(1))


def worst_error(rows: np.ndarray, query: np.ndarray) -> float:
    q = (query / np.linalg.norm(query.astype(np.float64))).astype(np.float32)
    norms = np.linalg.norm(rows.astype(np.float64), axis=1)
    exact = rows.astype(np.float64) @ q.astype(np.float64) / norms
    # fp32 accumulation, term by term, in the worst order we can cheaply produce (sequential)
    prod = (tf32(rows).astype(np.float32) * tf32(q)[None, :]).astype(np.float32)
    approx = np.cumsum(prod, axis=1, dtype=np.float32)[:, -1].astype(np.float64) / norms
    return float(np.abs(approx - exact).max())


def test_band_covers_twice_the_tf32_error():
    eps_budget = band() / 2
    assert eps_budget >= 2.0 ** -9 * (1 + 2.0 ** -11) + 300 * 2.0 ** -24      # the analytic bound, with the fp32 sum
    rng = np.random.default_rng(0)
    worst = 0.0
    # random gaussian rows and queries
    rows = rng.standard_normal((4000, 300)).astype(np.float32)
    worst = max(worst, worst_error(rows, rng.standard_normal(300).astype(np.float32)))
    # the synthetic table, a row against its own cluster
    table = synth.table_rows(np.arange(5000), seed=0)
    worst = max(worst, worst_error(table, table[17]))
    # adversarial: every mantissa has its 13 low bits set, every product is positive, the row is parallel to the query
    base = np.abs(rng.standard_normal(300)).astype(np.float32) + 0.5
    adv = (base.view(np.uint32) | np.uint32(0x1FFF)).view(np.float32)
    worst = max(worst, worst_error(np.stack([adv, adv * np.float32(3.0), adv[::-1].copy()]), adv))
    assert worst <= eps_budget, (worst, eps_budget)
    assert worst > 1e-4            # the adversarial case really exercises the truncation
