"""Regenerates tests/golden/scorer_kat.json (run from the repo root: python tests/golden/make_golden.py).

The reference cannot be imported in this container (gensim / aioredis / nltk missing, no
weights file), so these are NOT outputs of the reference: they are known-answer cases
authored from the semantics of src/backend.py:303-317 (SURVEY.md §8c lists them), with the
numeric columns produced by oracle/scorer_oracle.py -- the restatement that calls the same
scipy BLAS routines gensim binds.  The file pins the oracle against accidental change and
gives the CUDA tests fixed inputs; it does not pin the oracle to the reference
("parity unpinned").

Two sections:
  kat    : a 12-row hand-built table (dim 300) and string pairs whose results are known in
           closed form -- identical words, case folding, OOV on either side, orthogonal,
           anti-correlated, zero-norm, 3-4-5 triangle, quarter-overlap (cos 0.5).
  synth  : 96 (guess rows, answer row) cases on the seeded synthetic table
           (cassmantle_b200/synth.py, seed 0, 50_000 rows), N in {1, 4, 16}, with the
           oracle's fp32 result stored as hex so no decimal rounding is involved.
"""
from __future__ import annotations

import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from cassmantle_b200 import synth            # noqa: E402
from oracle import scorer_oracle as O        # noqa: E402

DIM = 300


def kat_table():
    """name -> sparse row description {index: value} or a ('fill', value, count) tuple."""
    rows = {
        "alpha": {0: 2.0},
        "beta": {1: 3.0},                         # orthogonal to alpha
        "gamma": {0: -1.0},                       # anti-parallel to alpha
        "delta": {0: 3.0, 1: 4.0},                # cos(alpha, delta) = 0.6
        "void": {},                               # zero norm: unitvec returns it unchanged
        "ones": ("fill", 1.0, 300),
        "quarter": ("fill", 1.0, 75),             # cos(ones, quarter) = 75/sqrt(300*75) = 0.5
        "twin1": ("fill", 0.37, 300),
        "twin2": ("fill", 0.37, 300),             # same direction as twin1 and ones
        "tiny": {5: 1e-20, 6: 1e-20},             # needs the fp64 norm: fp32 squares underflow
        "huge": {5: 1e20, 6: 1e20},               # fp32 squares overflow; cos(tiny, huge) = 1
        "mixed": {0: 1.0, 1: -2.0, 2: 0.5, 299: 4.0},
    }
    return rows


def dense(desc) -> np.ndarray:
    v = np.zeros(DIM, dtype=np.float32)
    if isinstance(desc, tuple):
        v[: desc[2]] = np.float32(desc[1])
    else:
        for i, x in desc.items():
            v[int(i)] = np.float32(x)
    return v


def main() -> None:
    rows = kat_table()
    keys = list(rows.keys())
    table = np.stack([dense(rows[k]) for k in keys])
    be = O.OracleBackend(O.OracleWordVectors(table, keys), min_score=0.01)

    pairs = [
        ("alpha", "alpha"), ("Alpha", "ALPHA"), ("zzz", "ZZZ"),        # equality before lookup
        ("alpha", "beta"), ("alpha", "gamma"), ("alpha", "delta"), ("delta", "alpha"),
        ("alpha", "void"), ("void", "alpha"), ("void", "beta"),
        ("ones", "quarter"), ("twin1", "twin2"), ("twin1", "ones"), ("tiny", "huge"),
        ("mixed", "alpha"), ("mixed", "delta"), ("mixed", "beta"), ("mixed", "ones"),
        ("nosuchword", "alpha"), ("alpha", "nosuchword"), ("Beta", "dELTA"), ("", "alpha"),
    ]
    cases = []
    for g, a in pairs:
        s = be.compute_score(g, a)
        cases.append({
            "input": g, "answer": a,
            "str": str(s),
            "f32_hex": float(np.float32(s)).hex(),
            "is_python_float": not isinstance(s, np.floating),
            "display": O.placeholder(str(s)),
            "centi": O.centi_score(str(s)),
        })

    # ---- seeded synthetic section ------------------------------------------------------
    V = 50_000
    T = synth.full_table(V, seed=0)
    syn = []
    for n_words in (1, 4, 16):
        idx, ln, ans, slot, ex = synth.make_guess_batch(32, n_words, V, n_answers=3,
                                                        seed=100 + n_words, ragged=n_words > 1)
        out = O.score_batch_faithful(T, idx, ln, ans, slot, ex, 0.01)
        for b in range(32):
            syn.append({"rows": idx[b].tolist(), "len": int(ln[b]),
                        "answer_row": int(ans[slot[b]]), "exact": int(ex[b]),
                        "f32_hex": float(out[b]).hex()})

    fmt = [{"f32_hex": float(np.float32(x)).hex(), "str": str(max(0.01, np.float32(x))),
            "display": O.placeholder(str(max(0.01, np.float32(x)))),
            "centi": O.centi_score(str(max(0.01, np.float32(x))))}
           for x in (0.4567891, 0.125, 0.1, 0.10000001, 0.99999994, 1.0, 0.5, 0.012345678,
                     -0.25, 0.0, 0.335, 0.045, 0.7071068, 0.33333334)]

    doc = {
        "provenance": "authored in this repo from src/backend.py:303-317 + gensim's published "
                      "similarity/unitvec algorithm; numeric values by oracle/scorer_oracle.py; "
                      "NOT reference outputs (reference not importable here) -- parity unpinned",
        "min_score": 0.01,
        "dim": DIM,
        "kat": {"keys": keys,
                "rows": {k: (list(v) if isinstance(v, tuple) else {str(i): x for i, x in v.items()})
                         for k, v in rows.items()},
                "cases": cases},
        "synth": {"table_seed": 0, "n_rows": V, "cases": syn},
        "format": fmt,
    }
    out_path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "scorer_kat.json")
    with open(out_path, "w") as f:
        json.dump(doc, f, indent=1)
    print("wrote", out_path, len(cases), "kat cases,", len(syn), "synthetic cases")


if __name__ == "__main__":
    main()
