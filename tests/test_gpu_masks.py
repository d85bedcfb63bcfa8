"""GPU parity tests of mask selection (cm_semantic_distances + cassmantle_b200/masks.py) against
the reference's own outputs (tests/golden/masks_golden.json) and the numpy restatement."""
import json
import os
import sys

import numpy as np
import pytest

from cassmantle_b200 import masks, synth
from cassmantle_b200._lib import CassMantleError, check
from cassmantle_b200.vocab import Vocabulary
from oracle import masks_oracle as M

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import toy_nltk                                     # noqa: E402

REL_TOL = 2e-6          # fp32 sqrt(dot) of ~300 terms, summed in a different order than BLAS sdot


@pytest.fixture(scope="module")
def mask_stack():
    from cassmantle_b200.engine import ScoringEngine
    with open(os.path.join(HERE, "golden", "masks_golden.json")) as f:
        g = json.load(f)
    keys = synth.make_vocab(g["vocab_rows"], g["seed"])
    table = synth.table_rows(np.arange(g["vocab_rows"]), g["seed"])
    eng = ScoringEngine.from_vectors(table)
    yield g, keys, table, eng, masks.MaskModel(eng, Vocabulary(keys))
    eng.close()


def _hex_to_f32(h):
    return np.frombuffer(bytes.fromhex(h), dtype=np.float32)[0]


def test_distances_and_selection_equal_the_reference_outputs(mask_stack):
    g, keys, table, eng, model = mask_stack
    for case in g["cases"]:
        want = np.array([_hex_to_f32(h) for h in case["distances_hex"]], dtype=np.float32)
        got = model.semantic_distances(case["filtered"])
        assert got.dtype == np.float32 and np.allclose(got, want, rtol=REL_TOL, atol=0)
        assert np.all(got[want == 0] == 0)                             # words outside the model: exactly 0
        for k, masks_want in case["masks"].items():
            words, got_masks = masks.select_descriptive_words(model, case["sentence"], int(k),
                                                              toy_nltk.word_tokenize, toy_nltk.pos_tag)
            assert words == case["words"] and got_masks == masks_want
        assert masks.construct_prompt_dict(model, case["sentence"], 2, toy_nltk.word_tokenize, toy_nltk.pos_tag) == \
            {"tokens": case["words"], "masks": case["masks"]["2"]}
    case = g["cases"][7]
    for w, h in zip(case["filtered"], case["distances_hex"]):
        assert float(masks.semantic_distance(w, case["filtered"], model)) == pytest.approx(float(_hex_to_f32(h)), rel=REL_TOL)


def test_entry_point_against_the_numpy_restatement(mask_stack):
    g, keys, table, eng, model = mask_stack
    rng = np.random.default_rng(9)
    lib = eng._lib
    for n, n_mean in ((1, 1), (2, 2), (7, 7), (64, 64), (300, 300), (4096, 4096), (10, 4), (33, 0), (5, 1)):
        rows = rng.integers(0, len(keys), n).astype(np.int32)
        rows[rng.random(n) < 0.2] = -1
        if n == 7:
            rows[:] = -1                                               # nothing in the model at all
        out = np.full(n, -5.0, dtype=np.float32)
        check(lib.cm_semantic_distances(eng._h, rows.ctypes.data, n, n_mean, out.ctypes.data))
        want = M.distances_from_rows(table, rows, n_mean)
        assert np.allclose(out, want, rtol=REL_TOL, atol=1e-6), (n, n_mean)
        assert np.all(out[rows < 0] == 0)
    out = np.zeros(4, np.float32)
    rows = np.array([0, 1, len(keys), 2], np.int32)                    # a row beyond the table = not in the model
    check(lib.cm_semantic_distances(eng._h, rows.ctypes.data, 4, 4, out.ctypes.data))
    assert out[2] == 0 and np.allclose(out, M.distances_from_rows(table, [0, 1, -1, 2]), rtol=REL_TOL)
    assert lib.cm_semantic_distances(eng._h, rows.ctypes.data, 0, 0, out.ctypes.data) == 0     # empty list: no-op
    for bad in ((4097, 1), (4, 5), (-1, 0)):
        with pytest.raises(CassMantleError):
            check(lib.cm_semantic_distances(eng._h, rows.ctypes.data, bad[0], bad[1], out.ctypes.data))
    with pytest.raises(CassMantleError):
        check(lib.cm_semantic_distances(eng._h, None, 4, 4, out.ctypes.data))
