"""CPU tests of mask selection: the oracle and this repo's host logic against
tests/golden/masks_golden.json -- outputs of the reference's own select_descriptive_words /
semantic_distance (src/utils.py:73-104), run with tests/golden/toy_nltk.py in nltk's place."""
import json
import os
import sys

import numpy as np
import pytest

from cassmantle_b200 import masks, synth
from cassmantle_b200.vocab import Vocabulary
from oracle import masks_oracle as M

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import toy_nltk                                     # noqa: E402


@pytest.fixture(scope="module")
def masks_golden():
    with open(os.path.join(HERE, "golden", "masks_golden.json")) as f:
        g = json.load(f)
    keys = synth.make_vocab(g["vocab_rows"], g["seed"])
    table = synth.table_rows(np.arange(g["vocab_rows"]), g["seed"])
    return g, keys, table


class ToyModel:
    def __init__(self, keys, table):
        self.key_to_index = {k: i for i, k in enumerate(keys)}
        self.table = table

    def __contains__(self, key):
        return key in self.key_to_index

    def __getitem__(self, key):
        return self.table[self.key_to_index[key]]


class OracleMaskModel(masks.MaskModel):
    """The product's MaskModel with the device call replaced by the numpy restatement (no GPU here)."""

    def __init__(self, keys, table):
        self.vocab, self.table = Vocabulary(keys), table

    def semantic_distances(self, word_list, extra=()):
        rows = self.vocab.lookup_many(list(word_list) + list(extra))
        return M.distances_from_rows(self.table, rows, len(word_list))


def _hex_to_f32(h):
    return np.frombuffer(bytes.fromhex(h), dtype=np.float32)[0]


def test_oracle_reproduces_the_reference_outputs(masks_golden):
    g, keys, table = masks_golden
    model = ToyModel(keys, table)
    assert len(g["cases"]) >= 25
    for case in g["cases"]:
        assert toy_nltk.word_tokenize(case["sentence"]) == case["words"]
        assert [list(t) for t in toy_nltk.pos_tag(case["words"])] == case["tagged"]
        want = np.array([_hex_to_f32(h) for h in case["distances_hex"]], dtype=np.float32)
        got = np.array([M.semantic_distance(w, case["filtered"], model) for w in case["filtered"]], dtype=np.float32)
        assert np.allclose(got, want, rtol=2e-6, atol=0)           # BLAS sdot may differ in the last bit across CPUs
        rows = [model.key_to_index.get(w, -1) for w in case["filtered"]]
        assert np.allclose(M.distances_from_rows(table, rows), want, rtol=2e-6, atol=0)
        for k, masks_want in case["masks"].items():
            words, got_masks = M.select_descriptive_words(model, case["sentence"], int(k), toy_nltk.word_tokenize,
                                                          toy_nltk.pos_tag)
            assert words == case["words"] and got_masks == masks_want


def test_host_logic_of_the_mirror_reproduces_the_reference_outputs(masks_golden):
    g, keys, table = masks_golden
    model = OracleMaskModel(keys, table)
    for case in g["cases"]:
        for k, masks_want in case["masks"].items():
            words, got = masks.select_descriptive_words(model, case["sentence"], int(k), toy_nltk.word_tokenize,
                                                        toy_nltk.pos_tag)
            assert words == case["words"] and got == masks_want
        d = masks.construct_prompt_dict(model, case["sentence"], 2, toy_nltk.word_tokenize, toy_nltk.pos_tag)
        assert d == {"tokens": case["words"], "masks": case["masks"]["2"]}
    # one word at a time, as src/utils.py:91 calls it; a word outside the model is 0
    case = g["cases"][5]
    want = [_hex_to_f32(h) for h in case["distances_hex"]]
    for w, x in zip(case["filtered"], want):
        assert masks.semantic_distance(w, case["filtered"], model) == pytest.approx(float(x), rel=2e-6)
    assert masks.semantic_distance("no-such-word", case["filtered"], model) == 0


def test_sentence_without_a_two_letter_token_raises_like_scikit_learn(masks_golden):
    g, keys, table = masks_golden
    for fn, model in ((M.select_descriptive_words, ToyModel(keys, table)),
                      (masks.select_descriptive_words, OracleMaskModel(keys, table))):
        with pytest.raises(ValueError, match="empty vocabulary"):
            fn(model, "a b , c !", 2, toy_nltk.word_tokenize, toy_nltk.pos_tag)


def test_nltk_is_only_needed_without_injected_callables(masks_golden):
    g, keys, table = masks_golden
    model = OracleMaskModel(keys, table)
    try:
        import nltk                                   # noqa: F401
    except ImportError:
        with pytest.raises(ImportError, match="tokenize="):
            masks.select_descriptive_words(model, "some sentence", 2)
