"""CPU test: the reference's OWN mask-selection code, run live, against the oracle restatement.

Runs only where /root/reference is mounted (the build container; skipped on the GPU box).
`src/utils.py` is imported unmodified with tests/golden/toy_nltk.py standing in for nltk -- the
one module it needs that this box lacks; numpy and scikit-learn are the real ones.  Fresh random
sentences every run (seeded), so this goes beyond the committed golden file."""
import os
import sys
import types

import numpy as np
import pytest

from cassmantle_b200 import synth
from oracle import masks_oracle as M
from tests.test_masks_oracle import ToyModel

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import toy_nltk                                     # noqa: E402

REF = "/root/reference"
pytestmark = pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "src")), reason="reference tree not mounted")


@pytest.fixture()
def reference_utils(monkeypatch):
    nltk = types.ModuleType("nltk")
    nltk.word_tokenize, nltk.pos_tag = toy_nltk.word_tokenize, toy_nltk.pos_tag
    monkeypatch.setitem(sys.modules, "nltk", nltk)
    monkeypatch.syspath_prepend(REF)
    for m in [k for k in sys.modules if k == "src" or k.startswith("src.")]:
        monkeypatch.delitem(sys.modules, m)
    from src import utils                             # the reference's file, untouched
    yield utils
    for m in [k for k in sys.modules if k == "src" or k.startswith("src.")]:
        sys.modules.pop(m, None)


def test_oracle_equals_the_reference_on_fresh_sentences(reference_utils):
    keys = synth.make_vocab(800, 3)
    table = synth.table_rows(np.arange(800), 3)
    model = ToyModel(keys, table)
    rng = np.random.default_rng(77)
    for trial in range(60):
        n = int(rng.integers(2, 30))
        words = [keys[int(i)] for i in rng.integers(0, 800, n)]
        if trial % 3 == 0:
            words[0] = words[0].upper()               # out of vocabulary (case-sensitive keys)
        if trial % 4 == 0 and n > 4:
            words[2] = words[1]
        sentence = " ".join(words) + ("!" if trial % 2 else ".")
        for k in (1, 2, 5):
            want = reference_utils.select_descriptive_words(model, sentence, k)
            got = M.select_descriptive_words(model, sentence, k, toy_nltk.word_tokenize, toy_nltk.pos_tag)
            assert got == want, (trial, k)
        filt = [w for w, p in toy_nltk.pos_tag(toy_nltk.word_tokenize(sentence)) if w.isalpha() and p in M.DESCRIPTIVE_TAGS]
        for w in filt:
            a = reference_utils.semantic_distance(w, filt, model)
            b = M.semantic_distance(w, filt, model)
            assert np.float32(a).tobytes() == np.float32(b).tobytes()      # same numpy calls, same machine
        assert reference_utils.construct_prompt_dict(model, sentence, 2) == \
            M.construct_prompt_dict(model, sentence, 2, toy_nltk.word_tokenize, toy_nltk.pos_tag)
    with pytest.raises(ValueError, match="empty vocabulary"):
        reference_utils.select_descriptive_words(model, "a b , c !", 2)
