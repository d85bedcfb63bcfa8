"""Regenerates tests/golden/masks_golden.json (run from the repo root, in the build container:
python tests/golden/make_masks_golden.py).

Outputs of the REFERENCE ITSELF: `select_descriptive_words` and `semantic_distance`
(/root/reference/src/utils.py:73-104) are imported unmodified and run here.  The one module the
reference needs and this box lacks, nltk, is replaced by tests/golden/toy_nltk.py (a deterministic
tokenizer and tagger -- see its header for why that does not weaken the check); numpy and
scikit-learn's TfidfVectorizer are the real ones.  The word vectors are rows of the seeded
synthetic table (cassmantle_b200/synth.py, seed 7, 500 rows) under the keys of
synth.make_vocab(500, seed 7), so the GPU tests can rebuild the same model on the box, where
/root/reference does not exist.

Stored per sentence: the tokens, the (word, tag) pairs, the filtered words, the reference's
fp32 semantic distance of each filtered word (hex, no decimal rounding), and the reference's
selected mask indices for num_words in {1, 2, 3}.
"""
from __future__ import annotations

import json
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

import toy_nltk                                   # noqa: E402
from cassmantle_b200 import synth                 # noqa: E402

V, SEED = 500, 7


class ToyModel:
    """What `model` is to src/utils.py: `word in model`, `model[word]` (gensim KeyedVectors)."""

    def __init__(self, keys, vectors):
        self.key_to_index = {k: i for i, k in enumerate(keys)}
        self.vectors = vectors

    def __contains__(self, key):
        return key in self.key_to_index

    def __getitem__(self, key):
        return self.vectors[self.key_to_index[key]]


def sentences(keys, rng):
    out = []
    for n_words in (3, 5, 8, 12, 20, 40):
        for _ in range(4):
            words = [keys[int(i)] for i in rng.integers(0, len(keys), n_words)]
            for j in range(0, n_words, 5):
                words[j] = words[j].capitalize()              # case-sensitive keys: now out of vocabulary
            if n_words >= 8:
                words[3] = words[1]                            # a repeated word: words.index() finds the first
                words[6] = "zzq" + words[6]                    # not in the model -> distance 0
                words[5] = words[5] + "9"                      # not alphabetic -> filtered out
            text = " ".join(words[:n_words // 2]) + ", " + " ".join(words[n_words // 2:]) + "."
            out.append(text)
    out.append("It's a " + keys[3] + " " + keys[3] + " " + keys[4] + "!")
    return out


def main():
    nltk = types.ModuleType("nltk")
    nltk.word_tokenize, nltk.pos_tag = toy_nltk.word_tokenize, toy_nltk.pos_tag
    sys.modules["nltk"] = nltk
    sys.path.insert(0, REF)
    from src import utils as ref                                # the reference's file, untouched

    keys = synth.make_vocab(V, SEED)
    vectors = synth.table_rows(np.arange(V), SEED)
    model = ToyModel(keys, vectors)
    rng = np.random.default_rng(2024)
    cases = []
    for text in sentences(keys, rng):
        words = toy_nltk.word_tokenize(text)
        tagged = toy_nltk.pos_tag(words)
        tags = ['JJ', 'RB', 'NN', 'NNS', 'JJR', 'JJS', 'RBR', 'RBS']          # src/utils.py:87
        filtered = [w for w, p in tagged if w.isalpha() and p in tags]
        dist = [ref.semantic_distance(w, filtered, model) for w in filtered]   # src/utils.py:91
        case = {"sentence": text, "words": words, "tagged": tagged, "filtered": filtered,
                "distances_hex": [np.float32(d).tobytes().hex() for d in dist], "masks": {}}
        for num_words in (1, 2, 3):
            got_words, masks = ref.select_descriptive_words(model, text, num_words)
            assert got_words == words
            case["masks"][str(num_words)] = [int(m) for m in masks]
        cases.append(case)
    with open(os.path.join(HERE, "masks_golden.json"), "w") as f:
        json.dump({"vocab_rows": V, "seed": SEED, "source": "src/utils.py:73-104 run under tests/golden/toy_nltk.py",
                   "cases": cases}, f, indent=1)
    print(f"{len(cases)} sentences, {sum(len(c['filtered']) for c in cases)} filtered words")


if __name__ == "__main__":
    main()
