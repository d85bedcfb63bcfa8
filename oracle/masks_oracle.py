"""CPU restatement of the reference's mask selection -- TEST INFRASTRUCTURE, not product code.

Only tests/ may import this.  It restates, call for call,

    semantic_distance(word, word_list, model)                  src/utils.py:73-79
    select_descriptive_words(model, sentence, num_words=2)     src/utils.py:81-104
    construct_prompt_dict(model, prompt, num_masked)           src/utils.py:106-111

with the two nltk calls (src/utils.py:83-84) passed in as functions, because nltk is not on this
box.  PARITY PINNED: tests/golden/masks_golden.json holds outputs of the reference's own
functions, imported unmodified and run in the build container with tests/golden/toy_nltk.py in
nltk's place (tests/golden/make_masks_golden.py); tests/test_masks_oracle.py checks this file
against them, and tests/test_reference_masks.py re-runs the reference live where it is mounted.

IDF (src/utils.py:93-96): `TfidfVectorizer().fit([sentence])` sees ONE document, so every
term's document frequency equals the number of documents and
idf = ln((1 + n) / (1 + df)) + 1 = 1.0 for all of them; `idf_dict.get(word, 1)` is 1 as well.
The weight is therefore exactly 1.0 and the score is the distance (x * 1.0 is exact in IEEE
arithmetic); the golden vectors, produced with the real scikit-learn, confirm it.
"""
from __future__ import annotations

from typing import Callable, Dict, List, Sequence, Tuple

import re

import numpy as np

DESCRIPTIVE_TAGS = ['JJ', 'RB', 'NN', 'NNS', 'JJR', 'JJS', 'RBR', 'RBS']     # src/utils.py:87


def semantic_distance(word: str, word_list: Sequence[str], model) -> float:
    """src/utils.py:73-79."""
    if word in model:
        word_vector = model[word]
        mean_vector = np.mean([model[w] for w in word_list if w in model], axis=0)
        return np.linalg.norm(word_vector - mean_vector)
    return 0


def select_descriptive_words(model, sentence: str, num_words: int, tokenize: Callable, pos_tag: Callable
                             ) -> Tuple[List[str], List[int]]:
    """src/utils.py:81-104 with the tokenizer and tagger injected."""
    words = tokenize(sentence)
    tagged_words = pos_tag(words)
    filtered_words = [word for word, pos in tagged_words if word.isalpha() and pos in DESCRIPTIVE_TAGS]
    distances = [semantic_distance(word, filtered_words, model) for word in filtered_words]
    # TfidfVectorizer().fit([sentence]) (src/utils.py:94): its default analyzer lower-cases and keeps
    # tokens of two or more word characters; with none left scikit-learn raises
    if not re.findall(r"(?u)\b\w\w+\b", sentence.lower()):
        raise ValueError("empty vocabulary; perhaps the documents only contain stop words")
    scores = [distances[i] * 1.0 for i in range(len(filtered_words))]          # idf == 1.0, see the header
    selected_indices = [words.index(filtered_words[i]) for i in np.argsort(scores)[-num_words:]]
    return words, sorted(selected_indices)


def construct_prompt_dict(model, prompt: str, num_masked: int, tokenize: Callable, pos_tag: Callable
                          ) -> Dict[str, list]:
    """src/utils.py:106-111."""
    words, masks = select_descriptive_words(model, prompt, num_masked, tokenize, pos_tag)
    return {'tokens': words, 'masks': masks}


def distances_from_rows(table: np.ndarray, rows: Sequence[int], n_mean: int = -1) -> np.ndarray:
    """The numeric part for row indices (-1 = not in the model): what cm_semantic_distances computes.
    The mean is taken over rows[:n_mean] (all of them by default)."""
    rows = np.asarray(rows, dtype=np.int64)
    head = rows if n_mean < 0 else rows[:n_mean]
    known = head[head >= 0]
    out = np.zeros(len(rows), dtype=np.float32)
    if known.size == 0:
        mean_vector = np.zeros(table.shape[1], dtype=np.float32)
    else:
        mean_vector = np.mean([table[r] for r in known], axis=0)
    for i, r in enumerate(rows):
        if r >= 0:
            out[i] = np.linalg.norm(table[r] - mean_vector)
    return out
