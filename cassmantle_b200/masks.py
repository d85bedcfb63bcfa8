"""Mask selection -- which words of the round's prompt become blanks.

Mirrors, name for name, src/utils.py:73-111:

    semantic_distance(word, word_list, model)
    select_descriptive_words(model, sentence, num_words=2)
    construct_prompt_dict(model, prompt, num_masked)

``model`` here is a :class:`MaskModel`: the vocabulary (case-sensitive, as gensim's ``word in
model`` is -- nothing is lower-cased on this path) plus the scoring engine whose resident table
holds the vectors.  The distances of all filtered words are computed in ONE device call
(C ABI ``cm_semantic_distances``) instead of one numpy mean per word.

The reference tokenises and tags with nltk (src/utils.py:83-84).  nltk is imported lazily and
only if no ``tokenize`` / ``pos_tag`` callables are passed; everything after those two calls is
independent of them and is what the parity tests pin against the reference's own function.

IDF: ``TfidfVectorizer().fit([sentence])`` (src/utils.py:93-96) is fitted on ONE document, which
makes every idf exactly 1.0 (ln((1+1)/(1+1)) + 1), and ``idf_dict.get(word, 1)`` is 1 too: the
score of a word IS its distance.  scikit-learn is therefore not imported; the one observable
effect it has -- ``ValueError`` for a sentence without any token of two or more word characters --
is reproduced.
"""
from __future__ import annotations

import re
from typing import Callable, Dict, List, Optional, Sequence, Tuple

import numpy as np

from ._lib import check
from .vocab import Vocabulary

DESCRIPTIVE_TAGS = ['JJ', 'RB', 'NN', 'NNS', 'JJR', 'JJS', 'RBR', 'RBS']      # src/utils.py:87
_TFIDF_TOKEN = re.compile(r"(?u)\b\w\w+\b")                                    # TfidfVectorizer's default


class MaskModel:
    """``word in model`` + the device table: what src/utils.py needs of the KeyedVectors object."""

    def __init__(self, engine, vocab: Vocabulary) -> None:
        self.engine, self.vocab = engine, vocab

    def __contains__(self, word: str) -> bool:
        return word in self.vocab

    def semantic_distances(self, word_list: Sequence[str], extra: Sequence[str] = ()) -> np.ndarray:
        """fp32: ``semantic_distance(w, word_list, model)`` for every w of ``word_list`` followed
        by every w of ``extra`` (words measured against the list's mean without joining it)."""
        words = list(word_list) + list(extra)
        rows = np.ascontiguousarray(self.vocab.lookup_many(words), dtype=np.int32)
        out = np.zeros(rows.size, dtype=np.float32)
        if rows.size:
            check(self.engine._lib.cm_semantic_distances(self.engine._h, rows.ctypes.data, int(rows.size),
                                                         len(word_list), out.ctypes.data))
        return out


def semantic_distance(word: str, word_list: Sequence[str], model: MaskModel):
    """src/utils.py:73-79 (one word; ``model.semantic_distances`` does a whole list per call)."""
    if word in model:
        return model.semantic_distances(word_list, extra=[word])[-1]
    return 0


def _nltk_calls() -> Tuple[Callable, Callable]:
    try:
        import nltk
    except ImportError as exc:                       # noqa: F841
        raise ImportError("nltk is not installed: pass tokenize= and pos_tag= "
                          "(the reference calls nltk.word_tokenize / nltk.pos_tag, src/utils.py:83-84)") from exc
    return nltk.word_tokenize, nltk.pos_tag


def select_descriptive_words(model: MaskModel, sentence: str, num_words: int = 2,
                             tokenize: Optional[Callable] = None, pos_tag: Optional[Callable] = None
                             ) -> Tuple[List[str], List[int]]:
    """src/utils.py:81-104: (tokens, sorted indices of the ``num_words`` most descriptive ones)."""
    if tokenize is None or pos_tag is None:
        nltk_tokenize, nltk_tag = _nltk_calls()
        tokenize, pos_tag = tokenize or nltk_tokenize, pos_tag or nltk_tag
    words = tokenize(sentence)
    tagged_words = pos_tag(words)
    filtered_words = [word for word, pos in tagged_words if word.isalpha() and pos in DESCRIPTIVE_TAGS]
    distances = model.semantic_distances(filtered_words)
    if not _TFIDF_TOKEN.findall(sentence.lower()):   # what TfidfVectorizer().fit([sentence]) raises
        raise ValueError("empty vocabulary; perhaps the documents only contain stop words")
    scores = [distances[i] * 1.0 for i in range(len(filtered_words))]          # idf == 1.0 (module docstring)
    selected_indices = [words.index(filtered_words[i]) for i in np.argsort(scores)[-num_words:]]
    return words, sorted(selected_indices)


def construct_prompt_dict(model: MaskModel, prompt: str, num_masked: int,
                          tokenize: Optional[Callable] = None, pos_tag: Optional[Callable] = None
                          ) -> Dict[str, list]:
    """src/utils.py:106-111."""
    words, masks = select_descriptive_words(model, prompt, num_masked, tokenize, pos_tag)
    return {'tokens': words, 'masks': masks}
