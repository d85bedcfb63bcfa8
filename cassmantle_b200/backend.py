"""Backend -- the scorer half of the reference's ``Backend`` class, on the GPU.

Mirrors src/backend.py:303-317 name for name:

    def compute_score(self, inputs: str, answer: str) -> float
    async def compute_scores(self, data: Dict[str, Dict[str, str]]) -> Dict[str, str]

so ``Server.compute_client_scores`` (src/server.py:63-76) and the browser
(static/script.js:362-411) run unchanged.  Return values keep the reference's types and
text: the Python float ``1.0`` for equal strings, the Python float ``min_score`` when the
word is unknown or the cosine does not beat the floor, otherwise a ``numpy.float32``
cosine; ``compute_scores`` returns ``str()`` of those.  Nothing here raises for an unknown
word (the reference swallows the KeyError, src/backend.py:308-309).

What changed underneath: ``self.wv.similarity`` is one batched CUDA launch over a table
replica resident in HBM (include/cassmantle_b200.h).  There is no CPU path.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

from ._lib import CM_MAX_ANSWERS
from .engine import ScoringEngine
from .vocab import Vocabulary


class Backend:
    def __init__(self, engine: ScoringEngine, vocab: Vocabulary, min_score: float = 0.1,
                 phrase_guesses: bool = False, max_words: int = 64) -> None:
        # min_score default as src/backend.py:22; Server overrides it to 0.01 (src/server.py:17)
        self.min_score = min_score
        self.engine = engine
        self.vocab = vocab
        self.num_masked = 2                       # src/backend.py:49
        # Extension switch (SURVEY.md §0 Track A): treat a guess as a phrase, mean-pooled
        # like gensim's n_similarity.  Off by default: the reference looks the whole string
        # up as one key, so "two words" is simply unknown there.
        self.phrase_guesses = phrase_guesses
        self.max_words = max_words
        self._answer_rows: List[int] = []
        self._answer_slot: Dict[int, int] = {}

    # ---- answers: registered with the engine, reused while the round lasts -----------
    def _slots_for(self, rows: Sequence[int]) -> List[int]:
        new = [r for r in dict.fromkeys(rows) if r not in self._answer_slot]
        if new:
            if len(self._answer_rows) + len(new) > CM_MAX_ANSWERS:
                self._answer_rows, self._answer_slot = [], {}
                new = list(dict.fromkeys(rows))
                if len(new) > CM_MAX_ANSWERS:
                    raise ValueError(f"more than {CM_MAX_ANSWERS} distinct answers in one batch")
            for r in new:
                self._answer_slot[r] = len(self._answer_rows)
                self._answer_rows.append(r)
        # a no-op unless the registered set changed -- here or behind our back (another user of
        # the engine, e.g. DeviceVocabulary.score_words, re-registers answers too)
        self.engine.set_answers(self._answer_rows)
        return [self._answer_slot[r] for r in rows]

    # ---- core: a list of (guess, answer) string pairs -> reference-typed scores ------
    def score_pairs(self, pairs: Sequence[Tuple[str, str]]) -> list:
        if not pairs:
            return []
        B = len(pairs)
        get = self.vocab.key_to_index.get
        guesses = [p[0].lower() for p in pairs]                       # backend.py:304
        answers = [p[1].lower() for p in pairs]
        exact = np.fromiter((g == a for g, a in zip(guesses, answers)), dtype=np.uint8, count=B)  # :305
        ans_rows = [get(a, -1) for a in answers]
        if not self.phrase_guesses:                                   # the reference: whole string = one key
            idx = np.fromiter((get(g, -1) for g in guesses), dtype=np.int32, count=B).reshape(B, 1)
            length = np.ones(B, dtype=np.int32)
        else:
            words = [g.split()[: self.max_words] for g in guesses]
            N = max(1, max(len(w) for w in words))
            idx = np.full((B, N), -1, dtype=np.int32)
            length = np.zeros(B, dtype=np.int32)
            for b, w in enumerate(words):
                length[b] = len(w)
                for k, t in enumerate(w):
                    idx[b, k] = get(t, -1)
        slots = np.asarray(self._slots_for(ans_rows), dtype=np.int32)
        scores, _ = self.engine.score_host(idx, length, slots, exact, self.min_score)
        floor32 = np.float32(self.min_score)
        keep = scores > floor32                  # the cosine wins max(min_score, score) only if greater
        # numpy.float32 like wv.similarity's result, or the Python floats 1.0 / min_score
        return [1.0 if e else (s if k else self.min_score)
                for e, k, s in zip(exact.tolist(), keep.tolist(), list(scores))]

    # ---- the reference's two methods ---------------------------------------------------
    def compute_score(self, inputs: str, answer: str) -> float:
        return self.score_pairs([(inputs, answer)])[0]

    async def compute_scores(self, data: Dict[str, Dict[str, str]]) -> Dict[str, str]:
        keys = list(data.keys())
        vals = self.score_pairs([(data[k]["input"], data[k]["answer"]) for k in keys])
        scores: Dict[str, str] = {}
        for key, score in zip(keys, vals):
            scores.update({key: str(score)})                          # backend.py:316
        return scores

    def most_similar(self, word: str, topn: int = 50) -> List[Tuple[str, float]]:
        """src/backend.py:297-301: ``self.wv.most_similar(word.lower(), topn=topn)``."""
        try:
            row = self.vocab.key_to_index[word.lower()]
        except KeyError as e:
            raise Exception("Word not in dictionary. ", e)
        rows, sims = self.engine.most_similar(row, topn + 1)      # gensim: topn + len(all_keys)
        keys = self.vocab.index_to_key
        result = [(keys[int(r)], float(s)) for r, s in zip(rows, sims) if int(r) != row]
        return result[:topn]

    def most_similar_many(self, words, topn: int = 50) -> List[List[Tuple[str, float]]]:
        """``[self.most_similar(w, topn) for w in words]`` with the table read once for all of them
        (what a micro-batcher holding several such calls would issue)."""
        rows = []
        for word in words:
            try:
                rows.append(self.vocab.key_to_index[word.lower()])
            except KeyError as e:
                raise Exception("Word not in dictionary. ", e)
        keys = self.vocab.index_to_key
        out = []
        for row, (got, sims) in zip(rows, self.engine.most_similar_batch(rows, topn + 1)):
            out.append([(keys[int(r)], float(s)) for r, s in zip(got, sims) if int(r) != row][:topn])
        return out

    def score_to_blur(self, score: float, min_blur: float = 0.0, max_blur: float = 15):
        return min_blur + (1 - score ** 2) * (max_blur - min_blur)    # backend.py:319-320
