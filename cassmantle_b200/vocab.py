"""Vocabulary: word -> table row, the ``key_to_index`` half of gensim's KeyedVectors.

This is the reference's whole "tokeniser" (SURVEY.md §8 a7): ``str.lower()`` on the guess
(src/backend.py:304) and one dict lookup of the complete string inside
``wv.similarity`` (src/backend.py:307).  A missing key is a KeyError there and a row of -1
here; the kernel turns -1 into ``min_score`` exactly as the ``except KeyError`` does.
Keys are case-sensitive, as in gensim; only lowercase forms are ever queried.
"""
from __future__ import annotations

from typing import Dict, Iterable, List, Sequence

import numpy as np

OOV = -1


class Vocabulary:
    def __init__(self, index_to_key: Sequence[str]) -> None:
        self.index_to_key: List[str] = list(index_to_key)
        self.key_to_index: Dict[str, int] = {}
        for i, k in enumerate(self.index_to_key):
            self.key_to_index.setdefault(k, i)          # first occurrence wins (gensim)

    def __len__(self) -> int:
        return len(self.index_to_key)

    def __contains__(self, key: str) -> bool:
        return key in self.key_to_index

    def lookup(self, key: str) -> int:
        return self.key_to_index.get(key, OOV)

    def lookup_many(self, keys: Iterable[str]) -> np.ndarray:
        get = self.key_to_index.get
        return np.fromiter((get(k, OOV) for k in keys), dtype=np.int32)


def pack_words(words: Sequence[str]):
    """UTF-8 bytes of `words`, concatenated, + int32 offsets -- the C ABI's string format.

    Text that is not plain ASCII is lower-cased here (Python's rules: the device only folds
    A-Z); ASCII words go through untouched, their case is folded on the device."""
    enc = [(w if w.isascii() else w.lower()).encode("utf-8") for w in words]
    offsets = np.zeros(len(enc) + 1, dtype=np.int32)
    if enc:
        np.cumsum([len(b) for b in enc], out=offsets[1:])
    return np.frombuffer(b"".join(enc) or b"\0", dtype=np.uint8), offsets


class DeviceVocabulary:
    """The same dictionary as :class:`Vocabulary`, resident in HBM (C ABI: cm_vocab_*).

    ``lookup_many`` lower-cases and looks the words up on the GPU; ``score_words`` runs the
    whole of ``Backend.compute_scores`` for one-word guesses there (src/backend.py:303-317).
    """

    def __init__(self, engine, index_to_key: Sequence[str]) -> None:
        import ctypes

        from . import _lib
        self._ct, self._lib, self._check, self.engine = ctypes, engine._lib, _lib.check, engine
        enc = [k.encode("utf-8") for k in index_to_key]
        offsets = np.zeros(len(enc) + 1, dtype=np.int64)
        np.cumsum([len(b) for b in enc], out=offsets[1:])
        blob = b"".join(enc) or b"\0"
        self._h = ctypes.c_void_p()
        self._check(self._lib.cm_vocab_create(ctypes.byref(self._h), engine._h, blob,
                                              offsets.ctypes.data, len(enc)))
        self.n_keys = len(enc)

    def close(self) -> None:
        if getattr(self, "_h", None) is not None and self._h.value:
            self._lib.cm_vocab_destroy(self._h)
            self._h = self._ct.c_void_p()

    def lookup_many(self, words: Sequence[str]) -> np.ndarray:
        data, offsets = pack_words(words)
        rows = np.empty(len(words), dtype=np.int32)
        self._check(self._lib.cm_vocab_lookup_host(self._h, data.ctypes.data, offsets.ctypes.data,
                                                   len(words), rows.ctypes.data))
        return rows

    def score_words(self, guesses: Sequence[str], answers: Sequence[str], answer_slot,
                    min_score: float) -> np.ndarray:
        """fp32 scores of guess i against answers[answer_slot[i]] (strings in, scores out)."""
        g, go = pack_words(guesses)
        a, ao = pack_words(answers)
        sl = np.ascontiguousarray(answer_slot, dtype=np.int32)
        out = np.empty(len(guesses), dtype=np.float32)
        self._check(self._lib.cm_score_words_host(
            self.engine._h, self._h, g.ctypes.data, go.ctypes.data, len(guesses), a.ctypes.data,
            ao.ctypes.data, len(answers), sl.ctypes.data, self._ct.c_float(np.float32(min_score)),
            out.ctypes.data, None))
        self.engine._answers = None            # the call re-registered the engine's answers
        return out
