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
