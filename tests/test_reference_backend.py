"""CPU test: the reference's OWN scorer methods, run live, against the oracle and this repo's mirror.

Runs only where /root/reference is mounted (the build container; skipped on the GPU box).
`src/backend.py` is imported unmodified; aioredis, gensim and nltk (absent here, none of them
executed by the methods under test) are empty stand-ins, and `self.wv` -- the gensim object -- is
the oracle's OracleWordVectors.  So this pins everything the reference itself writes on the hot
path (src/backend.py:297-320: lower(), equality first, KeyError -> min_score, max(), str(), the
"Word not in dictionary" exception, score_to_blur); what stays unpinned is only the arithmetic
inside gensim's `similarity` / `most_similar`, which the oracle restates from gensim's source.
"""
import asyncio
import os
import sys
import types

import numpy as np
import pytest

from cassmantle_b200.backend import Backend as Mirror
from cassmantle_b200.vocab import Vocabulary
from oracle import scorer_oracle as O
from tests.test_backend_host import OracleEngine
from tests.util import dense_kat_table

REF = "/root/reference"
pytestmark = pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "src")), reason="reference tree not mounted")


class WordVectors(O.OracleWordVectors):
    """+ most_similar with gensim's return type, for src/backend.py:297-301."""

    def most_similar(self, word, topn=10):
        keys = list(self.key_to_index)
        return [(keys[r], s) for r, s in O.most_similar_rows(self.vectors, self.key_to_index[word], topn)]


@pytest.fixture()
def reference_backend(golden, monkeypatch):
    for name in ("aioredis", "aioredis.exceptions", "gensim", "gensim.models", "nltk"):
        monkeypatch.setitem(sys.modules, name, types.ModuleType(name))
    sys.modules["aioredis.exceptions"].LockError = type("LockError", (Exception,), {})
    sys.modules["aioredis"].Redis = object
    sys.modules["gensim.models"].KeyedVectors = object
    monkeypatch.syspath_prepend(REF)
    for m in [k for k in sys.modules if k == "src" or k.startswith("src.")]:
        monkeypatch.delitem(sys.modules, m)
    from src.backend import Backend                    # the reference's file, untouched
    table, keys = dense_kat_table(golden), golden["kat"]["keys"]
    ref = object.__new__(Backend)                      # __init__ needs api_key.txt and the 3.6 GB file
    ref.min_score = 0.01                               # src/server.py:17
    ref.wv = WordVectors(table, keys)
    yield ref, table, keys
    for m in [k for k in sys.modules if k == "src" or k.startswith("src.")]:
        sys.modules.pop(m, None)


def _pairs(golden, keys):
    pairs = [(c["input"], c["answer"]) for c in golden["kat"]["cases"]]
    rng = np.random.default_rng(5)
    pool = list(keys) + [k.upper() for k in keys[:4]] + ["", " ", "xyzzy", "Alpha ", "ALPHA"]
    pairs += [(pool[int(i)], pool[int(j)]) for i, j in rng.integers(0, len(pool), (300, 2))]
    return pairs


def test_reference_compute_score_equals_oracle_and_mirror(reference_backend, golden):
    ref, table, keys = reference_backend
    oracle = O.OracleBackend(O.OracleWordVectors(table, keys), 0.01)
    mirror = Mirror(OracleEngine(table), Vocabulary(keys), min_score=0.01)
    for a, b in _pairs(golden, keys):
        want = ref.compute_score(a, b)                 # src/backend.py:303-310, the reference's own code
        for got in (oracle.compute_score(a, b), mirror.compute_score(a, b)):
            assert type(got) is type(want), (a, b)
            assert str(got) == str(want), (a, b)
    for c in golden["kat"]["cases"]:                   # and the authored known-answer strings hold for it
        assert str(ref.compute_score(c["input"], c["answer"])) == c["str"], c


def test_reference_compute_scores_equals_oracle_and_mirror(reference_backend, golden):
    ref, table, keys = reference_backend
    oracle = O.OracleBackend(O.OracleWordVectors(table, keys), 0.01)
    mirror = Mirror(OracleEngine(table), Vocabulary(keys), min_score=0.01)
    data = {f"k{i}": {"input": a, "answer": b} for i, (a, b) in enumerate(_pairs(golden, keys))}
    want = asyncio.run(ref.compute_scores(data))       # src/backend.py:312-317
    assert asyncio.run(oracle.compute_scores(data)) == want
    got = asyncio.run(mirror.compute_scores(data))
    assert got == want and list(got) == list(want)
    with pytest.raises(KeyError):                      # a malformed payload fails the same way
        asyncio.run(ref.compute_scores({"0": {"input": "alpha"}}))
    with pytest.raises(KeyError):
        asyncio.run(mirror.compute_scores({"0": {"input": "alpha"}}))


def test_reference_most_similar_and_score_to_blur(reference_backend):
    ref, table, keys = reference_backend
    mirror = Mirror(OracleEngine(table), Vocabulary(keys), min_score=0.01)
    for s in (0.0, 0.01, 0.3333, 0.99999994, 1.0, np.float32(0.4567891)):
        assert ref.score_to_blur(s) == mirror.score_to_blur(s)                 # src/backend.py:319-320
        assert ref.score_to_blur(s, 1.0, 9) == mirror.score_to_blur(s, 1.0, 9)
    with pytest.raises(Exception) as r1:
        ref.most_similar("No-Such-Word")                                       # src/backend.py:300-301
    with pytest.raises(Exception) as r2:
        mirror.most_similar("No-Such-Word")
    assert type(r1.value) is type(r2.value) is Exception and r1.value.args[0] == r2.value.args[0] == "Word not in dictionary. "
    want = ref.most_similar(keys[0].upper(), topn=3)                           # .lower() before the lookup
    assert [k for k, _ in want] == [keys[r] for r, _ in O.most_similar_rows(table, 0, 3)]
