"""CPU tests of the host-side lowering in cassmantle_b200/backend.py.

The product Backend only talks to a ScoringEngine (CUDA).  Here a stand-in engine that
answers with the oracle is injected so the string handling -- lower(), equality before
lookup, OOV -> -1, answer slots, result types and str() -- can be checked without a GPU.
"""
import asyncio

import numpy as np

from cassmantle_b200.backend import Backend
from cassmantle_b200.vocab import Vocabulary
from oracle import scorer_oracle as O
from tests.util import dense_kat_table


class OracleEngine:
    """Test double with ScoringEngine's two methods used by Backend."""

    def __init__(self, table):
        self.table = table
        self.answers = None
        self.calls = 0
        self.set_calls = 0

    def set_answers(self, rows, stream=0):
        if self.answers is not None and tuple(self.answers.tolist()) == tuple(int(r) for r in rows):
            return                                   # the real engine skips an unchanged set too
        self.answers = np.asarray(rows, dtype=np.int32)
        self.set_calls += 1

    def score_host(self, idx, length, slot, exact, min_score, want_centi=False):
        self.calls += 1
        return O.score_batch_faithful(self.table, idx, length, self.answers, slot, exact, min_score), None


def _pair(golden, **kw):
    table = dense_kat_table(golden)
    keys = golden["kat"]["keys"]
    eng = OracleEngine(table)
    return Backend(eng, Vocabulary(keys), min_score=0.01, **kw), \
        O.OracleBackend(O.OracleWordVectors(table, keys), 0.01), eng


def test_compute_score_types_and_values(golden):
    be, ref, _ = _pair(golden)
    for c in golden["kat"]["cases"]:
        got, want = be.compute_score(c["input"], c["answer"]), ref.compute_score(c["input"], c["answer"])
        assert type(got) is type(want), c
        assert str(got) == str(want) == c["str"], c


def test_compute_scores_matches_reference_format(golden):
    be, ref, eng = _pair(golden)
    data = {str(i): {"input": c["input"], "answer": c["answer"]}
            for i, c in enumerate(golden["kat"]["cases"])}
    got = asyncio.run(be.compute_scores(data))
    want = asyncio.run(ref.compute_scores(data))
    assert got == want and list(got) == list(want)
    assert eng.calls == 1                      # the whole dict is ONE device batch


def test_answers_registered_once_per_round(golden):
    be, _, eng = _pair(golden)
    for _ in range(5):
        be.compute_score("mixed", "alpha")
        be.compute_score("mixed", "delta")
    assert eng.set_calls == 2                  # alpha, then alpha+delta; then reused


def test_whole_string_is_one_key_by_default(golden):
    be, ref, _ = _pair(golden)
    assert be.compute_score("alpha delta", "alpha") == ref.compute_score("alpha delta", "alpha") == 0.01


def test_phrase_extension_matches_n_similarity(golden):
    be, ref, _ = _pair(golden, phrase_guesses=True)
    for g, a in [("alpha delta", "alpha"), ("mixed ones twin1", "quarter"), ("alpha nosuch", "alpha"),
                 ("Alpha", "alpha"), ("delta", "alpha")]:
        assert str(be.compute_score(g, a)) == str(ref.compute_score_phrase(g, a)), (g, a)


def test_empty_payload(golden):
    be, _, eng = _pair(golden)
    assert asyncio.run(be.compute_scores({})) == {}
    assert eng.calls == 0


def test_pack_words_format():
    from cassmantle_b200.vocab import pack_words
    data, off = pack_words(["ab", "", "CÉ", "d"])
    assert off.tolist() == [0, 2, 2, 5, 6] and off.dtype == np.int32
    assert bytes(data[:6]) == "abcéd".encode()            # non-ASCII text is lowered on the host
    data, off = pack_words([])
    assert off.tolist() == [0]


def test_lookup_oracle_first_occurrence_wins():
    rows = O.lookup_rows(["a", "B", "a", "b"], ["A", "b", "B", "c", ""])
    assert rows.tolist() == [0, 3, 3, -1, -1]


def test_most_similar_many_is_most_similar_per_word():
    """Backend.most_similar_many (one engine call for all words) against most_similar per word, with the oracle
    standing in for both engine calls: key lookup through lower(), the query word dropped, topn kept, errors."""
    from cassmantle_b200 import synth
    import pytest
    keys = synth.make_vocab(400, seed=0)
    table = synth.table_rows(np.arange(400), seed=0)

    class SimilarEngine:
        batch_calls = 0

        def most_similar(self, row, k):
            d = O.most_similar_dists(table, int(row))
            order = np.lexsort((np.arange(len(d)), -d))[:k]
            return order.astype(np.int32), d[order].astype(np.float32)

        def most_similar_batch(self, rows, k):
            self.batch_calls += 1
            return [self.most_similar(r, k) for r in rows]

    eng = SimilarEngine()
    be = Backend(eng, Vocabulary(keys), min_score=0.01)
    words = [keys[3].upper(), keys[77], keys[399]]
    many = be.most_similar_many(words, topn=7)
    assert eng.batch_calls == 1 and len(many) == 3
    for w, got in zip(words, many):
        assert got == be.most_similar(w, topn=7)
        assert len(got) == 7 and w.lower() not in [k for k, _ in got]
        assert all(isinstance(s, float) for _, s in got)
    assert be.most_similar_many([], topn=5) == []
    with pytest.raises(Exception) as exc:
        be.most_similar_many([keys[1], "no-such-word"], topn=5)
    assert exc.value.args[0] == "Word not in dictionary. "
