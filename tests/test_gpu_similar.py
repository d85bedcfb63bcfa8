"""GPU parity tests of most_similar (cm_most_similar) against the gensim restatement."""
import numpy as np
import pytest

from cassmantle_b200 import synth
from oracle import scorer_oracle as O
from tests.util import COS_TOL

pytestmark = pytest.mark.gpu


def _check(rows, sims, dists, k, query):
    """GPU top-k vs the oracle's full distance vector (ties / last-ulp swaps allowed)."""
    assert len(rows) == k and len(set(rows.tolist())) == k
    assert np.all(np.abs(sims - dists[rows]) <= COS_TOL)            # each cosine is right
    assert np.all(np.diff(sims) <= 0)                               # best first
    want = np.sort(dists[~np.isnan(dists)])[::-1][:k]
    assert np.all(np.abs(np.sort(sims)[::-1] - want) <= COS_TOL)    # and nothing better was missed
    assert query in rows                                            # cos(w, w) = 1 is always in the list


def test_most_similar_matches_oracle(small_engine, small_table):
    for q in (0, 123, 49_999, 31337):
        dists = O.most_similar_dists(small_table, q)
        for k in (1, 11, 51, 1024):
            rows, sims = small_engine.most_similar(q, k)
            _check(rows, sims, dists, k, q)


def test_backend_most_similar_is_the_reference_call(small_engine, small_table):
    from cassmantle_b200.backend import Backend
    from cassmantle_b200.vocab import Vocabulary
    keys = synth.make_vocab(50_000)
    be = Backend(small_engine, Vocabulary(keys), min_score=0.01)
    got = be.most_similar(keys[777].upper(), topn=50)               # .lower() as src/backend.py:299
    want = O.most_similar_rows(small_table, 777, 50)
    assert len(got) == 50 and all(isinstance(k, str) and isinstance(s, float) for k, s in got)
    assert keys[777] not in [k for k, _ in got]                      # the query word is dropped
    for (gk, gs), (wr, ws) in zip(got, want):
        assert abs(gs - ws) <= COS_TOL
    assert [k for k, _ in got][:10] == [keys[r] for r, _ in want][:10]   # clear-cut head of the list
    with pytest.raises(Exception) as ei:
        be.most_similar("no-such-word")
    assert "Word not in dictionary" in str(ei.value.args[0])


def test_most_similar_special_rows():
    from cassmantle_b200.engine import ScoringEngine
    rng = np.random.default_rng(5)
    vec = rng.standard_normal((5000, 300), dtype=np.float32)
    vec[7] = 0.0                  # zero norm -> 0/0 = NaN cosine: ranks last
    vec[8, 3] = np.nan            # NaN row
    vec[9] = vec[10] * 3.0        # same direction as row 10: cosine 1 (a tie with the query itself)
    with ScoringEngine.from_vectors(vec) as eng:
        dists = O.most_similar_dists(vec, 10)
        rows, sims = eng.most_similar(10, 20)
        assert set(rows[:2].tolist()) == {9, 10} and 7 not in rows and 8 not in rows
        assert np.all(np.abs(sims - dists[rows]) <= COS_TOL)
        rows, sims = eng.most_similar(10, 1024)
        assert 7 not in rows and 8 not in rows
        rows_all, _ = eng.most_similar(11, 1024)
        assert len(rows_all) == 1024
        # after the table changes the cached norms must be recomputed
        vec2 = vec.copy(); vec2[12] = vec2[11] * 0.5
        from cassmantle_b200 import _lib
        _lib.check(eng._lib.cm_engine_upload_rows(eng._h, vec2[12:13].ctypes.data, 12, 1))
        rows, sims = eng.most_similar(11, 3)
        assert set(rows[:2].tolist()) == {11, 12}


def test_full_size_most_similar():
    """3,000,000 x 300: one query = one pass over 3.6 GB; sampled check against the host."""
    from cassmantle_b200.engine import ScoringEngine
    V = synth.FULL_ROWS
    with ScoringEngine.synthetic(V, 300, seed=0) as eng:
        q = 2_718_281
        rows, sims = eng.most_similar(q, 51)
        assert rows[0] == q and abs(float(sims[0]) - 1.0) <= COS_TOL
        # the rest of the list must be rows of the query's cluster (cosine ~0.5 in the synthetic table)
        assert np.all(synth.cluster_of(rows, 0) == synth.cluster_of([q], 0)[0])
        sub = synth.table_rows(rows, 0)
        qv = synth.table_rows([q], 0)[0]
        for r, s, v in zip(rows, sims, sub):
            assert abs(float(s) - float(O.similarity(np.stack([qv, v]), 0, 1))) <= 2 * COS_TOL
        again, sims2 = eng.most_similar(q, 51)
        assert np.array_equal(rows, again) and np.array_equal(sims, sims2)
