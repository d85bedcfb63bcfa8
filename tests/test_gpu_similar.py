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


def test_most_similar_crowded_bin_takes_the_radix_fallback():
    """Thousands of rows with the same distance around the k-th place (duplicated rows): the
    fast path's one-bin sort cannot hold them, the call must fall back and still be exact."""
    from cassmantle_b200.engine import ScoringEngine
    rng = np.random.default_rng(11)
    vec = rng.standard_normal((20_000, 300), dtype=np.float32)
    vec[1000:6000] = vec[1000] * rng.uniform(0.5, 2.0, (5000, 1)).astype(np.float32)   # 5000 rows, cosine 1 with row 1000
    with ScoringEngine.from_vectors(vec) as eng:
        dists = O.most_similar_dists(vec, 1000)
        for k in (10, 1024):
            rows, sims = eng.most_similar(1000, k)
            assert len(set(rows.tolist())) == k and np.all((rows >= 1000) & (rows < 6000))
            assert np.all(np.abs(sims - dists[rows]) <= COS_TOL) and np.all(np.diff(sims) <= 0)
        # and the fast path is clean again afterwards (its counters were reset by the fallback call)
        d2 = O.most_similar_dists(vec, 7)
        rows, sims = eng.most_similar(7, 40)
        _check(rows, sims, d2, 40, 7)


@pytest.mark.parametrize("cap", ["0", "5"])
def test_most_similar_forced_fallback_matches_fast_path(small_table, cap):
    """CM_SIMILAR_BUCKET_CAP (read once per process) is a test hook, so this runs in a child."""
    import os, subprocess, sys, textwrap
    code = textwrap.dedent("""
        import numpy as np
        from cassmantle_b200 import synth
        from cassmantle_b200.engine import ScoringEngine
        with ScoringEngine.synthetic(50_000, 300, seed=3) as eng:
            for q, k in ((5, 1), (77, 51), (4242, 1024)):
                r, s = eng.most_similar(q, k)
                print(q, k, int(np.bitwise_xor.reduce(r.astype(np.int64) * 2654435761 % (1 << 31))), float(s.sum()))
    """)
    outs = []
    for env_cap in (None, cap):
        env = dict(os.environ)
        env.pop("CM_SIMILAR_BUCKET_CAP", None)
        if env_cap is not None:
            env["CM_SIMILAR_BUCKET_CAP"] = env_cap
        outs.append(subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True,
                                   cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__))), check=True).stdout)
    assert outs[0] == outs[1] and len(outs[0].splitlines()) == 3


@pytest.mark.parametrize("dim", [64, 128, 512, 1024])
def test_most_similar_other_dimensions(dim):
    """Dimensions other than 300 take the generic scan kernel; same selection after it."""
    from cassmantle_b200.engine import ScoringEngine
    rng = np.random.default_rng(dim)
    vec = rng.standard_normal((3001, dim), dtype=np.float32)
    with ScoringEngine.from_vectors(vec) as eng:
        for q, k in ((0, 5), (3000, 100)):
            rows, sims = eng.most_similar(q, k)
            _check(rows, sims, O.most_similar_dists(vec, q), k, q)


def _similar_child(env_extra, n_rows, queries):
    """most_similar in a child process (the scan variant is read once per process)."""
    import os, subprocess, sys, textwrap
    code = textwrap.dedent(f"""
        import numpy as np
        from cassmantle_b200.engine import ScoringEngine
        with ScoringEngine.synthetic({n_rows}, 300, seed=3) as eng:
            for q, k in {queries!r}:
                r, s = eng.most_similar(q, k)
                print(q, k, " ".join(map(str, r.tolist())), " ".join(v.hex() for v in s.astype(np.float64)))
    """)
    env = dict(os.environ)
    env.pop("CM_SIMILAR_GENERIC_SCAN", None)
    env.update(env_extra)
    return subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, check=True,
                          cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__)))).stdout.splitlines()


def _compare_scans(n_rows, queries):
    a = _similar_child({}, n_rows, queries)
    b = _similar_child({"CM_SIMILAR_GENERIC_SCAN": "1"}, n_rows, queries)
    assert len(a) == len(b) == len(queries)
    for la, lb in zip(a, b):
        fa, fb = la.split(" "), lb.split(" ")
        k = int(fa[1])
        assert fa[:2] == fb[:2]
        rows_a, rows_b = fa[2:2 + k], fb[2:2 + k]
        sims_a = np.array([float.fromhex(x) for x in fa[2 + k:]])
        sims_b = np.array([float.fromhex(x) for x in fb[2 + k:]])
        assert np.abs(sims_a - sims_b).max() <= 2 * COS_TOL         # each is within COS_TOL of the oracle
        assert fa[0] in rows_a and fa[0] in rows_b                    # the query itself is in both lists
        assert len(set(rows_a) ^ set(rows_b)) <= 4                   # at most a couple of last-ulp swaps at the k-th place


def test_16_lane_scan_agrees_with_the_generic_scan():
    _compare_scans(50_000, [(5, 1), (77, 51), (4242, 1024)])


def test_16_lane_scan_agrees_with_the_generic_scan_at_full_size():
    _compare_scans(synth.FULL_ROWS, [(2_718_281, 51), (7, 200)])


# ---- cm_most_similar_batch: many queries per pass over the table (tensor-core filter + exact re-score) ----
def _same_as_single(eng, queries, k):
    batch = eng.most_similar_batch(queries, k)
    assert len(batch) == len(queries)
    for q, (rows, sims) in zip(queries, batch):
        want_rows, want_sims = eng.most_similar(int(q), k)
        assert np.array_equal(sims, want_sims), (q, k)              # the same fp32 cosines, bit for bit
        inside = sims > sims[-1]                                    # entries tied at the k-th place may differ
        assert np.array_equal(rows[inside], want_rows[inside]), (q, k)
        assert len(set(rows.tolist())) == len(rows)


def test_batch_equals_single_queries(small_engine, small_table):
    rng = np.random.default_rng(5)
    for n_q, k in ((2, 11), (16, 51), (37, 1), (64, 11), (70, 11), (5, 1024)):
        queries = rng.integers(0, small_table.shape[0], size=n_q).astype(np.int32)
        _same_as_single(small_engine, queries, k)


def test_batch_matches_oracle(small_engine, small_table):
    queries = np.array([0, 123, 49_999, 31337, 123], dtype=np.int32)    # a repeated query is allowed
    for k in (11, 200):
        for q, (rows, sims) in zip(queries, small_engine.most_similar_batch(queries, k)):
            _check(rows, sims, O.most_similar_dists(small_table, int(q)), k, int(q))


def test_batch_other_widths_and_small_tables():
    from cassmantle_b200.engine import ScoringEngine
    rng = np.random.default_rng(11)
    for n_rows, dim in ((300, 300), (5000, 64), (4097, 128), (2000, 200), (1500, 320), (900, 324), (40, 300), (500, 16)):
        vec = rng.standard_normal((n_rows, dim)).astype(np.float32)
        vec[7] = vec[3]                                              # a duplicated row: an exact tie
        with ScoringEngine.from_vectors(vec) as eng:
            queries = np.array([3, 7, n_rows - 1, 0, 11], dtype=np.int32)
            for k in (1, 10, min(300, n_rows)):
                _same_as_single(eng, queries, k)
                for q, (rows, sims) in zip(queries, eng.most_similar_batch(queries, k)):
                    # rows 3 and 7 are copies: at k = 1 either of them is "the" best row
                    _check(rows, sims, O.most_similar_dists(vec, int(q)), k, 3 if int(q) == 7 and k == 1 else int(q))


def test_batch_special_rows_fall_back():
    """NaN rows, zero rows and a crowd of duplicates: candidates overflow or run short and the query is finished
    by the one-query path -- results still equal it."""
    from cassmantle_b200.engine import ScoringEngine
    rng = np.random.default_rng(3)
    vec = rng.standard_normal((20_000, 300)).astype(np.float32)
    vec[100] = 0.0
    vec[200, 5] = np.nan
    vec[1000:11_000] = vec[1000]                                     # 10,000 copies: more candidates than the cap
    with ScoringEngine.from_vectors(vec) as eng:
        queries = np.array([1000, 5, 100, 6, 200], dtype=np.int32)
        for k in (3, 50):
            batch = eng.most_similar_batch(queries, k)
            for q, (rows, sims) in zip(queries, batch):
                want_rows, want_sims = eng.most_similar(int(q), k)
                assert np.array_equal(sims, want_sims, equal_nan=True), (q, k)


def test_batch_argument_errors(small_engine):
    from cassmantle_b200._lib import CassMantleError
    with pytest.raises(CassMantleError):
        small_engine.most_similar_batch([0, 10**9], 5)
    with pytest.raises(CassMantleError):
        small_engine.most_similar_batch([0, 1], 0)
    assert small_engine.most_similar_batch([], 5) == []


def test_backend_most_similar_many(small_engine, small_table):
    from cassmantle_b200.backend import Backend
    from cassmantle_b200.vocab import Vocabulary
    keys = synth.make_vocab(small_table.shape[0])
    be = Backend(small_engine, Vocabulary(keys), min_score=0.01)
    words = [keys[5].upper(), keys[777], keys[4242]]
    many = be.most_similar_many(words, topn=20)
    assert many == [be.most_similar(w, topn=20) for w in words]
    with pytest.raises(Exception):
        be.most_similar_many(["no-such-word"], topn=5)
