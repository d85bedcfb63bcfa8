"""GPU parity tests of the device vocabulary (cm_vocab_*): bit-exact against a Python dict."""
import numpy as np
import pytest

from cassmantle_b200 import synth
from oracle import scorer_oracle as O
from tests.util import COS_TOL

pytestmark = pytest.mark.gpu
V = 50_000


@pytest.fixture(scope="module")
def vocab_stack(small_engine, small_table):
    from cassmantle_b200.vocab import DeviceVocabulary
    keys = synth.make_vocab(V)
    # the awkward corners of a real key list
    keys[10] = "Apple"                 # only reachable in this exact case -> never, queries are lowered
    keys[11] = "apple"
    keys[12] = keys[5]                 # duplicate key: the first occurrence (row 5) wins
    keys[13] = ""                      # the empty string is a legal dict key
    keys[14] = "x" * 200               # long key
    keys[15] = "café"             # non-ASCII key, already lower case
    keys[16] = "ÉCOLE"            # non-ASCII upper case: unreachable after lower()
    keys[17] = "new_york_times"
    keys[18] = "0th"
    dv = DeviceVocabulary(small_engine, keys)
    yield dv, keys
    dv.close()


def test_lookup_matches_python_dict(vocab_stack):
    dv, keys = vocab_stack
    rng = np.random.default_rng(0)
    words = []
    for _ in range(20_000):
        r = rng.random()
        k = keys[int(rng.integers(0, V))]
        if r < 0.5:
            words.append(k)
        elif r < 0.7:
            words.append(k.upper())
        elif r < 0.8:
            words.append(k.capitalize())
        elif r < 0.9:
            words.append(k + "q")                      # almost a key
        else:
            words.append("zz" + str(int(rng.integers(0, 10 ** 6))))
    words += ["Apple", "APPLE", "apple", keys[5], keys[5].upper(), "", "X" * 200, "x" * 199, "x" * 201,
              "café", "CAFÉ", "École", "école", "New_York_Times", "0TH", " ", "a b"]
    got = dv.lookup_many(words)
    want = O.lookup_rows(keys, words)
    assert np.array_equal(got, want)
    assert got[words.index("APPLE")] == 11 and got[words.index(keys[5].upper(), 20_000)] == 5
    assert got[words.index("", 20_000)] == 13
    assert dv.lookup_many([]).shape == (0,)


def test_lookup_device_entry_point(vocab_stack):
    import torch
    from cassmantle_b200 import _lib
    from cassmantle_b200.vocab import pack_words
    dv, keys = vocab_stack
    words = [keys[i].upper() for i in range(0, V, 7)] + ["nope"] * 5
    data, offsets = pack_words(words)
    dev = torch.device("cuda:0")
    d_b, d_o = torch.from_numpy(data.copy()).to(dev), torch.from_numpy(offsets).to(dev)
    d_r = torch.empty(len(words), dtype=torch.int32, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    _lib.check(dv._lib.cm_vocab_lookup_device(dv._h, d_b.data_ptr(), d_o.data_ptr(), len(words),
                                              d_r.data_ptr(), st or None))
    torch.cuda.synchronize()
    assert np.array_equal(d_r.cpu().numpy(), O.lookup_rows(keys, words))


def test_score_words_is_compute_scores(vocab_stack, small_table):
    """Strings in, scores out: the whole of Backend.compute_scores on the device."""
    dv, keys = vocab_stack
    ref = O.OracleBackend(O.OracleWordVectors(small_table, keys), 0.01)
    rng = np.random.default_rng(1)
    answers = [keys[123], keys[4567].upper(), "not-a-word"]
    guesses, slots = [], []
    for _ in range(3000):
        s = int(rng.integers(0, 3))
        r = rng.random()
        if r < 0.05:
            g = answers[s].swapcase()                  # equal after lower() -> 1.0 (even if unknown)
        elif r < 0.10:
            g = "qq" + str(int(rng.integers(0, 999)))
        else:
            g = keys[int(rng.integers(0, V))].capitalize()
        guesses.append(g)
        slots.append(s)
    got = dv.score_words(guesses, answers, slots, 0.01)
    for g, s, x in zip(guesses, slots, got):
        y = ref.compute_score(g, answers[s])
        if isinstance(y, float):                       # exact classes: 1.0 and the floor
            assert x == np.float32(y), (g, answers[s], x, y)
        else:
            assert abs(float(x) - float(y)) <= COS_TOL, (g, answers[s], x, y)
    # a second call with other answers must not see the first one's
    got2 = dv.score_words([keys[9], keys[9]], [keys[9], keys[8]], [0, 1], 0.01)
    assert got2[0] == np.float32(1.0)
    assert abs(float(got2[1]) - float(ref.compute_score(keys[9], keys[8]))) <= COS_TOL


def test_large_vocabulary_every_key_found():
    """1M keys: every key maps to its own row (or its first duplicate), bit for bit."""
    from cassmantle_b200.engine import ScoringEngine
    from cassmantle_b200.vocab import DeviceVocabulary
    n = 1_000_000
    keys = [f"w{i:x}" for i in range(n)]
    with ScoringEngine.synthetic(1000, 300, seed=0) as eng:     # the table is irrelevant here
        dv = DeviceVocabulary(eng, keys)
        got = dv.lookup_many(keys[::3] + [k.upper() for k in keys[1::5]] + ["w", "wg", "W" + "f" * 9])
        want = np.r_[np.arange(0, n, 3), np.arange(1, n, 5), [-1, -1, -1]].astype(np.int32)
        assert np.array_equal(got, want)
        dv.close()


def test_backend_and_device_vocabulary_share_an_engine(vocab_stack, small_engine, small_table):
    """Both register answers with the engine; interleaving them must not leave stale slots behind."""
    from cassmantle_b200.backend import Backend
    from cassmantle_b200.vocab import Vocabulary
    dv, keys = vocab_stack
    be = Backend(small_engine, Vocabulary(keys), min_score=0.01)
    ref = O.OracleBackend(O.OracleWordVectors(small_table, keys), 0.01)
    for i in range(3):
        a = float(be.compute_score(keys[100 + i], keys[200]))
        dv.score_words([keys[1]], [keys[300 + i]], [0], 0.01)          # re-registers other answers
        b = float(be.compute_score(keys[100 + i], keys[200]))
        want = float(ref.compute_score(keys[100 + i], keys[200]))
        assert a == b and abs(a - want) <= COS_TOL
