"""CPU tests: the oracle against its known-answer file, and the C restatement against numpy."""
import asyncio

import numpy as np
import pytest

from cassmantle_b200 import synth
from oracle import c_oracle, scorer_oracle as O
from tests.util import dense_kat_table

F32_EPS = 3e-7   # a few fp32 ulps at |x| <= 1: the BLAS dot order may differ between CPUs


def _backend(golden):
    table = dense_kat_table(golden)
    return O.OracleBackend(O.OracleWordVectors(table, golden["kat"]["keys"]), golden["min_score"])


def test_kat_exact_classes(golden):
    be = _backend(golden)
    for c in golden["kat"]["cases"]:
        s = be.compute_score(c["input"], c["answer"])
        assert (not isinstance(s, np.floating)) == c["is_python_float"], c
        if c["is_python_float"]:
            assert str(s) == c["str"], c          # "1.0" and "0.01" are exact classes
        else:
            assert abs(float(s) - float.fromhex(c["f32_hex"])) <= F32_EPS, c


def test_kat_closed_forms(golden):
    be = _backend(golden)
    assert be.compute_score("alpha", "delta") == np.float32(0.6)       # 3-4-5 triangle
    assert be.compute_score("Beta", "dELTA") == np.float32(0.8)
    assert abs(be.compute_score("ones", "quarter") - 0.5) < F32_EPS
    assert abs(be.compute_score("tiny", "huge") - 1.0) < F32_EPS       # needs the fp64 norm
    assert be.compute_score("alpha", "gamma") == 0.01                   # anti-parallel -> floor
    assert be.compute_score("alpha", "void") == 0.01                    # zero norm -> cos 0 -> floor
    assert be.compute_score("ZZZ", "zzz") == 1.0                        # equality precedes lookup
    assert be.compute_score("nosuchword", "alpha") == 0.01              # KeyError -> floor


def test_compute_scores_strings(golden):
    be = _backend(golden)
    data = {"3": {"input": "Alpha", "answer": "alpha"},
            "7": {"input": "alpha", "answer": "delta"},
            "1": {"input": "qqq", "answer": "delta"}}
    out = asyncio.run(be.compute_scores(data))
    assert list(out.keys()) == ["3", "7", "1"]
    assert out == {"3": "1.0", "7": "0.6", "1": "0.01"}
    assert O.client_result({"3": "1.0", "7": "1.0"})["won"] == 1
    assert O.client_result(out)["won"] == 0


def test_format_and_browser_rounding(golden):
    for c in golden["format"]:
        x = np.float32(float.fromhex(c["f32_hex"]))
        s = str(max(0.01, x))
        assert s == c["str"]
        assert O.placeholder(s) == c["display"]
        assert O.centi_score(s) == c["centi"]
    assert O.js_to_fixed2_of_percent("0.125") == "12.50"
    assert O.js_to_fixed2_of_percent("0.99999994") == "100.00"
    assert O.score_to_blur(1.0) == 0.0 and O.score_to_blur(0.0) == 15.0


def test_golden_synth_section(golden):
    sec = golden["synth"]
    T = synth.full_table(sec["n_rows"], seed=sec["table_seed"])
    for c in sec["cases"]:
        rows = np.asarray([c["rows"]], dtype=np.int32)
        got = O.score_batch_faithful(T, rows, np.asarray([c["len"]]), np.asarray([c["answer_row"]]),
                                     np.asarray([0]), np.asarray([c["exact"]], dtype=np.uint8), 0.01)[0]
        want = np.float32(float.fromhex(c["f32_hex"]))
        if want in (np.float32(0.01), np.float32(1.0)):
            assert got == want, c
        else:
            assert abs(float(got) - float(want)) <= F32_EPS, c


def test_n1_is_similarity_bit_for_bit():
    T = synth.full_table(2000, seed=3)
    for i, j in [(0, 1), (5, 1999), (17, 17), (100, 200)]:
        assert O.n_similarity(T, [i], [j]) == O.similarity(T, i, j)


def test_mean_is_in_order_fp32_sum():
    """numpy mean(axis=0) on fp32 rows == running fp32 sum then one fp32 divide."""
    T = synth.full_table(500, seed=1)
    rows = [3, 77, 401, 9, 250, 13, 499]
    acc = np.zeros(300, dtype=np.float32)
    for r in rows:
        acc = acc + T[r]
    assert np.array_equal(np.array([T[r] for r in rows]).mean(axis=0), acc / np.float32(len(rows)))


@pytest.mark.parametrize("seq,ragged", [(1, False), (4, True), (16, False), (33, True)])
def test_c_oracle_matches_numpy(seq, ragged):
    V = 20_000
    T = synth.full_table(V, seed=0)
    idx, ln, ans, slot, ex = synth.make_guess_batch(600, seq, V, 3, seed=seq, ragged=ragged)
    a = O.score_batch_faithful(T, idx, ln, ans, slot, ex, 0.01)
    v = O.score_batch_vectorised(T, idx, ln, ans, slot, ex, 0.01)
    c1 = c_oracle.score_batch(T, idx, ln, ans, slot, ex, 0.01, threads=1)
    c4 = c_oracle.score_batch(T, idx, ln, ans, slot, ex, 0.01, threads=4)
    assert np.array_equal(c1, c4)
    for other in (v, c1):
        assert np.all(other[ex.astype(bool)] == np.float32(1.0))
        assert np.abs(a - other).max() <= 1.5e-6
        assert ((a == np.float32(0.01)) != (other == np.float32(0.01))).sum() == 0


def test_oov_and_empty_guess_floor():
    T = synth.full_table(100, seed=0)
    idx = np.array([[5, -1, 7], [-1, -1, -1], [1, 2, 3], [4, 5, 6]], dtype=np.int32)
    ln = np.array([3, 0, 3, 2], dtype=np.int32)
    ans = np.array([9, -1], dtype=np.int32)
    slot = np.array([0, 0, 1, 0], dtype=np.int32)
    for fn in (O.score_batch_faithful, O.score_batch_vectorised,
               lambda *a: c_oracle.score_batch(*a)):
        out = fn(T, idx, ln, ans, slot, None, 0.01)
        assert out[0] == np.float32(0.01) and out[1] == np.float32(0.01) and out[2] == np.float32(0.01)
        assert out[3] >= np.float32(0.01)


def test_synth_generator_is_pinned():
    """The generator is part of the parity contract with the CUDA fill kernel: pin its bits."""
    r = synth.table_rows([0, 1, 2_999_999], seed=0)
    assert r.dtype == np.float32 and r.shape == (3, 300)
    sample = [float(r[0, 0]).hex(), float(r[1, 299]).hex(), float(r[2, 150]).hex()]
    assert sample == PINNED_SAMPLE, sample
    keys = synth.make_vocab(2000)
    assert len(set(keys)) == 2000 and all(k == k.lower() and k.isalnum() for k in keys)


PINNED_SAMPLE = ['0x1.75646a0000000p-6', '-0x1.070c6e0000000p-9', '-0x1.7195620000000p-2']


def test_c_synth_fill_is_bit_identical_to_numpy():
    a = c_oracle.synth_fill(70_001, 300, seed=0, threads=3)
    rows = np.r_[0:50, 30_000:30_050, 69_990:70_001]
    assert np.array_equal(a[rows].view(np.uint32), synth.table_rows(rows, 0).view(np.uint32))
    b = c_oracle.synth_fill(1000, 300, seed=9, threads=1)
    assert np.array_equal(b.view(np.uint32), synth.table_rows(np.arange(1000), 9).view(np.uint32))


# ---- properties of the restated scorer (hypothesis) -------------------------------------------
from hypothesis import given, settings, strategies as st   # noqa: E402


@settings(max_examples=60, deadline=None)
@given(st.integers(0, 1999), st.integers(0, 1999), st.floats(0.01, 100.0))
def test_similarity_properties(i, j, scale):
    T = _prop_table()
    s = float(O.similarity(T, i, j))
    assert -1.0 - 1e-6 <= s <= 1.0 + 1e-6
    assert s == float(O.similarity(T, j, i))                      # the fp32 dot is symmetric
    T2 = T.copy()
    T2[i] = (T2[i] * np.float32(scale)).astype(np.float32)        # cosine ignores the row's length
    assert abs(float(O.similarity(T2, i, j)) - s) <= 5e-7
    if i == j:
        assert abs(s - 1.0) <= 5e-7


@settings(max_examples=40, deadline=None)
@given(st.lists(st.integers(0, 1999), min_size=1, max_size=12), st.integers(0, 1999), st.floats(-0.5, 0.9))
def test_batch_scorers_agree_and_respect_the_floor(rows, ans, floor):
    T = _prop_table()
    idx = np.asarray([rows], dtype=np.int32)
    ln = np.asarray([len(rows)], dtype=np.int32)
    args = (T, idx, ln, np.asarray([ans], np.int32), np.asarray([0], np.int32), None, floor)
    a = O.score_batch_faithful(*args)[0]
    v = O.score_batch_vectorised(*args)[0]
    c = c_oracle.score_batch(*args)[0]
    f32 = np.float32(floor)
    assert a >= f32 and v >= f32 and c >= f32
    for other in (v, c):
        assert abs(float(a) - float(other)) <= 1.5e-6 or (a == f32) != (other == f32) and abs(float(max(a, other)) - floor) <= 1.5e-6


_PROP = {}


def _prop_table():
    if "t" not in _PROP:
        _PROP["t"] = synth.full_table(2000, seed=11, threads=2)
    return _PROP["t"]
