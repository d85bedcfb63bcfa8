"""Consumers of tests/golden/scorer_gensim_golden.json -- outputs of the reference's own scorer
methods running on the REAL gensim (generator: tests/golden/make_gensim_golden.py, which needs a
machine with gensim; this box has none).

  * file present: the oracle (CPU) and the CUDA path (-m gpu) are compared with it -- that is the
    pin the oracle's header asks for;
  * file absent: those tests skip with the reason, and only the dry run below executes: it drives
    the generator with the oracle standing in for gensim and feeds the result to the same
    consumers, so the kit itself (generation, file format, comparison) is known to work.
"""
import asyncio
import importlib.util
import json
import os

import numpy as np
import pytest

from oracle import scorer_oracle as O
from tests.util import COS_TOL

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden", "scorer_gensim_golden.json")
REF = "/root/reference"

_spec = importlib.util.spec_from_file_location("make_gensim_golden", os.path.join(HERE, "golden", "make_gensim_golden.py"))
kit = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(kit)


def _load(path=GOLDEN):
    if not os.path.exists(path):
        pytest.skip("tests/golden/scorer_gensim_golden.json not generated yet: run "
                    "tests/golden/make_gensim_golden.py on a machine that has gensim")
    with open(path) as f:
        return json.load(f)


def _same_score(got, want: dict, where):
    """`got` is a scorer result, `want` its golden encoding: same Python type; the exact classes
    (1.0, min_score) equal as text; a cosine within COS_TOL of gensim's."""
    assert type(got).__name__ == want["type"], (where, type(got).__name__, want)
    if want["type"] == "float":                     # the Python floats 1.0 and min_score
        assert str(got) == want["str"], (where, got, want)
    else:
        assert abs(float(got) - float.fromhex(want["hex"])) <= COS_TOL, (where, got, want)


def check_backend_against(doc: dict, backend, most_similar_fn):
    """`backend` has the reference's compute_score / compute_scores; most_similar_fn(word, topn)."""
    worst = 0.0
    for (a, b), want in zip(doc["pairs"], doc["scores"]):
        got = backend.compute_score(a, b)
        _same_score(got, want, (a, b))
        worst = max(worst, abs(float(got) - float.fromhex(want["hex"])))
    for req in doc["requests"]:
        got = asyncio.run(backend.compute_scores(req["data"]))
        assert list(got) == list(req["result"])
        for k, s in req["result"].items():
            if s in ("1.0", str(doc["min_score"])):
                assert got[k] == s, (k, got[k], s)
            else:
                assert abs(float(got[k]) - float(s)) <= COS_TOL, (k, got[k], s)
    for q in doc["most_similar"]:
        got = most_similar_fn(q["word"], q["topn"])
        want = [(k, float.fromhex(h)) for k, h in q["result"]]
        assert len(got) == len(want)
        for (gk, gs), (wk, ws) in zip(got, want):
            assert abs(gs - ws) <= COS_TOL or (np.isnan(gs) and np.isnan(ws)), (q["word"], gk, wk, gs, ws)
        # the same words, up to swaps between neighbours whose cosines are within the tolerance
        for i, (gk, _) in enumerate(got):
            if gk != want[i][0]:
                near = [wk for wk, ws in want if abs(ws - got[i][1]) <= 2 * COS_TOL]
                assert gk in near or i >= len(got) - 2, (q["word"], i, gk, want[i])
    if doc["most_similar_unknown_word"] is not None:
        with pytest.raises(Exception) as exc:
            most_similar_fn("No-Such-Word", 5)
        assert type(exc.value).__name__ == doc["most_similar_unknown_word"]["type"]
        assert exc.value.args[0] == doc["most_similar_unknown_word"]["arg0"]
    return worst


def _oracle_stack():
    keys, table = kit.build_table()
    wv = O.OracleWordVectors(table, keys)
    be = O.OracleBackend(wv, kit.MIN_SCORE)

    def most_similar(word, topn):
        try:
            row = wv.key_to_index[word.lower()]
        except KeyError as e:
            raise Exception("Word not in dictionary. ", e)
        return [(keys[r], s) for r, s in O.most_similar_rows(table, row, topn)]

    return be, most_similar


def test_oracle_against_real_gensim_outputs():
    doc = _load()
    assert not doc["dry_run"], "a --dry-run file pins nothing and must not be committed as the golden file"
    assert doc["versions"]["gensim"]
    be, ms = _oracle_stack()
    worst = check_backend_against(doc, be, ms)
    print(f"oracle vs gensim {doc['versions']['gensim']}: max |cos diff| {worst:.2e} over {len(doc['pairs'])} pairs")


@pytest.mark.gpu
def test_cuda_path_against_real_gensim_outputs():
    doc = _load()
    assert not doc["dry_run"]
    from cassmantle_b200.backend import Backend
    from cassmantle_b200.engine import ScoringEngine
    from cassmantle_b200.vocab import Vocabulary
    keys, table = kit.build_table()
    with ScoringEngine.from_vectors(table) as eng:
        be = Backend(eng, Vocabulary(keys), min_score=kit.MIN_SCORE)
        check_backend_against(doc, be, lambda w, topn: be.most_similar(w, topn=topn))


@pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "src")), reason="reference tree not mounted")
def test_kit_dry_run_round_trip(tmp_path):
    """The generator runs the reference's own methods (src/backend.py, unmodified) and the consumers
    accept its file -- with the oracle in gensim's place, so this proves the kit, not parity."""
    out = str(tmp_path / "dry.json")
    wv, versions = kit.oracle_word_vectors()
    doc = kit.generate(wv, REF, out, versions, dry_run=True)
    assert doc["dry_run"] and len(doc["pairs"]) > 3000 and len(doc["requests"]) == 200
    reread = _load(out)
    kinds = {s["str"] for s in reread["scores"] if s["type"] == "float"}
    assert kinds == {"1.0", "0.01"}                                  # both exact classes occur
    assert sum(s["type"] == "float32" for s in reread["scores"]) > 1000    # and plenty of real cosines
    be, ms = _oracle_stack()
    assert check_backend_against(reread, be, ms) == 0.0              # the oracle reproduces itself bit for bit


@pytest.mark.gpu
@pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "src")), reason="reference tree not mounted")
def test_kit_dry_run_against_cuda_path(tmp_path):
    out = str(tmp_path / "dry.json")
    wv, versions = kit.oracle_word_vectors()
    doc = kit.generate(wv, REF, out, versions, dry_run=True)
    from cassmantle_b200.backend import Backend
    from cassmantle_b200.engine import ScoringEngine
    from cassmantle_b200.vocab import Vocabulary
    keys, table = kit.build_table()
    with ScoringEngine.from_vectors(table) as eng:
        be = Backend(eng, Vocabulary(keys), min_score=kit.MIN_SCORE)
        check_backend_against(doc, be, lambda w, topn: be.most_similar(w, topn=topn))
