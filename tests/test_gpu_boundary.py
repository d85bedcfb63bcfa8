"""GPU tests of the reference-shaped Python boundary (Backend + MicroBatcher) on a real vocabulary."""
import asyncio

import numpy as np
import pytest

from cassmantle_b200 import synth
from oracle import scorer_oracle as O
from tests.util import COS_TOL

pytestmark = pytest.mark.gpu
V = 50_000


@pytest.fixture(scope="module")
def stack(small_engine, small_table):
    from cassmantle_b200.backend import Backend
    from cassmantle_b200.vocab import Vocabulary
    keys = synth.make_vocab(V)
    be = Backend(small_engine, Vocabulary(keys), min_score=0.01)
    ref = O.OracleBackend(O.OracleWordVectors(small_table, keys), 0.01)
    return be, ref, keys


def _requests(keys, n_req, rng):
    answers = [keys[123], keys[4567]]
    reqs = []
    for _ in range(n_req):
        d = {}
        for m, a in zip(("3", "11"), answers):
            r = rng.random()
            if r < 0.05:
                g = "zz-not-a-word"
            elif r < 0.10:
                g = a.upper()                       # equal after lower() -> "1.0"
            else:
                g = keys[int(rng.integers(0, V))].capitalize()
            d[m] = {"input": g, "answer": a}
        reqs.append(d)
    return reqs


def _same(got: dict, want: dict):
    assert list(got) == list(want)
    for k in got:
        if want[k] in ("1.0", "0.01"):
            assert got[k] == want[k]
        else:
            assert abs(float(got[k]) - float(want[k])) <= COS_TOL


def test_compute_scores_against_oracle_backend(stack):
    be, ref, keys = stack
    rng = np.random.default_rng(0)
    for d in _requests(keys, 50, rng):
        _same(asyncio.run(be.compute_scores(d)), asyncio.run(ref.compute_scores(d)))


def test_micro_batched_traffic(stack):
    """configs[1]-style: many concurrent 2-word requests coalesced into a few device batches."""
    from cassmantle_b200.server import MicroBatcher
    be, ref, keys = stack
    reqs = _requests(keys, 512, np.random.default_rng(1))

    async def main():
        mb = MicroBatcher(be, max_batch=512, max_delay_s=0.005)
        outs = await asyncio.gather(*[mb.compute_scores(r) for r in reqs])
        mb.close()
        return mb, outs

    mb, outs = asyncio.run(main())
    assert mb.batches <= 3 and mb.requests == 512
    for r, o in zip(reqs, outs):
        _same(o, asyncio.run(ref.compute_scores(r)))
    won = O.client_result(outs[0])
    assert set(won) == {"mean", "won"}
