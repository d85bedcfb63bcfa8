"""CPU tests of the request micro-batcher (cassmantle_b200/server.py).

The scorer injected here is the oracle Backend (test double); on a GPU box the same class wraps
cassmantle_b200.backend.Backend (tests/test_gpu_boundary.py)."""
import asyncio

import pytest

from cassmantle_b200.server import MicroBatcher
from oracle import scorer_oracle as O
from tests.util import dense_kat_table


class CountingScorer:
    def __init__(self, backend):
        self.backend, self.calls, self.sizes = backend, 0, []

    def score_pairs(self, pairs):
        self.calls += 1
        self.sizes.append(len(pairs))
        return [self.backend.compute_score(g, a) for g, a in pairs]


def _scorer(golden):
    table = dense_kat_table(golden)
    be = O.OracleBackend(O.OracleWordVectors(table, golden["kat"]["keys"]), 0.01)
    return CountingScorer(be), be


def test_concurrent_requests_share_one_batch(golden):
    sc, ref = _scorer(golden)
    cases = golden["kat"]["cases"]

    async def main():
        mb = MicroBatcher(sc, max_batch=10_000, max_delay_s=0.01)
        reqs = [{"0": {"input": c["input"], "answer": c["answer"]},
                 "5": {"input": c["answer"], "answer": c["input"]}} for c in cases]
        outs = await asyncio.gather(*[mb.compute_scores(r) for r in reqs])
        mb.close()
        return mb, reqs, outs

    mb, reqs, outs = asyncio.run(main())
    assert sc.calls == 1 and mb.batches == 1 and mb.requests == len(reqs)
    assert sc.sizes == [2 * len(reqs)]
    for r, o in zip(reqs, outs):
        assert o == asyncio.run(ref.compute_scores(r))       # own keys, own order, same strings
        assert list(o) == ["0", "5"]


def test_flush_on_max_batch_and_on_deadline(golden):
    sc, _ = _scorer(golden)

    async def main():
        mb = MicroBatcher(sc, max_batch=4, max_delay_s=0.5)
        one = {"1": {"input": "alpha", "answer": "delta"}}
        t = asyncio.get_running_loop().time()
        await asyncio.gather(*[mb.compute_scores(one) for _ in range(8)])   # 2 full batches, no wait
        fast = asyncio.get_running_loop().time() - t
        mb2 = MicroBatcher(sc, max_batch=1000, max_delay_s=0.05)
        t = asyncio.get_running_loop().time()
        out = await mb2.compute_scores(one)                                  # alone: waits the deadline
        slow = asyncio.get_running_loop().time() - t
        mb.close(); mb2.close()
        return fast, slow, out

    fast, slow, out = asyncio.run(main())
    assert sc.sizes[:2] == [4, 4]
    assert fast < 0.4 and 0.04 <= slow < 0.5
    assert out == {"1": "0.6"}


def test_bad_payload_fails_only_its_own_request(golden):
    sc, _ = _scorer(golden)

    async def main():
        mb = MicroBatcher(sc, max_batch=100, max_delay_s=0.01)
        good = mb.compute_scores({"1": {"input": "alpha", "answer": "alpha"}})
        bad = mb.compute_scores({"1": {"answer": "alpha"}})
        res = await asyncio.gather(good, bad, return_exceptions=True)
        mb.close()
        return res

    res = asyncio.run(main())
    assert res[0] == {"1": "1.0"}
    assert isinstance(res[1], KeyError)


def test_scorer_failure_reaches_every_caller_of_that_batch():
    class Boom:
        def score_pairs(self, pairs):
            raise RuntimeError("device lost")

    async def main():
        mb = MicroBatcher(Boom(), max_batch=100, max_delay_s=0.01)
        r = await asyncio.gather(*[mb.compute_scores({"1": {"input": "a", "answer": "b"}})
                                   for _ in range(3)], return_exceptions=True)
        empty = await mb.compute_scores({})
        mb.close()
        return r, empty

    r, empty = asyncio.run(main())
    assert all(isinstance(x, RuntimeError) for x in r)
    assert empty == {}


def test_loop_stays_responsive_while_scoring(golden):
    """The reference blocks its event loop in compute_scores; the batcher must not."""
    import time

    class Slow:
        def score_pairs(self, pairs):
            time.sleep(0.15)
            return [1.0] * len(pairs)

    async def main():
        mb = MicroBatcher(Slow(), max_batch=1, max_delay_s=0.0)
        ticks = 0

        async def ticker():
            nonlocal ticks
            for _ in range(20):
                await asyncio.sleep(0.005)
                ticks += 1

        t = asyncio.create_task(ticker())
        out = await mb.compute_scores({"1": {"input": "x", "answer": "y"}})
        done_ticks = ticks
        await t
        mb.close()
        return out, done_ticks

    out, done_ticks = asyncio.run(main())
    assert out == {"1": "1.0"} and done_ticks >= 10
