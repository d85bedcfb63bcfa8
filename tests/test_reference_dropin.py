"""CPU test: the reference's OWN server code drives this repo's scorer -- the drop-in claim.

Runs only where /root/reference is mounted (the build container; the GPU box has no copy, so it
is skipped there).  The reference's `src/server.py` is imported unmodified; the third-party
modules it cannot import here (aioredis, gensim, nltk -- none of them on the scoring path) are
replaced by empty stand-ins, Redis by an in-memory fake with the handful of commands
`Server.compute_client_scores` uses, and the two scorer methods of `Backend`
(src/backend.py:303-317) by `cassmantle_b200.backend.Backend`'s -- exactly the patch in
INTEGRATION.md §3.  The engine behind it is the oracle test double (no GPU here); on the GPU box
tests/test_gpu_boundary.py runs the same Backend on the CUDA engine.
"""
import asyncio
import json
import os
import sys
import types

import numpy as np
import pytest

from cassmantle_b200.backend import Backend as GpuScorer
from cassmantle_b200.vocab import Vocabulary
from oracle import scorer_oracle as O
from tests.test_backend_host import OracleEngine
from tests.util import dense_kat_table

REF = "/root/reference"
pytestmark = pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "src")), reason="reference tree not mounted")


class FakeRedis:
    """The Redis commands on the path of Server.compute_client_scores (src/server.py:58-94)."""

    def __init__(self):
        self.h = {}

    async def hget(self, name, key):
        v = self.h.get(name, {}).get(key)
        return None if v is None else (v if isinstance(v, bytes) else str(v).encode())

    async def hgetall(self, name):
        return {k.encode(): (v if isinstance(v, bytes) else str(v).encode()) for k, v in self.h.get(name, {}).items()}

    async def hset(self, name, key=None, value=None, mapping=None):
        d = self.h.setdefault(name, {})
        if mapping:
            d.update({str(k): v for k, v in mapping.items()})
        if key is not None:
            d[str(key)] = value

    async def expire(self, name, seconds):
        return True

    async def hincrby(self, name, key, n):
        d = self.h.setdefault(name, {})
        d[key] = int(d.get(key, 0)) + n


@pytest.fixture()
def reference_server(golden, monkeypatch):
    for name in ("aioredis", "aioredis.exceptions", "gensim", "gensim.models", "nltk"):
        monkeypatch.setitem(sys.modules, name, types.ModuleType(name))
    sys.modules["aioredis.exceptions"].LockError = type("LockError", (Exception,), {})
    sys.modules["aioredis"].Redis = object              # only used as a return annotation
    sys.modules["gensim.models"].KeyedVectors = object
    monkeypatch.syspath_prepend(REF)
    for m in [k for k in sys.modules if k == "src" or k.startswith("src.")]:
        monkeypatch.delitem(sys.modules, m)
    from src.server import Server                       # the reference's file, untouched
    srv = object.__new__(Server)                        # __init__ needs api_key.txt + the 3.6 GB file
    srv.min_score = 0.01                                # src/server.py:17
    srv.time_per_prompt = 600
    srv.redis_conn = FakeRedis()
    keys = golden["kat"]["keys"]
    table = dense_kat_table(golden)
    gpu = GpuScorer(OracleEngine(table), Vocabulary(keys), min_score=srv.min_score)
    srv.compute_score = gpu.compute_score
    srv.compute_scores = gpu.compute_scores
    ref = O.OracleBackend(O.OracleWordVectors(table, keys), 0.01)
    yield srv, ref
    for m in [k for k in sys.modules if k == "src" or k.startswith("src.")]:
        sys.modules.pop(m, None)


def test_reference_compute_client_scores_runs_on_this_scorer(reference_server):
    srv, ref = reference_server
    prompt = {"tokens": ["the", "alpha", "and", "the", "delta", "met", "mixed", "ones"], "masks": [1, 4]}

    async def play(inputs):
        await srv.redis_conn.hset("prompt", "current", json.dumps(prompt))
        await srv.reset_client("sess")                                  # src/server.py:34-40
        return await srv.compute_client_scores("sess", inputs)          # src/server.py:63-76

    got = asyncio.run(play({"1": "Delta", "4": "ALPHA"}))
    want = {"1": str(ref.compute_score("Delta", "alpha")), "4": str(ref.compute_score("ALPHA", "delta"))}
    assert {k: got[k] for k in ("1", "4")} == want == {"1": "0.6", "4": "0.6"}
    assert got["won"] == 0
    sess = srv.redis_conn.h["sess"]
    assert sess["1"] == "0.6" and sess["attempts"] == 1 and float(sess["max"]) == pytest.approx(0.6)

    got = asyncio.run(play({"1": "ALPHA", "4": "delta"}))               # both exact -> the round is won
    assert got == {"1": "1.0", "4": "1.0", "won": 1}
    got = asyncio.run(play({"1": "xyzzy", "4": "gamma"}))               # unknown word / orthogonal -> floor
    assert got == {"1": "0.01", "4": "0.01", "won": 0}

    # and the reference's own view of the session afterwards (src/server.py:96-123)
    async def view():
        await srv.redis_conn.hset("prompt", "current", json.dumps(prompt))
        return await srv.fetch_prompt_json("sess")
    out = asyncio.run(view())
    assert out["tokens"][1] == "*" and out["tokens"][4] == "*" and out["scores"]["1"] == "0.01"
