"""CPU tests of the multi-rank path: world_size-2 (and 3) gloo process groups."""
import os
import socket

import numpy as np
import pytest
import torch.multiprocessing as mp

from cassmantle_b200 import synth
from cassmantle_b200.dist import shard_bounds


def test_shard_bounds_partition():
    for n in (0, 1, 7, 4096, 65536, 65537):
        for world in (1, 2, 3, 8):
            spans = [shard_bounds(n, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_bounds(10, 2, 2)


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank: int, world: int, port: int, B: int, N: int, ret):
    import torch.distributed as dist
    from cassmantle_b200.dist import sharded_score
    from oracle import scorer_oracle as O
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    V = 5000
    table = synth.full_table(V, seed=0, threads=1)
    idx, ln, ans, slot, ex = synth.make_guess_batch(B, N, V, 2, seed=3, ragged=True)
    seen = []

    def score_fn(i, l, s, e, ms):                 # stand-in for engine.score_host on this rank
        seen.append(len(i))
        return O.score_batch_vectorised(table, i, l, ans, s, e, ms)

    out = sharded_score(score_fn, idx, ln, slot, ex, 0.01)
    want = O.score_batch_vectorised(table, idx, ln, ans, slot, ex, 0.01)
    lo, hi = shard_bounds(B, world, rank)
    ret[rank] = (bool(np.array_equal(out, want)), seen == [hi - lo])
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,B,N", [(2, 1001, 4), (2, 64, 1), (3, 10, 2), (2, 1, 1)])
def test_sharded_score_gloo(world, B, N):
    port = _free_port()
    ctx = mp.get_context("spawn")
    with ctx.Manager() as m:
        ret = m.dict()
        procs = [ctx.Process(target=_worker, args=(r, world, port, B, N, ret)) for r in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(120)
            assert p.exitcode == 0
        assert len(ret) == world
        for r in range(world):
            assert ret[r] == (True, True), (r, ret[r])
