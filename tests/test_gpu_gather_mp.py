"""GPU test (needs >= 2 GPUs, skipped otherwise): the NVLink score all-gather between real processes.

One process per GPU, CUDA-IPC handles exchanged over a gloo group, 25 gathered steps; every rank
must end up with every rank's scores of the last step, bit-identical to a local re-score."""
import os
import socket

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank: int, world: int, port: int, ret):
    import torch
    import torch.distributed as dist
    from cassmantle_b200 import synth
    from cassmantle_b200.dist import ScoreGather
    from cassmantle_b200.engine import ScoringEngine
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    V, B, N = 100_000, 3000, 6
    eng = ScoringEngine.synthetic(V, 300, seed=0, device=rank)
    ans = np.array([17, 4242], np.int32)
    eng.set_answers(ans)
    eng.synchronize()
    batches = [synth.make_guess_batch(B, N, V, 2, seed=50 + r, ragged=True) for r in range(world)]
    idx, ln, _, slot, ex = batches[rank]
    t = [torch.from_numpy(a).to(dev) for a in (idx, ln, slot, ex)]
    g = ScoreGather(eng, B, depth=4)
    g.exchange_handles()
    dist.barrier()
    stream = torch.cuda.Stream(device=dev)
    st = stream.cuda_stream
    for _ in range(25):
        g.score(t[0].data_ptr(), t[1].data_ptr(), B, N, t[2].data_ptr(), t[3].data_ptr(), 0.01, 0, st, flags=1)
    ptr = g.wait(st)
    stream.synchronize()
    got = g.as_tensor(ptr).cpu().numpy()
    ok = g.timed_out() == 0
    for r in range(world):
        i2, l2, _, s2, e2 = batches[r]
        want, _ = eng.score_host(i2, l2, s2, e2, 0.01)
        ok = ok and bool(np.array_equal(got[r, :B].view(np.uint32), want.view(np.uint32)))
    ret[rank] = ok
    dist.barrier()
    g.close()
    eng.close()
    dist.destroy_process_group()


def test_gather_between_processes():
    import torch
    import torch.multiprocessing as mp
    world = min(torch.cuda.device_count(), 4)
    if world < 2:
        pytest.skip("needs at least 2 GPUs")
    port = _free_port()
    ctx = mp.get_context("spawn")
    with ctx.Manager() as m:
        ret = m.dict()
        procs = [ctx.Process(target=_worker, args=(r, world, port, ret)) for r in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(180)
            assert p.exitcode == 0
        assert all(ret.get(r) is True for r in range(world)), dict(ret)
