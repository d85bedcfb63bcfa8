"""Developer probe (torchrun, one rank per GPU): us per 4096x16 step with and without the fused score
all-gather, for several ring depths (the peers are waited for with a lag of depth/2 steps)."""
import ctypes, os, sys
import numpy as np
import torch
import torch.distributed as dist
sys.path.insert(0, ".")
from cassmantle_b200 import _lib, synth
from cassmantle_b200.dist import ScoreGather
from cassmantle_b200.engine import ScoringEngine

rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
lib = _lib.load()
V, B, N, RING = synth.FULL_ROWS, 4096, 16, 32
eng = ScoringEngine.synthetic(V, 300, 0, local)
ans = np.random.default_rng(999).choice(V, size=2, replace=False).astype(np.int32)
eng.set_answers(ans); eng.synchronize()
bufs = []
for i in range(RING):
    idx, ln, _, slot, ex = synth.make_guess_batch(B, N, V, 2, seed=1000 + 7919 * rank + i, near_frac=0.0)
    bufs.append([torch.from_numpy(a).to(dev) for a in (idx, ln, slot, ex)])
out = torch.empty(B, dtype=torch.float32, device=dev)
stream = torch.cuda.Stream(device=dev); torch.cuda.set_stream(stream); st = stream.cuda_stream
ms = ctypes.c_float(np.float32(0.01))

def run(gather_h, flags, steps):
    arr = (_lib.cm_batch * (2 * RING))()
    for k in range(2 * RING):
        q = bufs[k % RING]
        arr[k] = _lib.cm_batch(q[0].data_ptr(), q[1].data_ptr(), q[2].data_ptr(), q[3].data_ptr(), out.data_ptr(), None, B, N)
    def go(n):
        i = 0
        while i < n:
            k = min(RING, n - i)
            rc = lib.cm_score_batch_list_device(eng._h, gather_h, arr, k, ms, flags, st)
            assert rc == 0, lib.cm_last_error()
            i += k
    go(200)
    dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream); go(steps); e1.record(stream)
    dist.barrier(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) * 1e3 / steps], dtype=torch.float64, device=dev)
    allt = [torch.zeros_like(t) for _ in range(world)]
    dist.all_gather(allt, t)
    return [round(float(x.item()), 2) for x in allt]

def regions(gather_h, flags, k, reps):
    """`reps` timed regions of exactly k steps, each behind a barrier, an aligning all-reduce and an L2-flushing read."""
    arr = (_lib.cm_batch * (2 * RING))()
    for j in range(2 * RING):
        q = bufs[j % RING]
        arr[j] = _lib.cm_batch(q[0].data_ptr(), q[1].data_ptr(), q[2].data_ptr(), q[3].data_ptr(), out.data_ptr(), None, B, N)
    flush = torch.zeros(512 << 20, dtype=torch.uint8, device=dev)
    sink = torch.zeros((), dtype=torch.int64, device=dev)
    al = torch.zeros(1, device=dev)
    outl = []
    for r in range(reps + 2):
        dist.barrier(); torch.cuda.synchronize()
        dist.all_reduce(al)
        sink.copy_(torch.count_nonzero(flush))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        rc = lib.cm_score_batch_list_device(eng._h, gather_h, arr, k, ms, flags, st)
        assert rc == 0
        e1.record(stream)
        dist.barrier(); torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) * 1e3 / k], dtype=torch.float64, device=dev)
        allt = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(allt, t)
        if r >= 2:
            outl.append([round(float(x.item()), 2) for x in allt])
    return outl

if os.environ.get("CM_PROBE_REGIONS"):
    g = ScoreGather(eng, B, depth=8)
    g.exchange_handles()
    dist.barrier()
    for row in regions(g._h, 1, 20, 10):
        if rank == 0: print(f"world {world}: 20-step region, us/step per rank {row} max {max(row)}", flush=True)
    g.wait(st); torch.cuda.synchronize(); dist.barrier(); g.close()
    for row in regions(None, 1, 20, 4):
        if rank == 0: print(f"world {world}: 20-step region, NO gather, us/step per rank {row} max {max(row)}", flush=True)
    eng.close(); dist.destroy_process_group(); sys.exit(0)

steps = 6000
res = run(None, 1, steps)
if rank == 0: print(f"world {world}: no gather: per rank {res} max {max(res)}", flush=True)
for depth in ((8,) if os.environ.get('CM_GATHER_TRACE') else (8, 16, 4)):
    for flags, tag in ((1, "wait lag depth/2"), (1 | _lib.CM_GATHER_DEFER_WAIT, "no peer wait in the ship")):
        g = ScoreGather(eng, B, depth=depth)
        g.exchange_handles()
        dist.barrier()
        res = run(g._h, flags, steps)
        g.wait(st); torch.cuda.synchronize()
        if rank == 0: print(f"world {world}: gather depth {depth}, {tag}: max {max(res)} us/step (timed out: {g.timed_out()})", flush=True)
        if rank == 0 and os.environ.get("CM_GATHER_TRACE") and flags == 1:
            tr = np.zeros(512, dtype=np.uint64)
            lib.cm_gather_trace(g._h, tr.ctypes.data)
            tr = tr.reshape(64, 8); tr = tr[np.argsort(tr[:, 0])]
            prev = None
            for row in tr[-10:]:
                b0 = int(row[1])
                print("  ship of step", int(row[0]), "start +%.2f us after the previous ship's start;" % ((b0 - prev) / 1e3 if prev else 0.0),
                      "own-order wait %.2f, staging+stores %.2f, fence %.2f, flags %.2f, peers %.2f us" % tuple(
                          (int(row[i + 1]) - int(row[i])) / 1e3 for i in range(1, 6)), flush=True)
                prev = b0
        dist.barrier()
        g.close()
eng.close()
dist.destroy_process_group()
