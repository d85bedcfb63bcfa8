"""Developer probe: CUDA-event timing of cm_score_batch_device on several shapes (not a test)."""
import ctypes
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from cassmantle_b200 import _lib, synth
from cassmantle_b200.engine import ScoringEngine

V, D = synth.FULL_ROWS, 300
shapes = [(4096, 16), (4096, 1), (512, 8), (65536, 16), (16384, 64), (65536, 1)]
REPS = (200, 2000, 20000)
import os
FLAGS = int(os.environ.get('CM_FLAGS', '0'))
args = [a for a in sys.argv[1:] if not a.startswith("reps=")]
for a in sys.argv[1:]:
    if a.startswith("reps="):
        REPS = tuple(int(x) for x in a[5:].split(","))
if args:
    shapes = [tuple(int(x) for x in s.split("x")) for s in args]
lib = _lib.load()
eng = ScoringEngine.synthetic(V, D, 0, 0)
dev = torch.device("cuda:0")
stream = torch.cuda.Stream(device=dev)
torch.cuda.set_stream(stream)
st = stream.cuda_stream
ms32 = ctypes.c_float(np.float32(0.01))
fn = lib.cm_score_batch_device
GATHER = int(os.environ.get('CM_PROBE_GATHER', '0'))
if GATHER:
    from cassmantle_b200.dist import ScoreGather
    gat = ScoreGather(eng, 65536, depth=4, rank=0, world=1)
    fg = lib.cm_score_batch_gather
    def fn(h, a, b, B, N, c, d, ms, o, ce, fl, st):
        return fg(h, gat._h, a, b, B, N, c, d, ms, o, ce, fl, st)
for B, N in shapes:
    ring = 8 if B * N > 1_000_000 else 64
    bufs = []
    ans = None
    for i in range(ring):
        idx, ln, a, slot, ex = synth.make_guess_batch(B, N, V, 2, seed=i, near_frac=0.0)
        ans = a if ans is None else ans
        bufs.append([torch.from_numpy(x).to(dev) for x in (idx, ln, slot, ex)])
    eng.set_answers(ans)
    out = torch.empty(B, dtype=torch.float32, device=dev)
    ptrs = [(b[0].data_ptr(), b[1].data_ptr(), b[2].data_ptr(), b[3].data_ptr()) for b in bufs]
    bytes_alg = B * (N * 1204 + 8)
    for reps in REPS:
        if reps * bytes_alg > 4e12:
            continue
        for i in range(20):
            p = ptrs[i % ring]
            fn(eng._h, p[0], p[1], B, N, p[2], p[3], ms32, out.data_ptr(), None, FLAGS, st)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record(stream)
        for i in range(reps):
            p = ptrs[i % ring]
            fn(eng._h, p[0], p[1], B, N, p[2], p[3], ms32, out.data_ptr(), None, FLAGS, st)
        e1.record(stream)
        t_enq = time.perf_counter() - t0
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / reps
        print(f"B={B} N={N} reps={reps}: {us:8.2f} us/launch  enqueue {t_enq / reps * 1e6:6.2f} us  "
              f"{bytes_alg / us / 1e3:8.1f} GB/s  frac {bytes_alg / us / 1e3 / 6539.9:.3f}", flush=True)
eng.close()
