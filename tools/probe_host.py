"""Developer probe: latency of cm_score_batch_host for pinned / pageable buffers (CM_HOST_MODE)."""
import ctypes, sys, time
import numpy as np, torch
sys.path.insert(0, ".")
from cassmantle_b200 import _lib, synth
from cassmantle_b200.engine import ScoringEngine
V = synth.FULL_ROWS
lib = _lib.load()
eng = ScoringEngine.synthetic(V, 300, 0, 0)
ms32 = ctypes.c_float(np.float32(0.01))
SHAPES = [tuple(int(x) for x in a.split('x')) for a in sys.argv[1:]] or [(4096, 16), (512, 8), (4096, 1), (64, 1), (65536, 16)]
for B, N in SHAPES:
    idx, ln, ans, slot, ex = synth.make_guess_batch(B, N, V, 2, seed=1, near_frac=0.0)
    eng.set_answers(ans)
    ref = None
    for pinned in (True, False):
        arrs = [torch.from_numpy(a) for a in (idx, ln, slot, ex)]
        out = torch.empty(B, dtype=torch.float32)
        if pinned:
            arrs = [a.pin_memory() for a in arrs]; out = out.pin_memory()
        a = [t.numpy() for t in arrs]; o = out.numpy()
        ts = []
        for i in range(300):
            t0 = time.perf_counter()
            rc = lib.cm_score_batch_host(eng._h, a[0].ctypes.data, a[1].ctypes.data, B, N, a[2].ctypes.data,
                                         a[3].ctypes.data, ms32, o.ctypes.data, None)
            ts.append(time.perf_counter() - t0)
            assert rc == 0
        if ref is None: ref = o.copy()
        assert np.array_equal(ref, o)
        ts = np.sort(np.array(ts[50:])) * 1e6
        print(f"B={B} N={N} pinned={pinned}: p50 {ts[len(ts)//2]:7.1f} us  p10 {ts[len(ts)//10]:7.1f}  p99 {ts[int(len(ts)*.99)]:7.1f}", flush=True)
eng.close()
