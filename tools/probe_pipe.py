"""Developer probe: per-step cost of the pipelined host path for several shapes and depths."""
import ctypes, sys, time
import numpy as np
sys.path.insert(0, ".")
from cassmantle_b200 import _lib, synth
from cassmantle_b200.engine import ScoringEngine, ScoringPipe
lib = _lib.load()
V = 3_000_000
eng = ScoringEngine.synthetic(V, 300, 0, 0)
ms32 = ctypes.c_float(np.float32(0.01))
import os
SHAPES = [tuple(int(x) for x in a.split('x')) for a in sys.argv[1:]] or [(64, 1), (4096, 1), (4096, 16)]
for (B, N) in SHAPES:
    for depth in [int(x) for x in os.environ.get('DEPTHS', '2,4,8').split(',')]:
        RING = 32
        pipe = ScoringPipe(eng, depth=RING, max_guesses=B, max_slots=N)
        ans = None
        for s in range(RING):
            idx, ln, a, slot, ex = synth.make_guess_batch(B, N, V, 2, seed=s, near_frac=0.0)
            ans = a if ans is None else ans
            v = pipe.buffers(s, B, N)
            v["guess_idx"][:] = idx; v["guess_len"][:] = ln; v["answer_slot"][:] = slot; v["exact"][:] = ex
        eng.set_answers(ans); eng.synchronize()
        cyc, sub, wait, ph = lib.cm_pipe_cycle, lib.cm_pipe_submit, lib.cm_pipe_wait, pipe._h
        K = 3000 if B * N < 500000 else 500
        for i in range(depth): sub(ph, i % RING, ms32, 0)
        t0 = time.perf_counter()
        for i in range(depth, depth + K):
            cyc(ph, (i - depth) % RING, i % RING, ms32, 0)
        dt = time.perf_counter() - t0
        for i in range(K, K + depth): wait(ph, i % RING)
        print(f"B={B} N={N} depth={depth}: {dt / K * 1e6:.2f} us per step", flush=True)
        pipe.close()
