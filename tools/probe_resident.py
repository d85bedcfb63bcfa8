"""Developer probe: latency of cm_score_batch_host with the resident kernel armed, with its trace."""
import ctypes, os, sys, time
import numpy as np
os.environ.setdefault("CM_RESIDENT_TRACE", "1")
sys.path.insert(0, ".")
from cassmantle_b200 import synth
from cassmantle_b200.engine import ScoringEngine

V = synth.FULL_ROWS
eng = ScoringEngine.synthetic(V, 300, 0, 0)
lib = eng._lib
shapes = [(512, 8), (64, 1), (2, 1), (1024, 1), (1024, 4)]
if len(sys.argv) > 1:
    shapes = [tuple(int(x) for x in a.split("x")) for a in sys.argv[1:]]
for B, N in shapes:
    idx, ln, ans, slot, ex = synth.make_guess_batch(B, N, V, 2, seed=4242)
    eng.set_answers(ans); eng.synchronize()
    out = np.empty(B, np.float32)
    ms = ctypes.c_float(np.float32(0.01))
    def call():
        lib.cm_score_batch_host(eng._h, idx.ctypes.data, ln.ctypes.data, B, N, slot.ctypes.data, ex.ctypes.data, ms, out.ctypes.data, None)
    for armed in (False, True):
        if armed: eng.arm_resident(idle_ms=100)
        ts = []
        for i in range(600):
            t0 = time.perf_counter(); call(); ts.append(time.perf_counter() - t0)
        ts = np.sort(np.array(ts[100:])) * 1e6
        line = f"{B}x{N} armed={armed}: p50 {ts[len(ts)//2]:.1f} us  p99 {ts[int(len(ts)*.99)]:.1f} us  min {ts[0]:.1f}"
        if armed:
            tr = np.zeros(7); lib.cm_engine_resident_trace(eng._h, tr.ctypes.data)
            line += (f" | device: flag +{(tr[1]-tr[0])/1e3:.2f} us, staged +{(tr[3]-tr[0])/1e3:.2f}, scored +{(tr[2]-tr[0])/1e3:.2f}"
                     f" | host: {tr[4]:.1f} us before ring, first result {tr[5]:.1f} us, all {tr[6]:.1f} us after ring")
            eng.disarm_resident()
        print(line, flush=True)
eng.close()
