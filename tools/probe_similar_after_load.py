"""Developer probe: is one-query most_similar slower right after a long scoring run (power cap)?"""
import ctypes, subprocess, sys, time
import numpy as np
import torch
sys.path.insert(0, ".")
from cassmantle_b200 import _lib, synth
from cassmantle_b200.engine import ScoringEngine

V = synth.FULL_ROWS
lib = _lib.load()
eng = ScoringEngine.synthetic(V, 300, 0, 0)
rng = np.random.default_rng(9)

def clocks():
    return subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,clocks.mem,power.draw", "--format=csv,noheader"],
                          capture_output=True, text=True).stdout.strip()

def time_similar(tag):
    ts = []
    for q in rng.integers(0, V, size=30):
        t = time.perf_counter(); eng.most_similar(int(q), 51); ts.append(time.perf_counter() - t)
    print(f"{tag}: median {np.median(ts) * 1e6:.0f} us  min {np.min(ts) * 1e6:.0f}  [{clocks()}]", flush=True)

eng.most_similar(1, 51)
time_similar("cold")
dev = torch.device("cuda:0")
B, N = 4096, 16
idx, ln, ans, slot, ex = synth.make_guess_batch(B, N, V, 2, seed=1, near_frac=0.0)
t = [torch.from_numpy(x).to(dev) for x in (idx, ln, slot, ex)]
eng.set_answers(ans)
out = torch.empty(B, dtype=torch.float32, device=dev)
ms = ctypes.c_float(np.float32(0.01))
t0 = time.perf_counter()
while time.perf_counter() - t0 < 3.0:
    for _ in range(2000):
        lib.cm_score_batch_device(eng._h, t[0].data_ptr(), t[1].data_ptr(), B, N, t[2].data_ptr(), t[3].data_ptr(), ms, out.data_ptr(), None, 1, None)
    torch.cuda.synchronize()
print("after 3 s of scoring:", clocks(), flush=True)
time_similar("hot")
time.sleep(1.0)
time_similar("1 s later")
eng.close()
