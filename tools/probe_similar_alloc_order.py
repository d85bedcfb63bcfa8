"""Developer probe: one-query most_similar when its scratch is allocated after other allocations."""
import sys, time
import numpy as np
import torch
sys.path.insert(0, ".")
from cassmantle_b200 import synth
from cassmantle_b200.engine import ScoringEngine

mode = sys.argv[1]
V = synth.FULL_ROWS
eng = ScoringEngine.synthetic(V, 300, 0, 0)
keep = []
dev = torch.device("cuda:0")
if "dev_small" in mode:
    keep += [torch.empty(300 * 1024, dtype=torch.uint8, device=dev) for _ in range(256)]
if "dev_big" in mode:
    keep += [torch.empty(512 << 20, dtype=torch.uint8, device=dev)]
    x = torch.empty(2 << 30, dtype=torch.uint8, device=dev); del x
if "pinned" in mode:
    keep += [torch.empty(64 << 20, dtype=torch.uint8, pin_memory=True) for _ in range(4)]
if "pipes" in mode:
    from cassmantle_b200.engine import ScoringPipe
    keep += [ScoringPipe(eng, depth=8), ScoringPipe(eng, depth=8, packed=True)]
if "resident" in mode:
    eng.arm_resident(idle_ms=100); eng.disarm_resident()
torch.cuda.synchronize()
rng = np.random.default_rng(1)
eng.most_similar(1, 51)
ts = []
for q in rng.integers(0, V, size=30):
    t = time.perf_counter(); eng.most_similar(int(q), 51); ts.append(time.perf_counter() - t)
print(f"{mode}: median {np.median(ts) * 1e6:.0f} us", flush=True)
