"""Developer probe: a few cm_most_similar_batch calls at full table size (for ncu launch lists / captures)."""
import sys
import numpy as np
sys.path.insert(0, ".")
from cassmantle_b200 import synth
from cassmantle_b200.engine import ScoringEngine

n_q = int(sys.argv[1]) if len(sys.argv) > 1 else 64
eng = ScoringEngine.synthetic(synth.FULL_ROWS, 300, 0, 0)
rng = np.random.default_rng(1)
for _ in range(3):
    got = eng.most_similar_batch(rng.integers(0, synth.FULL_ROWS, size=n_q).astype(np.int32), 51)
print(len(got), got[0][1][:3])
eng.close()
