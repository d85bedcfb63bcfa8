"""Measures the one-time cost of making the table resident (cm_engine_upload_rows) and checks it."""
import os, sys, time
import numpy as np
sys.path.insert(0, ".")
from cassmantle_b200 import synth
from cassmantle_b200.engine import ScoringEngine
from oracle import c_oracle
V = synth.FULL_ROWS
T = c_oracle.synth_fill(V, 300, 0, threads=os.cpu_count())
t0 = time.perf_counter(); eng = ScoringEngine.from_vectors(T); dt = time.perf_counter() - t0
print(f"engine create + upload of {T.nbytes/1e9:.1f} GB from pageable host memory: {dt:.2f} s = {T.nbytes/dt/1e9:.1f} GB/s")
from cassmantle_b200._lib import check
ts = []
for _ in range(3):                       # the upload alone, into the existing engine
    t0 = time.perf_counter()
    check(eng._lib.cm_engine_upload_rows(eng._h, T.ctypes.data, 0, V))
    ts.append(time.perf_counter() - t0)
print(f"cm_engine_upload_rows alone, best of 3: {min(ts):.2f} s = {T.nbytes/min(ts)/1e9:.1f} GB/s "
      f"(CM_UPLOAD_THREADS={os.environ.get('CM_UPLOAD_THREADS', 'default')})")
for r0 in (0, 1_234_567, V - 5):
    assert np.array_equal(eng.download_rows(r0, 5).view(np.uint32), T[r0:r0+5].view(np.uint32))
eng2 = ScoringEngine.synthetic(V, 300, 0)
assert np.array_equal(eng2.download_rows(2_000_000, 64).view(np.uint32), T[2_000_000:2_000_064].view(np.uint32))
print("uploaded table == device-filled synthetic table == host fill: ok")
