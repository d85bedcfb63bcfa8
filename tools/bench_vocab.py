"""Measures the device vocabulary lookup (cm_vocab_lookup_device) against the Python dict it replaces."""
import sys, time
import numpy as np, torch
sys.path.insert(0, ".")
from cassmantle_b200 import _lib
from cassmantle_b200.engine import ScoringEngine
from cassmantle_b200.vocab import DeviceVocabulary, Vocabulary, pack_words

n = 3_000_000
keys = [f"w{i:x}" for i in range(n)]
eng = ScoringEngine.synthetic(1000, 300, 0, 0)
t0 = time.perf_counter(); dv = DeviceVocabulary(eng, keys); t_build = time.perf_counter() - t0
rng = np.random.default_rng(0)
for B in (4096, 65536, 1_000_000):
    words = [keys[i].upper() if i % 3 else "zz%d" % i for i in rng.integers(0, n, B)]
    data, off = pack_words(words)
    dev = torch.device("cuda:0")
    d_b, d_o = torch.from_numpy(data.copy()).to(dev), torch.from_numpy(off).to(dev)
    d_r = torch.empty(B, dtype=torch.int32, device=dev)
    stream = torch.cuda.Stream(); st = stream.cuda_stream
    fn = dv._lib.cm_vocab_lookup_device
    for _ in range(5): fn(dv._h, d_b.data_ptr(), d_o.data_ptr(), B, d_r.data_ptr(), st)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 200 if B < 10 ** 6 else 20
    with torch.cuda.stream(stream):
        e0.record(stream)
        for _ in range(reps): fn(dv._h, d_b.data_ptr(), d_o.data_ptr(), B, d_r.data_ptr(), st)
        e1.record(stream)
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / reps
    pv = Vocabulary(keys) if B == 4096 else pv
    t0 = time.perf_counter(); rows = [pv.key_to_index.get(w.lower(), -1) for w in words]; t_py = time.perf_counter() - t0
    assert np.array_equal(d_r.cpu().numpy(), np.asarray(rows, dtype=np.int32))
    print(f"B={B}: device {us:.1f} us/launch = {B / us:.1f} M lookups/s ; python dict {B / t_py / 1e6:.2f} M lookups/s (1 core)", flush=True)
print(f"build (host C++ insert + upload, 3M keys): {t_build:.2f} s incl. Python encoding")
