"""Developer probe: capture a run of cm_score_batch_device launches in a CUDA graph and replay it."""
import ctypes, sys, time
import numpy as np, torch
sys.path.insert(0, ".")
from cassmantle_b200 import _lib, synth
from cassmantle_b200.engine import ScoringEngine
lib = _lib.load()
V = 3_000_000
eng = ScoringEngine.synthetic(V, 300, 0, 0)
dev = torch.device("cuda:0")
ms32 = ctypes.c_float(np.float32(0.01))
fn = lib.cm_score_batch_device
for (B, N) in [(4096, 1), (512, 8), (4096, 16)]:
    K = 32
    bufs, outs, wants = [], [], []
    ans = None
    for i in range(K):
        idx, ln, a, slot, ex = synth.make_guess_batch(B, N, V, 2, seed=i, near_frac=0.0)
        ans = a if ans is None else ans
        bufs.append([torch.from_numpy(x).to(dev) for x in (idx, ln, slot, ex)])
        outs.append(torch.zeros(B, dtype=torch.float32, device=dev))
    eng.set_answers(ans); eng.synchronize()
    for i in range(K):
        idx, ln, a, slot, ex = synth.make_guess_batch(B, N, V, 2, seed=i, near_frac=0.0)
        wants.append(eng.score_host(idx, ln, slot, ex, 0.01)[0])
    s = torch.cuda.Stream(device=dev)
    g = torch.cuda.CUDAGraph()
    for flags in (0, 1):
        try:
            with torch.cuda.stream(s):
                g = torch.cuda.CUDAGraph()
                g.capture_begin()
                st = torch.cuda.current_stream().cuda_stream
                for i in range(K):
                    b = bufs[i]
                    rc = fn(eng._h, b[0].data_ptr(), b[1].data_ptr(), B, N, b[2].data_ptr(), b[3].data_ptr(), ms32,
                            outs[i].data_ptr(), None, flags, st)
                    assert rc == 0, lib.cm_last_error()
                g.capture_end()
            for o in outs: o.zero_()
            g.replay(); torch.cuda.synchronize()
            ok = all(np.array_equal(outs[i].cpu().numpy().view(np.uint32), wants[i].view(np.uint32)) for i in range(K))
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            with torch.cuda.stream(s):
                e0.record(s)
                for _ in range(50): g.replay()
                e1.record(s)
            torch.cuda.synchronize()
            print(f"B={B} N={N} flags={flags}: graph of {K} launches replayed: ok={ok}, {e0.elapsed_time(e1)*1e3/50/K:.2f} us per step", flush=True)
        except Exception as exc:
            print(f"B={B} N={N} flags={flags}: capture failed: {exc!r}", flush=True)
            torch.cuda.synchronize()
