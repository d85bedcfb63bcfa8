"""GPU test of cm_score_batch_noalloc, the entry point for callers that may not allocate.

Some users of this library cannot let it call cudaMalloc: their GPU memory is owned by their
own allocator or fixed at start-up, or their request path cannot take cudaMalloc's device-wide
synchronisation.  For them the library offers one call that takes the table, the answers, the
batch and the outputs from the caller and keeps its only scratch -- the answers' unit vectors --
in a fixed 256 KiB array inside the library.

The test drives that call through the C ABI (ctypes) and compares it
  * with the ordinary engine path (cm_set_answers + cm_score_batch_device) on the same inputs:
    the two run the same kernels, so the results must agree bit for bit, and
  * with the CPU oracle directly, within the tolerance of the other parity tests
    (tests/util.py: COS_TOL), exact classes bit-exact.
It makes two calls in a row with different answers and batches, on one stream and WITHOUT a
host synchronisation between them -- the second must neither see what the first left in the
shared array nor overwrite it before the first has read it -- checks that the free device memory
is the same before and after the calls (nothing was allocated), that sizes beyond the fixed
maximum are refused with a clear error, and, where the box has a second GPU, repeats the
comparison there (the scratch address is looked up per device).

This file is named so that pytest collects it last: a run that stops at the first failure
has then already exercised every other test.  It also passes when run on its own.
"""
import ctypes

import numpy as np
import pytest

from cassmantle_b200 import synth
from oracle import scorer_oracle as O
from tests.util import COS_TOL, assert_scores_match

pytestmark = pytest.mark.gpu

CALLS = [  # (B, N, answers): the second call uses fewer answers and another shape
    (3000, 5, np.array([17, 29_999, -1, 1234, 8, 9, 10], np.int32)),
    (4096, 1, np.array([555, 2], np.int32)),
]
V, SEED = 30_000, 4


def _noalloc(lib, device, eng, t_ans, t, B, N, out, cen, stream):
    return lib.cm_score_batch_noalloc(
        device, eng.table_ptr, eng.n_rows, eng.dim, t_ans.data_ptr(), t_ans.numel(),
        t[0].data_ptr(), t[1].data_ptr(), B, N, t[2].data_ptr(), t[3].data_ptr(),
        ctypes.c_float(np.float32(0.01)), out.data_ptr(), cen.data_ptr(), stream)


def _free_bytes(torch, device):
    with torch.cuda.device(device):
        return torch.cuda.mem_get_info()[0]          # cudaMemGetInfo


def _run_on(device: int):
    import torch
    from cassmantle_b200 import _lib
    from cassmantle_b200.engine import ScoringEngine

    lib = _lib.load()
    dev = torch.device("cuda", device)
    table = synth.full_table(V, seed=SEED)
    with ScoringEngine.synthetic(V, 300, seed=SEED, device=device) as eng:
        stream = torch.cuda.Stream(device=dev)
        st = stream.cuda_stream
        # every buffer of both calls is created up front, on `stream`, so that the two calls under
        # test can be enqueued back to back with nothing between them
        probs = []
        with torch.cuda.stream(stream):
            for k, (B, N, ans) in enumerate(CALLS):
                idx, ln, _, slot, ex = synth.make_guess_batch(B, N, V, len(ans), seed=70 + k,
                                                              ragged=N > 1, table_seed=SEED)
                t = [torch.from_numpy(a).to(dev) for a in (idx, ln, slot, ex)]
                probs.append(dict(
                    B=B, N=N, ans=ans, host=(idx, ln, slot, ex), t=t, t_ans=torch.from_numpy(ans).to(dev),
                    want=torch.empty(B, dtype=torch.float32, device=dev),
                    want_c=torch.empty(B, dtype=torch.int32, device=dev),
                    got=torch.full((B,), -7.0, dtype=torch.float32, device=dev),
                    got_c=torch.full((B,), -7, dtype=torch.int32, device=dev)))
        # reference: the engine path (allocating) on the same device buffers
        for p in probs:
            t = p["t"]
            eng.set_answers(p["ans"], stream=st)
            eng.score_device(t[0].data_ptr(), t[1].data_ptr(), p["B"], p["N"], t[2].data_ptr(),
                             t[3].data_ptr(), 0.01, p["want"].data_ptr(), p["want_c"].data_ptr(), stream=st)
        # the first use on a device looks the scratch address up (and may load the module): do it
        # before measuring free memory, as a start-up call would
        p0 = probs[0]
        rc = _noalloc(lib, device, eng, p0["t_ans"], p0["t"], 1, p0["N"], p0["got"], p0["got_c"], st)
        assert rc == 0, lib.cm_last_error()
        stream.synchronize()
        p0["got"].fill_(-7.0)
        p0["got_c"].fill_(-7)
        torch.cuda.synchronize(dev)
        free_before = _free_bytes(torch, device)
        # the calls under test: no engine state, caller-owned buffers only, no host sync in between
        for p in probs:
            rc = _noalloc(lib, device, eng, p["t_ans"], p["t"], p["B"], p["N"], p["got"], p["got_c"], st)
            assert rc == 0, lib.cm_last_error()
        stream.synchronize()
        assert _free_bytes(torch, device) == free_before, "cm_score_batch_noalloc changed the free device memory"
        for p in probs:
            idx, ln, slot, ex = p["host"]
            g, w = p["got"].cpu().numpy(), p["want"].cpu().numpy()
            assert np.array_equal(g.view(np.uint32), w.view(np.uint32))      # same kernels, same bits
            assert np.array_equal(p["got_c"].cpu().numpy(), p["want_c"].cpu().numpy())
            oracle = O.score_batch_faithful(table, idx, ln, p["ans"], slot, ex, 0.01)
            assert_scores_match(g, oracle, 0.01, tol=COS_TOL)
            assert np.all(g[ex.astype(bool)] == np.float32(1.0))
            unknown_answer = (p["ans"][slot] < 0) & ~ex.astype(bool)
            assert np.all(g[unknown_answer] == np.float32(0.01))
        return lib, eng, probs, st, dev, torch


def test_noalloc_entry_point_matches_engine_path_and_oracle():
    lib, eng, probs, st, dev, torch = _run_on(0)


def test_noalloc_limits_are_refused():
    import torch
    from cassmantle_b200 import _lib
    from cassmantle_b200.engine import ScoringEngine

    lib = _lib.load()
    dev = torch.device("cuda:0")
    with ScoringEngine.synthetic(1000, 300, seed=1, device=0) as eng:
        st = torch.cuda.Stream(device=dev).cuda_stream
        B, N = 16, 1
        idx, ln, _, slot, ex = synth.make_guess_batch(B, N, 1000, 2, seed=1, table_seed=1)
        t = [torch.from_numpy(a).to(dev) for a in (idx, ln, slot, ex)]
        t_ans = torch.tensor([1, 2], dtype=torch.int32, device=dev)
        got = torch.empty(B, dtype=torch.float32, device=dev)
        got_c = torch.empty(B, dtype=torch.int32, device=dev)
        torch.cuda.synchronize(dev)
        # sizes beyond the fixed scratch are refused, with a message that says why
        many = torch.zeros(_lib.CM_NOALLOC_MAX_ANSWERS + 1, dtype=torch.int32, device=dev)
        rc = _noalloc(lib, 0, eng, many, t, B, N, got, got_c, st)
        assert rc == -1 and b"at most 64" in lib.cm_last_error()
        rc = lib.cm_score_batch_noalloc(0, eng.table_ptr, eng.n_rows, 2048, t_ans.data_ptr(), 2,
                                        t[0].data_ptr(), t[1].data_ptr(), B, N, t[2].data_ptr(), None,
                                        ctypes.c_float(0.01), got.data_ptr(), None, st)
        assert rc == -1 and b"dim" in lib.cm_last_error()
        # a device this box does not have is named as such
        rc = lib.cm_score_batch_noalloc(63, eng.table_ptr, eng.n_rows, 300, t_ans.data_ptr(), 2,
                                        t[0].data_ptr(), t[1].data_ptr(), B, N, t[2].data_ptr(), None,
                                        ctypes.c_float(0.01), got.data_ptr(), None, st)
        assert rc == -1 and b"device 63 not in" in lib.cm_last_error()
        # an empty batch is accepted and does nothing
        assert lib.cm_score_batch_noalloc(0, eng.table_ptr, eng.n_rows, 300, t_ans.data_ptr(), 2,
                                          None, None, 0, 1, None, None, ctypes.c_float(0.01),
                                          None, None, st) == 0


def test_noalloc_on_a_second_device():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs: the scratch address is cached per device")
    _run_on(1)
    _run_on(0)          # and the first device's cached address still holds afterwards
