"""GPU test of cm_score_batch_noalloc, the entry point for callers that may not allocate.

Some users of this library cannot let it call cudaMalloc: their GPU memory is owned by their
own allocator or fixed at start-up, or their request path cannot take cudaMalloc's device-wide
synchronisation.  For them the library offers one call that takes the table, the answers, the
batch and the outputs from the caller and keeps its only scratch -- the answers' unit vectors --
in a fixed 256 KiB array inside the library.

The test drives that call through the C ABI (ctypes) and compares it with the ordinary engine
path (cm_set_answers + cm_score_batch_device) on the same inputs: the two run the same kernels,
so the exact classes must agree bit for bit and the cosines within the tolerance used by the
other parity tests (tests/util.py: COS_TOL).  It makes two calls in a row with different
answers and batches -- the second must not see anything the first left in the shared array --
and checks that sizes beyond the fixed maximum are refused with a clear error.

This file is named so that pytest collects it last: a run that stops at the first failure
has then already exercised every other test.  It also passes when run on its own.
"""
import ctypes

import numpy as np
import pytest

from cassmantle_b200 import synth
from tests.util import COS_TOL, assert_scores_match

pytestmark = pytest.mark.gpu


def _noalloc(lib, eng, t_ans, t, B, N, out, cen, stream):
    return lib.cm_score_batch_noalloc(
        0, eng.table_ptr, eng.n_rows, eng.dim, t_ans.data_ptr(), t_ans.numel(),
        t[0].data_ptr(), t[1].data_ptr(), B, N, t[2].data_ptr(), t[3].data_ptr(),
        ctypes.c_float(np.float32(0.01)), out.data_ptr(), cen.data_ptr(), stream)


def test_noalloc_entry_point_matches_engine_path():
    import torch
    from cassmantle_b200 import _lib
    from cassmantle_b200.engine import ScoringEngine

    lib = _lib.load()
    dev = torch.device("cuda:0")
    V = 30_000
    with ScoringEngine.synthetic(V, 300, seed=4, device=0) as eng:
        stream = torch.cuda.Stream(device=dev)
        st = stream.cuda_stream
        calls = [  # (B, N, answers): the second call uses fewer answers and another shape
            (3000, 5, np.array([17, 29_999, -1, 1234, 8, 9, 10], np.int32)),
            (4096, 1, np.array([555, 2], np.int32)),
        ]
        for k, (B, N, ans) in enumerate(calls):
            idx, ln, _, slot, ex = synth.make_guess_batch(B, N, V, len(ans), seed=70 + k,
                                                          ragged=N > 1, table_seed=4)
            t = [torch.from_numpy(a).to(dev) for a in (idx, ln, slot, ex)]
            t_ans = torch.from_numpy(ans).to(dev)
            # reference: the engine path (allocating) on the same device buffers
            want = torch.empty(B, dtype=torch.float32, device=dev)
            want_c = torch.empty(B, dtype=torch.int32, device=dev)
            eng.set_answers(ans, stream=st)
            eng.score_device(t[0].data_ptr(), t[1].data_ptr(), B, N, t[2].data_ptr(), t[3].data_ptr(),
                             0.01, want.data_ptr(), want_c.data_ptr(), stream=st)
            # the call under test: no engine state, caller-owned buffers only
            got = torch.full((B,), -7.0, dtype=torch.float32, device=dev)
            got_c = torch.full((B,), -7, dtype=torch.int32, device=dev)
            rc = _noalloc(lib, eng, t_ans, t, B, N, got, got_c, st)
            assert rc == 0, lib.cm_last_error()
            stream.synchronize()
            g, w = got.cpu().numpy(), want.cpu().numpy()
            assert_scores_match(g, w, 0.01, tol=COS_TOL)
            assert np.array_equal(g.view(np.uint32), w.view(np.uint32))      # same kernels, same bits
            assert np.array_equal(got_c.cpu().numpy(), want_c.cpu().numpy())
            assert np.all(g[ex.astype(bool)] == np.float32(1.0))
            unknown_answer = (ans[slot] < 0) & ~ex.astype(bool)
            assert np.all(g[unknown_answer] == np.float32(0.01))

        # sizes beyond the fixed scratch are refused, with a message that says why
        many = torch.zeros(_lib.CM_NOALLOC_MAX_ANSWERS + 1, dtype=torch.int32, device=dev)
        rc = _noalloc(lib, eng, many, t, B, N, got, got_c, st)
        assert rc == -1 and b"at most 64" in lib.cm_last_error()
        rc = lib.cm_score_batch_noalloc(0, eng.table_ptr, eng.n_rows, 2048, t_ans.data_ptr(), 2,
                                        t[0].data_ptr(), t[1].data_ptr(), B, N, t[2].data_ptr(), None,
                                        ctypes.c_float(0.01), got.data_ptr(), None, st)
        assert rc == -1 and b"dim" in lib.cm_last_error()
        # an empty batch is accepted and does nothing
        assert lib.cm_score_batch_noalloc(0, eng.table_ptr, eng.n_rows, 300, t_ans.data_ptr(), 2,
                                          None, None, 0, 1, None, None, ctypes.c_float(0.01),
                                          None, None, st) == 0
