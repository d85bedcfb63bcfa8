"""GPU parity tests: the CUDA path, called through the C ABI, against the oracle.

Bar (stated): the exact classes are bit-exact -- 1.0 for equal strings, min_score for unknown
words / empty guesses / cosines that do not beat the floor; cosines within COS_TOL = 2e-6
absolute of the oracle (north_star allows 1e-3).  The synthetic table fill is bit-exact.
"""
import asyncio
import ctypes

import numpy as np
import pytest

from cassmantle_b200 import synth
from oracle import c_oracle, scorer_oracle as O
from tests.util import COS_TOL, assert_scores_match, compact_problem, dense_kat_table

pytestmark = pytest.mark.gpu


# ---------------------------------------------------------------------------------------
# table residency
# ---------------------------------------------------------------------------------------
def test_synthetic_fill_is_bit_exact(small_engine, small_table):
    for row0, n in [(0, 64), (12345, 257), (49_990, 10)]:
        got = small_engine.download_rows(row0, n)
        assert np.array_equal(got.view(np.uint32), small_table[row0:row0 + n].view(np.uint32))


def test_upload_download_roundtrip():
    from cassmantle_b200.engine import ScoringEngine
    rng = np.random.default_rng(0)
    vec = rng.standard_normal((1000, 300), dtype=np.float32)
    with ScoringEngine.from_vectors(vec, chunk_rows=333) as eng:
        assert np.array_equal(eng.download_rows(0, 1000), vec)
        assert np.array_equal(eng.download_rows(998, 2), vec[998:])


# ---------------------------------------------------------------------------------------
# known answers through the reference-shaped Python boundary
# ---------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def kat_backend(golden):
    from cassmantle_b200.backend import Backend
    from cassmantle_b200.engine import ScoringEngine
    from cassmantle_b200.vocab import Vocabulary
    table = dense_kat_table(golden)
    eng = ScoringEngine.from_vectors(table)
    yield Backend(eng, Vocabulary(golden["kat"]["keys"]), min_score=golden["min_score"])
    eng.close()


def test_kat_compute_score(kat_backend, golden):
    for c in golden["kat"]["cases"]:
        s = kat_backend.compute_score(c["input"], c["answer"])
        assert (not isinstance(s, np.floating)) == c["is_python_float"], c
        if c["is_python_float"]:
            assert str(s) == c["str"], c
        else:
            assert abs(float(s) - float.fromhex(c["f32_hex"])) <= COS_TOL, c


def test_kat_exact_strings(kat_backend):
    data = {"3": {"input": "Alpha", "answer": "alpha"},
            "7": {"input": "alpha", "answer": "delta"},
            "9": {"input": "Beta", "answer": "dELTA"},
            "1": {"input": "qqq", "answer": "delta"},
            "2": {"input": "alpha", "answer": "gamma"},
            "4": {"input": "void", "answer": "alpha"}}
    out = asyncio.run(kat_backend.compute_scores(data))
    assert out == {"3": "1.0", "7": "0.6", "9": "0.8", "1": "0.01", "2": "0.01", "4": "0.01"}
    assert list(out) == ["3", "7", "9", "1", "2", "4"]


def test_kat_fp64_norm_path(kat_backend):
    # fp32 squares of 1e-20 underflow and of 1e20 overflow; gensim's double nrm2 does not care
    assert abs(float(kat_backend.compute_score("tiny", "huge")) - 1.0) <= COS_TOL
    assert abs(float(kat_backend.compute_score("ones", "quarter")) - 0.5) <= COS_TOL


def test_golden_synth_cases(small_engine, golden):
    cases = golden["synth"]["cases"]
    for n_slots in (1, 4, 16):
        sel = [c for c in cases if len(c["rows"]) == n_slots]
        idx = np.asarray([c["rows"] for c in sel], dtype=np.int32)
        ln = np.asarray([c["len"] for c in sel], dtype=np.int32)
        ans = sorted({c["answer_row"] for c in sel})
        slot = np.asarray([ans.index(c["answer_row"]) for c in sel], dtype=np.int32)
        ex = np.asarray([c["exact"] for c in sel], dtype=np.uint8)
        want = np.asarray([float.fromhex(c["f32_hex"]) for c in sel], dtype=np.float32)
        small_engine.set_answers(ans)
        got, _ = small_engine.score_host(idx, ln, slot, ex, 0.01)
        assert_scores_match(got, want, 0.01)


# ---------------------------------------------------------------------------------------
# seeded batches against the numpy oracle
# ---------------------------------------------------------------------------------------
@pytest.mark.parametrize("B,N,ragged", [
    (1, 1, False), (7, 1, False), (64, 1, False), (512, 8, False), (513, 8, True),
    (4096, 1, False), (4096, 16, False), (1000, 16, True), (300, 33, True), (129, 64, True),
    (50, 100, True), (4096, 16, True), (2500, 40, True),     # >= 32768 rows: the 16-lane paired kernel
    (3, 4096, True), (9, 4096, False),                       # CM_MAX_SEQ word slots per guess
])
def test_batch_matches_oracle(small_engine, small_table, B, N, ragged):
    idx, ln, ans, slot, ex = synth.make_guess_batch(B, N, 50_000, n_answers=3, seed=B * 131 + N,
                                                    ragged=ragged)
    small_engine.set_answers(ans)
    got, centi = small_engine.score_host(idx, ln, slot, ex, 0.01, want_centi=True)
    want = O.score_batch_faithful(small_table, idx, ln, ans, slot, ex, 0.01) if B * N <= 70_000 \
        else c_oracle.score_batch(small_table, idx, ln, ans, slot, ex, 0.01, threads=4)
    assert_scores_match(got, want, 0.01)
    assert np.all(got[ex.astype(bool)] == np.float32(1.0))
    # hundredths of a percent, as the browser rounds them (static/script.js:300); the stated
    # rounding boundary: score*1e4 within 2e-3 of a half-integer may round either way
    x = got.astype(np.float64) * 1e4
    safe = np.abs((x % 1.0) - 0.5) > 2e-3
    floor32 = np.float32(0.01)
    strs = ["1.0" if ex[b] else (str(got[b]) if got[b] > floor32 else "0.01") for b in range(B)]
    want_centi = np.asarray([O.centi_score(s) for s in strs])
    assert np.array_equal(centi[safe], want_centi[safe])


def test_device_entry_point_equals_host_entry_point(small_engine):
    import torch
    B, N = 2048, 16
    idx, ln, ans, slot, ex = synth.make_guess_batch(B, N, 50_000, 2, seed=5, ragged=True)
    small_engine.set_answers(ans)
    want, want_c = small_engine.score_host(idx, ln, slot, ex, 0.01, want_centi=True)
    dev = torch.device("cuda:0")
    t = [torch.from_numpy(a).to(dev) for a in (idx, ln, slot, ex)]
    out = torch.empty(B, dtype=torch.float32, device=dev)
    cen = torch.empty(B, dtype=torch.int32, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    small_engine.score_device(t[0].data_ptr(), t[1].data_ptr(), B, N, t[2].data_ptr(),
                              t[3].data_ptr(), 0.01, out.data_ptr(), cen.data_ptr(), stream=st)
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy().view(np.uint32), want.view(np.uint32))
    assert np.array_equal(cen.cpu().numpy(), want_c)
    # exact == NULL means "no exact matches"
    small_engine.score_device(t[0].data_ptr(), t[1].data_ptr(), B, N, t[2].data_ptr(), 0, 0.01,
                              out.data_ptr(), 0, stream=st)
    torch.cuda.synchronize()
    want2, _ = small_engine.score_host(idx, ln, slot, None, 0.01)
    assert np.array_equal(out.cpu().numpy().view(np.uint32), want2.view(np.uint32))


def test_batch_list_entry_point_equals_single_calls(small_engine):
    """cm_score_batch_list_device: k batches of different shapes in one C-ABI call."""
    import ctypes
    import torch
    from cassmantle_b200 import _lib
    lib = _lib.load()
    dev = torch.device("cuda:0")
    st = torch.cuda.current_stream().cuda_stream
    shapes = [(700, 16), (4096, 1), (33, 3), (2048, 8)]
    keep, wants = [], []
    arr = (_lib.cm_batch * len(shapes))()
    ans = None
    for k, (B, N) in enumerate(shapes):
        idx, ln, a, slot, ex = synth.make_guess_batch(B, N, 50_000, 2, seed=300 + k, ragged=N > 1)
        if ans is None:
            ans = a
            small_engine.set_answers(ans)
        wants.append(small_engine.score_host(idx, ln, slot, ex, 0.01, want_centi=True))
        t = [torch.from_numpy(x).to(dev) for x in (idx, ln, slot, ex)]
        out = torch.full((B,), -3.0, dtype=torch.float32, device=dev)
        cen = torch.full((B,), -3, dtype=torch.int32, device=dev)
        keep.append((t, out, cen))
        arr[k] = _lib.cm_batch(t[0].data_ptr(), t[1].data_ptr(), t[2].data_ptr(), t[3].data_ptr(),
                               out.data_ptr(), cen.data_ptr(), B, N)
    torch.cuda.synchronize()
    rc = lib.cm_score_batch_list_device(small_engine._h, None, arr, len(shapes), ctypes.c_float(np.float32(0.01)),
                                        _lib.CM_SCORE_OVERLAP_PREVIOUS, st)
    assert rc == 0, lib.cm_last_error()
    torch.cuda.synchronize()
    for (t, out, cen), (want, want_c) in zip(keep, wants):
        assert np.array_equal(out.cpu().numpy().view(np.uint32), want.view(np.uint32))
        assert np.array_equal(cen.cpu().numpy(), want_c)
    # an error in the middle stops the list and reports which call failed, as that call would
    arr[1].n_slots = 0
    assert lib.cm_score_batch_list_device(small_engine._h, None, arr, len(shapes), ctypes.c_float(0.01), 0, st) == -1
    assert b"n_slots" in lib.cm_last_error()
    assert lib.cm_score_batch_list_device(small_engine._h, None, None, 0, ctypes.c_float(0.01), 0, st) == 0


@pytest.mark.parametrize("B,N", [(4096, 16), (512, 8), (3000, 1), (100, 100), (2500, 40), (1, 1)])
def test_packed_24_bit_indices_equal_32_bit_indices(small_engine, B, N):
    """cm_score_batch_device_idx24: 3-byte indices (0xFFFFFF = unknown word), same bits out."""
    import ctypes
    import torch
    from cassmantle_b200 import _lib
    from cassmantle_b200.engine import pack_idx24
    lib = _lib.load()
    dev = torch.device("cuda:0")
    idx, ln, ans, slot, ex = synth.make_guess_batch(B, N, 50_000, 3, seed=B + N, ragged=N > 1, oov_frac=0.2)
    idx[0, 0] = 49_999                                        # the last row of the table
    small_engine.set_answers(ans)
    want, want_c = small_engine.score_host(idx, ln, slot, ex, 0.01, want_centi=True)
    packed = pack_idx24(idx)
    assert packed.shape == (B, N, 3) and packed[idx < 0].min(initial=255) == 255
    t = [torch.from_numpy(a).to(dev) for a in (packed, ln, slot, ex)]
    out = torch.full((B,), -3.0, dtype=torch.float32, device=dev)
    cen = torch.full((B,), -3, dtype=torch.int32, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    rc = lib.cm_score_batch_device_idx24(small_engine._h, t[0].data_ptr(), t[1].data_ptr(), B, N, t[2].data_ptr(),
                                         t[3].data_ptr(), ctypes.c_float(np.float32(0.01)), out.data_ptr(),
                                         cen.data_ptr(), 0, st)
    assert rc == 0, lib.cm_last_error()
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy().view(np.uint32), want.view(np.uint32))
    assert np.array_equal(cen.cpu().numpy(), want_c)


@pytest.mark.parametrize("route", ["inplace", "copy"])
def test_pipe_with_packed_indices(small_engine, monkeypatch, route):
    """cm_pipe_set_idx24: the slots carry 3 bytes per index on both routes; scores as the 32-bit pipe's."""
    from cassmantle_b200.engine import ScoringPipe, pack_idx24
    monkeypatch.setenv("CM_PIPE_ROUTE", route)
    ans = np.array([7, 4242, 31337], np.int32)
    small_engine.set_answers(ans)
    small_engine.synchronize()
    pipe = ScoringPipe(small_engine, depth=3, max_guesses=4096, max_slots=16, packed=True)
    shapes = [(4096, 16), (100, 1), (777, 5), (4096, 12), (1, 16)]
    for rep in range(2):
        for i, (B, N) in enumerate(shapes):
            idx, ln, _, slot, ex = synth.make_guess_batch(B, N, 50_000, 3, seed=600 + i, ragged=N > 1, oov_frac=0.1)
            want = small_engine.score_host(idx, ln, slot, ex, 0.01, want_centi=True)
            s_i = i % 3
            v = pipe.buffers(s_i, B, N)
            assert v["guess_idx"].dtype == np.uint8 and v["guess_idx"].shape == (B, N, 3)
            v["guess_idx"][:] = pack_idx24(idx)
            v["guess_len"][:] = ln
            v["answer_slot"][:] = slot
            v["exact"][:] = ex
            v["scores"][:] = -1.0
            pipe.submit(s_i, 0.01, want_centi=True)
            pipe.wait(s_i)
            assert np.array_equal(v["scores"].view(np.uint32), want[0].view(np.uint32)), (route, B, N)
            assert np.array_equal(v["centi"], want[1])
    pipe.close()


def test_overlapped_launches_match_serial(small_engine):
    """CM_SCORE_OVERLAP_PREVIOUS: back-to-back independent batches may overlap on the GPU;
    every batch must still come out bit-identical to a serial run, including right after
    the answers change (the kernel waits for its predecessor before it reads them)."""
    import torch
    from cassmantle_b200._lib import CM_SCORE_OVERLAP_PREVIOUS
    dev = torch.device("cuda:0")
    stream = torch.cuda.Stream(device=dev)
    B, N, K = 1500, 6, 40
    batches, want = [], []
    answers = [np.array([11, 22], np.int32), np.array([333, 4444, 5], np.int32)]
    for i in range(K):
        idx, ln, _, slot, ex = synth.make_guess_batch(B, N, 50_000, 2, seed=900 + i, ragged=True)
        ans = answers[(i // 10) % 2]
        small_engine.set_answers(ans)
        want.append(small_engine.score_host(idx, ln, slot, ex, 0.01)[0])
        batches.append((ans, [torch.from_numpy(a).to(dev) for a in (idx, ln, slot, ex)]))
    outs = [torch.zeros(B, dtype=torch.float32, device=dev) for _ in range(K)]
    torch.cuda.synchronize()
    with torch.cuda.stream(stream):
        st = stream.cuda_stream
        for i, (ans, t) in enumerate(batches):
            small_engine.set_answers(ans, stream=st)          # no-op unless the round changed
            small_engine.score_device(t[0].data_ptr(), t[1].data_ptr(), B, N, t[2].data_ptr(),
                                      t[3].data_ptr(), 0.01, outs[i].data_ptr(), 0, stream=st,
                                      flags=CM_SCORE_OVERLAP_PREVIOUS)
    torch.cuda.synchronize()
    for i in range(K):
        assert np.array_equal(outs[i].cpu().numpy().view(np.uint32), want[i].view(np.uint32)), i


def test_host_entry_point_from_several_threads(small_engine):
    """ctypes releases the GIL: concurrent callers of the blocking entry point take turns on the
    engine's staging block and every one gets its own scores back."""
    import threading
    ans = np.array([42, 4242], np.int32)
    small_engine.set_answers(ans)
    small_engine.synchronize()
    jobs = []
    for t in range(6):
        idx, ln, _, slot, ex = synth.make_guess_batch(700 + 13 * t, 1 + t % 4, 50_000, 2, seed=800 + t, ragged=True)
        jobs.append((idx, ln, slot, ex, small_engine.score_host(idx, ln, slot, ex, 0.01)[0]))
    errors = []

    def worker(job):
        idx, ln, slot, ex, want = job
        for _ in range(40):
            got, _ = small_engine.score_host(idx, ln, slot, ex, 0.01)
            if not np.array_equal(got.view(np.uint32), want.view(np.uint32)):
                errors.append("mismatch")

    threads = [threading.Thread(target=worker, args=(j,)) for j in jobs]
    for th in threads:
        th.start()
    for th in threads:
        th.join()
    assert not errors


def test_device_entry_point_in_a_cuda_graph(small_engine):
    """A run of scoring launches (programmatic-dependent-launch edges included) can be captured
    once and replayed: the launch-bound small shapes drop from ~4-5 us to ~2.8 us per step."""
    import torch
    from cassmantle_b200._lib import CM_SCORE_OVERLAP_PREVIOUS
    dev = torch.device("cuda:0")
    B, N, K = 512, 1, 12
    ans = np.array([5, 77], np.int32)
    small_engine.set_answers(ans)
    small_engine.synchronize()
    bufs, outs, wants = [], [], []
    for i in range(K):
        idx, ln, _, slot, ex = synth.make_guess_batch(B, N, 50_000, 2, seed=300 + i)
        wants.append(small_engine.score_host(idx, ln, slot, ex, 0.01)[0])
        bufs.append([torch.from_numpy(a).to(dev) for a in (idx, ln, slot, ex)])
        outs.append(torch.zeros(B, dtype=torch.float32, device=dev))
    s = torch.cuda.Stream(device=dev)
    torch.cuda.synchronize()
    with torch.cuda.stream(s):
        g = torch.cuda.CUDAGraph()
        g.capture_begin()
        st = torch.cuda.current_stream().cuda_stream
        for i in range(K):
            b = bufs[i]
            small_engine.score_device(b[0].data_ptr(), b[1].data_ptr(), B, N, b[2].data_ptr(), b[3].data_ptr(),
                                      0.01, outs[i].data_ptr(), 0, stream=st, flags=CM_SCORE_OVERLAP_PREVIOUS)
        g.capture_end()
    for rep in range(3):
        for o in outs:
            o.zero_()
        g.replay()
        torch.cuda.synchronize()
        for i in range(K):
            assert np.array_equal(outs[i].cpu().numpy().view(np.uint32), wants[i].view(np.uint32)), (rep, i)


PIPE_ROUTES = {"default": None, "copy-engine": ("0", "0", None), "copy-in": ("0", "100000000", None),
               "in-place": ("100000000", "100000000", "inplace"), "mid-size-by-copy": (None, None, "copy")}


def _pipe_route(monkeypatch, route):
    """The pipe picks its route (kernel reads/writes the pinned block in place, or the copy
    engine moves it) from the batch size; the thresholds are read at cm_pipe_create."""
    for name in ("CM_PIPE_DIRECT_IN_MAX", "CM_PIPE_DIRECT_OUT_MAX", "CM_PIPE_ROUTE"):
        monkeypatch.delenv(name, raising=False)
    if PIPE_ROUTES[route] is not None:
        for name, v in zip(("CM_PIPE_DIRECT_IN_MAX", "CM_PIPE_DIRECT_OUT_MAX", "CM_PIPE_ROUTE"), PIPE_ROUTES[route]):
            if v is not None:
                monkeypatch.setenv(name, v)


@pytest.mark.parametrize("route", list(PIPE_ROUTES))
def test_pipe_slot_reuse_never_sees_a_previous_batch(small_engine, monkeypatch, route):
    """One slot pair reused 40 times with a different batch each time: a kernel that reads the
    page-locked block in place must see the bytes written since the previous launch."""
    from cassmantle_b200.engine import ScoringPipe
    _pipe_route(monkeypatch, route)
    small_engine.set_answers(np.array([11, 222, 3333], np.int32))
    small_engine.synchronize()
    pipe = ScoringPipe(small_engine, depth=2, max_guesses=1024, max_slots=4)
    want = {}
    for i in range(40):
        s = i % 2
        if i >= 2:
            pipe.wait(s)
            got = pipe.buffers(s, 1024, 4)
            assert np.array_equal(got["scores"].view(np.uint32), want[i - 2][0].view(np.uint32)), i - 2
            assert np.array_equal(got["centi"], want[i - 2][1]), i - 2
        idx, ln, _, slot, ex = synth.make_guess_batch(1024, 4, 50_000, 3, seed=7000 + i, ragged=True)
        want[i] = small_engine.score_host(idx, ln, slot, ex, 0.01, want_centi=True)
        v = pipe.buffers(s, 1024, 4)
        v["guess_idx"][:] = idx; v["guess_len"][:] = ln; v["answer_slot"][:] = slot; v["exact"][:] = ex
        v["scores"][:] = -1.0
        pipe.submit(s, 0.01, want_centi=True)
    for i in (38, 39):
        pipe.wait(i % 2)
        assert np.array_equal(pipe.buffers(i % 2, 1024, 4)["scores"].view(np.uint32), want[i][0].view(np.uint32))
    pipe.close()


def test_pipe_route_tuning_keeps_results_identical(small_engine, monkeypatch):
    """Mid-size batches (64 KiB < inputs <= 640 KiB): the pipe times the in-place route against the
    copy engine and switches between them as it runs.  1400 submits walk through every phase
    (start-up, both trial windows, hold); each result must equal the blocking call's."""
    from cassmantle_b200.engine import ScoringPipe
    _pipe_route(monkeypatch, "default")
    small_engine.set_answers(np.array([5, 50, 500], np.int32))
    small_engine.synchronize()
    B, N, depth, n_batches = 2048, 16, 4, 6
    batches, want = [], []
    for i in range(n_batches):
        idx, ln, _, slot, ex = synth.make_guess_batch(B, N, 50_000, 3, seed=8100 + i, ragged=True)
        batches.append((idx, ln, slot, ex))
        want.append(small_engine.score_host(idx, ln, slot, ex, 0.01)[0])
    pipe = ScoringPipe(small_engine, depth=depth, max_guesses=B, max_slots=N)
    which = {}
    for i in range(1400):
        s = i % depth
        if i >= depth:
            pipe.wait(s)
            got = pipe.buffers(s, B, N)["scores"]
            assert np.array_equal(got.view(np.uint32), want[which[s]].view(np.uint32)), i
        b = (i * 5 + i // 7) % n_batches                   # no fixed slot <-> batch pairing
        v = pipe.buffers(s, B, N)
        v["guess_idx"][:] = batches[b][0]; v["guess_len"][:] = batches[b][1]
        v["answer_slot"][:] = batches[b][2]; v["exact"][:] = batches[b][3]
        v["scores"][:] = -1.0
        pipe.submit(s, 0.01)
        which[s] = b
    for s in range(depth):
        pipe.wait(s)
        assert np.array_equal(pipe.buffers(s, B, N)["scores"].view(np.uint32), want[which[s]].view(np.uint32))
    pipe.close()


@pytest.mark.parametrize("route", list(PIPE_ROUTES))
def test_pipe_matches_host_entry_point(small_engine, monkeypatch, route):
    """cm_pipe_*: slots filled in place, several in flight, results bit-identical to the
    synchronous host call; shapes may differ from slot to slot and slots are reusable."""
    from cassmantle_b200._lib import CassMantleError
    from cassmantle_b200.engine import ScoringPipe
    _pipe_route(monkeypatch, route)
    ans = np.array([7, 4242, 31337], np.int32)
    small_engine.set_answers(ans)
    small_engine.synchronize()
    pipe = ScoringPipe(small_engine, depth=3, max_guesses=2048, max_slots=16)
    shapes = [(2048, 16), (100, 1), (777, 5), (2048, 1), (1, 16), (640, 8), (333, 3)]
    inflight = {}
    want = {}
    for i, (B, N) in enumerate(shapes):
        s = i % 3
        if s in inflight:
            pipe.wait(s)
            j, v = inflight.pop(s)
            assert np.array_equal(v["scores"].view(np.uint32), want[j][0].view(np.uint32)), j
            assert np.array_equal(v["centi"], want[j][1]), j
        idx, ln, _, slot, ex = synth.make_guess_batch(B, N, 50_000, 3, seed=500 + i, ragged=N > 1)
        want[i] = small_engine.score_host(idx, ln, slot, ex, 0.01, want_centi=True)
        v = pipe.buffers(s, B, N)
        v["guess_idx"][:] = idx
        v["guess_len"][:] = ln
        v["answer_slot"][:] = slot
        v["exact"][:] = ex
        v["scores"][:] = -1.0
        pipe.submit(s, 0.01, want_centi=True)
        inflight[s] = (i, v)
        with pytest.raises(CassMantleError):
            pipe.buffers(s, B, N)              # the slot is busy until it has been waited for
    for s, (j, v) in inflight.items():
        pipe.wait(s)
        assert np.array_equal(v["scores"].view(np.uint32), want[j][0].view(np.uint32)), j
    # cycle = wait(older slot) + submit(filled slot) in one call
    idx, ln, _, slot, ex = synth.make_guess_batch(64, 2, 50_000, 3, seed=999)
    w0 = small_engine.score_host(idx, ln, slot, ex, 0.01)[0]
    for s in (0, 1):
        v = pipe.buffers(s, 64, 2)
        v["guess_idx"][:] = idx; v["guess_len"][:] = ln; v["answer_slot"][:] = slot; v["exact"][:] = ex
    pipe.submit(0, 0.01)
    pipe.cycle(0, 1, 0.01)
    assert np.array_equal(pipe.buffers(0, 64, 2)["scores"].view(np.uint32), w0.view(np.uint32))
    pipe.wait(1)
    assert np.array_equal(pipe.buffers(1, 64, 2)["scores"].view(np.uint32), w0.view(np.uint32))
    # cycle_many: k consecutive slots (modulo the depth) per call
    for s in (0, 1, 2):
        v = pipe.buffers(s, 64, 2)
        v["guess_idx"][:] = idx; v["guess_len"][:] = ln; v["answer_slot"][:] = slot; v["exact"][:] = ex
        v["scores"][:] = -1.0
    pipe.cycle_many(-1, 2, 2, 0.01)            # submit slots 2 and 0
    pipe.cycle_many(2, 1, 1, 0.01)             # wait 2, submit 1
    assert np.array_equal(pipe.buffers(2, 64, 2)["scores"].view(np.uint32), w0.view(np.uint32))
    pipe.cycle_many(0, -1, 2, 0.01)            # wait 0 and 1
    for s in (0, 1):
        assert np.array_equal(pipe.buffers(s, 64, 2)["scores"].view(np.uint32), w0.view(np.uint32))
    with pytest.raises(CassMantleError):
        pipe.cycle_many(0, 1, 4, 0.01)         # more than the pipe has slots
    with pytest.raises(CassMantleError):
        pipe.buffers(0, 4096, 16)              # beyond max_guesses
    with pytest.raises(CassMantleError):
        pipe.buffers(5, 16, 16)                # no such slot
    pipe.close()


def test_fused_gather_single_gpu_emulation(small_engine):
    """cm_gather_*: W ranks emulated on ONE GPU, one after the other on one stream.  The waits
    for the peers are deferred (CM_GATHER_DEFER_WAIT, cm_gather_flush) so no kernel ever waits
    for one that has not been launched.  A step is shipped by the NEXT launch of its rank (or
    by flush/wait): every rank must end up with every rank's slice, step after step, with
    more steps than ring slots."""
    import torch
    from cassmantle_b200._lib import CM_GATHER_DEFER_WAIT
    from cassmantle_b200.dist import ScoreGather, shard_bounds
    dev = torch.device("cuda:0")
    W, B, N = 3, 1000, 4
    idx, ln, ans, slot, ex = synth.make_guess_batch(B, N, 50_000, 2, seed=31, ragged=True)
    small_engine.set_answers(ans)
    want, _ = small_engine.score_host(idx, ln, slot, ex, 0.01)
    per = -(-B // W)
    gs = [ScoreGather(small_engine, per, depth=2, rank=r, world=W) for r in range(W)]
    for r in range(W):
        for q in range(W):
            if q != r:
                gs[r].link_local(q, gs[q])
    stream = torch.cuda.Stream(device=dev)
    st = stream.cuda_stream
    t = [torch.from_numpy(a).to(dev) for a in (idx, ln, slot, ex)]
    local = torch.empty(B, dtype=torch.float32, device=dev)
    for step in range(5):                      # more steps than ring slots: buffers are reused
        for rep in range(1 + step % 2):            # odd steps: two launches, the first shipped by the second
            for r in range(W):
                lo, hi = shard_bounds(B, W, r)
                gs[r].score(t[0][lo:].data_ptr(), t[1][lo:].data_ptr(), hi - lo, N, t[2][lo:].data_ptr(),
                            t[3][lo:].data_ptr(), 0.01, local[lo:].data_ptr(), st, flags=CM_GATHER_DEFER_WAIT)
        for r in range(W):
            gs[r].flush(st)                    # every rank ships before anybody waits
        for r in range(W):
            ptr = gs[r].wait(st)
            stream.synchronize()
            got = gs[r].as_tensor(ptr).cpu().numpy()
            for q in range(W):
                lo, hi = shard_bounds(B, W, q)
                assert np.array_equal(got[q, : hi - lo].view(np.uint32), want[lo:hi].view(np.uint32))
            assert gs[r].timed_out() == 0
    assert np.array_equal(local.cpu().numpy().view(np.uint32), want.view(np.uint32))
    # an empty slice still publishes its step
    gs2 = [ScoreGather(small_engine, 8, depth=2, rank=r, world=2) for r in range(2)]
    gs2[0].link_local(1, gs2[1]); gs2[1].link_local(0, gs2[0])
    gs2[0].score(t[0].data_ptr(), t[1].data_ptr(), 8, N, t[2].data_ptr(), t[3].data_ptr(), 0.01,
                 local.data_ptr(), st, flags=CM_GATHER_DEFER_WAIT)
    gs2[1].score(0, 0, 0, N, 0, 0, 0.01, 0, st, flags=CM_GATHER_DEFER_WAIT)
    gs2[0].flush(st)
    p = gs2[1].wait(st)
    stream.synchronize()
    assert np.array_equal(gs2[1].as_tensor(p).cpu().numpy()[0, :8].view(np.uint32), want[:8].view(np.uint32))
    assert gs2[1].timed_out() == 0
    # a world of one waits only for itself: the in-kernel wait is safe on a single stream
    g1 = ScoreGather(small_engine, B, depth=2, rank=0, world=1)
    for _ in range(7):                         # ships ride in the following launches
        g1.score(t[0].data_ptr(), t[1].data_ptr(), B, N, t[2].data_ptr(), t[3].data_ptr(), 0.01,
                 local.data_ptr(), st, flags=1)
    p1 = g1.wait(st)                           # the last step is shipped by the wait itself
    assert p1 == g1.result()
    stream.synchronize()
    assert np.array_equal(g1.as_tensor(p1).cpu().numpy()[0, :B].view(np.uint32), want.view(np.uint32))
    assert g1.timed_out() == 0
    for g in gs + gs2 + [g1]:
        g.close()


# ---------------------------------------------------------------------------------------
# edge cases the reference's semantics define
# ---------------------------------------------------------------------------------------
def test_edge_cases(small_engine, small_table):
    V = 50_000
    idx = np.array([[5, -1, 7, 8],          # OOV inside the used part        -> floor
                    [5, 6, -1, -1],         # OOV only in the padding         -> cosine
                    [-1, -1, -1, -1],       # empty guess                     -> floor
                    [V, 1, 2, 3],           # row beyond the table = unknown  -> floor
                    [9, 9, 9, 9],           # the answer row itself           -> ~1
                    [1, 2, 3, 4],           # unknown answer                  -> floor
                    [1, 2, 3, 4],           # slot out of range               -> floor
                    [1, 2, 3, 4]],          # exact flag wins over everything -> 1.0
                   dtype=np.int32)
    ln = np.array([4, 2, 0, 4, 4, 4, 4, 4], dtype=np.int32)
    ans = np.array([9, -1], dtype=np.int32)
    slot = np.array([0, 0, 0, 0, 0, 1, 7, 1], dtype=np.int32)
    ex = np.array([0, 0, 0, 0, 0, 0, 0, 1], dtype=np.uint8)
    small_engine.set_answers(ans)
    got, _ = small_engine.score_host(idx, ln, slot, ex, 0.01)
    f = np.float32(0.01)
    assert got[0] == f and got[2] == f and got[3] == f and got[5] == f and got[6] == f
    assert got[7] == np.float32(1.0)
    assert abs(float(got[4]) - 1.0) <= COS_TOL
    want1 = O.n_similarity(small_table, [5, 6], [9])
    assert abs(float(got[1]) - max(0.01, float(want1))) <= COS_TOL
    # empty batch is a no-op
    e, _ = small_engine.score_host(np.zeros((0, 4), np.int32), np.zeros(0, np.int32),
                                   np.zeros(0, np.int32), None, 0.01)
    assert e.shape == (0,)


def test_nan_and_zero_rows_floor():
    from cassmantle_b200.engine import ScoringEngine
    vec = np.ones((4, 300), dtype=np.float32)
    vec[1] = 0.0
    vec[2, 7] = np.nan
    vec[3] = -1.0
    with ScoringEngine.from_vectors(vec) as eng:
        eng.set_answers([0])
        got, _ = eng.score_host(np.array([[0], [1], [2], [3]], np.int32), None,
                                np.zeros(4, np.int32), None, 0.01)
    assert abs(float(got[0]) - 1.0) <= COS_TOL
    assert got[1] == np.float32(0.01)      # zero norm: unitvec leaves it, cos 0
    assert got[2] == np.float32(0.01)      # max(min_score, nan) is min_score in Python
    assert got[3] == np.float32(0.01)      # anti-parallel


def test_min_score_is_a_parameter(small_engine, small_table):
    idx, ln, ans, slot, ex = synth.make_guess_batch(512, 1, 50_000, 2, seed=77)
    small_engine.set_answers(ans)
    for ms in (0.1, 0.01, 0.5, -1.0):
        got, _ = small_engine.score_host(idx, ln, slot, ex, ms)
        want = O.score_batch_faithful(small_table, idx, ln, ans, slot, ex, ms)
        assert_scores_match(got, want, ms)


def test_many_answers(small_engine, small_table):
    from cassmantle_b200._lib import CM_MAX_ANSWERS
    rng = np.random.default_rng(3)
    ans = rng.integers(0, 50_000, CM_MAX_ANSWERS).astype(np.int32)
    ans[5] = -1
    idx = rng.integers(0, 50_000, (3000, 1)).astype(np.int32)
    slot = rng.integers(0, CM_MAX_ANSWERS, 3000).astype(np.int32)
    small_engine.set_answers(ans)
    got, _ = small_engine.score_host(idx, None, slot, None, 0.01)
    want = c_oracle.score_batch(small_table, idx, np.ones(3000, np.int32), ans, slot, None, 0.01)
    assert_scores_match(got, want, 0.01)


def test_other_dims():
    """vector_size is a table property, not a constant: any multiple of 4 up to 1024."""
    from cassmantle_b200.engine import ScoringEngine
    rng = np.random.default_rng(11)
    for dim in (4, 64, 128, 132, 512, 1024):
        vec = rng.standard_normal((500, dim), dtype=np.float32)
        idx = rng.integers(0, 500, (200, 5)).astype(np.int32)
        ln = rng.integers(1, 6, 200).astype(np.int32)
        ans = np.array([3, 4], np.int32)
        slot = rng.integers(0, 2, 200).astype(np.int32)
        with ScoringEngine.from_vectors(vec) as eng:
            eng.set_answers(ans)
            got, _ = eng.score_host(idx, ln, slot, None, 0.01)
        want = O.score_batch_faithful(vec, idx, ln, ans, slot, None, 0.01)
        assert_scores_match(got, want, 0.01)


def test_abi_rejects_bad_arguments(small_engine):
    from cassmantle_b200 import _lib
    lib = _lib.load()
    h = small_engine._h
    one = np.zeros(1, np.int32)
    f = np.zeros(1, np.float32)
    assert lib.cm_score_batch_host(h, one.ctypes.data, one.ctypes.data, 1, 0, one.ctypes.data,
                                   None, 0.01, f.ctypes.data, None) == -1
    assert lib.cm_score_batch_host(h, None, one.ctypes.data, 1, 1, one.ctypes.data,
                                   None, 0.01, f.ctypes.data, None) == -1
    assert lib.cm_set_answers(h, one.ctypes.data, 0, None) == -1
    assert lib.cm_set_answers(h, one.ctypes.data, _lib.CM_MAX_ANSWERS + 1, None) == -1
    assert lib.cm_engine_upload_rows(h, f.ctypes.data, 49_999, 2) == -1
    assert b"outside" in lib.cm_last_error()
    h2 = ctypes.c_void_p()
    assert lib.cm_engine_create(ctypes.byref(h2), 99, 10, 300) == -1


# ---------------------------------------------------------------------------------------
# BASELINE.json's full sizes: size-independent properties + a sampled oracle check
# ---------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def full_engine():
    from cassmantle_b200.engine import ScoringEngine
    eng = ScoringEngine.synthetic(synth.FULL_ROWS, 300, seed=0, device=0)
    yield eng
    eng.close()


@pytest.mark.parametrize("B,N", [(512, 8), (4096, 16), (65536, 16), (16384, 64), (65536, 1),
                                 (300_001, 1)])     # the last one takes the persistent one-word kernel
def test_full_size_configs(full_engine, B, N):
    V = synth.FULL_ROWS
    idx, ln, ans, slot, ex = synth.make_guess_batch(B, N, V, n_answers=3, seed=B + N)
    full_engine.set_answers(ans)
    got, _ = full_engine.score_host(idx, ln, slot, ex, 0.01)
    again, _ = full_engine.score_host(idx, ln, slot, ex, 0.01)
    assert np.array_equal(got.view(np.uint32), again.view(np.uint32))          # idempotent
    assert np.all(got >= np.float32(0.01)) and np.all(got <= np.float32(1.0 + 1e-6))
    assert np.all(got[ex.astype(bool)] == np.float32(1.0))
    # a guess made only of the answer's own row scores ~1; doubling a guess' words (same mean
    # direction) leaves its score unchanged to rounding
    if N >= 2:
        half = N // 2
        dup = idx.copy()
        dup[:, half:2 * half] = dup[:, :half]
        dl = np.full(B, 2 * half, np.int32)
        hl = np.full(B, half, np.int32)
        a, _ = full_engine.score_host(dup, dl, slot, None, 0.01)
        b, _ = full_engine.score_host(idx, hl, slot, None, 0.01)
        # a guess that has an OOV word in its first half floors in both; one with the OOV word only
        # in the second half floors in neither after the overwrite
        assert np.abs(a - b).max() <= 4e-6
    own = np.repeat(ans[slot][:, None], N, axis=1).astype(np.int32)
    s, _ = full_engine.score_host(own, ln, slot, None, 0.01)
    assert np.abs(s - 1.0).max() <= COS_TOL
    # sampled oracle check on the rows this batch touches (the host never builds 3.6 GB)
    pick = np.random.default_rng(0).choice(B, size=min(B, 1024), replace=False)
    table, ridx, rans = compact_problem(idx[pick], ans)
    want = c_oracle.score_batch(table, ridx, ln[pick], rans, slot[pick], ex[pick], 0.01, threads=4)
    assert_scores_match(got[pick], want, 0.01)
