"""GPU tests of the resident micro-batch kernel (cm_engine_arm_resident): a blocking host call on a
small batch goes through a page-locked mailbox instead of a launch; the scores must be
bit-identical to the launch path and within tolerance of the oracle, across changing answers,
idle exits and re-launches (BASELINE configs[1]; src/backend.py:312-317 is what it serves)."""
import threading
import time

import numpy as np
import pytest

from cassmantle_b200 import synth
from oracle import scorer_oracle as O
from tests.util import COS_TOL, assert_scores_match

pytestmark = pytest.mark.gpu
V = 50_000

# within the mailbox limits (<= 1024 guesses, <= 4096 words) and, the last three, beyond them (launch path)
SHAPES = [(512, 8), (64, 1), (1024, 1), (1024, 4), (1, 1), (1, 16), (333, 3), (7, 64), (1, 4096), (149, 27),
          (2048, 1), (4, 4096), (1000, 16)]


def _launches(lib):
    return int(lib.cm_kernel_launch_count())


def test_resident_path_equals_launch_path_and_oracle(small_table):
    from cassmantle_b200.engine import ScoringEngine
    with ScoringEngine.synthetic(V, 300, seed=0, device=0) as eng:
        lib = eng._lib
        cases = []
        for k, (B, N) in enumerate(SHAPES):
            idx, ln, ans, slot, ex = synth.make_guess_batch(B, N, V, 1 + k % 3, seed=900 + k, ragged=N > 1)
            eng.set_answers(ans)
            want = eng.score_host(idx, ln, slot, ex, 0.01, want_centi=True)          # launch path
            cases.append((idx, ln, ans, slot, ex, want))
        eng.arm_resident(idle_ms=200)
        for rep in range(3):
            for k, (idx, ln, ans, slot, ex, (want, want_c)) in enumerate(cases):
                eng.set_answers(ans)                                            # answers change between requests
                n0 = _launches(lib)
                got, got_c = eng.score_host(idx, ln, slot, ex, 0.01, want_centi=True)
                assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), (rep, idx.shape)
                assert np.array_equal(got_c, want_c)
                assert _launches(lib) - n0 == (0 if k < 10 else 1), idx.shape    # no scoring launch inside the limits
        idx, ln, ans, slot, ex, (want, _) = cases[0]
        eng.set_answers(ans)
        got, _ = eng.score_host(idx, ln, slot, None, 0.01)                       # exact == NULL
        eng.disarm_resident()
        want2, _ = eng.score_host(idx, ln, slot, None, 0.01)
        assert np.array_equal(got.view(np.uint32), want2.view(np.uint32))
        assert_scores_match(want, O.score_batch_faithful(small_table, idx, ln, ans, slot, ex, 0.01), 0.01, tol=COS_TOL)


def test_nan_floor_takes_the_launch_path():
    from cassmantle_b200.engine import ScoringEngine
    with ScoringEngine.synthetic(V, 300, seed=0, device=0) as eng:
        idx, ln, ans, slot, ex = synth.make_guess_batch(64, 2, V, 2, seed=77)
        eng.set_answers(ans)
        eng.arm_resident()
        n0 = _launches(eng._lib)
        got, _ = eng.score_host(idx, ln, slot, ex, float("nan"))                 # the sentinel protocol needs a real floor
        assert _launches(eng._lib) - n0 == 1 and got.shape == (64,)


def test_idle_exit_relaunch_and_table_change():
    """The kernel leaves after its idle limit and the next call brings it back; a table upload stops
    it (it reads the table through the non-coherent path) and later calls see the new rows."""
    from cassmantle_b200.engine import ScoringEngine
    rng = np.random.default_rng(4)
    vec = rng.standard_normal((4000, 300), dtype=np.float32)
    with ScoringEngine.from_vectors(vec) as eng:
        lib = eng._lib
        idx, ln, ans, slot, ex = synth.make_guess_batch(256, 2, 4000, 2, seed=5)
        eng.set_answers(ans)
        eng.arm_resident(idle_ms=5)
        a, _ = eng.score_host(idx, ln, slot, ex, 0.01)
        assert_scores_match(a, O.score_batch_faithful(vec, idx, ln, ans, slot, ex, 0.01), 0.01, tol=COS_TOL)
        for _ in range(4):
            time.sleep(0.03)                                                    # > idle limit: it has left
            n0 = _launches(lib)
            b, _ = eng.score_host(idx, ln, slot, ex, 0.01)
            assert _launches(lib) - n0 == 1                                      # the re-launch of the resident kernel
            assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
        vec2 = vec[::-1].copy()
        lib.cm_engine_upload_rows(eng._h, vec2.ctypes.data, 0, 4000)
        eng._answers = None
        eng.set_answers(ans)
        c, _ = eng.score_host(idx, ln, slot, ex, 0.01)
        assert_scores_match(c, O.score_batch_faithful(vec2, idx, ln, ans, slot, ex, 0.01), 0.01, tol=COS_TOL)
        # other work on the device carries on while it is resident
        import torch
        x = torch.ones(1 << 20, device="cuda:0")
        assert float((x * 2).sum().item()) == 2.0 * (1 << 20)


def test_resident_calls_from_several_threads_and_clean_shutdown():
    from cassmantle_b200.engine import ScoringEngine
    eng = ScoringEngine.synthetic(V, 300, seed=0, device=0)
    ans = np.array([11, 2222], np.int32)
    eng.set_answers(ans)
    batches = [synth.make_guess_batch(300 + 10 * i, 1 + i % 4, V, 2, seed=40 + i, ragged=True) for i in range(6)]
    wants = [eng.score_host(b[0], b[1], b[3], b[4], 0.01)[0] for b in batches]
    eng.arm_resident(idle_ms=50)
    errors = []

    def worker(i):
        try:
            for _ in range(40):
                got, _ = eng.score_host(batches[i][0], batches[i][1], batches[i][3], batches[i][4], 0.01)
                if not np.array_equal(got.view(np.uint32), wants[i].view(np.uint32)):
                    errors.append(i)
        except Exception as exc:      # noqa: BLE001
            errors.append(repr(exc))

    ts = [threading.Thread(target=worker, args=(i,)) for i in range(6)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert not errors
    t0 = time.perf_counter()
    eng.close()                                                                # stops the resident kernel at once
    assert time.perf_counter() - t0 < 0.04


def test_batched_most_similar_beside_the_resident_kernel():
    """The tensor-core scan of cm_most_similar_batch takes all of an SM's shared memory, so it cannot share SMs with
    the resident kernel: the call stops that kernel (it never idles out here: idle_ms = 0) and the next blocking score
    call brings it back -- no hang, same results on both sides."""
    from cassmantle_b200.engine import ScoringEngine
    with ScoringEngine.synthetic(V, 300, seed=0, device=0) as eng:
        idx, ln, ans, slot, ex = synth.make_guess_batch(256, 4, V, 2, seed=77, ragged=True)
        eng.set_answers(ans)
        want, _ = eng.score_host(idx, ln, slot, ex, 0.01)
        queries = np.array([5, 1234, 49_999, 17], dtype=np.int32)
        want_sim = [eng.most_similar(int(q), 11) for q in queries]
        eng.arm_resident(idle_ms=0)
        for _ in range(3):
            got, _ = eng.score_host(idx, ln, slot, ex, 0.01)              # resident path
            assert np.array_equal(got, want)
            t0 = time.perf_counter()
            batch = eng.most_similar_batch(queries, 11)
            assert time.perf_counter() - t0 < 2.0
            for (rows, sims), (wr, ws) in zip(batch, want_sim):
                assert np.array_equal(sims, ws) and np.array_equal(rows, wr)
        got, _ = eng.score_host(idx, ln, slot, ex, 0.01)
        assert np.array_equal(got, want)
