"""ScoringEngine -- Python owner of one cm_engine (one GPU, one resident table replica).

Takes the place of the ``KeyedVectors`` object the reference keeps in ``self.wv``
(src/backend.py:45) for the one thing the scorer uses it for: ``similarity`` on rows.
All arithmetic happens in the CUDA library behind include/cassmantle_b200.h.
"""
from __future__ import annotations

import ctypes
from typing import Optional, Sequence, Tuple

import numpy as np

from . import _lib
from ._lib import CassMantleError, check  # noqa: F401  (re-exported)


def _ptr(a: Optional[np.ndarray]) -> Optional[int]:
    return None if a is None else a.ctypes.data


class ScoringEngine:
    def __init__(self, n_rows: int, dim: int = 300, device: int = 0) -> None:
        self._lib = _lib.load()
        self._h = ctypes.c_void_p()
        check(self._lib.cm_engine_create(ctypes.byref(self._h), device, n_rows, dim))
        self.n_rows, self.dim, self.device = int(n_rows), int(dim), int(device)
        self._answers: Optional[Tuple[int, ...]] = None

    # ---- lifecycle -------------------------------------------------------------------
    def close(self) -> None:
        if getattr(self, "_h", None) is not None and self._h.value:
            self._lib.cm_engine_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self) -> None:  # pragma: no cover - best effort
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self) -> "ScoringEngine":
        return self

    def __exit__(self, *exc) -> None:
        self.close()

    @classmethod
    def from_vectors(cls, vectors: np.ndarray, device: int = 0,
                     chunk_rows: int = 1 << 20) -> "ScoringEngine":
        """Upload a host matrix (e.g. gensim's mmap'd ``wv.vectors``) chunk by chunk."""
        if vectors.ndim != 2 or vectors.dtype != np.float32:
            raise ValueError("vectors must be a 2-D float32 array")
        eng = cls(vectors.shape[0], vectors.shape[1], device)
        for lo in range(0, vectors.shape[0], chunk_rows):
            blk = np.ascontiguousarray(vectors[lo:lo + chunk_rows])
            check(eng._lib.cm_engine_upload_rows(eng._h, blk.ctypes.data, lo, blk.shape[0]))
        return eng

    @classmethod
    def synthetic(cls, n_rows: int, dim: int = 300, seed: int = 0, device: int = 0) -> "ScoringEngine":
        eng = cls(n_rows, dim, device)
        check(eng._lib.cm_engine_fill_synthetic(eng._h, seed))
        return eng

    def download_rows(self, row0: int, n: int) -> np.ndarray:
        out = np.empty((n, self.dim), dtype=np.float32)
        check(self._lib.cm_engine_download_rows(self._h, out.ctypes.data, row0, n))
        return out

    @property
    def table_ptr(self) -> int:
        p = ctypes.c_void_p()
        check(self._lib.cm_engine_table(self._h, ctypes.byref(p), None, None))
        return int(p.value)

    @property
    def stream(self) -> int:
        p = ctypes.c_void_p()
        check(self._lib.cm_engine_stream(self._h, ctypes.byref(p)))
        return int(p.value or 0)

    def synchronize(self) -> None:
        check(self._lib.cm_engine_synchronize(self._h))

    def arm_resident(self, idle_ms: int = 0) -> None:
        """Keep a small kernel resident so that ``score_host`` on micro-batches (<= 1024 guesses,
        <= 4096 words) needs no launch: the batch goes through a page-locked mailbox.  The kernel
        leaves after ``idle_ms`` (0 = 20 ms) without work and is re-launched on demand."""
        check(self._lib.cm_engine_arm_resident(self._h, int(idle_ms)))

    def disarm_resident(self) -> None:
        check(self._lib.cm_engine_disarm_resident(self._h))

    # ---- answers ---------------------------------------------------------------------
    def _stream(self, stream: Optional[int]) -> int:
        # None = the engine's own stream; an int (0 included: CUDA's default stream) is used as is
        return self.stream if stream is None else int(stream)

    def set_answers(self, rows: Sequence[int], stream: Optional[int] = None) -> None:
        """Register the round's hidden words (table rows, -1 = unknown word)."""
        key = tuple(int(r) for r in rows)
        if key == self._answers:
            return
        arr = np.asarray(key, dtype=np.int32)
        check(self._lib.cm_set_answers(self._h, arr.ctypes.data, arr.size, self._stream(stream) or None))
        self._answers = key

    def set_answers_device(self, d_rows: int, n: int, stream: Optional[int] = None) -> None:
        check(self._lib.cm_set_answers_device(self._h, d_rows, n, self._stream(stream) or None))
        self._answers = None

    @property
    def answers(self) -> Optional[Tuple[int, ...]]:
        return self._answers

    # ---- scoring ---------------------------------------------------------------------
    def score_host(self, guess_idx: np.ndarray, guess_len: Optional[np.ndarray],
                   answer_slot: np.ndarray, exact: Optional[np.ndarray], min_score: float,
                   want_centi: bool = False) -> Tuple[np.ndarray, Optional[np.ndarray]]:
        """Score a batch held in host memory; returns (fp32 scores, int32 hundredths|None)."""
        gi = np.ascontiguousarray(guess_idx, dtype=np.int32)
        if gi.ndim == 1:
            gi = gi[:, None]
        B, N = gi.shape
        gl = (np.ones(B, dtype=np.int32) if guess_len is None
              else np.ascontiguousarray(guess_len, dtype=np.int32))
        sl = np.ascontiguousarray(answer_slot, dtype=np.int32)
        ex = None if exact is None else np.ascontiguousarray(exact, dtype=np.uint8)
        if gl.shape != (B,) or sl.shape != (B,) or (ex is not None and ex.shape != (B,)):
            raise ValueError("guess_len / answer_slot / exact must have one entry per guess")
        scores = np.empty(B, dtype=np.float32)
        centi = np.empty(B, dtype=np.int32) if want_centi else None
        if B:
            check(self._lib.cm_score_batch_host(
                self._h, gi.ctypes.data, gl.ctypes.data, B, max(N, 1), sl.ctypes.data, _ptr(ex),
                ctypes.c_float(np.float32(min_score)), scores.ctypes.data, _ptr(centi)))
        return scores, centi

    def most_similar(self, row: int, k: int) -> Tuple[np.ndarray, np.ndarray]:
        """The k rows with the highest cosine to ``row`` (itself included), best first."""
        rows = np.empty(k, dtype=np.int32)
        sims = np.empty(k, dtype=np.float32)
        n = ctypes.c_int(0)
        check(self._lib.cm_most_similar(self._h, int(row), int(k), rows.ctypes.data, sims.ctypes.data,
                                        ctypes.byref(n)))
        return rows[: n.value], sims[: n.value]

    def most_similar_batch_arrays(self, rows, k: int):
        """``most_similar(row, k)`` for every row of ``rows`` in one pass over the table (C ABI
        cm_most_similar_batch): (rows[Q, k], sims[Q, k], n[Q]) -- query q's results are the first n[q] entries of row q,
        identical to the one-query results."""
        q = np.ascontiguousarray(rows, dtype=np.int32).reshape(-1)
        out_rows = np.empty((q.size, k), dtype=np.int32)
        out_sims = np.empty((q.size, k), dtype=np.float32)
        n = np.zeros(q.size, dtype=np.int32)
        check(self._lib.cm_most_similar_batch(self._h, q.ctypes.data, int(q.size), int(k), out_rows.ctypes.data,
                                              out_sims.ctypes.data, n.ctypes.data))
        return out_rows, out_sims, n

    def most_similar_batch(self, rows, k: int):
        """The same as a list of (rows, sims) pairs, one per query (views of the two result arrays)."""
        out_rows, out_sims, n = self.most_similar_batch_arrays(rows, k)
        if k > 0 and bool((n == k).all()):
            return list(zip(out_rows, out_sims))
        return [(out_rows[i, : n[i]], out_sims[i, : n[i]]) for i in range(len(n))]

    def score_device(self, d_guess_idx: int, d_guess_len: int, n_guesses: int, n_slots: int,
                     d_answer_slot: int, d_exact: int, min_score: float, d_scores: int,
                     d_centi: int = 0, stream: Optional[int] = None, flags: int = 0) -> None:
        """Asynchronous scoring on buffers already resident in HBM (raw device addresses)."""
        check(self._lib.cm_score_batch_device(
            self._h, d_guess_idx, d_guess_len, n_guesses, n_slots, d_answer_slot,
            d_exact or None, ctypes.c_float(np.float32(min_score)), d_scores, d_centi or None,
            flags, self._stream(stream) or None))


def pack_idx24(idx: np.ndarray) -> np.ndarray:
    """int32 row numbers [..., N] (-1 = not in the vocabulary) -> uint8 [..., N, 3], little-endian,
    0xFFFFFF for -1: the packed index form of ``ScoringPipe(packed=True)`` / cm_pipe_set_idx24."""
    a = np.asarray(idx, dtype=np.int32)
    u = np.ascontiguousarray(np.where(a < 0, -1, a)).view(np.uint32)     # -1 -> 0xFFFFFFFF -> low 3 bytes 0xFFFFFF
    return np.ascontiguousarray(u[..., None].view(np.uint8)[..., :3])


class ScoringPipe:
    """Ring of page-locked batch slots on one engine (C ABI: cm_pipe_*).

    ``buffers(slot, B, N)`` returns numpy views INTO the slot's pinned block; fill them in
    place, ``submit(slot, min_score)``, carry on with other slots, then ``wait(slot)`` and read
    ``scores`` from the same views.  Several slots in flight overlap the input DMA of one batch
    with the scoring of another.
    """

    def __init__(self, engine: ScoringEngine, depth: int = 4, max_guesses: int = 4096,
                 max_slots: int = 16, packed: bool = False) -> None:
        self._lib = engine._lib
        self.engine = engine
        self._h = ctypes.c_void_p()
        check(self._lib.cm_pipe_create(ctypes.byref(self._h), engine._h, depth, max_guesses, max_slots))
        self.depth = depth
        self.packed = bool(packed)      # 3-byte indices in the slots: ``guess_idx`` views are uint8 [B, N, 3]
        if self.packed:
            check(self._lib.cm_pipe_set_idx24(self._h, 1))

    def close(self) -> None:
        if getattr(self, "_h", None) is not None and self._h.value:
            self._lib.cm_pipe_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self) -> None:  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass

    def buffers(self, slot: int, n_guesses: int, n_slots: int) -> dict:
        ptrs = [ctypes.c_void_p() for _ in range(6)]
        check(self._lib.cm_pipe_buffers(self._h, slot, n_guesses, n_slots,
                                        *[ctypes.byref(p) for p in ptrs]))

        def view(p, ctype, shape):
            n = int(np.prod(shape))
            return np.ctypeslib.as_array((ctype * n).from_address(p.value)).reshape(shape)

        B, N = n_guesses, n_slots
        return {
            "guess_idx": (view(ptrs[0], ctypes.c_uint8, (B, N, 3)) if self.packed
                          else view(ptrs[0], ctypes.c_int32, (B, N))),
            "guess_len": view(ptrs[1], ctypes.c_int32, (B,)),
            "answer_slot": view(ptrs[2], ctypes.c_int32, (B,)),
            "exact": view(ptrs[3], ctypes.c_uint8, (B,)),
            "scores": view(ptrs[4], ctypes.c_float, (B,)),
            "centi": view(ptrs[5], ctypes.c_int32, (B,)),
        }

    def submit(self, slot: int, min_score: float, want_centi: bool = False) -> None:
        check(self._lib.cm_pipe_submit(self._h, slot, ctypes.c_float(np.float32(min_score)),
                                       1 if want_centi else 0))

    def wait(self, slot: int) -> None:
        check(self._lib.cm_pipe_wait(self._h, slot))

    def cycle(self, wait_slot: int, submit_slot: int, min_score: float, want_centi: bool = False) -> None:
        """``wait(wait_slot)`` (skipped when negative) then ``submit(submit_slot)`` in one call."""
        check(self._lib.cm_pipe_cycle(self._h, wait_slot, submit_slot, ctypes.c_float(np.float32(min_score)),
                                      1 if want_centi else 0))

    def cycle_many(self, wait_first: int, submit_first: int, count: int, min_score: float,
                   want_centi: bool = False) -> None:
        """``count`` consecutive slots in one call: wait ``wait_first + i``, submit ``submit_first + i``
        (modulo the depth); a negative index skips the waits / the submits."""
        check(self._lib.cm_pipe_cycle_many(self._h, wait_first, submit_first, count,
                                           ctypes.c_float(np.float32(min_score)), 1 if want_centi else 0))
