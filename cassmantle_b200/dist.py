"""Batch sharding across GPUs.

Every guess is independent, so the only thing that shards is the batch: rank r of W scores the
contiguous slice ``shard_bounds(B, W, r)`` against its own full table replica, and -- where a
caller needs all B scores on every rank, as north_star states -- the fp32 scores (4 bytes per
guess) are all-gathered.  There is no other exchange on the path.  ``torch.distributed`` is
plumbing: NCCL over NVLink on GPUs, gloo in the CPU tests.
"""
from __future__ import annotations

from typing import Callable, Optional, Tuple

import numpy as np


def shard_bounds(n: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous, near-equal slices: the first ``n % world`` ranks get one extra guess."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad world/rank")
    base, extra = divmod(n, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def sharded_score(score_fn: Callable[..., np.ndarray], guess_idx: np.ndarray,
                  guess_len: np.ndarray, answer_slot: np.ndarray, exact: Optional[np.ndarray],
                  min_score: float, group=None, device=None) -> np.ndarray:
    """Score this rank's slice with ``score_fn`` and all-gather: every rank returns all B scores.

    ``score_fn(idx, len, slot, exact, min_score) -> fp32[len(idx)]`` is
    ``lambda *a: engine.score_host(*a)[0]`` in production.  Slices are padded to a common
    length for the collective (``all_gather_into_tensor`` wants equal shards).
    """
    import torch
    import torch.distributed as dist

    B = guess_idx.shape[0]
    if not (dist.is_available() and dist.is_initialized()):
        return score_fn(guess_idx, guess_len, answer_slot, exact, min_score)
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    lo, hi = shard_bounds(B, world, rank)
    local = score_fn(guess_idx[lo:hi], guess_len[lo:hi], answer_slot[lo:hi],
                     None if exact is None else exact[lo:hi], min_score)
    per = -(-B // world) if B else 0
    send = torch.zeros(per, dtype=torch.float32, device=device)
    if hi > lo:
        send[: hi - lo] = torch.from_numpy(np.ascontiguousarray(local, dtype=np.float32)).to(send.device)
    recv = torch.empty(per * world, dtype=torch.float32, device=device)
    if per:
        dist.all_gather_into_tensor(recv, send, group=group)
    recv = recv.cpu().numpy().reshape(world, per) if per else np.zeros((world, 0), np.float32)
    out = np.empty(B, dtype=np.float32)
    for r in range(world):
        a, b = shard_bounds(B, world, r)
        out[a:b] = recv[r, : b - a]
    return out


class _DevicePtr:
    """Minimal __cuda_array_interface__ carrier so torch/cupy can view a raw device pointer."""

    def __init__(self, ptr: int, shape, typestr: str = "<f4") -> None:
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr,
                                         "data": (int(ptr), False), "version": 3}


class ScoreGather:
    """Score all-gather over NVLink peer memory (C ABI: cm_gather_*, csrc/cm_gather.cu).

    ``score`` is one kernel launch: it scores this step into a staging slot and, with one extra
    CTA, ships the PREVIOUS step into every rank's receive buffer through peer mappings
    (publishing a step flag and waiting for every rank's flag) while the other CTAs gather rows.
    ``wait(stream)`` ships the latest step if nothing has taken it along yet, waits for all
    ranks and returns the device address of the complete buffer.  ``torch.distributed`` only
    carries the 64-byte IPC handles once, at construction.
    """

    def __init__(self, engine, max_guesses_per_rank: int, depth: int = 4, group=None,
                 rank: Optional[int] = None, world: Optional[int] = None) -> None:
        import ctypes

        from . import _lib
        self._ct, self._lib, self.engine = ctypes, engine._lib, engine
        self._check = _lib.check
        if rank is None or world is None:
            import torch.distributed as dist
            rank, world = dist.get_rank(group), dist.get_world_size(group)
        self.rank, self.world = rank, world
        self._h = ctypes.c_void_p()
        self._check(self._lib.cm_gather_create(ctypes.byref(self._h), engine._h, rank, world,
                                               max_guesses_per_rank, depth))
        self.stride = int(self._lib.cm_gather_stride(self._h))

    def exchange_handles(self, group=None) -> None:
        """All ranks swap IPC handles of their receive blocks (one all_gather_object)."""
        import torch.distributed as dist
        buf = self._ct.create_string_buffer(64)
        self._check(self._lib.cm_gather_export(self._h, buf))
        handles = [None] * self.world
        dist.all_gather_object(handles, bytes(buf.raw), group=group)
        for r, h in enumerate(handles):
            if r != self.rank:
                self._check(self._lib.cm_gather_import(self._h, r, self._ct.c_char_p(h)))

    def block_ptr(self) -> int:
        p = self._ct.c_void_p()
        self._check(self._lib.cm_gather_block(self._h, self._ct.byref(p), None))
        return int(p.value)

    def link_local(self, peer_rank: int, other: "ScoreGather") -> None:
        """Same-process linking (tests / single-process multi-GPU with peer access enabled)."""
        self._check(self._lib.cm_gather_import_ptr(self._h, peer_rank, other.block_ptr()))

    def score(self, d_guess_idx: int, d_guess_len: int, n_guesses: int, n_slots: int,
              d_answer_slot: int, d_exact: int, min_score: float, d_scores: int,
              stream: int, flags: int = 0) -> None:
        self._check(self._lib.cm_score_batch_gather(
            self.engine._h, self._h, d_guess_idx or None, d_guess_len or None, n_guesses, n_slots,
            d_answer_slot or None, d_exact or None, self._ct.c_float(np.float32(min_score)),
            d_scores or None, None, flags, stream or None))

    def result(self) -> int:
        """Device address of the receive buffer of the latest gathered launch ([world, stride])."""
        p = self._ct.c_void_p()
        self._check(self._lib.cm_gather_result(self._h, self._ct.byref(p)))
        return int(p.value)

    def flush(self, stream: int) -> None:
        """Ship the latest step now, without waiting for the peers (single-GPU emulation)."""
        self._check(self._lib.cm_gather_flush(self._h, stream or None))

    def wait(self, stream: int) -> int:
        """Make ``stream`` wait for the latest gathered step; returns its buffer's device address."""
        p = self._ct.c_void_p()
        self._check(self._lib.cm_gather_wait(self._h, self._ct.byref(p), stream or None))
        return int(p.value)

    def as_tensor(self, ptr: int):
        import torch
        return torch.as_tensor(_DevicePtr(ptr, (self.world, self.stride)),
                               device=f"cuda:{self.engine.device}")

    def timed_out(self) -> int:
        v = self._ct.c_int(0)
        self._check(self._lib.cm_gather_status(self._h, self._ct.byref(v)))
        return int(v.value)

    def close(self) -> None:
        if self._h is not None and self._h.value:
            self._lib.cm_gather_destroy(self._h)
            self._h = self._ct.c_void_p()
