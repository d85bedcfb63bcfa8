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
