"""Deterministic synthetic word-vector table, vocabulary and guess batches.

The reference loads ``word2vec-google-news-300`` (download_model.py:9-10,
src/backend.py:45): a 3,000,000 x 300 fp32 table whose rows are NOT unit-norm.
Neither that file nor a network exists on the build or GPU boxes, so every test
and benchmark here runs on a seeded stand-in with the properties that matter to
the scorer: fp32, row-major, non-unit row norms, and a cluster structure so that
cosines spread over (-0.2, 0.8) instead of collapsing around zero.

The generator is a pure integer hash followed by two correctly-rounded fp32
multiplies, so the CUDA fill kernel (csrc/cm_scorer.cu: cm_synth_fill_table)
and this numpy code produce the SAME BITS for any (seed, row, col).  That lets a
3.6 GB table be created in HBM in about a millisecond while the host side
materialises only the rows it wants to check.

Value definition (all integer arithmetic is modulo 2**32):

    mix32(x)      : x^=x>>16; x*=0x7feb352d; x^=x>>15; x*=0x846ca68b; x^=x>>16
    bytesum(h)    : sum of the four bytes of h, minus 510      (in [-510, 510])
    noise(r, c)   : bytesum(mix32((r*D + c) ^ mix32(seed ^ 0x9e3779b9)))
    cluster(r)    : mix32(r ^ mix32(seed ^ 0x85ebca6b)) & 1023
    centre(k, c)  : bytesum(mix32((k*D + c) ^ mix32(seed ^ 0xc2b2ae35)))
    scale(r)      : 0.5f + 4.5f * ((mix32(r*0x9e3779b1 + seed) >> 8) * 2**-24)
    table[r, c]   : float(noise + centre(cluster(r), c)) * (scale(r) * INV_UNIT)

Row 0 <= r < V with r*D + c < 2**32 (true for 3M x 300).
"""
from __future__ import annotations

from concurrent.futures import ThreadPoolExecutor
from typing import List, Sequence, Tuple

import numpy as np

DIM = 300                      # word2vec-google-news-300 (download_model.py:9)
FULL_ROWS = 3_000_000          # rows of the real table
INV_UNIT = np.float32(1.0 / 3620.0)   # 147.8 (bytesum std) * sqrt(2) * sqrt(300): |row| ~ scale(r)
N_CLUSTERS = 1024

_U32 = np.uint32


def mix32(x: np.ndarray) -> np.ndarray:
    """lowbias32 integer finaliser on uint32 arrays (wraps modulo 2**32)."""
    x = np.asarray(x, dtype=np.uint32).copy()
    x ^= x >> _U32(16)
    x *= _U32(0x7FEB352D)
    x ^= x >> _U32(15)
    x *= _U32(0x846CA68B)
    x ^= x >> _U32(16)
    return x


def _mix32_scalar(x: int) -> int:
    x &= 0xFFFFFFFF
    x ^= x >> 16
    x = (x * 0x7FEB352D) & 0xFFFFFFFF
    x ^= x >> 15
    x = (x * 0x846CA68B) & 0xFFFFFFFF
    x ^= x >> 16
    return x


def _bytesum(h: np.ndarray) -> np.ndarray:
    s = (h & _U32(0xFF)) + ((h >> _U32(8)) & _U32(0xFF)) \
        + ((h >> _U32(16)) & _U32(0xFF)) + (h >> _U32(24))
    return s.astype(np.int32) - np.int32(510)


def row_scale(rows: np.ndarray, seed: int) -> np.ndarray:
    """Per-row fp32 magnitude in [0.5, 5): rows are deliberately not unit-norm."""
    rows = np.asarray(rows, dtype=np.uint32)
    with np.errstate(over="ignore"):
        h = mix32(rows * _U32(0x9E3779B1) + _U32(seed & 0xFFFFFFFF))
    u = (h >> _U32(8)).astype(np.float32) * np.float32(2.0 ** -24)
    return np.float32(0.5) + np.float32(4.5) * u


def table_rows(rows: Sequence[int] | np.ndarray, seed: int = 0, dim: int = DIM) -> np.ndarray:
    """fp32[len(rows), dim] of the synthetic table, bit-identical to the CUDA fill."""
    rows = np.asarray(rows, dtype=np.int64).reshape(-1)
    if rows.size and (rows.min() < 0 or (int(rows.max()) + 1) * dim > 2 ** 32):
        raise ValueError("row index outside the generator's 32-bit element range")
    r32 = rows.astype(np.uint32)
    cols = np.arange(dim, dtype=np.uint32)[None, :]
    k_noise = _U32(_mix32_scalar(seed ^ 0x9E3779B9))
    k_clu = _U32(_mix32_scalar(seed ^ 0x85EBCA6B))
    k_cen = _U32(_mix32_scalar(seed ^ 0xC2B2AE35))
    with np.errstate(over="ignore"):
        noise = _bytesum(mix32((r32[:, None] * _U32(dim) + cols) ^ k_noise))
        clu = mix32(r32 ^ k_clu) & _U32(N_CLUSTERS - 1)
        centre = _bytesum(mix32((clu[:, None] * _U32(dim) + cols) ^ k_cen))
    coef = row_scale(r32, seed) * INV_UNIT                      # one fp32 rounding
    return (noise + centre).astype(np.float32) * coef[:, None]  # one fp32 rounding


def full_table(n_rows: int, seed: int = 0, dim: int = DIM, threads: int = 8,
               chunk: int = 65536) -> np.ndarray:
    """Materialise the whole table on the host (the CPU baseline gathers from it)."""
    out = np.empty((n_rows, dim), dtype=np.float32)

    def fill(lo: int) -> None:
        hi = min(lo + chunk, n_rows)
        out[lo:hi] = table_rows(np.arange(lo, hi), seed, dim)

    with ThreadPoolExecutor(max_workers=max(1, threads)) as ex:
        list(ex.map(fill, range(0, n_rows, chunk)))
    return out


# --------------------------------------------------------------------------- #
# Vocabulary: lowercase pseudo-words.  The reference only ever queries
# lowercase keys (src/backend.py:304) made of [a-z0-9] (static/script.js:422).
# --------------------------------------------------------------------------- #
_ONSETS = ["b", "c", "d", "f", "g", "h", "j", "k", "l", "m", "n", "p", "r", "s", "t", "v",
           "w", "z", "br", "ch", "cl", "dr", "fl", "gr", "pl", "pr", "sh", "sl", "st", "th", "tr", ""]
_VOWELS = ["a", "e", "i", "o", "u", "ai", "ea", "ee", "io", "oo", "ou", "y"]
_CODAS = ["", "", "n", "r", "s", "t", "l", "m", "d", "ng", "st", "ck"]


def make_vocab(n_rows: int, seed: int = 0) -> List[str]:
    """``n_rows`` distinct lowercase keys; key i maps to table row i (key_to_index)."""
    keys: List[str] = []
    seen = set()
    i = 0
    while len(keys) < n_rows:
        h = _mix32_scalar(i * 0x9E3779B1 + seed + 0x51ED27)
        n_syl = 1 + (h & 3) % 3 + ((h >> 2) & 1)
        w = ""
        for s in range(n_syl):
            g = _mix32_scalar(h + 0x632BE5AB * (s + 1))
            w += _ONSETS[g % len(_ONSETS)] + _VOWELS[(g >> 8) % len(_VOWELS)] \
                + _CODAS[(g >> 16) % len(_CODAS)]
        if w in seen:
            w = f"{w}{i}"
        seen.add(w)
        keys.append(w)
        i += 1
    return keys


def cluster_of(rows: np.ndarray, seed: int = 0) -> np.ndarray:
    """Cluster id (0..1023) of each table row; rows of one cluster have cosine ~0.5."""
    k_clu = _U32(_mix32_scalar(seed ^ 0x85EBCA6B))
    return (mix32(np.asarray(rows, dtype=np.uint32) ^ k_clu) & _U32(N_CLUSTERS - 1)).astype(np.int32)


def make_guess_batch(n_guesses: int, seq: int, n_rows: int, n_answers: int = 2,
                     seed: int = 0, oov_frac: float = 0.05, exact_frac: float = 0.01,
                     ragged: bool = False, near_frac: float = 0.3, table_seed: int = 0
                     ) -> Tuple[np.ndarray, np.ndarray, np.ndarray, np.ndarray, np.ndarray]:
    """Index-form guess batch (what the vocabulary lookup hands the kernel).

    Returns (guess_idx int32[B,seq] with -1 = out-of-vocabulary, guess_len int32[B],
    answer_rows int32[A], answer_slot int32[B], exact uint8[B]).  About ``oov_frac`` of
    the guesses contain one OOV word (-> min_score, src/backend.py:308-309) and about
    ``exact_frac`` are flagged as string-equal to the answer (-> 1.0, src/backend.py:305).
    Guess rows are uniform over the table (the worst case for any cache), except that
    ``near_frac`` of the guesses take their FIRST word from the answer's cluster so the
    scores are not all floored at ``min_score``.
    """
    rng = np.random.default_rng(seed)
    idx = rng.integers(0, n_rows, size=(n_guesses, seq), dtype=np.int64).astype(np.int32)
    if ragged:
        length = rng.integers(1, seq + 1, size=n_guesses).astype(np.int32)
    else:
        length = np.full(n_guesses, seq, dtype=np.int32)
    oov = rng.random(n_guesses) < oov_frac
    pos = (rng.integers(0, 1 << 30, size=n_guesses) % length).astype(np.int64)
    idx[np.nonzero(oov)[0], pos[oov]] = -1
    idx[np.arange(seq)[None, :] >= length[:, None]] = -1      # padding slots
    answers = rng.choice(n_rows, size=n_answers, replace=False).astype(np.int32)
    slot = rng.integers(0, n_answers, size=n_guesses).astype(np.int32)
    exact = (rng.random(n_guesses) < exact_frac).astype(np.uint8)
    if near_frac > 0:
        scan = np.arange(min(n_rows, 1 << 22), dtype=np.uint32)
        clu = cluster_of(scan, table_seed)
        near = rng.random(n_guesses) < near_frac
        for a in range(n_answers):
            members = np.nonzero(clu == cluster_of(answers[a:a + 1], table_seed)[0])[0]
            sel = np.nonzero(near & (slot == a) & (idx[:, 0] >= 0))[0]
            if members.size and sel.size:
                idx[sel, 0] = members[rng.integers(0, members.size, size=sel.size)]
    return idx, length, answers, slot, exact
