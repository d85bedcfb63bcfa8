"""ctypes binding of oracle/scorer_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT.

Same import rule as oracle/scorer_oracle.py: tests/, smoke() and bench.py's CPU legs only.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from typing import Optional

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liboracle_scorer.so")
_lib = None


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "scorer_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B" if force else "-s"])
    return _SO


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(build())
        _lib.oracle_score_batch.restype = ctypes.c_int
        _lib.oracle_score_batch.argtypes = [
            ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_int,
            ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
            ctypes.c_float, ctypes.c_void_p, ctypes.c_int]
        _lib.oracle_similarity.restype = ctypes.c_float
        _lib.oracle_similarity.argtypes = [ctypes.c_void_p, ctypes.c_int,
                                           ctypes.c_int64, ctypes.c_int64]
        _lib.oracle_synth_fill.restype = ctypes.c_int
        _lib.oracle_synth_fill.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_int,
                                           ctypes.c_uint32, ctypes.c_int]
    return _lib


def synth_fill(n_rows: int, dim: int = 300, seed: int = 0, threads: int = 8) -> np.ndarray:
    """Host copy of the seeded synthetic table (bit-identical to cassmantle_b200/synth.py)."""
    out = np.empty((n_rows, dim), dtype=np.float32)
    if lib().oracle_synth_fill(out.ctypes.data, n_rows, dim, seed, threads) != 0:
        raise ValueError("oracle_synth_fill rejected its arguments")
    return out


def score_batch(table: np.ndarray, guess_idx: np.ndarray, guess_len: np.ndarray,
                answer_rows: np.ndarray, answer_slot: np.ndarray,
                exact: Optional[np.ndarray], min_score: float, threads: int = 1) -> np.ndarray:
    assert table.dtype == np.float32 and table.flags.c_contiguous
    gi = np.ascontiguousarray(guess_idx, dtype=np.int32)
    gl = np.ascontiguousarray(guess_len, dtype=np.int32)
    ar = np.ascontiguousarray(answer_rows, dtype=np.int32)
    sl = np.ascontiguousarray(answer_slot, dtype=np.int32)
    ex = None if exact is None else np.ascontiguousarray(exact, dtype=np.uint8)
    B, N = gi.shape
    out = np.empty(B, dtype=np.float32)
    rc = lib().oracle_score_batch(
        table.ctypes.data, table.shape[1], gi.ctypes.data, B, N, gl.ctypes.data,
        ar.ctypes.data, sl.ctypes.data, None if ex is None else ex.ctypes.data,
        ctypes.c_float(np.float32(min_score)), out.ctypes.data, threads)
    if rc != 0:
        raise ValueError("oracle_score_batch rejected its arguments")
    return out


def similarity(table: np.ndarray, i: int, j: int) -> np.float32:
    return np.float32(lib().oracle_similarity(table.ctypes.data, table.shape[1], i, j))
