"""ctypes binding of include/cassmantle_b200.h -- the stub a reference maintainer would add.

The reference is Python (src/backend.py), so its foreign-function binding for the scorer
is ctypes over the C ABI.  There is deliberately NO fallback: if the shared library has
not been built, or there is no sm_100 GPU, the calls raise.
"""
from __future__ import annotations

import ctypes
import os
import re
from typing import List

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libcassmantle_b200.so")
HEADER_PATH = os.path.join(os.path.dirname(HERE), "include", "cassmantle_b200.h")

CM_OK = 0
CM_MAX_DIM = 1024
CM_MAX_ANSWERS = 4096
CM_MAX_SEQ = 4096
CM_SCORE_OVERLAP_PREVIOUS = 0x1
CM_NOALLOC_MAX_ANSWERS = 64
CM_GATHER_DEFER_WAIT = 0x2


class CassMantleError(RuntimeError):
    """A C-ABI call returned a negative cm_status."""

    def __init__(self, code: int, message: str) -> None:
        super().__init__(f"cassmantle_b200 error {code}: {message}")
        self.code = code


_c = ctypes
_P = _c.c_void_p
_SIGNATURES = {
    "cm_last_error": (_c.c_char_p, []),
    "cm_version": (_c.c_char_p, []),
    "cm_kernel_launch_count": (_c.c_int64, []),
    "cm_engine_create": (_c.c_int, [_c.POINTER(_P), _c.c_int, _c.c_int64, _c.c_int]),
    "cm_engine_destroy": (_c.c_int, [_P]),
    "cm_engine_upload_rows": (_c.c_int, [_P, _P, _c.c_int64, _c.c_int64]),
    "cm_engine_download_rows": (_c.c_int, [_P, _P, _c.c_int64, _c.c_int64]),
    "cm_engine_fill_synthetic": (_c.c_int, [_P, _c.c_uint32]),
    "cm_engine_flush_l2": (_c.c_int, [_P, _P]),
    "cm_engine_table": (_c.c_int, [_P, _c.POINTER(_P), _c.POINTER(_c.c_int64), _c.POINTER(_c.c_int)]),
    "cm_engine_stream": (_c.c_int, [_P, _c.POINTER(_P)]),
    "cm_engine_synchronize": (_c.c_int, [_P]),
    "cm_set_answers": (_c.c_int, [_P, _P, _c.c_int, _P]),
    "cm_set_answers_device": (_c.c_int, [_P, _P, _c.c_int, _P]),
    "cm_score_batch_device": (_c.c_int, [_P, _P, _P, _c.c_int, _c.c_int, _P, _P, _c.c_float,
                                         _P, _P, _c.c_uint32, _P]),
    "cm_engine_arm_resident": (_c.c_int, [_P, _c.c_int]),
    "cm_engine_disarm_resident": (_c.c_int, [_P]),
    "cm_engine_resident_trace": (_c.c_int, [_P, _P]),
    "cm_score_batch_device_idx24": (_c.c_int, [_P, _P, _P, _c.c_int, _c.c_int, _P, _P, _c.c_float,
                                               _P, _P, _c.c_uint32, _P]),
    "cm_pipe_set_idx24": (_c.c_int, [_P, _c.c_int]),
    "cm_pipe_create": (_c.c_int, [_c.POINTER(_P), _P, _c.c_int, _c.c_int, _c.c_int]),
    "cm_pipe_destroy": (_c.c_int, [_P]),
    "cm_pipe_depth": (_c.c_int, [_P]),
    "cm_pipe_buffers": (_c.c_int, [_P, _c.c_int, _c.c_int, _c.c_int] + [_c.POINTER(_P)] * 6),
    "cm_pipe_submit": (_c.c_int, [_P, _c.c_int, _c.c_float, _c.c_int]),
    "cm_pipe_wait": (_c.c_int, [_P, _c.c_int]),
    "cm_pipe_cycle": (_c.c_int, [_P, _c.c_int, _c.c_int, _c.c_float, _c.c_int]),
    "cm_pipe_cycle_many": (_c.c_int, [_P, _c.c_int, _c.c_int, _c.c_int, _c.c_float, _c.c_int]),
    "cm_pipe_probe_copies": (_c.c_int, [_P, _c.c_int, _c.c_int, _c.POINTER(_c.c_double)]),
    "cm_vocab_create": (_c.c_int, [_c.POINTER(_P), _P, _P, _P, _c.c_int64]),
    "cm_vocab_destroy": (_c.c_int, [_P]),
    "cm_vocab_size": (_c.c_int, [_P, _c.POINTER(_c.c_int64), _c.POINTER(_c.c_int64)]),
    "cm_vocab_lookup_device": (_c.c_int, [_P, _P, _P, _c.c_int, _P, _P]),
    "cm_vocab_lookup_host": (_c.c_int, [_P, _P, _P, _c.c_int, _P]),
    "cm_score_words_host": (_c.c_int, [_P, _P, _P, _P, _c.c_int, _P, _P, _c.c_int, _P, _c.c_float, _P, _P]),
    "cm_most_similar": (_c.c_int, [_P, _c.c_int32, _c.c_int, _P, _P, _c.POINTER(_c.c_int)]),
    "cm_most_similar_batch": (_c.c_int, [_P, _P, _c.c_int, _c.c_int, _P, _P, _P]),
    "cm_semantic_distances": (_c.c_int, [_P, _P, _c.c_int, _c.c_int, _P]),
    "cm_blur_create": (_c.c_int, [_c.POINTER(_P), _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_int]),
    "cm_blur_destroy": (_c.c_int, [_P]),
    "cm_blur_set_image": (_c.c_int, [_P, _P, _c.c_int, _c.c_int]),
    "cm_blur_batch_host": (_c.c_int, [_P, _P, _c.c_int, _P]),
    "cm_blur_batch_pinned": (_c.c_int, [_P, _P, _c.c_int, _c.POINTER(_P)]),
    "cm_blur_batch_jpeg": (_c.c_int, [_P, _P, _c.c_int, _c.c_int, _P, _c.c_int64, _P]),
    "cm_blur_batch_jpeg_pinned": (_c.c_int, [_P, _P, _c.c_int, _c.c_int, _c.POINTER(_P), _c.POINTER(_P)]),
    "cm_gather_create": (_c.c_int, [_c.POINTER(_P), _P, _c.c_int, _c.c_int, _c.c_int, _c.c_int]),
    "cm_gather_destroy": (_c.c_int, [_P]),
    "cm_gather_export": (_c.c_int, [_P, _P]),
    "cm_gather_import": (_c.c_int, [_P, _c.c_int, _P]),
    "cm_gather_import_ptr": (_c.c_int, [_P, _c.c_int, _P]),
    "cm_gather_block": (_c.c_int, [_P, _c.POINTER(_P), _c.POINTER(_c.c_int64)]),
    "cm_gather_stride": (_c.c_int, [_P]),
    "cm_score_batch_gather": (_c.c_int, [_P, _P, _P, _P, _c.c_int, _c.c_int, _P, _P, _c.c_float,
                                         _P, _P, _c.c_uint32, _P]),
    "cm_score_batch_list_device": (_c.c_int, [_P, _P, _P, _c.c_int, _c.c_float, _c.c_uint32, _P]),
    "cm_gather_result": (_c.c_int, [_P, _c.POINTER(_P)]),
    "cm_gather_flush": (_c.c_int, [_P, _P]),
    "cm_gather_wait": (_c.c_int, [_P, _c.POINTER(_P), _P]),
    "cm_gather_trace": (_c.c_int, [_P, _P]),
    "cm_gather_status": (_c.c_int, [_P, _c.POINTER(_c.c_int)]),
    "cm_score_batch_noalloc": (_c.c_int, [_c.c_int, _P, _c.c_int64, _c.c_int, _P, _c.c_int, _P, _P,
                                          _c.c_int, _c.c_int, _P, _P, _c.c_float, _P, _P, _P]),
    "cm_score_batch_host": (_c.c_int, [_P, _P, _P, _c.c_int, _c.c_int, _P, _P, _c.c_float,
                                       _P, _P]),
}

class cm_batch(ctypes.Structure):
    """struct cm_batch of include/cassmantle_b200.h (one entry of cm_score_batch_list_device)."""
    _fields_ = [("guess_idx", _P), ("guess_len", _P), ("answer_slot", _P), ("exact", _P),
                ("scores", _P), ("centi", _P), ("n_guesses", _c.c_int32), ("n_slots", _c.c_int32)]


_lib = None


def header_symbols() -> List[str]:
    """Every function name declared in include/cassmantle_b200.h."""
    with open(HEADER_PATH) as f:
        text = re.sub(r"/\*.*?\*/", "", f.read(), flags=re.S)
    return sorted(set(re.findall(r"\b(cm_[a-z0-9_]+)\s*\(", text)))


def load() -> ctypes.CDLL:
    """Load the CUDA library; raises (never falls back) when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build the sm_100a extension first "
            "(python -m cassmantle_b200.build, or __graft_entry__.build()). "
            "cassmantle_b200 has no CPU fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in _SIGNATURES.items():
        fn = getattr(lib, name)        # AttributeError = header and library disagree
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int) -> None:
    if rc != CM_OK:
        raise CassMantleError(rc, load().cm_last_error().decode("utf-8", "replace"))
