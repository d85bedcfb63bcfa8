"""cassmantle_b200 -- B200-native guess scorer behind CassMantle's Backend interface.

Scope: the data-parallel hot path of the reference, Backend.compute_score /
compute_scores (src/backend.py:303-317).  The arithmetic runs in hand-written sm_100a
CUDA behind the C ABI in include/cassmantle_b200.h; importing this package does not load
the library, creating a ScoringEngine does, and it raises if the library is not built.
"""
from .vocab import Vocabulary  # noqa: F401

__all__ = ["Vocabulary", "ScoringEngine", "Backend"]
__version__ = "0.1.0"


def __getattr__(name):
    if name == "ScoringEngine":
        from .engine import ScoringEngine
        return ScoringEngine
    if name == "Backend":
        from .backend import Backend
        return Backend
    raise AttributeError(name)
