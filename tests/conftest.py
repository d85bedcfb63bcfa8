import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def golden():
    import json
    with open(os.path.join(ROOT, "tests", "golden", "scorer_kat.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def small_engine():
    """50k-row synthetic table on cuda:0 (same seed/size as the golden 'synth' section)."""
    from cassmantle_b200.engine import ScoringEngine
    eng = ScoringEngine.synthetic(50_000, 300, seed=0, device=0)
    yield eng
    eng.close()


@pytest.fixture(scope="session")
def small_table():
    from cassmantle_b200 import synth
    return synth.full_table(50_000, seed=0)
