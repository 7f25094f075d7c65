import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture
def ljs():
    import learning_jax_sharding_b200 as mod
    return mod


@pytest.fixture
def cpu_checker(ljs):
    """Route per-shard compute of CPU-resident shards to the oracle-backed checker (tests only)."""
    from tests import _cpu_checker
    ljs.runtime.set_local_compute_for_testing(_cpu_checker)
    yield _cpu_checker
    ljs.runtime.set_local_compute_for_testing(None)
