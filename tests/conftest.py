import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def engine():
    """One GpuConsensus context on cuda:0 for the whole GPU test session."""
    from consensuscruncher_b200.engine import GpuConsensus
    eng = GpuConsensus(0)
    yield eng
    eng.close()
