import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu under gpurun)")


@pytest.fixture(scope="session")
def engine():
    """The product engine on cuda:0.  Fails loudly (no skip, no fallback) when the library or GPU is missing."""
    import laser_slam_b200 as zh
    eng = zh.Engine(0)
    yield eng
    eng.close()
