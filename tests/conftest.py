import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu)")


def pytest_collection_modifyitems(config, items):
    # `-m gpu` on a box without CUDA should fail loudly, not silently pass.
    pass


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN
