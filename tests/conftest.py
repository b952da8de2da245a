import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    """The CPU oracle (test infrastructure; compiled on demand)."""
    from oracle import oracle as O
    O.lib()
    return O


@pytest.fixture(scope="session")
def product():
    """The product package; fails loudly when the native libraries are missing."""
    import octree_mg_b200 as om
    om.libs()
    return om
