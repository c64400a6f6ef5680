import os
import sys

import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if REPO not in sys.path:
    sys.path.insert(0, REPO)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def sclx():
    """the product binding; building it here if the in-tree library is missing"""
    lib = os.path.join(REPO, "scalix_b200", "lib", "libscalix_b200.so")
    if not os.path.exists(lib):
        import __graft_entry__

        __graft_entry__.build()
    import scalix_b200

    return scalix_b200


@pytest.fixture(scope="session")
def oracle():
    import oracle as o

    return o
