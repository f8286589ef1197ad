import os
import sys

import pytest

# one hardware queue per stream: the logical-rank exchange tests run many streams in one process
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
os.environ.setdefault("CUDA_MODULE_LOADING", "EAGER")  # no kernel is loaded (a device-wide sync) in mid-pipeline

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


@pytest.fixture(scope="session")
def built():
    """Make sure the product library and the oracle are built (no-op when up to date)."""
    import __graft_entry__ as g
    g.build(quiet=True)
    return True


@pytest.fixture(scope="session")
def oracle(built):
    from tests import oracle_lib
    return oracle_lib.Oracle()


@pytest.fixture(scope="session")
def hj(built):
    import htm_hashjoin_b200 as pkg
    return pkg
