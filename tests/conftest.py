import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, 'tests', 'golden', 'r2p1d_oracle.npz')


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a B200 (sm_100a) GPU; run with -m gpu')


@pytest.fixture(scope='session')
def golden():
    import numpy as np
    return np.load(GOLDEN)


@pytest.fixture(scope='session')
def lib():
    """The C-ABI library; building it is __graft_entry__.build()'s job, not the tests'."""
    from rnb_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        import __graft_entry__
        __graft_entry__.build()
    return _lib.load()
