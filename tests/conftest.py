import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN_DIR = os.path.join(ROOT, 'tests', 'golden')


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box with -m gpu)')


@pytest.fixture(scope='session')
def built_library():
    """Build (if stale) and load libofb200.so; never skipped - a missing library is a failure."""
    import __graft_entry__
    __graft_entry__.build()
    from opticalflow_b200 import _lib
    return _lib.load()


@pytest.fixture(scope='session')
def cuda_device(built_library):
    import torch
    assert torch.cuda.is_available(), 'gpu-marked test run without a CUDA device'
    return torch.device('cuda', 0)
