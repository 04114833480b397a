import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA (sm_100) device; run with -m gpu on the B200 box')


@pytest.fixture(scope='session')
def c_oracle():
    from oracle import c_oracle as C
    C.build()
    return C
