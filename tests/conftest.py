import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, 'oracle')):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box)')


@pytest.fixture(scope='session')
def built_lib():
    """Make sure libkgp_b200.so exists (compiles with nvcc; no GPU needed)."""
    import __graft_entry__ as g
    g.build()
    from kgp_b200 import _cabi
    return _cabi.lib()
