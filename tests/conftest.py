import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'oracle'))
GOLDEN = os.path.join(ROOT, 'tests', 'golden')


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box with -m gpu)')


def pytest_collection_modifyitems(config, items):
    import torch
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason='no CUDA device')
    for item in items:
        if 'gpu' in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope='session')
def pkg():
    import pkgpath
    return pkgpath.load_package()


@pytest.fixture(scope='session')
def oracle():
    import geotx_oracle
    return geotx_oracle


@pytest.fixture(scope='session')
def knn_c():
    """ctypes handle on the plain-C k-NN oracle (built on demand with gcc)."""
    import ctypes
    import __graft_entry__ as ge
    lib = ctypes.CDLL(ge.build_oracle())
    lib.geotx_oracle_knn.restype = None
    return lib


@pytest.fixture(scope='session')
def golden():
    import numpy as np

    def load(name):
        return np.load(os.path.join(GOLDEN, name), allow_pickle=False)
    return load
