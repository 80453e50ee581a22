"""Shared test plumbing.  `-m "not gpu"` tests run on a CPU-only box; `-m gpu` tests are the parity
tests proper and call the CUDA path through the C ABI."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.dirname(os.path.abspath(__file__))):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    import gs_oracle
    gs_oracle.build()
    return gs_oracle


@pytest.fixture(scope="session")
def native():
    """The C-ABI library; builds it in-tree if it is missing (nvcc cross-compiles without a GPU)."""
    from mvs_gaussian_splatting_b200 import _native
    if not os.path.exists(_native.LIB_PATH):
        _native.build()
    return _native


@pytest.fixture(scope="session")
def cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from mvs_gaussian_splatting_b200 import _native
    _native.lib()  # fail loudly if the extension is missing on a GPU box
    return torch.device("cuda:0")
