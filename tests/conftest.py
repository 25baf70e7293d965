import importlib
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")
PKG_NAME = "2012-04_stanford_cme213_b200"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pkg():
    return importlib.import_module(PKG_NAME)


def golden(name):
    return np.load(os.path.join(GOLDEN, name))


@pytest.fixture(scope="session")
def cme():
    """The package (ctypes binding).  Loading fails loudly if the shared library is not built."""
    p = pkg()
    p.load()
    return p


@pytest.fixture(scope="session")
def ctx(cme):
    import torch
    assert torch.cuda.is_available(), "gpu test collected without a CUDA device"
    torch.cuda.set_device(0)
    c = cme.Context(0)
    yield c
    c.close()


# ---- device helpers for the gpu tests (torch is only the allocator here) ------------------------------
def dev(a):
    """numpy array -> CUDA tensor with the same bytes (uint32/uint64 travel as int32/int64)."""
    import torch
    a = np.ascontiguousarray(a)
    if a.dtype == np.uint32:
        a = a.view(np.int32)
    elif a.dtype == np.uint64:
        a = a.view(np.int64)
    elif a.dtype == np.bool_:
        a = a.view(np.uint8)
    return torch.from_numpy(a).cuda()


def dev_empty(n, dtype):
    import torch
    tdt = {np.float32: torch.float32, np.float64: torch.float64, np.int32: torch.int32, np.uint32: torch.int32,
           np.int64: torch.int64, np.uint64: torch.int64, np.uint8: torch.uint8}[np.dtype(dtype).type]
    return torch.empty(int(n), dtype=tdt, device="cuda")


def host(t, dtype=None):
    import torch
    torch.cuda.synchronize()
    a = t.cpu().numpy()
    return a.view(dtype) if dtype is not None else a
