import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def golden():
    import numpy as np

    def load(name):
        return dict(np.load(os.path.join(GOLDEN, name + ".npz")))
    return load


@pytest.fixture(scope="session")
def c_oracle():
    """ctypes handle of oracle/libdmnerf_oracle.so (built on demand with gcc)."""
    import ctypes as C
    import subprocess
    so = os.path.join(ROOT, "oracle", "libdmnerf_oracle.so")
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle")], check=True)   # no-op when up to date
    lib = C.CDLL(so)
    p, i64, i32 = C.c_void_p, C.c_int64, C.c_int
    lib.oracle_sample_pdf.argtypes = [p, p, p, i64, i64, i32, i32, p, p]
    lib.oracle_sample_pdf.restype = None
    lib.oracle_composite.argtypes = [p, p, p, i64, i32, i32, i32, p, p, p, p]
    lib.oracle_composite.restype = None
    return lib
