import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """The checkers (oracle/) and the product library are built in-tree; build what is missing."""
    if not os.path.exists(os.path.join(ROOT, "oracle", "libiblb_oracle.so")):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "oracle"])
    if not os.path.exists(os.path.join(ROOT, "cuda_iblb_11_b200", "libiblb.so")):
        from cuda_iblb_11_b200.build import build
        build()
    yield


def has_gpu():
    try:
        import ctypes
        cudart = ctypes.CDLL("libcudart.so")
    except OSError:
        try:
            import torch
            return torch.cuda.is_available()
        except Exception:
            return False
    n = ctypes.c_int(0)
    return cudart.cudaGetDeviceCount(ctypes.byref(n)) == 0 and n.value > 0
