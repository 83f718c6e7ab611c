import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


def _have_gpu():
    try:
        import ctypes
        lib = ctypes.CDLL("libcudart.so.12")
        n = ctypes.c_int(0)
        return lib.cudaGetDeviceCount(ctypes.byref(n)) == 0 and n.value > 0
    except OSError:
        try:
            import torch
            return torch.cuda.is_available()
        except Exception:
            return False


@pytest.fixture(scope="session")
def have_gpu():
    return _have_gpu()


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Build the oracle (always) and the CUDA library (when missing)."""
    import oracle_lib
    oracle_lib.build()
    from raytrace_b200 import build as b
    b.build()
