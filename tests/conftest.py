import os
import sys

import pytest

os.environ.setdefault("HDDM_B200_DEBUG_KNOBS", "1")  # the A/B parity tests switch evaluators (include/hddm_b200.h)
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _has_gpu():
    try:
        import ctypes
        from hddm_b200 import _lib
        n = ctypes.c_int(0)
        return _lib.lib().hddm_b200_device_count(ctypes.byref(n)) == 0 and n.value > 0
    except Exception:
        return False


@pytest.fixture(scope="session")
def gpu():
    if not _has_gpu():
        pytest.fail("a test marked gpu ran without a usable CUDA device")
    return True


@pytest.fixture(scope="session")
def golden_dir():
    return os.path.join(ROOT, "tests", "golden")
