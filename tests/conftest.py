import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    # the oracle is plain C++: (re)build it wherever the tests run; the reference .so is only
    # rebuilt where /root/reference exists (oracle/Makefile), elsewhere the prebuilt one is used.
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "liboracle.so"])


def has_gpu():
    try:
        import antpolarcodes_b200 as ap
        return ap.lib().polar_cuda_device_count() > 0
    except Exception:
        return False


@pytest.fixture(scope="session")
def gpu():
    if not has_gpu():
        pytest.skip("no CUDA device")
    return True
