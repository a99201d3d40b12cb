"""The C++ drop-in layer (antpolarcodes_b200/csrc/host/polar_cuda_plugin.h): subclasses of the reference's own
Decoder / Encoder base classes over the C ABI.  tests/cpp/plugin_test.cpp drives the reference's CPU classes and
the CUDA subclasses through the same base-class pointers and compares them frame by frame.

The driver is compiled in the dev container (it needs the reference's headers and oracle/_ref/libpolar_ref.so,
the unmodified reference library); the binary travels to the GPU box."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "tests", "cpp", "_build", "plugin_test")


def _build():
    if os.path.isdir("/root/reference/include/polarcode"):
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "tests", "cpp")])
    if not os.path.exists(BIN):
        pytest.skip("plugin_test was not built (needs the reference headers, dev container only)")


def test_plugin_header_builds_and_links():
    """The header compiles against the reference's headers and the driver resolves every symbol of
    libpolar_ref.so and libpolar_cuda.so; without a GPU it stops at the device check (exit code 77)."""
    _build()
    r = subprocess.run([BIN, "--quick"], capture_output=True, text=True, timeout=300)
    assert r.returncode in (0, 77), r.stdout[-2000:] + r.stderr[-2000:]


@pytest.mark.gpu
def test_plugin_matches_reference_classes(gpu):
    _build()
    r = subprocess.run([BIN], capture_output=True, text=True, timeout=900)
    print(r.stdout[-3000:])
    assert r.returncode == 0, r.stdout[-4000:] + r.stderr[-2000:]
    assert "PASSED" in r.stdout
