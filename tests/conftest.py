import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _ensure_oracle():
    so = os.path.join(ROOT, "oracle", "liboracle.so")
    src = os.path.join(ROOT, "oracle", "uhdr_oracle.c")
    if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "oracle"], check=True, capture_output=True)


@pytest.fixture(scope="session")
def orc():
    """Our plain-C restatement (the oracle)."""
    _ensure_oracle()
    from oracle.pyoracle import CpuImpl

    return CpuImpl.oracle()


@pytest.fixture(scope="session")
def ref():
    """The unmodified reference hot path compiled into oracle/_ref (skips when not built)."""
    from oracle.pyoracle import CpuImpl

    if not CpuImpl.have_reference():
        if os.path.exists("/root/reference/lib/src/gainmapmath.cpp"):
            subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "ref"], check=True, capture_output=True)
        else:
            pytest.skip("oracle/_ref/libuhdr_ref.so not built (no /root/reference here)")
    return CpuImpl.reference()


@pytest.fixture(scope="session")
def ref_lut():
    from oracle.pyoracle import CpuImpl

    if not CpuImpl.have_reference(lut=True):
        pytest.skip("oracle/_ref/libuhdr_ref_lut.so not built")
    return CpuImpl.reference(lut=True)


@pytest.fixture(scope="session")
def uh():
    """The product: the C ABI through its Python mirror. Fails loudly if the library is missing."""
    from libultrahdr_dev_b200.api import JpegR

    return JpegR()
