import ctypes
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def emu():
    """The kernels compiled for the host SIMT emulator (tests/_emu/, test infrastructure)."""
    from graspqp_b200 import _lib, build
    path = build.build_emu()
    return _lib.bind(ctypes.CDLL(path), strict=False)


@pytest.fixture(scope="session")
def oracle_built():
    subprocess.check_call(["bash", os.path.join(ROOT, "oracle", "build_ref.sh")], stdout=subprocess.DEVNULL)
    return True


def ptr(a):
    return a.ctypes.data if a is not None else None
