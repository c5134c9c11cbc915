"""Shared fixtures.  `-m "not gpu"` tests need no GPU; `-m gpu` tests call the CUDA library through its C ABI."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from hobbyraytracer_b200 import abi  # noqa: E402

ORACLE_SO = os.path.join(ROOT, "oracle", "_build", "liboracle.so")
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libhrt_ref.so")
PRODUCT_SO = os.path.join(ROOT, "hobbyraytracer_b200", "csrc", "libhrt_b200.so")
GOLDEN = os.path.join(ROOT, "tests", "golden")
CACHE = os.path.join(ROOT, "tests", "_cache")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _make(target):
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle"), target], check=True,
                   stdout=subprocess.DEVNULL)


@pytest.fixture(scope="session")
def orc():
    """The CPU restatement oracle (oracle/oracle.cpp), built on demand with g++."""
    src = os.path.join(ROOT, "oracle", "oracle.cpp")
    if not os.path.exists(ORACLE_SO) or os.path.getmtime(ORACLE_SO) < os.path.getmtime(src):
        _make("oracle")
    return abi.Lib(ORACLE_SO, "orc_")


@pytest.fixture(scope="session")
def ref():
    """The unmodified reference behind oracle/ref_harness.cpp.  Built here when /root/reference exists; on the GPU
    box the prebuilt library travels with the snapshot.  Tests that need it skip when neither is available."""
    if not os.path.exists(REF_SO):
        if os.path.isdir("/root/reference/hobbyraytracer"):
            _make("ref")
        else:
            pytest.skip("oracle/_ref/libhrt_ref.so not built and /root/reference absent")
    return abi.Lib(REF_SO, "ref_")


@pytest.fixture(scope="session")
def hrt():
    """The product: CUDA library behind the C ABI.  Fails loudly if it is not built."""
    lib = abi.Lib(PRODUCT_SO, "hrt_")
    return lib


@pytest.fixture(scope="session")
def cache_dir():
    os.makedirs(CACHE, exist_ok=True)
    return CACHE
