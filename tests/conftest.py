import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def restatement():
    """oracle/sart_oracle.c (the C restatement); built on demand, it needs only gcc."""
    from oracle import pyoracle
    if not os.path.exists(pyoracle.ORACLE_SO):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "libsart_oracle.so"])
    return pyoracle.Restatement()


@pytest.fixture(scope="session")
def reference():
    """The reference's own CPU solver (oracle/_ref); prebuilt where /root/reference existed, else skipped."""
    from oracle import pyoracle
    if not pyoracle.Reference.available():
        pytest.skip("oracle/_ref/libsart_ref.so not built (needs /root/reference at build time)")
    return pyoracle.Reference()


@pytest.fixture(scope="session")
def golden():
    return np.load(os.path.join(ROOT, "tests", "golden", "golden_v1.npz"))


@pytest.fixture(scope="session")
def sart():
    """The product: ctypes view of libsart_b200.so (fails loudly when the library is not built)."""
    from __graft_entry__ import load_package
    return load_package()
