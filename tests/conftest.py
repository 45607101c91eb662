import os
import sys

import numpy as np
import pytest

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def _ensure_oracle_built():
    from oracle import checker
    if not checker.available("oracle"):
        import subprocess
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "oracle"])


@pytest.fixture(scope="session")
def oracle():
    """The plain-C restatement (always available: built from committed sources)."""
    _ensure_oracle_built()
    from oracle.checker import Checker
    return Checker("oracle")


@pytest.fixture(scope="session")
def ref():
    """The compiled reference (only where /root/reference was available to build it)."""
    from oracle import checker
    if not checker.available("ref"):
        pytest.skip("oracle/_ref/libcabac_ref.so not built (reference tree absent)")
    return checker.Checker("ref")


@pytest.fixture(scope="session")
def engine():
    import entropy_coding_b200 as E
    eng = E.CabacEngine(0)
    yield eng
    eng.close()


def load_golden(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))
