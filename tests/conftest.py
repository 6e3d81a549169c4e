import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle():
    from _oracle import Oracle
    return Oracle()


@pytest.fixture(scope="session")
def ref():
    """The compiled reference (oracle/_ref); skipped where it was not built."""
    from _oracle import RefLib
    if not RefLib.available("avx") and not os.path.isdir("/root/reference/src/lib"):
        pytest.skip("oracle/_ref not built and /root/reference absent")
    return RefLib("avx")


@pytest.fixture(scope="session")
def nb():
    import nbody_b200
    nbody_b200.lib()
    return nbody_b200


def load_golden(name):
    path = os.path.join(ROOT, "tests", "golden", name)
    return np.load(path)


@pytest.fixture(scope="session")
def golden():
    return load_golden
