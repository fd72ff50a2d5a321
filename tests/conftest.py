import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

PKG_NAME = "sensor-fusion-pose-estimation-challenge_b200"
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_pkg():
    """The package directory name has hyphens; import it by name through importlib."""
    return importlib.import_module(PKG_NAME)


def load_sub(name):
    return importlib.import_module(PKG_NAME + "." + name)


@pytest.fixture(scope="session")
def pkg():
    return load_pkg()


@pytest.fixture(scope="session")
def synth():
    return load_sub("synth")


def golden(name):
    import numpy as np
    return np.load(os.path.join(GOLDEN, name + ".npz"))
