import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def load_pkg():
    """The package directory is named after the reference repo and carries a hyphen."""
    return importlib.import_module("ollama-operator_b200")


@pytest.fixture(scope="session")
def pkg():
    return load_pkg()


@pytest.fixture(scope="session")
def oracle():
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle as o
    o.lib()
    return o


@pytest.fixture(scope="session")
def golden():
    import numpy as np
    return np.load(os.path.join(ROOT, "tests", "golden", "quant_kat.npz"))


@pytest.fixture(scope="session")
def model_dir(tmp_path_factory):
    d = os.environ.get("B2_MODEL_DIR") or str(tmp_path_factory.mktemp("models"))
    os.makedirs(d, exist_ok=True)
    return d
