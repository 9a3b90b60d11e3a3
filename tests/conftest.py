import importlib
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def bdpt():
    """The product package (ctypes over libbdpt.so). Builds the library if it is missing."""
    pkg = importlib.import_module("bidirectional-path-tracer_b200")
    if not os.path.exists(pkg.LIB_PATH):
        pkg.build_library()
    return pkg


@pytest.fixture(scope="session")
def port():
    from oracle.oracle import Oracle
    return lambda: Oracle("port")


@pytest.fixture(scope="session")
def scenes():
    from tools import testscenes as ts
    return {
        "cornell_spheres": lambda: ts.cornell(True, (8, 12)),
        "cornell_box": lambda: ts.cornell(False),
        "unit_room": lambda: ts.unit_room(seed=1, n_blobs=6),
        "sphere_grid": lambda: ts.sphere_grid(6, 9, 10),   # BASELINE.json config 5 in miniature
    }


def load_golden(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))


@pytest.fixture(scope="session")
def golden():
    return load_golden


@pytest.fixture(scope="session")
def gpu_ctx(bdpt):
    ctx = bdpt.Context(0)   # raises without a CUDA device: no CPU fallback
    yield ctx
    ctx.close()
