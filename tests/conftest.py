import importlib
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")
REFERENCE_DIR = "/root/reference/Machine Learning for Predicting Heart Failure Progression"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on a B200)")


@pytest.fixture(scope="session")
def pkg():
    return importlib.import_module("machine-learning-replications_b200")


def _load(pkg, name):
    with open(os.path.join(GOLDEN, f"{name}_model.blob"), "rb") as fh:
        blob = fh.read()
    golden = dict(np.load(os.path.join(GOLDEN, f"golden_{name}.npz")))
    if name == "published":
        specs = pkg.synth.schema_specs()
    else:
        cand = pkg.synth.candidate_specs()
        specs = [cand[c] for c in np.load(os.path.join(GOLDEN, "retrained_columns.npy"))]
    return {"name": name, "blob": blob, "arrays": pkg.export.read_blob(blob), "golden": golden, "specs": specs}


@pytest.fixture(scope="session", params=["published", "retrained"])
def case(request, pkg):
    return _load(pkg, request.param)


@pytest.fixture(scope="session")
def published(pkg):
    return _load(pkg, "published")


@pytest.fixture(scope="session")
def retrained(pkg):
    return _load(pkg, "retrained")


@pytest.fixture(scope="session")
def cuda_lib(pkg):
    """Build (if needed) and load libhfe.so; fails loudly when it cannot be had."""
    import __graft_entry__ as entry
    entry.build()
    return pkg.capi.load_library()
