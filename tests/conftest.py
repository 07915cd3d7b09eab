"""pytest configuration: the `gpu` marker and shared fixtures.

`pytest -m "not gpu"` runs here on CPU (oracle vs golden vectors, host logic, the C-ABI library
loading and exporting its symbols, gloo world_size-2 logic); `pytest -m gpu` runs on a B200 and
calls the CUDA path through the C ABI, checking it against the oracle.
"""
import importlib
import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200, sm_100a)")


@pytest.fixture(scope="session")
def pkg():
    return importlib.import_module("tensorflow-distributed-clustering_b200")


@pytest.fixture(scope="session")
def golden_index():
    with open(os.path.join(GOLDEN, "index.json")) as f:
        return json.load(f)


def load_golden(name, pkg, index):
    """Fixture arrays + the inputs (stored, or regenerated from the seed and SHA-256 checked)."""
    import hashlib

    meta = index[name]
    z = dict(np.load(os.path.join(GOLDEN, name + ".npz")))
    if "X" not in z:
        if meta["kind"] == "blobs":
            X, _ = pkg.datagen.blobs_numpy(meta["N"], meta["D"], meta["K"], meta["seed"])
        else:
            raise AssertionError("large fixture of unknown kind")
        if hashlib.sha256(X.tobytes()).hexdigest() != meta["x_sha256"]:
            pytest.skip("NumPy generator stream differs from the one the fixture was made with")
        z["X"] = X
    return z, meta


@pytest.fixture(scope="session")
def cuda_device():
    import torch

    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda", 0)
