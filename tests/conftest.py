"""Test configuration: registers the ``gpu`` marker and puts the package + oracle on sys.path.

CPU (``-m "not gpu"``): oracle vs the reference's golden vectors, host logic, C-ABI exports.
GPU (``-m gpu``): parity of the CUDA path, called through the C-ABI, against the oracle.
"""

import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "lightning-thunder_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)
sys.dont_write_bytecode = True


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    import torch

    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def golden():
    import torch

    path = os.path.join(ROOT, "tests", "golden", "regions.pt")
    return torch.load(path, map_location="cpu", weights_only=True)["cases"]
