"""Test configuration.  `-m "not gpu"`: oracle vs golden vectors / dense second opinion, host logic, C-ABI
exports (no compute).  `-m gpu`: parity of the CUDA path (through the C ABI) against the oracle."""
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def liblgk():
    """the built C-ABI library (built on demand here; shipped prebuilt to the GPU box)"""
    from lesion_gnn_b200 import _lib
    if not os.path.exists(_lib.LIBRARY):
        subprocess.run(["make", "-C", os.path.join(ROOT, "lesion_gnn_b200", "csrc"), "-j", "8"], check=True)
    return _lib.lib()


@pytest.fixture(scope="session")
def dev(liblgk):
    return torch.device("cuda:0")
