"""pytest configuration: markers and import paths.

`-m "not gpu"` runs everywhere (oracle vs golden vectors, host logic, C-ABI
symbol checks, gloo multi-process tests); `-m gpu` needs a B200 and calls the
CUDA path through the C-ABI.
"""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "vllm-i64_b200")
for p in (ROOT, PKG, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200, sm_100a)")


@pytest.fixture(scope="session")
def golden():
    import numpy as np

    def load(name):
        path = os.path.join(GOLDEN, name)
        return np.load(path) if name.endswith(".npz") else np.load(path)

    return load
