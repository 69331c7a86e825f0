"""pytest configuration: the ``gpu`` marker + shared fixtures.

``-m "not gpu"``: oracle vs golden vectors, host logic, C-ABI symbol check (no GPU needed).
``-m gpu``: parity tests proper — the CUDA path through the C ABI against the oracle.
"""

from __future__ import annotations

import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


def _has_gpu() -> bool:
    try:
        import torch

        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def golden():
    def load(name):
        return np.load(os.path.join(GOLDEN, name), allow_pickle=False)

    return load


@pytest.fixture(scope="session")
def built_lib():
    """Make sure the CUDA library and the C oracle are built (nvcc cross-compiles without a GPU)."""
    from tygrit_b200 import build

    return build.build_cuda(), build.build_oracle()
