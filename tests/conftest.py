"""pytest configuration: the ``gpu`` marker and shared fixtures.

``-m "not gpu"`` runs on the CPU-only build container (oracle vs golden vectors, host logic, C-ABI
symbol checks); ``-m gpu`` runs the parity tests proper on a B200 through the C ABI.
"""
from __future__ import annotations

import glob
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


def golden_system_files():
    return sorted(glob.glob(os.path.join(GOLDEN, "system_*.npz")))


def golden_laplace_files():
    return sorted(glob.glob(os.path.join(GOLDEN, "laplace_*.npz")))


def load_system(path):
    from padne_b200.flat import FlatSystem
    d = np.load(path)
    return FlatSystem.from_npz_dict(d, name=os.path.basename(path)), d


@pytest.fixture(scope="session")
def ctx():
    """One pdn context on cuda:0 for the whole session (gpu tests only)."""
    from padne_b200 import capi
    c = capi.Context(0)
    yield c
    c.close()
