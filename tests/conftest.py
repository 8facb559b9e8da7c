"""pytest configuration: `gpu` marker, repo root on sys.path, shared fixtures."""

import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def params():
    from oracle import scoring
    return scoring.load_params()


@pytest.fixture(scope="session")
def golden():
    return np.load(os.path.join(GOLDEN, "scoring_golden.npz"))
