"""pytest configuration: the `gpu` marker and shared fixtures."""

import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

GOLDEN_DIR = Path(__file__).resolve().parent / "golden"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "reference: needs /root/reference (build container only)")


@pytest.fixture(scope="session")
def chirp_golden():
    return np.load(GOLDEN_DIR / "chirp_1s_60fps.npz")


@pytest.fixture(scope="session")
def claps_golden():
    return np.load(GOLDEN_DIR / "claps_known_answers.npz")
