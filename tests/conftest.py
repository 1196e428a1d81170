"""pytest configuration: registers the ``gpu`` marker and puts the repo root on sys.path."""
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


GOLDEN_DIR = ROOT / "tests" / "golden"


def golden_cases():
    return sorted(p.stem for p in GOLDEN_DIR.glob("*.npz"))


@pytest.fixture(scope="session")
def golden():
    import numpy as np

    def load(name):
        return dict(np.load(GOLDEN_DIR / f"{name}.npz", allow_pickle=False))
    return load
