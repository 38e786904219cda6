import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def read_pgm(path):
    d = Path(path).read_bytes()
    parts = d.split(b"\n", 3)
    assert parts[0] == b"P5"
    w, h = map(int, parts[1].split())
    return np.frombuffer(parts[3], np.uint8).reshape(h, w).copy()


@pytest.fixture(scope="session")
def fixtures():
    return {n: read_pgm(ROOT / "tests/data" / f"{n}.pgm") for n in ("EuRoC_l", "EuRoC_r", "TUM_l", "TUM_r")}


@pytest.fixture(scope="session")
def golden():
    g = {}
    for n in ("EuRoC_l", "EuRoC_r", "TUM_l", "TUM_r"):
        g[n] = np.load(ROOT / "tests/golden" / f"extract_{n}.npz")
    for p in ("EuRoC", "TUM"):
        g["stereo_" + p] = np.load(ROOT / "tests/golden" / f"stereo_{p}.npz")
    return g


BF = 47.90639384423901
B = BF / 435.2046959714599
