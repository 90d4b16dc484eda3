import json
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
for p in (ROOT, GOLDEN):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden():
    with open(os.path.join(GOLDEN, "golden.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle as o

    o.port()  # builds libflrl_oracle.so on first use
    return o


def flrl_import():
    """The product package (its directory name is not a Python identifier)."""
    import importlib

    return importlib.import_module("fl-rl-compression-mpi_b200")


@pytest.fixture(scope="session")
def flrl():
    return flrl_import()


def make_fl_mix(n, seed=42):
    """SURVEY.md §8d input 2: per-frame width U{1..8}; bytes U[0,256) & mask; byte 0 of each
    frame carries the top bit so the frame width is exactly w."""
    import numpy as np

    rng = np.random.default_rng(seed)
    frames = (n + 127) // 128
    w = rng.integers(1, 9, size=frames).astype(np.uint16)
    data = rng.integers(0, 256, size=frames * 128, dtype=np.uint16) & ((1 << np.repeat(w, 128)) - 1)
    data[::128] |= 1 << (w - 1)
    return data[:n].astype(np.uint8)
