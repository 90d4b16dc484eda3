"""Deterministic recipes of the golden inputs.  Shared by the
generator (make_golden.py, runs the reference binary) and the tests (which run
the oracle / the CUDA path on the same bytes)."""
import os

import numpy as np


def _mixed(n=8229):
    """frame f has width w = 1 + f % 8; v = ((i*2654435761) >> 7) & ((1<<w)-1);
    the first element of each frame carries the top bit so the width is exact."""
    i = np.arange(n, dtype=np.uint64)
    w = (1 + (i // 128) % 8).astype(np.uint64)
    v = ((i * np.uint64(2654435761)) >> np.uint64(7)) & ((np.uint64(1) << w) - np.uint64(1))
    first = (i % 128) == 0
    v = np.where(first, v | (np.uint64(1) << (w - np.uint64(1))), v)
    return v.astype(np.uint8).tobytes()


INPUTS = {
    "empty": lambda d: b"",
    "one_zero": lambda d: b"\x00",
    "one_ff": lambda d: b"\xff",
    "mul7_127": lambda d: bytes((i * 7) % 256 for i in range(127)),
    "mul7_mod128_128": lambda d: bytes((i * 7) % 128 for i in range(128)),
    "mul7_mod64_129": lambda d: bytes((i * 7) % 64 for i in range(129)),
    "zeros_1000": lambda d: bytes(1000),
    "mixed_8229": lambda d: _mixed(),
    "sample_bmp": lambda d: open(os.path.join(d, "sample_1280x853.bmp"), "rb").read(),
}


def make_input(name, golden_dir):
    return INPUTS[name](golden_dir)
