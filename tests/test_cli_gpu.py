"""The `compress` CLI's GPU methods on a B200: files byte-identical with the CPU methods' and the reference binary's,
and cross-decodable in both directions.  -m gpu."""
import os
import subprocess

import numpy as np
import pytest

from conftest import make_fl_mix
from golden_inputs import make_input

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
CLI = os.path.join(ROOT, "fl-rl-compression-mpi_b200", "compress")
REF_CLI = os.path.join(ROOT, "oracle", "_ref", "compress_ref")


def _run(*args, cli=CLI):
    r = subprocess.run([cli, *args], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "[ERROR]" not in r.stderr, r.stderr
    return r


@pytest.mark.parametrize("which", ["bmp", "mix", "empty", "tiny"])
def test_cli_fl_gpu_equals_cpu_and_reference(which, tmp_path):
    data = {"bmp": make_input("sample_bmp", GOLDEN), "mix": make_fl_mix((5 << 20) + 77, seed=5).tobytes(), "empty": b"",
            "tiny": b"\x07"}[which]
    src = tmp_path / "in"
    src.write_bytes(data)
    g, c, m = tmp_path / "g.fl", tmp_path / "c.fl", tmp_path / "m.fl"
    r = _run("c", "fl", str(src), str(g))
    if data:
        assert '[TIMER] Step: "Compression (kernels)"' in r.stdout
    _run("c", "fl-cpu", str(src), str(c))
    _run("c", "fl-nccl", str(src), str(m))  # every GPU of the box (1 here is fine)
    assert g.read_bytes() == c.read_bytes() == m.read_bytes()
    if os.path.exists(REF_CLI):
        t = tmp_path / "t.fl"
        _run("c", "fl-cpu", str(src), str(t), cli=REF_CLI)
        assert t.read_bytes() == g.read_bytes()
        if data:
            back = tmp_path / "back_ref"
            _run("d", "fl-cpu", str(g), str(back), cli=REF_CLI)  # the reference decodes our GPU file
            assert back.read_bytes() == data
    for method in ("fl", "fl-cpu", "fl-nccl"):
        back = tmp_path / f"back_{method}"
        _run("d", method, str(c), str(back))
        assert back.read_bytes() == data


@pytest.mark.parametrize("which", ["bmp", "runs", "empty"])
def test_cli_rl_gpu_equals_cpu(which, tmp_path):
    rng = np.random.default_rng(1)
    data = {"bmp": make_input("sample_bmp", GOLDEN),
            "runs": np.repeat(rng.integers(0, 5, size=40000).astype(np.uint8), rng.geometric(1 / 300.0, size=40000)).tobytes(),
            "empty": b""}[which]
    src = tmp_path / "in"
    src.write_bytes(data)
    g, c = tmp_path / "g.rl", tmp_path / "c.rl"
    _run("c", "rl", str(src), str(g))
    _run("c", "rl-cpu", str(src), str(c))
    assert g.read_bytes() == c.read_bytes()
    for method in ("rl", "rl-cpu"):
        back = tmp_path / f"back_{method}"
        _run("d", method, str(g), str(back))
        assert back.read_bytes() == data
