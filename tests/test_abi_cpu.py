"""CPU-side checks of the boundary: the C-ABI library loads and exports every symbol include/flrl.h declares, the
CLI's CPU methods are bit-exact with the oracle and the reference binary's golden files, the CLI contract
(usage, exit codes, error lines) matches the reference, and the GPU entry points fail loudly without a device."""
import hashlib
import os
import re
import subprocess

import numpy as np
import pytest

from golden_inputs import INPUTS, make_input

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
CLI = os.path.join(ROOT, "fl-rl-compression-mpi_b200", "compress")
REF_CLI = os.path.join(ROOT, "oracle", "_ref", "compress_ref")


def test_library_exports_every_declared_symbol(flrl):
    hdr = open(os.path.join(ROOT, "include", "flrl.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    names = sorted(set(re.findall(r"\b(flrl_[a-z0-9_]+)\s*\(", hdr)))
    assert len(names) >= 25
    missing = [n for n in names if not hasattr(flrl.lib, n)]
    assert not missing, missing


@pytest.mark.parametrize("name", list(INPUTS))
def test_fl_cpu_method_matches_golden(name, golden, flrl, oracle):
    e = next(x for x in golden["fl"] if x["name"] == name)
    data = make_input(name, GOLDEN)
    bits, values = flrl.fl_cpu_compress(data)
    blob = oracle.fl_container(len(data), bits, values)
    assert hashlib.sha256(blob).hexdigest() == e["sha256"]
    assert flrl.fl_cpu_decompress(len(data), bits, values).tobytes() == data


def test_cpu_methods_match_oracle_random(flrl, oracle):
    rng = np.random.default_rng(21)
    for n in (1, 2, 127, 128, 129, 255, 256, 1000, 4096, 65537, 300001):
        w = rng.integers(1, 9, size=(n + 127) // 128)
        data = (rng.integers(0, 256, size=n, dtype=np.uint16) & ((1 << np.repeat(w, 128)[:n]) - 1)).astype(np.uint8)
        b, v = flrl.fl_cpu_compress(data)
        b0, v0 = oracle.fl_compress(data)
        assert np.array_equal(b, b0) and np.array_equal(v, v0), n
        runs = np.repeat(rng.integers(0, 3, size=n).astype(np.uint8), rng.geometric(1 / 150.0, size=n))[:n]
        c, x = flrl.rl_cpu_compress(runs)
        c0, x0 = oracle.rl_compress(runs)
        assert np.array_equal(c, c0) and np.array_equal(x, x0), n
        assert np.array_equal(flrl.rl_cpu_decompress(c, x), runs)


def test_cpu_decompress_rejects_malformed(flrl, oracle):
    data = np.arange(1000, dtype=np.uint8)
    b, v = oracle.fl_compress(data)
    bad = b.copy()
    bad[2] = 9
    with pytest.raises(flrl.FlrlError):
        flrl.fl_cpu_decompress(data.size, bad, v)
    with pytest.raises(flrl.FlrlError):
        flrl.fl_cpu_decompress(data.size, b, v[:10])
    with pytest.raises(flrl.FlrlError):
        flrl.rl_cpu_decompress(np.array([3, 0, 2], np.uint8), np.array([1, 2, 3], np.uint8))


def _run(*args):
    return subprocess.run([CLI, *args], capture_output=True, text=True, timeout=120)


def test_cli_usage_and_exit_codes(tmp_path):
    """src/args_parser.cu:8-68: argc != 5, unknown op or method -> usage on stderr, exit 1."""
    for bad in ([], ["c", "fl-cpu", "a"], ["x", "fl-cpu", "a", "b"], ["c", "zip", "a", "b"]):
        r = _run(*bad)
        assert r.returncode == 1 and r.stderr.startswith("USAGE: ") and "operation - c (compress) or d (decompress)" in r.stderr
    # runtime errors: "[ERROR]: ..." on stderr, exit code 0 (src/main.cu:95-99)
    r = _run("c", "fl-cpu", str(tmp_path / "missing"), str(tmp_path / "out"))
    assert r.returncode == 0 and r.stderr.strip() == "[ERROR]: [FileIO] Cannot open file"
    r = _run("d", "rl-cpu", str(tmp_path / "missing"), str(tmp_path / "out"))
    assert r.returncode == 0 and r.stderr.strip() == "[ERROR]: [FileIO] Cannot open file"


@pytest.mark.parametrize("name", list(INPUTS))
def test_cli_fl_cpu_files_are_the_reference_files(name, golden, tmp_path):
    """`compress c fl-cpu` writes the byte-identical container the reference binary wrote (tests/golden)."""
    e = next(x for x in golden["fl"] if x["name"] == name)
    src, dst, back = tmp_path / "in", tmp_path / "out.fl", tmp_path / "back"
    src.write_bytes(make_input(name, GOLDEN))
    r = _run("c", "fl-cpu", str(src), str(dst))
    assert r.returncode == 0 and r.stderr == "" and '[TIMER] Step: "Load data from file"' in r.stdout
    assert hashlib.sha256(dst.read_bytes()).hexdigest() == e["sha256"]
    r = _run("d", "fl-cpu", str(dst), str(back))
    assert r.returncode == 0 and r.stderr == ""
    assert back.read_bytes() == src.read_bytes()


def test_cli_rl_cpu_round_trip_and_container(tmp_path, oracle):
    data = make_input("sample_bmp", GOLDEN)
    src, dst, back = tmp_path / "in", tmp_path / "out.rl", tmp_path / "back"
    src.write_bytes(data)
    assert _run("c", "rl-cpu", str(src), str(dst)).returncode == 0
    c, v = oracle.rl_compress(data)
    assert dst.read_bytes() == oracle.rl_container(len(data), c, v)
    assert _run("d", "rl-cpu", str(dst), str(back)).returncode == 0
    assert back.read_bytes() == data


@pytest.mark.skipif(not os.path.exists(REF_CLI), reason="oracle/_ref/compress_ref not built")
def test_cli_cross_decodes_with_reference_binary(tmp_path):
    """Files written by either implementation decode byte-identically with the other (north_star)."""
    data = make_input("sample_bmp", GOLDEN)
    src = tmp_path / "in"
    src.write_bytes(data)
    ours, theirs = tmp_path / "ours.fl", tmp_path / "theirs.fl"
    assert _run("c", "fl-cpu", str(src), str(ours)).returncode == 0
    subprocess.run([REF_CLI, "c", "fl-cpu", str(src), str(theirs)], check=True, capture_output=True, timeout=120)
    assert ours.read_bytes() == theirs.read_bytes()
    a, b = tmp_path / "a", tmp_path / "b"
    subprocess.run([REF_CLI, "d", "fl-cpu", str(ours), str(a)], check=True, capture_output=True, timeout=120)
    assert _run("d", "fl-cpu", str(theirs), str(b)).returncode == 0
    assert a.read_bytes() == data and b.read_bytes() == data


def test_gpu_methods_have_no_cpu_fallback(tmp_path, flrl):
    """Without a CUDA device the GPU entry points fail with the CUDA error; they never route to the CPU codec."""
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    src = tmp_path / "in"
    src.write_bytes(b"abc" * 100)
    r = _run("c", "fl", str(src), str(tmp_path / "o"))
    assert r.returncode == 0 and r.stderr.startswith("[ERROR]: CUDA Error") and not (tmp_path / "o").exists()
    with pytest.raises(flrl.FlrlError):
        flrl.fl_gpu_compress(b"abc" * 100)
    with pytest.raises(flrl.FlrlError):
        flrl.rl_gpu_compress(b"abc" * 100)
