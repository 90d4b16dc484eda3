"""Pins the oracle (oracle/flrl_oracle.c) before anything trusts it:
against the golden vectors the REFERENCE BINARY produced (tests/golden, SURVEY.md
§8c) and — when oracle/_ref/libflref.so travelled with the repo — against the
reference's own cpuCompress/cpuDecompress on random inputs.  CPU only."""
import hashlib
import os

import numpy as np
import pytest

from golden_inputs import INPUTS, make_input

GOLDEN = os.path.dirname(os.path.abspath(__file__)) + "/golden"


@pytest.mark.parametrize("name", list(INPUTS))
def test_fl_golden_vectors(name, golden, oracle):
    e = next(x for x in golden["fl"] if x["name"] == name)
    data = make_input(name, GOLDEN)
    assert hashlib.sha256(data).hexdigest() == e["input_sha256"]
    bits, values = oracle.fl_compress(data)
    assert [len(data), bits.size, values.size] == e["header"]
    blob = oracle.fl_container(len(data), bits, values)
    assert len(blob) == e["file_len"]
    assert hashlib.sha256(blob).hexdigest() == e["sha256"]
    if "file" in e:
        assert blob == open(os.path.join(GOLDEN, e["file"]), "rb").read()
    # decode side, from the reference-written bytes
    n, b, v = oracle.fl_container_parse(blob)
    back = oracle.fl_decompress(b, v, n)
    assert back.tobytes() == data


def test_fl_bmp_width_histogram(oracle):
    """width histogram of the sample BMP {4:1, 5:157, 6:1796, 7:4366, 8:19272}."""
    bits, _ = oracle.fl_compress(make_input("sample_bmp", GOLDEN))
    hist = {int(k): int(v) for k, v in zip(*np.unique(bits, return_counts=True))}
    assert hist == {4: 1, 5: 157, 6: 1796, 7: 4366, 8: 19272}


@pytest.mark.parametrize("e", ["plan_encode", "plan_decode", "plan_split_256"])
def test_rl_plan_examples(e, golden, oracle):
    g = next(x for x in golden["rl"] if x["name"] == e)
    data = np.array(g["input"], dtype=np.uint8)
    counts, values = oracle.rl_compress(data)
    assert counts.tolist() == g["counts"] and values.tolist() == g["values"]
    assert oracle.rl_decompress(np.array(g["counts"], np.uint8), np.array(g["values"], np.uint8)).tolist() == g["input"]


def test_rl_edge_cases(oracle):
    c, v = oracle.rl_compress(b"")
    assert c.size == 0 and v.size == 0
    assert oracle.rl_decompress(c, v).size == 0
    for run in (1, 254, 255, 256, 509, 510, 511, 765, 766, 100000):
        c, v = oracle.rl_compress(bytes([9]) * run + b"\x01")
        full, rem = divmod(run, 255)
        exp = [255] * full + ([rem] if rem else []) + [1]
        assert c.tolist() == exp
        assert v.tolist() == [9] * (len(exp) - 1) + [1]
        assert oracle.rl_decompress(c, v).tobytes() == bytes([9]) * run + b"\x01"


def test_rl_roundtrip_random(oracle):
    rng = np.random.default_rng(7)
    for mean in (1.5, 4, 64, 700):
        lens = rng.geometric(1.0 / mean, size=4000)
        vals = rng.integers(0, 256, size=lens.size, dtype=np.uint8)
        data = np.repeat(vals, lens)
        c, v = oracle.rl_compress(data)
        assert c.min() >= 1
        assert int(c.astype(np.uint64).sum()) == data.size
        assert np.array_equal(oracle.rl_decompress(c, v), data)


@pytest.mark.skipif(not os.path.exists(os.path.join(os.path.dirname(GOLDEN), "..", "oracle", "_ref", "libflref.so")),
                    reason="oracle/_ref not built (needs /root/reference at build time)")
def test_port_matches_reference_cpu_codec(oracle, capfd):
    """The restatement vs the reference's own fl_cpu.cu, on ragged random inputs."""
    rng = np.random.default_rng(11)
    for n in (1, 2, 127, 128, 129, 255, 256, 1000, 4096, 65537, 300001):
        w = rng.integers(1, 9, size=(n + 127) // 128)
        data = (rng.integers(0, 256, size=n, dtype=np.uint16) & ((1 << np.repeat(w, 128)[:n]) - 1)).astype(np.uint8)
        b0, v0 = oracle.ref_fl_compress(data)
        b1, v1 = oracle.fl_compress(data)
        assert np.array_equal(b0, b1) and np.array_equal(v0, v1), n
        assert np.array_equal(oracle.ref_fl_decompress(b0, v0, n), data)
        assert np.array_equal(oracle.fl_decompress(b0, v0, n), data)
    capfd.readouterr()  # swallow the reference's [TIMER] lines


def test_fl_shard_concat_identity(oracle):
    """SURVEY.md §0.5 / src/file_io.cu:46-51: shards cut at multiples of 128 bytes compress
    independently and concatenate to the single-stream output."""
    rng = np.random.default_rng(5)
    n = 100_000 + 77
    data = (rng.integers(0, 256, size=n, dtype=np.uint8) >> rng.integers(0, 8, size=n).astype(np.uint8))
    bits, values = oracle.fl_compress(data)
    for shards in (2, 3, 8):
        per = (n // (128 * shards)) * 128
        cuts = [i * per for i in range(shards)] + [n]
        parts = [oracle.fl_compress(data[cuts[i]:cuts[i + 1]]) for i in range(shards)]
        assert np.array_equal(np.concatenate([p[0] for p in parts]), bits)
        assert np.array_equal(np.concatenate([p[1] for p in parts]), values)
