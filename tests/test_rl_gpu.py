"""RL parity on the GPU: the CUDA path through the C ABI (include/flrl.h) against the oracle
(oracle/flrl_oracle.c, written from IMPLEMENTATION-PLAN.md:81-179), bit-exact.  Run with -m gpu."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _runs(rng, n, mean, alphabet=256):
    lens = rng.geometric(1.0 / mean, size=int(n / mean * 1.2) + 64)
    while lens.sum() < n:
        lens = np.concatenate([lens, rng.geometric(1.0 / mean, size=lens.size)])
    vals = rng.integers(0, alphabet, size=lens.size).astype(np.uint8)
    return np.repeat(vals, lens)[:n].copy()


def _check(flrl, oracle, data):
    data = np.ascontiguousarray(data, dtype=np.uint8)
    c_ref, v_ref = oracle.rl_compress(data)
    counts, values = flrl.rl_encode_device(flrl.to_device(data))
    assert counts.numel() == c_ref.size, (data.size, counts.numel(), c_ref.size)
    c, v = counts.cpu().numpy(), values.cpu().numpy()
    if not np.array_equal(c, c_ref):
        bad = int(np.flatnonzero(c != c_ref)[0])
        raise AssertionError(f"counts differ at pair {bad}: {c[bad:bad + 4]} vs {c_ref[bad:bad + 4]} (n={data.size})")
    assert np.array_equal(v, v_ref), data.size
    if data.size:
        out = flrl.rl_decode_device(flrl.to_device(c_ref), flrl.to_device(v_ref), data.size)
        o = out.cpu().numpy()
        assert o.size == data.size
        if not np.array_equal(o, data):
            bad = int(np.flatnonzero(o != data)[0])
            raise AssertionError(f"decode differs at byte {bad} of {data.size}")


@pytest.mark.parametrize("e", ["plan_encode", "plan_decode", "plan_split_256"])
def test_rl_plan_examples_gpu(e, golden, oracle, flrl):
    """IMPLEMENTATION-PLAN.md:87-89, :158-160, :125."""
    g = next(x for x in golden["rl"] if x["name"] == e)
    data = np.array(g["input"], dtype=np.uint8)
    counts, values = flrl.rl_encode_device(flrl.to_device(data))
    assert counts.cpu().tolist() == g["counts"] and values.cpu().tolist() == g["values"]
    out = flrl.rl_decode_device(flrl.to_device(np.array(g["counts"], np.uint8)), flrl.to_device(np.array(g["values"], np.uint8)),
                                len(g["input"]))
    assert out.cpu().tolist() == g["input"]


def test_rl_empty(flrl):
    import torch

    counts, values = flrl.rl_encode_device(torch.empty(0, dtype=torch.uint8, device="cuda"))
    assert counts.numel() == 0 and values.numel() == 0
    out = flrl.rl_decode_device(counts, values, 0)
    assert out.numel() == 0


@pytest.mark.parametrize("run", [1, 2, 127, 128, 129, 254, 255, 256, 509, 510, 511, 765, 766, 4095, 4096, 4097, 4096 + 255, 8191, 8192, 8193,
                                 8192 + 255, 16383, 16384, 16385, 16384 + 255, 32768 + 254, 100000])
def test_rl_long_run_split(run, oracle, flrl):
    """Greedy 255-split, also across row (128 B), warp (4 KiB) and tile (16 KiB) boundaries."""
    _check(flrl, oracle, np.frombuffer(bytes([9]) * run + b"\x01", dtype=np.uint8))
    _check(flrl, oracle, np.frombuffer(b"\x07\x07\x03" + bytes([9]) * run, dtype=np.uint8))


@pytest.mark.parametrize("n", [1, 2, 15, 16, 17, 31, 32, 33, 127, 128, 129, 255, 256, 257, 4095, 4096, 4097, 8191, 8192, 8193, 8192 + 256,
                               8192 * 2 - 1, 16384, 16385, 16384 + 127, 16384 + 128, 16384 * 2 - 255, 16384 * 2 - 1, 8192 * 3 + 77, 65537,
                               300001, (1 << 20) + 5])
@pytest.mark.parametrize("mean", [1.3, 4, 64, 700])
def test_rl_ragged_sizes(n, mean, oracle, flrl):
    _check(flrl, oracle, _runs(np.random.default_rng(n + int(mean * 10)), n, mean))


def test_rl_constant_buffers(oracle, flrl):
    """One run over many tiles: the run-start look-back must carry the phase across every tile."""
    for n in (8192 * 5, 8192 * 40 + 3, 255 * 8192, 255 * 8192 + 1, (1 << 22) + 11):
        _check(flrl, oracle, np.zeros(n, dtype=np.uint8))
    d = np.zeros(8192 * 9 + 100, dtype=np.uint8)
    d[8192 * 4 + 17] = 1  # one interruption in the middle
    _check(flrl, oracle, d)


def test_rl_no_runs_and_small_alphabet(oracle, flrl):
    rng = np.random.default_rng(1)
    _check(flrl, oracle, (np.arange(100003) % 251).astype(np.uint8))  # every byte is its own pair
    _check(flrl, oracle, rng.integers(0, 256, size=(1 << 20) + 3, dtype=np.uint8))
    _check(flrl, oracle, _runs(rng, 1 << 20, 3000, alphabet=2))  # long runs, many 255-splits
    _check(flrl, oracle, _runs(rng, 1 << 20, 300, alphabet=2))


@pytest.mark.parametrize("mean", [3, 500, 5000])
def test_rl_shards_concatenate(mean, oracle, flrl):
    """include/flrl.h shard semantics: prev_run + <=255 bytes of look-ahead make shard outputs concatenate to the
    single-stream output, wherever the cuts fall."""
    rng = np.random.default_rng(mean)
    data = _runs(rng, 400_000, mean, alphabet=3)
    c_ref, v_ref = oracle.rl_compress(data)
    cuts = [0, 1, 77, 8192, 100_001, 250_000, 399_999, data.size]
    cs, vs = [], []
    for a, b in zip(cuts[:-1], cuts[1:]):
        prev = 0
        while a - prev - 1 >= 0 and data[a - prev - 1] == data[a]:
            prev += 1
        ahead = min(data.size, b + 255)
        d = flrl.to_device(data[a:ahead])
        c, v = flrl.rl_encode_device(d, n=b - a, prev_run=prev)
        cs.append(c.cpu().numpy())
        vs.append(v.cpu().numpy())
    assert np.array_equal(np.concatenate(cs), c_ref)
    assert np.array_equal(np.concatenate(vs), v_ref)


def test_rl_decode_rejects_malformed(oracle, flrl):
    data = _runs(np.random.default_rng(5), 50_000, 20)
    c, v = oracle.rl_compress(data)
    with pytest.raises(flrl.FlrlError):
        flrl.rl_decode_device(flrl.to_device(c), flrl.to_device(v), data.size - 1)  # output too small
    bad = c.copy()
    bad[123] = 0
    with pytest.raises(flrl.FlrlError):
        flrl.rl_decode_device(flrl.to_device(bad), flrl.to_device(v), data.size)


@pytest.mark.parametrize("mean", [4, 64, 4096])
def test_rl_full_size_properties(mean, oracle, flrl):
    """BASELINE.json config 3 at full size (1 GiB): device round trip, sum(counts) == n, counts in 1..255, no two
    adjacent pairs mergeable; the whole pair stream is compared with the oracle."""
    import torch

    import bench

    n = 1 << 30
    d_in = bench.gen_rl_device(torch, n, mean)
    enc = flrl.RLEncoder(n)
    enc.launch(d_in)
    counts, values = enc.result()
    assert int(counts.sum(dtype=torch.int64).item()) == n
    assert int(counts.min().item()) >= 1
    same = values[1:] == values[:-1]
    assert bool((counts[:-1][same] == 255).all()), "adjacent equal-valued pairs must come from a 255-split"
    dec = flrl.RLDecoder(n, counts.numel())
    dec.launch(counts, values)
    assert torch.equal(dec.result(), d_in)
    c_ref, v_ref = oracle.rl_compress(d_in.cpu().numpy())  # EVERY pair against the oracle (the C port: seconds per GiB)
    assert counts.numel() == c_ref.size
    assert np.array_equal(counts.cpu().numpy(), c_ref)
    assert np.array_equal(values.cpu().numpy(), v_ref)


@pytest.mark.parametrize("mean", [1, 2])
def test_rl_dense_data_at_scale(mean, oracle, flrl):
    """Dense data (random bytes, mean run 2) at 256 MiB: every 16 KiB tile has more pair boundaries than the encoder's ring
    of staged pairs holds, so all of them go through its second pass, and the decoder takes its short-run (start-bit)
    expand everywhere.  Every pair is compared with the oracle."""
    import torch

    import bench

    n = 1 << 28
    d_in = torch.randint(0, 256, (n,), dtype=torch.uint8, device="cuda", generator=torch.Generator(device="cuda").manual_seed(5)) \
        if mean == 1 else bench.gen_rl_device(torch, n, 2)
    enc = flrl.RLEncoder(n)
    enc.launch(d_in)
    counts, values = enc.result()
    c_ref, v_ref = oracle.rl_compress(d_in.cpu().numpy())
    assert counts.numel() == c_ref.size
    assert np.array_equal(counts.cpu().numpy(), c_ref)
    assert np.array_equal(values.cpu().numpy(), v_ref)
    dec = flrl.RLDecoder(n, counts.numel())
    dec.launch(counts, values)
    assert torch.equal(dec.result(), d_in)
