"""Host layer of the C ABI (include/flrl.h: flrl_*_gpu_compress / _decompress, HOST buffers) against the oracle,
with pageable and pinned buffers, one GPU and — when the box has them — several (NCCL size all-gather).  -m gpu."""
import numpy as np
import pytest

from conftest import make_fl_mix

pytestmark = pytest.mark.gpu


def _rl_data(n, mean, seed):
    rng = np.random.default_rng(seed)
    lens = rng.geometric(1.0 / mean, size=int(n / mean * 1.3) + 64)
    while lens.sum() < n:
        lens = np.concatenate([lens, lens])
    return np.repeat(rng.integers(0, 4, size=lens.size).astype(np.uint8), lens)[:n].copy()


def _ngpus():
    import torch

    return [g for g in (1, 2, 4, 8) if g <= torch.cuda.device_count()]


@pytest.mark.parametrize("n", [0, 1, 127, 128, 129, 16384, 100_003, (1 << 22) + 77, (80 << 20) + 13])
def test_fl_host_pageable(n, oracle, flrl):
    """malloc-owned outputs, pageable input: the reference's gpuCompress / gpuDecompress shape (fl_gpu.cu:289-645)."""
    data = make_fl_mix(n, seed=n + 1)
    bits, values = flrl.fl_gpu_compress(data)
    b_ref, v_ref = oracle.fl_compress(data)
    assert np.array_equal(bits, b_ref) and np.array_equal(values, v_ref)
    back = flrl.fl_gpu_decompress(n, b_ref, v_ref)
    assert np.array_equal(back, data)


@pytest.mark.parametrize("n", [0, 1, 255, 256, 8193, 100_003, (1 << 22) + 77, (70 << 20) + 5])
@pytest.mark.parametrize("mean", [3, 900])
def test_rl_host_pageable(n, mean, oracle, flrl):
    data = _rl_data(n, mean, n + mean)
    counts, values = flrl.rl_gpu_compress(data)
    c_ref, v_ref = oracle.rl_compress(data)
    assert np.array_equal(counts, c_ref) and np.array_equal(values, v_ref)
    back = flrl.rl_gpu_decompress(c_ref, v_ref, expected_size=n)
    assert np.array_equal(back, data)


@pytest.mark.parametrize("kind", ["fl", "rl"])
def test_host_pinned_round_trip(kind, oracle, flrl):
    n = (24 << 20) + 128 * 3 + 5
    data = make_fl_mix(n, seed=3) if kind == "fl" else _rl_data(n, 40, 4)
    for g in _ngpus():
        hc = flrl.HostCodec(kind, n, ngpus=g)
        hc.load_input(data)
        hc.compress()
        meta, payload = hc.compressed()
        m_ref, p_ref = (oracle.fl_compress if kind == "fl" else oracle.rl_compress)(data)
        assert np.array_equal(meta, m_ref) and np.array_equal(payload, p_ref), (kind, g)
        hc.decompress()
        assert hc.check_round_trip(), (kind, g)
        k_ms, t_ms = hc.timing()
        assert 0 < k_ms <= t_ms


@pytest.mark.parametrize("kind,mean", [("fl", 0), ("rl", 40), ("rl", 3), ("rl", 3_000_000)])
def test_host_chunk_pipeline(kind, mean, oracle, flrl):
    """The single-GPU chunk pipeline (copy-in / kernel / copy-out overlapped, pinned buffers) with small chunks: chunk
    outputs must concatenate to the single-stream output (FL: 128-aligned chunks; RL: prev_run + look-ahead, runs
    longer than a chunk carried over whole chunks)."""
    flrl.lib.flrl_set_pipe_chunk(1 << 20)
    try:
        n = (11 << 20) + 128 * 5 + 77
        data = make_fl_mix(n, seed=5) if kind == "fl" else _rl_data(n, mean, 6)
        m_ref, p_ref = (oracle.fl_compress if kind == "fl" else oracle.rl_compress)(data)
        for g in _ngpus():  # every shard runs its own pipeline; shard 0 delivers as it goes, the others after the size exchange
            for peer in ((0, 1) if g > 1 else (0,)):
                flrl.lib.flrl_set_concat_mode(peer)
                hc = flrl.HostCodec(kind, n, ngpus=g)
                hc.load_input(data)
                hc.compress()
                meta, payload = hc.compressed()
                assert np.array_equal(meta, m_ref) and np.array_equal(payload, p_ref), (kind, g, peer)
                hc.decompress()
                assert hc.check_round_trip(), (kind, g, peer)
                k_ms, t_ms = hc.timing()
                assert 0 < k_ms <= t_ms
                per = hc.timing_per_gpu()
                assert len(per) == g and all(t > 0 and bi > 0 and bo > 0 for _, t, bi, bo in per)
    finally:
        flrl.lib.flrl_set_pipe_chunk(0)
        flrl.lib.flrl_set_concat_mode(0)


def test_host_pipeline_rejects_malformed(oracle, flrl):
    """A container whose widths are out of range must be refused before any chunk is copied: a width of 255 would make
    a chunk's value bytes exceed the pipeline's staging buffer (pinned buffers, several chunks)."""
    flrl.lib.flrl_set_pipe_chunk(1 << 20)
    try:
        n = (5 << 20) + 300
        data = make_fl_mix(n, seed=11)
        hc = flrl.HostCodec("fl", n, ngpus=1)
        hc.load_input(data)
        hc.compress()
        meta, payload = hc.compressed()
        meta[12345] = 255
        with pytest.raises(flrl.FlrlError):
            hc.decompress()
        meta[12345] = 0
        with pytest.raises(flrl.FlrlError):
            hc.decompress()
        hc.compress()
        hc.s1.value -= 4096  # values shorter than the widths imply
        with pytest.raises(flrl.FlrlError):
            hc.decompress()
        rc = flrl.HostCodec("rl", n, ngpus=1)
        rc.load_input(_rl_data(n, 5, 12))
        rc.compress()
        counts, _ = rc.compressed()
        counts[777] = 0
        with pytest.raises(flrl.FlrlError):
            rc.decompress()
    finally:
        flrl.lib.flrl_set_pipe_chunk(0)


def test_host_release_and_reuse(oracle, flrl):
    """flrl_release() frees every cached buffer, stream and event (the pipeline's too); the next call rebuilds them."""
    n = (3 << 20) + 5
    data = make_fl_mix(n, seed=13)
    flrl.lib.flrl_set_pipe_chunk(1 << 20)
    try:
        hc = flrl.HostCodec("fl", n, ngpus=1)
        hc.load_input(data)
        hc.compress()
        flrl.lib.flrl_release()
        assert flrl.lib.flrl_prepare(1) == 0  # optional: contexts, kernel modules (and NCCL for several GPUs) up front
        hc.compress()
        hc.decompress()
        assert hc.check_round_trip()
    finally:
        flrl.lib.flrl_set_pipe_chunk(0)


@pytest.mark.parametrize("peer", [0, 1])
def test_multi_gpu_concat_identity(peer, oracle, flrl):
    """Shards + NCCL size all-gather + concatenation == single-stream output, both concat modes."""
    gs = [g for g in _ngpus() if g > 1]
    if not gs:
        pytest.skip("needs >= 2 GPUs")
    flrl.lib.flrl_set_concat_mode(peer)
    try:
        n = (9 << 20) + 1000
        fl = make_fl_mix(n, seed=8)
        rl = _rl_data(n, 700, 9)  # long runs straddle the shard cuts
        for g in gs:
            b, v = flrl.fl_gpu_compress(fl, ngpus=g)
            b_ref, v_ref = oracle.fl_compress(fl)
            assert np.array_equal(b, b_ref) and np.array_equal(v, v_ref), g
            assert np.array_equal(flrl.fl_gpu_decompress(n, b, v, ngpus=g), fl), g
            c, x = flrl.rl_gpu_compress(rl, ngpus=g)
            c_ref, x_ref = oracle.rl_compress(rl)
            assert np.array_equal(c, c_ref) and np.array_equal(x, x_ref), g
            assert np.array_equal(flrl.rl_gpu_decompress(c, x, expected_size=n, ngpus=g), rl), g
    finally:
        flrl.lib.flrl_set_concat_mode(0)


def test_host_rejects_malformed(oracle, flrl):
    data = make_fl_mix(50_000, seed=2)
    b, v = oracle.fl_compress(data)
    bad = b.copy()
    bad[3] = 11
    with pytest.raises(flrl.FlrlError):
        flrl.fl_gpu_decompress(data.size, bad, v)
    with pytest.raises(flrl.FlrlError):
        flrl.fl_gpu_decompress(data.size, b, v[: v.size // 2])
    with pytest.raises(flrl.FlrlError):
        flrl.fl_gpu_decompress(data.size, b[:-1], v)
