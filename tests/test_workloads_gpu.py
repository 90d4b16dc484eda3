"""Parity on the data shapes of BASELINE.json configs[3] and configs[4] (SURVEY.md §8d inputs 4 and 5) at sizes the oracle
finishes in seconds: BMP-like 24-bit image stack and the mixed-entropy stream, FL and RL, device layer and host layer.  -m gpu."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def bmp_like_stack(images: int) -> np.ndarray:
    """1280x853 24-bit images behind a 138-byte header (BITMAPV5, bfOffBits = 138): channel value
    clamp(A_c x/W + B_c y/H + 8 sin(2 pi t/64) + noise), noise ~ U{-n..n}, n cycling {0, 1, 3} per image; seed = image index."""
    W, H = 1280, 853
    out = []
    x = np.arange(W, dtype=np.float32)[None, :, None] / W
    y = np.arange(H, dtype=np.float32)[:, None, None] / H
    t = (np.arange(W, dtype=np.float32)[None, :, None] + np.arange(H, dtype=np.float32)[:, None, None])
    A = np.array([200, 90, 30], dtype=np.float32)[None, None, :]
    B = np.array([20, 120, 210], dtype=np.float32)[None, None, :]
    for i in range(images):
        rng = np.random.default_rng(i)
        nz = (0, 1, 3)[i % 3]
        noise = rng.integers(-nz, nz + 1, size=(H, W, 3)) if nz else 0
        img = np.clip(A * x + B * y + 8 * np.sin(2 * np.pi * t / 64) + noise, 0, 255).astype(np.uint8)
        hdr = np.zeros(138, dtype=np.uint8)
        hdr[0:2] = (0x42, 0x4D)
        hdr[10] = 138
        out += [hdr, img.reshape(-1)[: 3275520]]
    return np.concatenate(out)


def mixed_entropy_stream(segments: int, seg: int = 1 << 20) -> np.ndarray:
    """1 MiB segments cycling: zeros, constant runs (mean 4096), narrow noise (w in 1..4), geometric runs (mean 64),
    full-entropy bytes; seed = segment index."""
    out = []
    for i in range(segments):
        rng = np.random.default_rng(i)
        k = i % 5
        if k == 0:
            s = np.zeros(seg, dtype=np.uint8)
        elif k in (1, 3):
            mean = 4096 if k == 1 else 64
            lens = rng.geometric(1.0 / mean, size=seg // mean * 2 + 64)
            s = np.repeat(rng.integers(0, 256, size=lens.size).astype(np.uint8), lens)
            while s.size < seg:
                s = np.concatenate([s, s])
            s = s[:seg]
        elif k == 2:
            w = rng.integers(1, 5, size=seg // 128)
            s = (rng.integers(0, 256, size=seg) & ((1 << np.repeat(w, 128)) - 1)).astype(np.uint8)
        else:
            s = rng.integers(0, 256, size=seg, dtype=np.uint8)
        out.append(s)
    return np.concatenate(out)


DATA = {"bmp_stack": lambda: bmp_like_stack(3), "mixed_stream": lambda: mixed_entropy_stream(11)}


@pytest.mark.parametrize("which", list(DATA))
def test_fl_device_parity(which, oracle, flrl):
    data = DATA[which]()
    bits, values = flrl.fl_encode_device(flrl.to_device(data))
    b_ref, v_ref = oracle.fl_compress(data)
    assert np.array_equal(bits.cpu().numpy(), b_ref) and np.array_equal(values.cpu().numpy(), v_ref)
    out = flrl.fl_decode_device(flrl.to_device(b_ref), flrl.to_device(v_ref), data.size)
    assert np.array_equal(out.cpu().numpy(), data)


@pytest.mark.parametrize("which", list(DATA))
def test_rl_device_parity(which, oracle, flrl):
    data = DATA[which]()
    counts, values = flrl.rl_encode_device(flrl.to_device(data))
    c_ref, v_ref = oracle.rl_compress(data)
    assert np.array_equal(counts.cpu().numpy(), c_ref) and np.array_equal(values.cpu().numpy(), v_ref)
    out = flrl.rl_decode_device(flrl.to_device(c_ref), flrl.to_device(v_ref), data.size)
    assert np.array_equal(out.cpu().numpy(), data)


@pytest.mark.parametrize("kind", ["fl", "rl"])
def test_host_layer_mixed_stream_chunked(kind, oracle, flrl):
    """configs[4] shape through the host layer's chunk pipeline (2 MiB chunks: every segment kind meets a chunk cut),
    on one GPU and on every GPU count the box has."""
    import torch

    flrl.lib.flrl_set_pipe_chunk(2 << 20)
    try:
        data = mixed_entropy_stream(11)[: (11 << 20) - 4321]
        m_ref, p_ref = (oracle.fl_compress if kind == "fl" else oracle.rl_compress)(data)
        for g in [g for g in (1, 2, 4, 8) if g <= torch.cuda.device_count()]:
            hc = flrl.HostCodec(kind, data.size, ngpus=g)
            hc.load_input(data)
            hc.compress()
            meta, payload = hc.compressed()
            assert np.array_equal(meta, m_ref) and np.array_equal(payload, p_ref), g
            hc.decompress()
            assert hc.check_round_trip(), g
    finally:
        flrl.lib.flrl_set_pipe_chunk(0)
