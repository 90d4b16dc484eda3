"""FL parity on the GPU: the CUDA path through the C ABI (include/flrl.h) against the oracle,
bit-exact.  Needs a B200: run with -m gpu."""
import hashlib
import os

import numpy as np
import pytest

from conftest import make_fl_mix
from golden_inputs import INPUTS, make_input

pytestmark = pytest.mark.gpu
GOLDEN = os.path.dirname(os.path.abspath(__file__)) + "/golden"


def _roundtrip(flrl, oracle, data):
    import torch

    data = np.ascontiguousarray(data, dtype=np.uint8)
    n = data.size
    d_in = flrl.to_device(data)
    bits, values = flrl.fl_encode_device(d_in)
    b_ref, v_ref = oracle.fl_compress(data)
    assert bits.numel() == b_ref.size and values.numel() == v_ref.size, (n, bits.numel(), b_ref.size, values.numel(), v_ref.size)
    assert np.array_equal(bits.cpu().numpy(), b_ref), n
    assert np.array_equal(values.cpu().numpy(), v_ref), n
    if n:
        out = flrl.fl_decode_device(flrl.to_device(b_ref), flrl.to_device(v_ref), n)
        assert np.array_equal(out.cpu().numpy(), data), n
    torch.cuda.synchronize()
    return bits, values


@pytest.mark.parametrize("name", list(INPUTS))
def test_fl_golden_vectors_gpu(name, golden, oracle, flrl):
    """The reference binary's own outputs (tests/golden)."""
    e = next(x for x in golden["fl"] if x["name"] == name)
    data = np.frombuffer(make_input(name, GOLDEN), dtype=np.uint8)
    bits, values = _roundtrip(flrl, oracle, data)
    blob = oracle.fl_container(data.size, bits.cpu().numpy(), values.cpu().numpy())
    assert len(blob) == e["file_len"]
    assert hashlib.sha256(blob).hexdigest() == e["sha256"]


@pytest.mark.parametrize("n", [1, 2, 15, 16, 127, 128, 129, 255, 256, 257, 1000, 16383, 16384, 16385, 16384 + 128,
                               16384 * 2 - 1, 16384 * 3 + 77, 65537, 300001, 1 << 20, (1 << 22) + 5])
def test_fl_ragged_sizes(n, oracle, flrl):
    _roundtrip(flrl, oracle, make_fl_mix(n, seed=n))


@pytest.mark.parametrize("w", range(0, 9))
def test_fl_uniform_width(w, oracle, flrl):
    """Every frame the same width (w = 0: all-zero input, width clamps to 1)."""
    rng = np.random.default_rng(w)
    n = 16384 * 5 + 128 * 3 + 17
    data = (rng.integers(0, 256, size=n, dtype=np.uint16) & ((1 << w) - 1)).astype(np.uint8)
    _roundtrip(flrl, oracle, data)


def test_fl_random_bytes(oracle, flrl):
    rng = np.random.default_rng(3)
    _roundtrip(flrl, oracle, rng.integers(0, 256, size=(1 << 21) + 3, dtype=np.uint8))


def test_fl_decode_rejects_malformed(oracle, flrl):
    data = make_fl_mix(50000, seed=9)
    b, v = oracle.fl_compress(data)
    bad = b.copy()
    bad[7] = 9
    with pytest.raises(flrl.FlrlError):
        flrl.fl_decode_device(flrl.to_device(bad), flrl.to_device(v), data.size)
    with pytest.raises(flrl.FlrlError):
        flrl.fl_decode_device(flrl.to_device(b), flrl.to_device(v[: v.size // 2]), data.size)
    with pytest.raises(flrl.FlrlError):
        flrl.fl_decode_device(flrl.to_device(b[:-1]), flrl.to_device(v), data.size)


def test_fl_full_size_properties(oracle, flrl):
    """BASELINE.json config 2 at full size (1 GiB): round trip on the device, widths equal the
    generator's, value size equals 16*sum(widths); all of it is compared with the reference's CPU codec."""
    import torch

    n = 1 << 30
    g = torch.Generator(device="cuda").manual_seed(42)
    w = torch.randint(1, 9, (n // 128,), device="cuda", generator=g, dtype=torch.int32)
    data = torch.randint(0, 256, (n // 128, 128), device="cuda", generator=g, dtype=torch.int32)
    data &= ((1 << w) - 1)[:, None]
    data[:, 0] |= 1 << (w - 1)
    d_in = data.to(torch.uint8).reshape(-1)
    del data
    enc = flrl.FLEncoder(n)
    enc.launch(d_in)
    bits, values = enc.result(n)
    assert torch.equal(bits.to(torch.int32), w)
    assert values.numel() == int(w.sum().item()) * 16
    dec = flrl.FLDecoder(n)
    dec.launch(bits, values, n)
    assert torch.equal(dec.result(n), d_in)
    # EVERY byte against the reference itself: its cpuCompress (oracle/_ref, compiled in place from the unmodified sources)
    # — or the port when that is not built — on all host threads over 128-byte aligned chunks, whose outputs concatenate
    # to the single-stream output.  A prefix compare would miss a once-per-100k-tiles hazard.
    import os
    import threading

    host = d_in.cpu().numpy()
    bits_h, values_h = bits.cpu().numpy(), values.cpu().numpy()
    comp = oracle.ref_fl_compress if oracle.have_ref() else oracle.fl_compress
    T = max(1, len(os.sched_getaffinity(0)))
    per = (n // (128 * 4 * T)) * 128
    cuts = list(range(0, n, per)) + [n]
    res = [None] * (len(cuts) - 1)

    def job(t):
        for i in range(t, len(res), T):
            res[i] = comp(host[cuts[i]:cuts[i + 1]])

    saved = os.dup(1)  # the reference prints [TIMER] lines with printf
    dn = os.open(os.devnull, os.O_WRONLY)
    os.dup2(dn, 1)
    try:
        ts = [threading.Thread(target=job, args=(t,)) for t in range(T)]
        [t.start() for t in ts]
        [t.join() for t in ts]
    finally:
        os.dup2(saved, 1)
        os.close(saved)
        os.close(dn)
    assert np.array_equal(bits_h, np.concatenate([r[0] for r in res])), "widths differ from the reference's"
    v_ref = np.concatenate([r[1] for r in res])
    assert values_h.size == v_ref.size and np.array_equal(values_h, v_ref), "values differ from the reference's"
