"""The four kernels are persistent (a CTA takes tiles from an atomic counter and waits for earlier tiles).  Several of them
sharing one GPU from different streams must neither deadlock nor disturb each other: every stream's output is checked
against the oracle.  Needs a B200: run with -m gpu."""
import numpy as np
import pytest

from conftest import make_fl_mix

pytestmark = pytest.mark.gpu


def _rl_runs(n, mean, seed):
    rng = np.random.default_rng(seed)
    lens = rng.geometric(1.0 / mean, size=max(n // mean * 2, 16))
    vals = rng.integers(0, 256, size=lens.size, dtype=np.uint8)
    return np.repeat(vals, lens)[:n].copy()


@pytest.mark.timeout(120)
def test_streams_share_the_gpu(oracle, flrl):
    import torch

    n = 48 << 20  # each launch wants the whole GPU (a few thousand tiles, a grid as large as fits)
    fl_in = [make_fl_mix(n, seed=s) for s in (1, 2)]
    rl_in = [_rl_runs(n, 40, 3), _rl_runs(n, 3000, 4)]
    streams = [torch.cuda.Stream() for _ in range(4)]
    d_fl = [flrl.to_device(a) for a in fl_in]
    d_rl = [flrl.to_device(a) for a in rl_in]
    fl_enc = [flrl.FLEncoder(n) for _ in fl_in]
    rl_enc = [flrl.RLEncoder(n) for _ in rl_in]
    torch.cuda.synchronize()
    for rep in range(3):  # three waves, all four streams busy at once
        for i in range(2):
            with torch.cuda.stream(streams[i]):
                fl_enc[i].launch(d_fl[i])
            with torch.cuda.stream(streams[2 + i]):
                rl_enc[i].launch(d_rl[i])
    torch.cuda.synchronize()
    fl_out = [e.result(n) for e in fl_enc]
    rl_out = [e.result() for e in rl_enc]
    for i in range(2):
        b_ref, v_ref = oracle.fl_compress(fl_in[i])
        assert np.array_equal(fl_out[i][0].cpu().numpy(), b_ref)
        assert np.array_equal(fl_out[i][1].cpu().numpy(), v_ref)
        c_ref, w_ref = oracle.rl_compress(rl_in[i])
        assert np.array_equal(rl_out[i][0].cpu().numpy(), c_ref)
        assert np.array_equal(rl_out[i][1].cpu().numpy(), w_ref)

    # the decoders, the same way
    fl_dec = [flrl.FLDecoder(n) for _ in fl_in]
    rl_dec = [flrl.RLDecoder(n) for _ in rl_in]
    torch.cuda.synchronize()
    for rep in range(3):
        for i in range(2):
            with torch.cuda.stream(streams[i]):
                fl_dec[i].launch(fl_out[i][0], fl_out[i][1], n)
            with torch.cuda.stream(streams[2 + i]):
                rl_dec[i].launch(rl_out[i][0], rl_out[i][1])
    torch.cuda.synchronize()
    for i in range(2):
        assert np.array_equal(fl_dec[i].result(n).cpu().numpy(), fl_in[i])
        assert np.array_equal(rl_dec[i].result().cpu().numpy(), rl_in[i])
