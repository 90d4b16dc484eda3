"""flrl_fl_decode_static (include/flrl.h): the FL decoder with no allocation on its path — its scratch is a static device
array inside the library.  Checked through the C ABI against the existing entry point flrl_fl_decode, bit-exact (the
tolerance of every FL test here), on two calls in a row with different inputs (the scratch keeps the first call's
contents).  Collected last on purpose.  -m gpu."""
import numpy as np
import pytest

from conftest import make_fl_mix

pytestmark = pytest.mark.gpu


def _decode_static(flrl, torch, bits, values, n):
    out = torch.full((max((n + 15) // 16 * 16, 16),), 0xA5, dtype=torch.uint8, device="cuda")  # the caller's buffers, as for flrl_fl_decode
    status = torch.full((1,), 77, dtype=torch.int32, device="cuda")
    rc = flrl.lib.flrl_fl_decode_static(bits.data_ptr(), bits.numel(), values.data_ptr(), values.numel(), out.data_ptr(), n,
                                        status.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert rc == 0, flrl.lib.flrl_last_error()
    torch.cuda.synchronize()
    return out[:n], int(status.item())


def test_fl_decode_static_matches_fl_decode_two_calls_in_a_row(flrl):
    import torch

    # the second call is smaller and different: stale descriptors / tile offsets of the first must not leak into it
    for n, seed in (((5 << 20) + 128 * 3 + 77, 1), ((2 << 20) + 41, 2)):
        data = make_fl_mix(n, seed=seed)
        d_in = flrl.to_device(data)
        bits, values = flrl.fl_encode_device(d_in)
        want = flrl.fl_decode_device(bits, values, n)  # existing entry point: flrl_fl_decode with a caller workspace
        got, status = _decode_static(flrl, torch, bits, values, n)
        assert status == 0
        assert torch.equal(got, want), n
        assert np.array_equal(got.cpu().numpy(), data), n


def test_fl_decode_static_reports_malformed_and_rejects_oversize(flrl):
    import torch

    n = (1 << 20) + 5
    data = make_fl_mix(n, seed=3)
    bits, values = flrl.fl_encode_device(flrl.to_device(data))
    bad = bits.clone()
    bad[100] = 9
    _, status = _decode_static(flrl, torch, bad, values, n)
    assert status != 0  # same status word as flrl_fl_decode
    got, status = _decode_static(flrl, torch, bits, values, n)  # and the next call is clean again
    assert status == 0 and np.array_equal(got.cpu().numpy(), data)
    too_big = flrl.STATIC_MAX_BYTES + 128
    rc = flrl.lib.flrl_fl_decode_static(bits.data_ptr(), (too_big + 127) // 128, values.data_ptr(), values.numel(), got.data_ptr(), too_big,
                                        0, torch.cuda.current_stream().cuda_stream)
    assert rc == -3 and b"FLRL_STATIC_MAX_BYTES" in flrl.lib.flrl_last_error()  # FLRL_E_CAPACITY, nothing launched
