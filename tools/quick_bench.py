"""Scratch device-resident timing of the FL/RL kernels (development aid; bench.py is the contract)."""
import importlib
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
flrl = importlib.import_module("fl-rl-compression-mpi_b200")


def gen_fl(n, seed=42):
    g = torch.Generator(device="cuda").manual_seed(seed)
    w = torch.randint(1, 9, (n // 128,), device="cuda", generator=g, dtype=torch.int32)
    data = torch.randint(0, 256, (n // 128, 128), device="cuda", generator=g, dtype=torch.int32)
    data &= ((1 << w) - 1)[:, None]
    data[:, 0] |= 1 << (w - 1)
    return data.to(torch.uint8).reshape(-1), w


def timeit(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return min(ts), sum(ts) / len(ts)


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 30
    d_in, w = gen_fl(n)
    enc = flrl.FLEncoder(n)
    enc.launch(d_in)
    bits, values = enc.result(n)
    ok = torch.equal(bits.to(torch.int32), w) and values.numel() == int(w.sum().item()) * 16
    alg = n + n // 128 + values.numel()
    best, avg = timeit(lambda: enc.launch(d_in))
    print(f"fl encode n={n} ok={ok} best {best:.3f} ms avg {avg:.3f} ms  {n/best/1e6:.1f} GB/s uncompressed, "
          f"{alg/best/1e6:.1f} GB/s algorithmic ({alg/best/1e6/6546.9:.3f} of measured HBM)")
    dec = flrl.FLDecoder(n)
    dec.launch(bits, values, n)
    ok = torch.equal(dec.result(n), d_in)
    best, avg = timeit(lambda: dec.launch(bits, values, n))
    print(f"fl decode n={n} ok={ok} best {best:.3f} ms avg {avg:.3f} ms  {n/best/1e6:.1f} GB/s uncompressed, "
          f"{alg/best/1e6:.1f} GB/s algorithmic ({alg/best/1e6/6546.9:.3f} of measured HBM)")
    # copy roofline on this very box
    dst = torch.empty_like(d_in)
    best, _ = timeit(lambda: dst.copy_(d_in))
    print(f"torch copy {2*n/best/1e6:.1f} GB/s")


if __name__ == "__main__":
    main()
