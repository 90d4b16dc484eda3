"""Scratch device-resident timing of the RL kernels (development aid; bench.py is the contract).
usage: quick_rl.py [n] [means...]   env: FLRL_LIB_PATH selects a variant build of libflrl.so"""
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
flrl = importlib.import_module("fl-rl-compression-mpi_b200")
import bench  # generators

PEAK = 6530.6


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


def expected_pairs(d):
    n = d.numel()
    b = torch.ones(n, dtype=torch.bool, device=d.device)
    b[1:] = d[1:] != d[:-1]
    st = torch.nonzero(b).flatten()
    L = torch.diff(torch.cat([st, torch.tensor([n], device=d.device)]))
    return int(((L + 254) // 255).sum().item())


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 30
    means = [int(x) for x in sys.argv[2:]] or [4, 64, 4096]
    tag = os.environ.get("FLRL_LIB_PATH", "default").split("/")[-1]
    for mean in means:
        d = bench.gen_rl_device(torch, n, mean)
        enc, dec = flrl.RLEncoder(n), flrl.RLDecoder(n)
        enc.launch(d)
        counts, vals = enc.result()
        P = counts.numel()
        ok_p = P == expected_pairs(d)
        ok_rt = torch.equal(torch.repeat_interleave(vals, counts.to(torch.int64)), d)
        alg = n + 2 * P
        best, avg = timeit(lambda: enc.launch(d))
        print(f"[{tag}] rl{mean} encode ok={ok_p and ok_rt} best {best:.3f} avg {avg:.3f} ms  {n/best/1e6:.0f} GB/s unc, frac {alg/best/1e6/PEAK:.3f}", flush=True)
        dec.launch(counts, vals)
        ok = torch.equal(dec.result(), d)
        best, avg = timeit(lambda: dec.launch(counts, vals))
        print(f"[{tag}] rl{mean} decode ok={ok} best {best:.3f} avg {avg:.3f} ms  {n/best/1e6:.0f} GB/s unc, frac {alg/best/1e6/PEAK:.3f}", flush=True)
        del d, enc, dec, counts, vals


if __name__ == "__main__":
    main()
