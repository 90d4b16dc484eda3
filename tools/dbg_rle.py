"""Scratch: RL encoder diagnostics on the GPU (development aid)."""
import importlib, os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
flrl = importlib.import_module("fl-rl-compression-mpi_b200")
from oracle import oracle as o
import bench

def case(data, tag):
    data = np.ascontiguousarray(data, dtype=np.uint8)
    c_ref, v_ref = o.rl_compress(data)
    enc = flrl.RLEncoder(data.size)
    enc.counts.fill_(77); enc.values.fill_(77)
    enc.launch(flrl.to_device(data))
    c, v = enc.result()
    c, v = c.cpu().numpy(), v.cpu().numpy()
    ok = c.size == c_ref.size and np.array_equal(c, c_ref) and np.array_equal(v, v_ref)
    msg = f"{tag}: n={data.size} P={c.size}/{c_ref.size} ok={ok}"
    if not ok and c.size == c_ref.size:
        bc = np.flatnonzero(c != c_ref); bv = np.flatnonzero(v != v_ref)
        msg += f" bad counts {bc.size} first {bc[:5]} got {c[bc[:5]]} want {c_ref[bc[:5]]}; bad values {bv.size} first {bv[:5]} got {v[bv[:5]]} want {v_ref[bv[:5]]}"
    print(msg, flush=True)

for run in [16383, 16384, 16385, 16384 + 255, 32768 + 254, 100000]:
    case(np.frombuffer(bytes([9]) * run + b"\x01", dtype=np.uint8), f"run{run}a")
    case(np.frombuffer(b"\x07\x07\x03" + bytes([9]) * run, dtype=np.uint8), f"run{run}b")
rng = np.random.default_rng(1)
def runs(n, mean, alphabet=256):
    lens = rng.geometric(1.0 / mean, size=int(n / mean * 1.2) + 64)
    vals = rng.integers(0, alphabet, size=lens.size).astype(np.uint8)
    return np.repeat(vals, lens)[:n].copy()
for n in [16384 * 2 - 1, 65537, 300001, (1 << 20) + 5, (1 << 22) + 11]:
    for mean in [1.3, 4, 64, 700, 5000]:
        case(runs(n, mean), f"n{n}m{mean}")
case(np.zeros((1 << 22) + 11, np.uint8), "zeros")
for lg in [24, 26, 28, 30]:
    for mean in [4096, 64, 4]:
        n = 1 << lg
        d = bench.gen_rl_device(torch, n, mean)
        enc = flrl.RLEncoder(n)
        torch.cuda.synchronize()
        t = time.time()
        enc.launch(d)
        torch.cuda.synchronize()
        counts, vals = enc.result()
        ok = torch.equal(torch.repeat_interleave(vals, counts.to(torch.int64)), d)
        print(f"2^{lg} mean {mean}: first launch {1e3*(time.time()-t):.2f} ms, P={counts.numel()} roundtrip ok={ok}", flush=True)
        ts = []
        for _ in range(5):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); enc.launch(d); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
        print(f"   best {min(ts):.3f} ms -> {(n + 2*counts.numel())/min(ts)/1e6:.0f} GB/s algorithmic", flush=True)
        del d, enc, counts, vals
