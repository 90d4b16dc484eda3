"""Scratch device-resident timing of the FL kernels (development aid).  env: FLRL_LIB_PATH selects a variant build."""
import importlib, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
flrl = importlib.import_module("fl-rl-compression-mpi_b200")
import bench
PEAK = bench.peaks()[0]
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 30
tag = os.environ.get("FLRL_LIB_PATH", "default").split("/")[-1]
d, w = bench.gen_fl_device(torch, n)
enc, dec = flrl.FLEncoder(n), flrl.FLDecoder(n)
enc.launch(d)
bits, values = enc.result(n)
dec.launch(bits, values, n)
ok = torch.equal(dec.result(n), d)
alg = n + bits.numel() + values.numel()
def timeit(fn, iters=10):
    for _ in range(3): fn()
    torch.cuda.synchronize(); ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    return min(ts), sum(ts) / len(ts)
for name, fn in (("encode", lambda: enc.launch(d)), ("decode", lambda: dec.launch(bits, values, n))):
    best, avg = timeit(fn)
    print(f"[{tag}] fl {name} ok={ok} best {best:.4f} avg {avg:.4f} ms  frac {alg/best/1e6/PEAK:.3f}", flush=True)
