"""Development aid: per-tile phase timeline of rl_encode_kernel (needs a -DFLRL_TRACE variant build:
tools/build_variant.sh trace -DFLRL_TRACE; FLRL_LIB_PATH=.../variants/libflrl_trace.so python tools/trace_rle.py [mean])."""
import ctypes as C
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
flrl = importlib.import_module("fl-rl-compression-mpi_b200")
import bench

mean = int(sys.argv[1]) if len(sys.argv) > 1 else 64
n = 1 << 30
d = bench.gen_rl_device(torch, n, mean)
enc = flrl.RLEncoder(n)
for _ in range(3):
    enc.launch(d)
torch.cuda.synchronize()
tiles = n // 16384 + 2
buf = torch.zeros(tiles * 16, dtype=torch.int64, device="cuda")
flrl.lib.flrl_debug_set_rle_trace.argtypes = [C.c_void_p]
assert flrl.lib.flrl_debug_set_rle_trace(buf.data_ptr()) == 0
enc.launch(d)
torch.cuda.synchronize()
t = buf.reshape(tiles, 16)[1000:60000].double()
names = {0: "sv: loads issued", 8: "c: first data", 9: "c: phase1 done", 1: "sv: LB#1 done", 10: "c: phase2 done",
         3: "sv: p2 seen", 12: "c: compaction done", 14: "c: tile done (+drain)", 5: "base: slot known"}
t0 = t[:, 0]
for k in (0, 8, 9, 1, 10, 3, 12, 14, 5):
    dt = (t[:, k] - t0) / 1000.0
    print(f"{names[k]:22s} mean {dt.mean().item():7.2f} us  p10 {dt.quantile(0.1).item():7.2f}  p90 {dt.quantile(0.9).item():7.2f}")
starts = torch.sort(t0).values
print("tile start rate: %.1f tiles/us" % ((starts.numel() - 1) / ((starts[-1] - starts[0]).item() / 1000.0)))
