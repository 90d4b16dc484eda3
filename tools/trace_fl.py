"""Development aid: per-tile phase timeline of fl_encode_kernel (library built with EXTRA=-DFLRL_TRACE)."""
import ctypes as C
import importlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
flrl = importlib.import_module("fl-rl-compression-mpi_b200")
from quick_bench import gen_fl  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 30
d_in, w = gen_fl(n)
enc = flrl.FLEncoder(n)
for _ in range(3):
    enc.launch(d_in)
torch.cuda.synchronize()
tiles = (n // 128 + 127) // 128
lead = min(int(os.environ.get("FLRL_FL_LEAD", 2048)), tiles)
ctas = tiles + lead
buf = torch.zeros(ctas * 12, dtype=torch.int64, device="cuda")
flrl.lib.flrl_debug_set_trace.argtypes = [C.c_void_p]
assert flrl.lib.flrl_debug_set_trace(buf.data_ptr()) == 0
enc.launch(d_in)
torch.cuda.synchronize()
flrl.lib.flrl_debug_set_trace(None)
t = buf.cpu().numpy().reshape(ctas, 12).astype(np.float64)[lead:tiles - 1]  # CTAs that do both passes
t0 = t[:, 0].min()
print(f"kernel span {(t[:, 7].max() - t0) / 1e3:.1f} us over {tiles} tiles, lead {lead}")


def stat(x, label):
    print(f"{label:28s} mean {x.mean():8.0f} ns  p50 {np.percentile(x, 50):8.0f}  p90 {np.percentile(x, 90):8.0f}  max {x.max():8.0f}")


stat(t[:, 1] - t[:, 0], "start->P loaded (L2)")
stat(t[:, 2] - t[:, 1], "P loaded->offsets")
stat(t[:, 5] - t[:, 2], "offsets->packed")
stat(t[:, 3] - t[:, 5], "packed->W loaded (DRAM)")
stat(t[:, 3] - t[:, 0], "start->W loaded (DRAM)")
stat(t[:, 4] - t[:, 3], "W loaded->agg posted")
stat(t[:, 6] - t[:, 0], "start->lookback done")
stat(t[:, 7] - t[:, 0], "control warp lifetime")
stat(t[:, 4] - t[:, 0], "pack warps lifetime")
mid = (t0 + t[:, 7].max()) / 2
print("alive at mid-kernel:", int(((t[:, 0] <= mid) & (np.maximum(t[:, 7], t[:, 4]) >= mid)).sum()))
