import importlib, os, sys, torch
sys.path.insert(0, "/root/repo")
import bench
flrl = importlib.import_module("fl-rl-compression-mpi_b200")
n = 1 << 30
d = bench.gen_rl_device(torch, n, 64)
enc = flrl.RLEncoder(n)
def t(label):
    for _ in range(3): enc.launch(d)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(10): enc.launch(d)
    b.record(); torch.cuda.synchronize()
    print(label, round(a.elapsed_time(b) / 10, 4), "ms", flush=True)
t("dbg=" + os.environ.get("FLRL_RLE_DBG", "0") + " ctas=" + os.environ.get("FLRL_RLE_CTAS", "3"))
