#!/usr/bin/env python
"""bench.py — the driver's measurement contract for the FL/RL codec hot path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--workload all|fl|rl4|rl64|rl4096|stack16g|stream64g]

One "step" = one pass of the hot path over one batch of synthetic input resident in HBM: compress the whole
workload, then decompress the result.  The headline (`value`, `ms_per_step`, `config.workload`) is BASELINE.json
configs[1]: FL over 1 GiB of bytes whose per-frame width is uniform in 1..8.  With the default `--workload all` the
three configs[2] workloads (RL, 1 GiB, geometric runs of mean 4 / 64 / 4096) are measured in the same run and reported
under `detail.workloads`; `roofline` is the kernel furthest below the HBM roofline over all four.
`value` = uncompressed bytes through the codec per second = 2 * N * n_gpus / step time (a step compresses N bytes and
decompresses N bytes on every GPU), GB/s, device timed with CUDA events, max over ranks.  `e2e` = the same round trip
through the host layer of the C ABI (flrl_*_gpu_compress / _decompress: HOST buffers in, HOST buffers out, copies
inside the timed region); at N > 1 rank 0 makes ONE call with ngpus = N on the N shards (the product's own multi-GPU
path: shards, NCCL all-gather of the sizes, every GPU delivers at its offset of the merged container in host memory).

N > 1 (torchrun, one rank per GPU), default workloads: every rank owns one shard of the same shape (weak scaling).
Inside the step the ranks all-gather the per-shard compressed sizes (NCCL, one u64 per GPU) and scan them into output
offsets on the device: the merged container = the shards where they are + this offset table, which is how the product
delivers (each GPU writes at its offset of the host buffer / file over its own PCIe link).  Gathering the payloads
into ONE GPU's memory (the reference's gpuNCCLCompress ends like that, fl_gpu.cu:221-238) is bound by that GPU's
NVLink ingress; it is timed as a second step variant and reported as `detail.concat_in_step_ms` /
`detail.value_with_concat_to_rank0`, not hidden.

`--workload stack16g | stream64g` (BASELINE.json configs[3] / configs[4]): STRONG scaling — the total size is fixed
(16 / 64 GiB, `--total-bytes` overrides), cut across the ranks with the reference's shard rule (file_io.cu:46-51), FL
and RL both, same step.  Device generators follow SURVEY.md §8(d) inputs 4 and 5 (tests/test_workloads_gpu.py holds the
numpy statement used for parity at small sizes).

`--impl reference` times the reference's own CPU implementation (oracle/_ref, built from the unmodified sources by
oracle/Makefile; the oracle port for RL, which the reference lacks) on all host cores over a bounded sample.  This
file is one of the few allowed users of oracle/.
"""
from __future__ import annotations

import argparse
import importlib
import json
import math
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

GIB = 1 << 30
WORKLOADS = {
    "fl": "configs[1]: FL 1 GiB synthetic bytes, per-frame bit-width uniform 1..8",
    "rl4": "configs[2]: RL 1 GiB synthetic bytes, geometric run lengths mean 4",
    "rl64": "configs[2]: RL 1 GiB synthetic bytes, geometric run lengths mean 64",
    "rl4096": "configs[2]: RL 1 GiB synthetic bytes, geometric run lengths mean 4096",
}
BIG = {
    "stack16g": ("configs[3]: FL+RL on a 16 GiB synthetic BMP-like image stack (24-bit pixels, gradients + noise)", 16 * GIB),
    "stream64g": ("configs[4]: FL+RL on a 64 GiB sharded synthetic mixed-entropy stream", 64 * GIB),
}
# committed `ncu --set full` captures that hold dram__bytes_read.sum + dram__bytes_write.sum per launch (1 GiB workloads)
NCU_FILES = {"fl": "profiles/r02_fl_ncu_full.csv", "rl4": "profiles/r02_rl4_ncu_full.csv", "rl64": "profiles/r02_rl64_ncu_full.csv",
             "rl4096": "profiles/r02_rl4096_ncu_full.csv"}


def ncu_traffic(workload, kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch (bytes) of `kernel` from the committed `ncu --set full`
    summary of `workload` (profiles/, written by tools/ncu_summary.py from a 1 GiB capture), or None."""
    import csv

    path = os.path.join(ROOT, NCU_FILES.get(workload, ""))
    if not os.path.isfile(path):
        return None
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    try:
        with open(path) as f:
            rows = list(csv.reader(f))
        hdr, units = rows[0], rows[1]
        ir, iw, ik = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum"), hdr.index("Kernel Name")
        it = hdr.index("gpu__time_duration.sum")
        best = None
        for r in rows[2:]:
            if kernel in r[ik]:
                tot = float(r[ir]) * scale[units[ir]] + float(r[iw]) * scale[units[iw]]
                if best is None or float(r[it]) > best[0]:  # the launch that did the work (RL encode: pass 0, not the empty pass 1)
                    best = (float(r[it]), tot)
        return int(best[1]) if best else None
    except (ValueError, KeyError, IndexError):
        return None


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


# ------------------------------------------------------------------------------- clocks
class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.index), "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for ln in self.proc.stdout:
            self.lines.append(ln)

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.t.join(timeout=2)
        sm, mx, reasons = [], None, set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx = float(f[2])
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


# --------------------------------------------------------------------------- generators
def gen_fl_device(torch, n, seed=42):
    """SURVEY.md §8d input 2 on the device: width w_f ~ U{1..8}; bytes U[0,256) & mask; byte 0 of each frame
    carries the top bit so the frame's width is exactly w_f."""
    g = torch.Generator(device="cuda").manual_seed(seed)
    frames = n // 128
    w = torch.randint(1, 9, (frames,), device="cuda", generator=g, dtype=torch.int32)
    out = torch.empty(n, dtype=torch.uint8, device="cuda")
    step = 1 << 21  # frames per slab: bounds the int32 temporaries
    for f0 in range(0, frames, step):
        f1 = min(frames, f0 + step)
        d = torch.randint(0, 256, (f1 - f0, 128), device="cuda", generator=g, dtype=torch.int32)
        d &= ((1 << w[f0:f1]) - 1)[:, None]
        d[:, 0] |= 1 << (w[f0:f1] - 1)
        out[f0 * 128:f1 * 128] = d.to(torch.uint8).reshape(-1)
    return out, w


def gen_rl_device(torch, n, mean, seed=None):
    """SURVEY.md §8d input 3 on the device: run lengths ~ Geom(1/mean), every run's value differs from the
    previous one ((prev + 1 + U{0..254}) mod 256), truncated to n bytes."""
    g = torch.Generator(device="cuda").manual_seed(mean if seed is None else seed)
    runs = int(n / mean * 1.05) + 1024
    while True:
        u = torch.rand(runs, device="cuda", generator=g, dtype=torch.float64).clamp_(1e-300, 1.0)
        import math
        L = (torch.log(u) / math.log1p(-1.0 / mean)).floor_().to(torch.int64) + 1 if mean > 1 else torch.ones(
            runs, dtype=torch.int64, device="cuda")
        ends = torch.cumsum(L, 0)
        if int(ends[-1].item()) >= n:
            break
        runs *= 2
    k = int(torch.searchsorted(ends, torch.tensor([n], device="cuda"), right=False).item()) + 1
    ends = ends[:k].clamp_(max=n)
    inc = torch.randint(1, 256, (k,), device="cuda", generator=g, dtype=torch.int64)
    vals = (torch.cumsum(inc, 0) % 256).to(torch.uint8)
    starts = torch.cat([torch.zeros(1, dtype=torch.int64, device="cuda"), ends[:-1]])
    out = torch.repeat_interleave(vals, ends - starts)
    assert out.numel() == n
    return out


def gen_narrow_device(torch, n, g, wmax=4):
    """Narrow noise: per-frame width w ~ U{1..wmax}, bytes U[0,256) & ((1 << w) - 1)."""
    frames = (n + 127) // 128
    w = torch.randint(1, wmax + 1, (frames,), device="cuda", generator=g, dtype=torch.int32)
    out = torch.empty(frames * 128, dtype=torch.uint8, device="cuda")
    step = 1 << 21
    for f0 in range(0, frames, step):
        f1 = min(frames, f0 + step)
        d = torch.randint(0, 256, (f1 - f0, 128), device="cuda", generator=g, dtype=torch.int32)
        d &= ((1 << w[f0:f1]) - 1)[:, None]
        out[f0 * 128:f1 * 128] = d.to(torch.uint8).reshape(-1)
    return out[:n]


def gen_stream_device(torch, b0, n, seg=1 << 20):
    """Bytes [b0, b0 + n) of SURVEY.md §8d input 5 on the device: 1 MiB segments cycling zeros / constant runs (mean
    4096) / narrow noise (w in 1..4) / geometric runs (mean 64) / full-entropy bytes.  All segments of one kind inside
    the range are drawn in one batch (seeded by kind and first segment), so this is the device analogue of the numpy
    statement in tests/test_workloads_gpu.py, not the same bytes."""
    assert b0 % seg == 0
    s0, nseg = b0 // seg, (n + seg - 1) // seg
    out = torch.empty(nseg * seg, dtype=torch.uint8, device="cuda")
    view = out.view(nseg, seg)
    for kind in range(5):
        first = (kind - s0) % 5
        idx = list(range(first, nseg, 5))
        for j0 in range(0, len(idx), 2048):  # slabs of at most 2 GiB of one kind: bounds the generators' temporaries
            sel = idx[j0:j0 + 2048]
            c = len(sel)
            g = torch.Generator(device="cuda").manual_seed(1000 * (s0 + sel[0]) + kind)
            if kind == 0:
                batch = torch.zeros(c * seg, dtype=torch.uint8, device="cuda")
            elif kind == 1:
                batch = gen_rl_device(torch, c * seg, 4096, seed=7 + s0 + sel[0])
            elif kind == 2:
                batch = gen_narrow_device(torch, c * seg, g)
            elif kind == 3:
                batch = gen_rl_device(torch, c * seg, 64, seed=11 + s0 + sel[0])
            else:
                batch = torch.randint(0, 256, (c * seg,), device="cuda", generator=g, dtype=torch.uint8)
            view[sel[0]:sel[-1] + 1:5] = batch.view(c, seg)
            del batch
    return out[:n]


IMG_W, IMG_H, IMG_HDR = 1280, 853, 138
IMG_PIX = 3275520
IMG_BYTES = IMG_HDR + IMG_PIX  # 3,275,658: the size of example_input/sample_1280x853.bmp


def gen_stack_device(torch, b0, n, batch=32):
    """Bytes [b0, b0 + n) of SURVEY.md §8d input 4 on the device: a stack of 1280x853 24-bit images behind 138-byte
    headers; channel value clamp(A_c x/W + B_c y/H + 8 sin(2 pi (x+y)/64) + noise), noise ~ U{-k..k}, k cycling {0, 1, 3}
    per image (flat bands for RL, dark regions for FL)."""
    i0, i1 = b0 // IMG_BYTES, (b0 + n + IMG_BYTES - 1) // IMG_BYTES
    x = torch.arange(IMG_W, dtype=torch.float32, device="cuda")[None, :, None]
    y = torch.arange(IMG_H, dtype=torch.float32, device="cuda")[:, None, None]
    A = torch.tensor([200.0, 90.0, 30.0], device="cuda")[None, None, :]
    B = torch.tensor([20.0, 120.0, 210.0], device="cuda")[None, None, :]
    base = (A * x / IMG_W + B * y / IMG_H + 8 * torch.sin(2 * math.pi * (x + y) / 64)).reshape(-1)[:IMG_PIX]
    hdr = torch.zeros(IMG_HDR, dtype=torch.uint8, device="cuda")
    hdr[0], hdr[1], hdr[10] = 0x42, 0x4D, IMG_HDR
    out = torch.empty((i1 - i0) * IMG_BYTES, dtype=torch.uint8, device="cuda")
    rows = out.view(i1 - i0, IMG_BYTES)
    rows[:, :IMG_HDR] = hdr
    g = torch.Generator(device="cuda").manual_seed(i0)
    for j0 in range(i0, i1, batch):
        j1 = min(i1, j0 + batch)
        k = torch.tensor([(0, 1, 3)[j % 3] for j in range(j0, j1)], device="cuda", dtype=torch.int16)[:, None]
        r = torch.randint(0, 7, (j1 - j0, IMG_PIX), device="cuda", generator=g, dtype=torch.int16)
        noise = torch.where(k == 3, r - 3, torch.where(k == 1, r % 3 - 1, torch.zeros_like(r)))
        rows[j0 - i0:j1 - i0, IMG_HDR:] = (base[None, :] + noise).clamp_(0, 255).to(torch.uint8)
        del r, noise
    lo = b0 - i0 * IMG_BYTES
    return out[lo:lo + n]


def run_before_device(torch, d, a):
    """Number of bytes equal to d[a] immediately before position a (the prev_run of a shard cut at a)."""
    if a == 0:
        return 0
    v, look = d[a], 1 << 16
    while True:
        lo = max(0, a - look)
        same = (d[lo:a] == v).flip(0)
        r = int(same.to(torch.int8).cummin(0).values.sum(dtype=torch.int64).item())
        if r < a - lo or lo == 0:
            return r
        look *= 16


# ------------------------------------------------------------------------------ our arm
class Work:
    """Device-resident round trip of one shard on the current device through the C ABI's device layer.  The shard is
    encoded and decoded in pieces of at most `piece` bytes (one piece for the 1 GiB workloads): FL pieces are cut at
    multiples of 128 bytes, RL pieces carry prev_run and 255 bytes of look-ahead, so piece outputs concatenate to the
    single-stream output (include/flrl.h shard semantics)."""

    def __init__(self, flrl, torch, kind, d_in, piece=None):
        self.flrl, self.torch, self.kind, self.d_in = flrl, torch, kind, d_in
        n = self.n = d_in.numel()
        piece = n if not piece else min(n, piece - piece % 128)
        self.cuts = [(a, min(n, a + piece)) for a in range(0, n, piece)]
        pmax = max(b - a for a, b in self.cuts)
        self.kernels = ("fl_encode_kernel", "fl_decode_kernel") if kind == "fl" else ("rl_encode_kernel", "rl_decode_kernel")
        self.enc = flrl.FLEncoder(pmax) if kind == "fl" else flrl.RLEncoder(pmax)
        self.dec = flrl.FLDecoder(pmax) if kind == "fl" else flrl.RLDecoder(pmax)
        self.prev = [0 if kind == "fl" else run_before_device(torch, d_in, a) for a, _ in self.cuts]
        self.total_size = torch.zeros(1, dtype=torch.int64, device="cuda")
        # first pass: per-piece compressed sizes (the decoder's launch arguments) and the round-trip check
        self.sizes, self.csize, self.alg_bytes = [], 0, 0
        for i, (a, b) in enumerate(self.cuts):
            self._encode(i)
            if kind == "fl":
                bits, values = self.enc.result(b - a)
                self.sizes.append((bits.numel(), values.numel()))
                self.csize += bits.numel() + values.numel()
                self.alg_bytes += (b - a) + bits.numel() + values.numel()  # N + F + valuesSize, both directions
            else:
                counts, _ = self.enc.result()
                self.sizes.append((counts.numel(),))
                self.csize += 2 * counts.numel()
                self.alg_bytes += (b - a) + 2 * counts.numel()  # N + 2P
            self._decode(i)
            out = self.dec.result(b - a) if kind == "fl" else self.dec.result()
            assert torch.equal(out, d_in[a:b]), f"{kind} round trip failed (piece {i})"

    def _encode(self, i):
        a, b = self.cuts[i]
        if self.kind == "fl":
            self.enc.launch(self.d_in[a:b])
        else:
            self.enc.launch(self.d_in[a:min(self.n, b + 255)], n=b - a, prev_run=self.prev[i])

    def _decode(self, i):
        a, b = self.cuts[i]
        if self.kind == "fl":
            nb, nv = self.sizes[i]
            self.dec.launch(self.enc.bits[:nb], self.enc.values[:nv], b - a)
        else:
            (p,) = self.sizes[i]
            self.dec.launch(self.enc.counts[:p], self.enc.values[:p])

    def step(self, exchange=None, events=None):
        """compress the shard, [exchange the sizes], decompress it; pieces back to back."""
        single = len(self.cuts) == 1  # one launch per shard: the kernel's own size word is the shard's size (no extra launches)
        if not single:
            self.total_size.zero_()
        for i in range(len(self.cuts)):
            if events:
                events[i][0].record()
            self._encode(i)
            if events:
                events[i][1].record()
            if not single:
                self.total_size += self.enc.size
            if exchange and i == len(self.cuts) - 1:
                exchange()
            if events:
                events[i][2].record()
            self._decode(i)
            if events:
                events[i][3].record()

    def size_tensor(self):
        return self.enc.size if len(self.cuts) == 1 else self.total_size

    def payload(self):
        return self.enc.values if self.kind == "fl" else self.enc.counts


def make_input(torch, workload, n, rank, world, total):
    """The rank's shard of the workload on the device + (begin, length) of the shard in the whole workload."""
    sharding = importlib.import_module("fl-rl-compression-mpi_b200.sharding")
    if workload == "fl":
        return gen_fl_device(torch, n, seed=42 + rank)[0], (rank * n, n)
    if workload.startswith("rl"):
        mean = int(workload[2:])
        return gen_rl_device(torch, n, mean, seed=mean + 7919 * rank), (rank * n, n)
    b0, m = sharding.shard_range(total, world, rank)  # the reference's rule, file_io.cu:46-51
    gen = gen_stack_device if workload == "stack16g" else gen_stream_device
    d = gen(torch, b0, m)
    if d.data_ptr() % 16:  # the device layer wants 16-byte aligned buffers (a shard of the stack starts inside an image)
        d = d.clone()
    return d, (b0, m)


def pcie_probe(torch, devices):
    """Pinned host <-> device copy bandwidth of this box, both directions at once on every device of `devices`
    (GB/s per GPU and direction): the ceiling of any host-buffer round trip (tools/pcie_ceiling.cu is the full table)."""
    n, reps = 128 << 20, 4
    st = []
    for dv in devices:
        with torch.cuda.device(dv):
            st.append((torch.empty(n, dtype=torch.uint8).pin_memory(), torch.empty(n, dtype=torch.uint8).pin_memory(),
                       torch.empty(n, dtype=torch.uint8, device=f"cuda:{dv}"), torch.empty(n, dtype=torch.uint8, device=f"cuda:{dv}"),
                       torch.cuda.Stream(device=dv), torch.cuda.Stream(device=dv)))

    def go():
        for hi, ho, di, do, s0, s1 in st:
            for _ in range(reps):
                with torch.cuda.stream(s0):
                    di.copy_(hi, non_blocking=True)
                with torch.cuda.stream(s1):
                    ho.copy_(do, non_blocking=True)
        for dv in devices:
            torch.cuda.synchronize(dv)

    go()
    t0 = time.perf_counter()
    go()
    return n * reps / (time.perf_counter() - t0) / 1e9


def e2e_round_trip(flrl, torch, kind, host_in, ngpus, steps, warmup):
    """The same metric through the host layer of the C ABI: pinned HOST buffers in, HOST buffers out, H2D / D2H inside
    the timed region.  host_in: a HostCodec whose input is loaded.  Returns (seconds per step, h2d bytes, d2h bytes, per-GPU figures)."""
    hc = host_in
    for _ in range(max(1, min(warmup, 2))):
        hc.compress()
        hc.decompress()
    assert hc.check_round_trip(), "host-layer round trip failed"
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        hc.compress()
        hc.decompress()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / steps
    comp = hc.compressed_bytes()
    return dt, hc.n + comp, comp + hc.n, hc.timing_per_gpu()


def measure(args, torch, dist, flrl, sharding, workload, kind, rank, local, world, headline):
    """Times `args.steps` steps of one workload / codec on every rank; returns rank-0's dict (None elsewhere)."""
    big = workload in BIG
    total = args.total_bytes or (BIG[workload][1] if big else args.bytes * world)
    d_in, (b0, n) = make_input(torch, workload, args.bytes, rank, world, total)
    work = Work(flrl, torch, kind, d_in, piece=args.piece_bytes if big else None)
    sizes = torch.zeros(world, dtype=torch.int64, device="cuda")
    offsets = torch.zeros(world, dtype=torch.int64, device="cuda")

    def exchange():
        """N > 1, inside every timed step: the ONLY exchange of the path — all-gather the per-shard compressed sizes
        (one u64 per GPU, NCCL) and turn them into output offsets by an exclusive scan, on the device, no host sync."""
        sharding.exchange_sizes(dist, work.size_tensor(), sizes, offsets)  # the logic tests/test_sharding_gloo.py runs over gloo

    ex = exchange if world > 1 else None
    ev = lambda: torch.cuda.Event(enable_timing=True)
    for _ in range(args.warmup):
        work.step(ex)
    torch.cuda.synchronize()
    ke = [[(ev(), ev(), ev(), ev()) for _ in work.cuts] for _ in range(args.steps)]
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    sampler = ClockSampler(local)
    if rank == 0 and headline:
        # nvidia-smi delivers one sample per 100 ms and the timed region is a few ms: keep the same load running
        # (untimed, no collective) until the sampler has produced samples, then enter the timed region without a gap.
        sampler.start()
        t_wait = time.perf_counter()
        while len(sampler.lines) < 4 and time.perf_counter() - t_wait < 3.0:
            work.step(None)
            torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t_begin, t_end = ev(), ev()
    t_begin.record()
    for e in ke:
        work.step(ex, e)
    t_end.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    clocks = sampler.stop() if rank == 0 and headline else None
    total_ms = t_begin.elapsed_time(t_end)
    enc_ms = sum(a.elapsed_time(b) for e in ke for a, b, _, _ in e) / args.steps
    dec_ms = sum(c.elapsed_time(d) for e in ke for _, _, c, d in e) / args.steps
    stats = torch.tensor([total_ms, enc_ms, dec_ms, float(work.alg_bytes), float(work.csize)], dtype=torch.float64, device="cuda")
    if world > 1:
        mx = stats.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        dist.all_reduce(stats, op=dist.ReduceOp.SUM)
        total_ms, enc_ms, dec_ms = mx[:3].tolist()
        alg_all, csize_all = stats[3].item(), stats[4].item()
    else:
        alg_all, csize_all = float(work.alg_bytes), float(work.csize)
    ms_per_step = total_ms / args.steps

    # the same step with the payloads gathered into rank 0's memory (the reference's merged-on-rank-0 result)
    concat_ms = concat_step_ms = None
    if world > 1 and len(work.cuts) == 1 and not args.no_concat:
        gather_buf = torch.empty(int(sum(sizes.tolist())) + 16, dtype=torch.uint8, device="cuda") if rank == 0 else None
        pay = 1 if kind == "fl" else 2  # FL: values (the widths' places are known a priori); RL: counts and values

        def concat():
            for _ in range(pay):
                sharding.concat_to_rank0(dist, work.payload(), sizes, rank, gather_buf)

        work.step(ex)
        concat()  # warm-up (NCCL p2p channels)
        torch.cuda.synchronize()
        dist.barrier()
        c0, c1, c2 = ev(), ev(), ev()
        c0.record()
        for _ in range(3):
            concat()
        c1.record()
        for _ in range(3):
            work.step(ex)
            concat()
        c2.record()
        torch.cuda.synchronize()
        t = torch.tensor([c0.elapsed_time(c1) / 3, c1.elapsed_time(c2) / 3], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        concat_ms, concat_step_ms = round(t[0].item(), 4), round(t[1].item(), 4)
        del gather_buf

    # end to end through the host layer
    e2e = None
    if not args.no_e2e and (headline or world == 1):
        e2e_n = n if not big else min(n, args.e2e_bytes // world)
        steps = max(1, min(args.steps, 3 if big else 5))
        hc = flrl.HostCodec(kind, e2e_n, ngpus=1)
        hc.load_input(d_in[:e2e_n])
        dt, h2d, d2h, _ = e2e_round_trip(flrl, torch, kind, hc, 1, steps, args.warmup)
        del hc
        per_rank = None
        if world > 1:
            t = torch.tensor([dt], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            per_rank = round(2 * e2e_n * world / t.item() / 1e9, 3)
            # ONE call of the product's multi-GPU API from rank 0 over all shards while the other ranks wait on the host
            # (gloo barrier: their GPUs stay idle).  The shards travel to rank 0's pinned input buffer first, untimed:
            # this is "the file in host memory".
            gathered = [torch.empty(e2e_n, dtype=torch.uint8, device="cuda") for _ in range(world)] if rank == 0 else None
            dist.gather(d_in[:e2e_n].contiguous(), gathered, dst=0)
            torch.cuda.synchronize()
            if rank == 0:
                hcN = flrl.HostCodec(kind, e2e_n * world, ngpus=world)
                for r in range(world):
                    hcN.src.array[r * e2e_n:(r + 1) * e2e_n] = gathered[r].cpu().numpy()
                del gathered
                dt, h2d, d2h, per_gpu = e2e_round_trip(flrl, torch, kind, hcN, world, steps, args.warmup)
                del hcN
            dist.barrier(group=args.host_group)
        if rank == 0:
            api = "flrl_%s_gpu_compress_into + flrl_%s_gpu_decompress_into (include/flrl.h), pinned host buffers, ngpus=%d" % (kind, kind, world)
            e2e = {"value": round(2 * e2e_n * world / dt / 1e9, 3), "unit": "GB/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                   "api": api, "bytes": e2e_n * world}
            if per_rank is not None:
                e2e["per_rank_calls_value"] = per_rank  # N processes, each calling the 1-GPU host API on its shard
            if headline:
                bw = pcie_probe(torch, list(range(world)))
                e2e["pcie_duplex_gbs_per_gpu_per_direction"] = round(bw, 1)
                # a round trip moves max(in, out) = the uncompressed bytes over each GPU's link twice: ceiling = that bandwidth x GPUs
                e2e["frac_of_pcie"] = round(e2e["value"] / (bw * world), 3)
            if big and e2e_n < n:
                e2e["note"] = f"first {e2e_n} bytes of every shard (pinned host memory bounds the e2e sample)"
    if rank != 0:
        return None
    peak, _ = peaks()
    nbytes_all = total if big else args.bytes * world
    return {
        "workload": workload, "codec": kind, "bytes": int(nbytes_all), "compressed_bytes": int(csize_all),
        "value": round(2 * nbytes_all / (ms_per_step * 1e-3) / 1e9, 3), "ms_per_step": round(ms_per_step, 4),
        "compress_ms": round(enc_ms, 4), "decompress_ms": round(dec_ms, 4),
        "compress_gbs": round(nbytes_all / world / enc_ms / 1e6, 2), "decompress_gbs": round(nbytes_all / world / dec_ms / 1e6, 2),
        "compress_frac_of_hbm": round(alg_all / world / enc_ms / 1e6 / peak, 4),
        "decompress_frac_of_hbm": round(alg_all / world / dec_ms / 1e6 / peak, 4),
        "algorithmic_bytes_per_gpu": int(alg_all / world), "pieces_per_gpu": len(work.cuts), "kernels": work.kernels,
        "concat_to_rank0_ms": concat_ms, "concat_in_step_ms": concat_step_ms,
        "value_with_concat_to_rank0": round(2 * nbytes_all / (concat_step_ms * 1e-3) / 1e9, 3) if concat_step_ms else None,
        "e2e": e2e, "clocks": clocks, "gpu_launches": 2 * args.steps * len(work.cuts),
    }


def run_ours(args):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    torch.cuda.set_device(local)
    args.host_group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        args.host_group = dist.new_group(backend="gloo")  # host-side barrier: waiting ranks leave their GPUs idle
    flrl = importlib.import_module("fl-rl-compression-mpi_b200")  # raises if libflrl.so is missing: no fallback
    sharding = importlib.import_module("fl-rl-compression-mpi_b200.sharding")
    big = args.workload in BIG
    if big:
        plan = [(args.workload, "fl"), (args.workload, "rl")]
    elif args.workload == "all":
        plan = [("fl", "fl"), ("rl4", "rl"), ("rl64", "rl"), ("rl4096", "rl")]
    else:
        plan = [(args.workload, args.workload[:2])]
    results = []
    for i, (w, kind) in enumerate(plan):
        r = measure(args, torch, dist, flrl, sharding, w, kind, rank, local, world, headline=(i == 0))
        results.append(r)
        torch.cuda.empty_cache()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    peak, peak_src = peaks()
    head = results[0]
    # the kernel furthest below the HBM roofline over everything measured in this run
    worst = None
    for r in results:
        for k, frac, ms in ((r["kernels"][0], r["compress_frac_of_hbm"], r["compress_ms"]), (r["kernels"][1], r["decompress_frac_of_hbm"], r["decompress_ms"])):
            if worst is None or frac < worst[2]:
                worst = (r, k, frac, ms)
    wr, wk, wfrac, wms = worst
    achieved = wr["algorithmic_bytes_per_gpu"] / (wms * 1e-3) / 1e9
    label = BIG[args.workload][0] if big else WORKLOADS[head["workload"]]
    par = "1 GPU"
    if world > 1:
        par = ("contiguous shards of ONE workload (reference rule, file_io.cu:46-51), " if big else "one shard of the same shape per GPU, ") + \
              "NCCL all-gather of the shard sizes + device-side offset scan in every step; merged container = shards in place + offsets " \
              "(gathering into one GPU: detail.concat_in_step_ms)"
    line = {
        "metric": "FL/RL compress & decompress GB/s (uncompressed)",
        "value": head["value"], "unit": "GB/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": head["ms_per_step"],
        "higher_is_better": True, "scaling": "strong" if big else "weak", "vs_baseline": None,
        "dtype": "u8", "data": "synthetic",
        "config": {"workload": label, "codec": head["codec"], "bytes_total": head["bytes"], "compressed_bytes_total": head["compressed_bytes"],
                   "step": "compress the workload then decompress it (device-resident)",
                   "l2": "inputs larger than L2 (>= 1 GiB per pass vs 126 MB)", "parallelism": par,
                   "concat_inside_step": False},
        "detail": {"compress_gbs": head["compress_gbs"], "decompress_gbs": head["decompress_gbs"], "compress_ms": head["compress_ms"],
                   "decompress_ms": head["decompress_ms"], "compress_frac_of_hbm": head["compress_frac_of_hbm"],
                   "decompress_frac_of_hbm": head["decompress_frac_of_hbm"], "concat_to_rank0_ms": head["concat_to_rank0_ms"],
                   "concat_in_step_ms": head["concat_in_step_ms"], "value_with_concat_to_rank0": head["value_with_concat_to_rank0"],
                   "workloads": {(r["workload"] if not big else r["workload"] + "/" + r["codec"]): {k: v for k, v in r.items() if k not in ("clocks",)} for r in results}},
        "roofline": {"bound": "hbm", "kernel": wk, "workload": wr["workload"], "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                     "frac": round(achieved / peak, 4), "traffic": ncu_traffic(wr["workload"], wk) if wr["bytes"] == GIB * world else None,
                     "traffic_source": NCU_FILES.get(wr["workload"]),
                     "algorithmic_bytes": wr["algorithmic_bytes_per_gpu"], "launch_ms": wms, "peak_source": peak_src},
        "e2e": head["e2e"],
        "gpu_launches": head["gpu_launches"],
        "clocks": head["clocks"],
    }
    if not args.no_cpu_baseline and world == 1 and not big:
        line["cpu_baseline"] = cpu_baseline("fl" if head["workload"] == "fl" else head["workload"], threads=1, budget_s=12.0)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


# ------------------------------------------------------- CPU baseline / reference arm
class _Quiet:
    """The reference prints [TIMER] lines with printf: keep fd 1 clean for the JSON line."""

    def __enter__(self):
        sys.stdout.flush()
        self.saved = os.dup(1)
        dn = os.open(os.devnull, os.O_WRONLY)
        os.dup2(dn, 1)
        os.close(dn)

    def __exit__(self, *a):
        os.dup2(self.saved, 1)
        os.close(self.saved)


def gen_fl_host(n, seed=42):
    import numpy as np

    rng = np.random.default_rng(seed)
    frames = (n + 127) // 128
    w = rng.integers(1, 9, size=frames).astype(np.uint16)
    d = rng.integers(0, 256, size=frames * 128, dtype=np.uint16) & ((1 << np.repeat(w, 128)) - 1)
    d[::128] |= 1 << (w - 1)
    return d[:n].astype(np.uint8)


def gen_rl_host(n, mean, seed=None):
    import numpy as np

    rng = np.random.default_rng(mean if seed is None else seed)
    runs = int(n / mean * 1.1) + 1024
    L = rng.geometric(1.0 / mean, size=runs)
    while L.sum() < n:
        L = np.concatenate([L, rng.geometric(1.0 / mean, size=runs)])
    vals = (np.cumsum(rng.integers(1, 256, size=L.size)) % 256).astype(np.uint8)
    return np.repeat(vals, L)[:n].copy()


def cpu_codec(workload):
    """(compress(data) -> state, decompress(state) -> bytes, kind).  FL: the reference's own cpuCompress /
    cpuDecompress compiled in place (oracle/_ref) when present, else the oracle port.  RL: the oracle port
    (the reference ships no RL code)."""
    from oracle import oracle as o

    if workload == "fl":
        if o.have_ref():
            return (lambda d: (d.size,) + o.ref_fl_compress(d)), (lambda s: o.ref_fl_decompress(s[1], s[2], s[0])), "reference"
        return (lambda d: (d.size,) + o.fl_compress(d)), (lambda s: o.fl_decompress(s[1], s[2], s[0])), "port"
    return (lambda d: o.rl_compress(d)), (lambda s: o.rl_decompress(s[0], s[1])), "port"


def host_sample(workload, n, seed):
    return gen_fl_host(n, seed) if workload == "fl" else gen_rl_host(n, int(workload[2:]), seed)


def cpu_round_trip(workload, chunks, threads):
    """One compress + decompress of every chunk, `threads` chunks at a time (ctypes releases the GIL).
    Chunks are 128-byte aligned, so FL outputs concatenate to the single-stream output."""
    comp, decomp, kind = cpu_codec(workload)
    ok = [True] * len(chunks)

    def job(i):
        back = decomp(comp(chunks[i]))
        ok[i] = back.size == chunks[i].size and bool((back == chunks[i]).all())

    t0 = time.perf_counter()
    with _Quiet():
        if threads == 1:
            for i in range(len(chunks)):
                job(i)
        else:
            ts = [threading.Thread(target=job, args=(i,)) for i in range(len(chunks))]
            for t in ts:
                t.start()
            for t in ts:
                t.join()
    dt = time.perf_counter() - t0
    assert all(ok), "CPU baseline round trip failed"
    return dt, kind


def cpu_baseline(workload, threads=1, budget_s=12.0):
    """Bounded sample of the same workload on the host cores (rank 0, N = 1 only)."""
    probe = host_sample(workload, 4 << 20, 1)
    dt, kind = cpu_round_trip(workload, [probe], 1)
    per_byte = dt / probe.size
    n = int(min(512 << 20, max(8 << 20, budget_s / per_byte)))
    n -= n % 128
    data = host_sample(workload, n, 2)
    dt, kind = cpu_round_trip(workload, [data], 1)
    return {"value": round(2 * n / dt / 1e9, 4), "unit": "GB/s", "cores": threads, "kind": kind,
            "sample": f"{n} bytes of the same workload generator, 1 compress + 1 decompress, {dt:.2f} s, "
                      f"host has {os.cpu_count()} logical CPUs"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = max(1, len(os.sched_getaffinity(0)))
    per = args.ref_chunk_bytes - args.ref_chunk_bytes % 128
    chunks = [host_sample(args.workload, per, 100 + i) for i in range(threads)]
    kind = None
    for _ in range(args.warmup):
        _, kind = cpu_round_trip(args.workload, chunks, threads)
    t = 0.0
    for _ in range(args.steps):
        dt, kind = cpu_round_trip(args.workload, chunks, threads)
        t += dt
    n = per * threads
    v = round(2 * n * args.steps / t / 1e9, 4)
    line = {
        "impl": "reference",
        "metric": "FL/RL compress & decompress GB/s (uncompressed)", "value": v, "unit": "GB/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(t / args.steps * 1e3, 3),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": WORKLOADS[args.workload], "step": "compress a sample then decompress it, on the host cores",
                   "sample_bytes_per_step": n},
        "cpu_baseline": {"value": v, "unit": "GB/s", "cores": threads, "kind": kind,
                         "sample": f"{threads} chunks x {per} bytes per step (one 128-aligned chunk per thread; the reference "
                                   f"codec itself is single-threaded), same workload generator"},
        "e2e": {"value": v, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--workload", choices=["all"] + list(WORKLOADS) + list(BIG), default="all")
    ap.add_argument("--bytes", type=int, default=GIB, help="uncompressed bytes per GPU of the 1 GiB workloads")
    ap.add_argument("--total-bytes", type=int, default=0, help="stack16g / stream64g: total size over all GPUs (default 16 / 64 GiB)")
    ap.add_argument("--piece-bytes", type=int, default=4 * GIB, help="stack16g / stream64g: bytes per kernel launch")
    ap.add_argument("--e2e-bytes", type=int, default=4 * GIB, help="stack16g / stream64g: bytes of the end-to-end sample (pinned host memory)")
    ap.add_argument("--ref-chunk-bytes", type=int, default=16 << 20)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-concat", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "ours" and args.warmup < 3:
        args.warmup = 3  # timing rule: W >= 3
    if args.impl == "reference":
        if args.workload in BIG or args.workload == "all":
            args.workload = "fl"  # the reference arm times the headline workload
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
