#!/usr/bin/env python
"""bench.py — the driver's measurement contract for the FL/RL codec hot path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload fl|rl4|rl64|rl4096]

One "step" = one pass of the hot path over one batch of synthetic input resident in HBM:
compress the whole workload, then decompress the result (BASELINE.json configs[1]: FL over
1 GiB of bytes whose per-frame width is uniform in 1..8; configs[2] for the rl* workloads).
`value` = uncompressed bytes through the codec per second = 2 * N * n_gpus / step time
(a step compresses N bytes and decompresses N bytes on every GPU), GB/s, device timed with
CUDA events, max over ranks.  `e2e` = the same round trip through the host layer of the C
ABI (flrl_*_gpu_compress / _decompress: HOST buffers in, HOST buffers out, copies inside
the timed region).  `roofline` is for the slower of the two kernels of the step.

N > 1 (torchrun, one rank per GPU): every rank owns one shard of the same shape (weak
scaling); the only exchange on the path is the all-gather of the per-shard compressed sizes
(NCCL, one u64 per GPU) from which every rank computes its output offset on the device.  The
device-resident concatenation to rank 0 (NCCL send/recv over NVLink) is timed separately
(detail.concat_to_rank0_ms): end to end each GPU delivers its payload to the host at its offset.

`--impl reference` times the reference's own CPU implementation (oracle/_ref, built from the
unmodified sources by oracle/Makefile; the oracle port for RL, which the reference lacks) on
all host cores over a bounded sample.  This file is one of the few allowed users of oracle/.
"""
from __future__ import annotations

import argparse
import importlib
import json
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


# ------------------------------------------------------------------------------ our arm
class FLWork:
    """Device-resident FL round trip on the current device (through the C ABI's device layer)."""

    kernels = ("fl_encode_kernel", "fl_decode_kernel")

    def __init__(self, flrl, torch, n, rank):
        self.flrl, self.torch, self.n = flrl, torch, n
        self.d_in, w = gen_fl_device(torch, n, seed=42 + rank)
        self.enc, self.dec = flrl.FLEncoder(n), flrl.FLDecoder(n)
        self.enc.launch(self.d_in)
        self.bits, self.values = self.enc.result(n)
        assert torch.equal(self.bits.to(torch.int32), w), "FL widths differ from the generator's"
        self.csize = self.values.numel()
        self.alg_bytes = n + self.bits.numel() + self.csize  # N + F + valuesSize, both directions
        self.dec.launch(self.bits, self.values, n)
        assert torch.equal(self.dec.result(n), self.d_in), "FL round trip failed"

    def encode(self):
        self.enc.launch(self.d_in)

    def decode(self):
        self.dec.launch(self.bits, self.values, self.n)

    def size_tensor(self):
        return self.enc.size

    def payload(self):
        return self.enc.values


class RLWork:
    kernels = ("rl_encode_kernel", "rl_decode_kernel")

    def __init__(self, flrl, torch, n, rank, mean):
        self.flrl, self.torch, self.n = flrl, torch, n
        self.d_in = gen_rl_device(torch, n, mean, seed=mean + 7919 * rank)
        self.enc, self.dec = flrl.RLEncoder(n), flrl.RLDecoder(n)
        self.enc.launch(self.d_in)
        self.counts, self.vals = self.enc.result()
        self.pairs = self.counts.numel()
        self.csize = 2 * self.pairs
        self.alg_bytes = n + 2 * self.pairs  # N + 2P
        self.dec.launch(self.counts, self.vals)
        assert torch.equal(self.dec.result(), self.d_in), "RL round trip failed"

    def encode(self):
        self.enc.launch(self.d_in)

    def decode(self):
        self.dec.launch(self.counts, self.vals)

    def size_tensor(self):
        return self.enc.size

    def payload(self):
        return self.enc.counts


def make_work(flrl, torch, workload, n, rank):
    if workload == "fl":
        return FLWork(flrl, torch, n, rank)
    return RLWork(flrl, torch, n, rank, int(workload[2:]))


def e2e_round_trip(flrl, torch, workload, work, steps, warmup):
    """The same metric through the host layer of the C ABI: pinned HOST buffers in, HOST buffers out,
    H2D / D2H inside the timed region.  Returns (seconds per step, h2d bytes, d2h bytes) or None."""
    if not hasattr(flrl, "HostCodec"):
        return None
    hc = flrl.HostCodec(workload[:2], work.n)
    hc.load_input(work.d_in)  # device -> pinned host, untimed: this is the "file in host memory"
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
    return dt, work.n + comp, comp + work.n


def run_ours(args):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    flrl = importlib.import_module("fl-rl-compression-mpi_b200")  # raises if libflrl.so is missing: no fallback
    sharding = importlib.import_module("fl-rl-compression-mpi_b200.sharding")
    n = args.bytes
    work = make_work(flrl, torch, args.workload, n, rank)

    sizes = torch.zeros(world, dtype=torch.int64, device="cuda")
    gather_buf = None
    if world > 1 and rank == 0:
        gather_buf = torch.empty(world * ((n + 15) // 16 * 16), dtype=torch.uint8, device="cuda")

    offsets = torch.zeros(world, dtype=torch.int64, device="cuda")

    def exchange():
        """N > 1, inside every timed step: the ONLY exchange of the path — all-gather the per-shard compressed sizes
        (one u64 per GPU, NCCL) and turn them into output offsets by an exclusive scan, on the device, no host sync."""
        sharding.exchange_sizes(dist, work.size_tensor(), sizes, offsets)  # the logic tests/test_sharding_gloo.py runs over gloo

    def concat_to_rank0():
        """The device-resident merge of the reference's gpuNCCLCompress (fl_gpu.cu:221-238): every rank sends its payload
        to its offset in rank 0's buffer (NCCL send/recv over NVLink).  It serialises on rank 0's ingress link, so it
        is measured on its own (detail.concat_to_rank0_ms) and is not part of the timed step: end to end, every GPU
        delivers its payload straight to its offset in the host buffer instead (host layer, e2e)."""
        sharding.concat_to_rank0(dist, work.payload(), sizes, rank, gather_buf)

    def step():
        work.encode()
        if world > 1:
            exchange()
        work.decode()

    def step_local():  # no collective: rank 0 alone runs it while waiting for clock samples
        work.encode()
        work.decode()

    ev = lambda: torch.cuda.Event(enable_timing=True)
    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    # per-kernel launch durations, measured live on the launching stream (torch's current stream)
    ke = [(ev(), ev(), ev(), ev()) for _ in range(args.steps)]
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    sampler = ClockSampler(local)
    if rank == 0:
        # nvidia-smi delivers one sample per 100 ms and the timed region is a few ms: keep the same load running
        # (untimed) until the sampler has produced samples, then enter the timed region without a gap.
        sampler.start()
        t_wait = time.perf_counter()
        while len(sampler.lines) < 4 and time.perf_counter() - t_wait < 3.0:
            step_local()
            torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t_begin, t_end = ev(), ev()
    t_begin.record()
    for a, b, c, d in ke:
        a.record()
        work.encode()
        b.record()
        if world > 1:
            exchange()
        c.record()
        work.decode()
        d.record()
    t_end.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    clocks = sampler.stop() if rank == 0 else None
    total_ms = t_begin.elapsed_time(t_end)
    enc_ms = sum(a.elapsed_time(b) for a, b, _, _ in ke) / args.steps
    dec_ms = sum(c.elapsed_time(d) for _, _, c, d in ke) / args.steps
    if world > 1:
        t = torch.tensor([total_ms, enc_ms, dec_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms, enc_ms, dec_ms = t.tolist()
    ms_per_step = total_ms / args.steps
    concat_ms = None
    if world > 1:
        exchange()
        concat_to_rank0()  # warm-up (NCCL p2p channels)
        torch.cuda.synchronize()
        dist.barrier()
        c0, c1 = ev(), ev()
        c0.record()
        for _ in range(3):
            concat_to_rank0()
        c1.record()
        torch.cuda.synchronize()
        t = torch.tensor([c0.elapsed_time(c1) / 3], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        concat_ms = round(t.item(), 4)

    e2e = None
    if not args.no_e2e:
        r = e2e_round_trip(flrl, torch, args.workload, work, max(1, min(args.steps, 5)), args.warmup)
        if r is not None:
            dt, h2d, d2h = r
            if world > 1:
                t = torch.tensor([dt], dtype=torch.float64, device="cuda")
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                dt = t.item()
            e2e = {"value": round(2 * n * world / dt / 1e9, 3), "unit": "GB/s", "h2d_bytes_per_step": h2d,
                   "d2h_bytes_per_step": d2h, "api": "flrl_%s_gpu_compress + flrl_%s_gpu_decompress (include/flrl.h), pinned host buffers"
                   % (args.workload[:2], args.workload[:2])}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    peak, peak_src = peaks()
    slow = 0 if enc_ms >= dec_ms else 1
    k_ms = (enc_ms, dec_ms)[slow]
    achieved = work.alg_bytes / (k_ms * 1e-3) / 1e9
    line = {
        "metric": "FL/RL compress & decompress GB/s (uncompressed)",
        "value": round(2 * n * world / (ms_per_step * 1e-3) / 1e9, 3),
        "unit": "GB/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms_per_step, 4),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8", "data": "synthetic",
        "config": {"workload": WORKLOADS[args.workload], "bytes_per_gpu": n, "compressed_bytes_per_gpu": int(work.csize),
                   "step": "compress N bytes then decompress them, per GPU",
                   "l2": "inputs larger than L2 (1 GiB per pass vs 126 MB)",
                   "parallelism": "contiguous shards, one per GPU; NCCL all-gather of the shard sizes + device-side offset scan in every step" if world > 1 else "1 GPU"},
        "detail": {"compress_gbs": round(n / enc_ms / 1e6, 2), "decompress_gbs": round(n / dec_ms / 1e6, 2),
                   "compress_ms": round(enc_ms, 4), "decompress_ms": round(dec_ms, 4), "concat_to_rank0_ms": concat_ms,
                   "compress_frac_of_hbm": round(work.alg_bytes / enc_ms / 1e6 / peak, 4),
                   "decompress_frac_of_hbm": round(work.alg_bytes / dec_ms / 1e6 / peak, 4)},
        "roofline": {"bound": "hbm", "kernel": work.kernels[slow], "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                     "frac": round(achieved / peak, 4), "traffic": TRAFFIC.get((args.workload, work.kernels[slow])) if n == GIB else None,
                     "algorithmic_bytes": int(work.alg_bytes), "launch_ms": round(k_ms, 4), "peak_source": peak_src},
        "e2e": e2e,
        "gpu_launches": 2 * args.steps,
        "clocks": clocks,
    }
    if not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(args.workload, threads=1, budget_s=12.0)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


# dram__bytes_read.sum + dram__bytes_write.sum per launch (bytes) from the committed `ncu --set full` captures
# (profiles/r01_run8_fl_ncu_full.csv for FL, r01_run7_*_ncu_full.csv for RL; 1 GiB per launch, the bench's default size)
TRAFFIC = {
    ("fl", "fl_encode_kernel"): 1_643_363_000, ("fl", "fl_decode_kernel"): 1_640_467_000,
    ("rl4", "rl_encode_kernel"): 1_578_695_000, ("rl4", "rl_decode_kernel"): 1_558_388_000,
    ("rl64", "rl_encode_kernel"): 1_110_706_000, ("rl64", "rl_decode_kernel"): 1_054_236_000,
    ("rl4096", "rl_encode_kernel"): 1_087_804_000, ("rl4096", "rl_decode_kernel"): 1_028_774_000,
}


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
    ap.add_argument("--workload", choices=list(WORKLOADS), default="fl")
    ap.add_argument("--bytes", type=int, default=GIB, help="uncompressed bytes per GPU (default: the 1 GiB of the config)")
    ap.add_argument("--ref-chunk-bytes", type=int, default=16 << 20)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "ours" and args.warmup < 3:
        args.warmup = 3  # timing rule: W >= 3
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
