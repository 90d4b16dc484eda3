"""File-to-file timing of the CLI (BASELINE.json configs[4]: end to end including file I/O): the mixed-entropy stream is
written to a tmpfs file, `compress c|d <method>` runs on it as a user would, wall-clock timed, output compared with the
input; the reference binary (oracle/_ref/compress_ref: fl-cpu on the host cores, fl = its own GPU kernels) runs on a
bounded sample of the same file.  usage: python tools/bench_files.py [--bytes N] [--dir /dev/shm] [--out file.json]"""
import argparse, json, os, re, shutil, subprocess, sys, time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
CLI = os.path.join(ROOT, "fl-rl-compression-mpi_b200", "compress")
REF = os.path.join(ROOT, "oracle", "_ref", "compress_ref")


def run(cmd, env=None):
    t0 = time.perf_counter()
    p = subprocess.run(cmd, capture_output=True, text=True, env=env)
    dt = time.perf_counter() - t0
    timers = {m.group(1): float(m.group(2)) for m in re.finditer(r'\[TIMER\] Step: "([^"]+)", Time: ([0-9.]+) ms', p.stdout)}
    err = [ln for ln in (p.stdout + p.stderr).splitlines() if "[ERROR]" in ln]
    return dt, timers, err, p.returncode


def same(a, b):
    return subprocess.run(["cmp", "-s", a, b]).returncode == 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--bytes", type=int, default=0)
    ap.add_argument("--dir", default="/dev/shm")
    ap.add_argument("--out", default="")
    ap.add_argument("--ref-bytes", type=int, default=256 << 20)
    ap.add_argument("--ref-gpu-bytes", type=int, default=1 << 30)
    args = ap.parse_args()
    import torch
    import bench

    ngpu = torch.cuda.device_count()
    free = shutil.disk_usage(args.dir).free
    n = args.bytes or min(16 << 30, free // 5)
    n -= n % (1 << 20)
    d = os.path.join(args.dir, "flrl_files")
    os.makedirs(d, exist_ok=True)
    src, comp, back = (os.path.join(d, x) for x in ("in.bin", "comp.bin", "back.bin"))
    with open(src, "wb") as f:
        for b0 in range(0, n, 1 << 30):
            f.write(bench.gen_stream_device(torch, b0, min(1 << 30, n - b0)).cpu().numpy().tobytes())
    res = {"medium": f"tmpfs ({args.dir})", "bytes": n, "gpus": ngpu, "workload": "configs[4] mixed-entropy stream (bench.gen_stream_device)",
           "host_cpus": os.cpu_count(), "runs": []}
    methods = [("fl", 1, {}), ("rl", 1, {})]
    if ngpu > 1:
        methods += [("fl-nccl", ngpu, {}), ("rl", ngpu, {"FLRL_NGPUS": str(ngpu)})]
    for m, g, extra in methods:
        env = dict(os.environ, **extra)
        for rep in range(2):  # the second run has the page cache, the driver context cache and the tmpfs pages warm
            tc, timc, ec, _ = run([CLI, "c", m, src, comp], env)
            td, timd, ed, _ = run([CLI, "d", m, comp, back], env)
        ok = not ec and not ed and same(src, back)
        res["runs"].append({"method": m, "gpus": g, "ok": ok, "compress_s": round(tc, 3), "decompress_s": round(td, 3),
                            "compress_gbs_file_to_file": round(n / tc / 1e9, 2), "decompress_gbs_file_to_file": round(n / td / 1e9, 2),
                            "compressed_bytes": os.path.getsize(comp) if os.path.exists(comp) else None,
                            "timers_compress_ms": timc, "timers_decompress_ms": timd, "errors": ec + ed})
        print(json.dumps(res["runs"][-1]), flush=True)
    # cross-decode: a file written by our GPU path must decode with the reference's CPU path and vice versa (a bounded sample)
    if os.path.exists(REF):
        m = min(n, args.ref_bytes)
        small, small_c, small_b, ours_c = (os.path.join(d, x) for x in ("small.bin", "small.ref", "small.back", "small.ours"))
        with open(src, "rb") as f, open(small, "wb") as g:
            g.write(f.read(m))
        tc, timc, ec, _ = run([REF, "c", "fl-cpu", small, small_c])
        td, timd, ed, _ = run([REF, "d", "fl-cpu", small_c, small_b])
        run([CLI, "c", "fl", small, ours_c])
        res["reference_fl_cpu"] = {"bytes": m, "compress_s": round(tc, 3), "decompress_s": round(td, 3),
                                   "compress_gbs_file_to_file": round(m / tc / 1e9, 3), "decompress_gbs_file_to_file": round(m / td / 1e9, 3),
                                   "round_trip_ok": same(small, small_b), "our_file_identical_to_reference_file": same(ours_c, small_c),
                                   "timers_compress_ms": timc, "timers_decompress_ms": timd, "cores": 1}
        print(json.dumps(res["reference_fl_cpu"]), flush=True)
        # the reference's own GPU kernels on this box (compiled for sm_100a, unmodified): `c fl` / `d fl`
        m = min(n, args.ref_gpu_bytes)
        with open(src, "rb") as f, open(small, "wb") as g:
            g.write(f.read(m))
        tc, timc, ec, _ = run([REF, "c", "fl", small, small_c])
        td, timd, ed, _ = run([REF, "d", "fl", small_c, small_b])
        run([CLI, "c", "fl", small, ours_c])
        res["reference_fl_gpu"] = {"bytes": m, "compress_s": round(tc, 3), "decompress_s": round(td, 3), "round_trip_ok": same(small, small_b),
                                   "reference_gpu_file_identical_to_ours": same(ours_c, small_c), "timers_compress_ms": timc,
                                   "timers_decompress_ms": timd, "errors": ec + ed}
        print(json.dumps(res["reference_fl_gpu"]), flush=True)
    shutil.rmtree(d, ignore_errors=True)
    if args.out:
        with open(args.out, "w") as f:
            json.dump(res, f, indent=1)


if __name__ == "__main__":
    main()
