"""The N > 1 host logic on CPU: world_size 2 over gloo (no GPU).  Each rank compresses its shard with the library's
CPU codec, the ranks exchange ONLY the compressed sizes, rank 0 concatenates — the result must be the single-stream
output byte for byte."""
import importlib
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def _worker(rank, world, port, kind, n, q):
    import torch
    import torch.distributed as dist

    flrl = importlib.import_module("fl-rl-compression-mpi_b200")
    sharding = importlib.import_module("fl-rl-compression-mpi_b200.sharding")
    from conftest import make_fl_mix

    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        if kind == "fl":
            data = make_fl_mix(n, seed=11)
        else:
            rng = np.random.default_rng(12)
            lens = rng.geometric(1 / 300.0, size=n // 100)
            data = np.repeat(rng.integers(0, 256, size=lens.size).astype(np.uint8), lens)[:n].copy()
        b0, ln = sharding.shard_range(n, world, rank, 128)
        if kind == "rl":
            # a plain per-shard encode concatenates exactly when every cut is a run start: move the cut forward to one
            cuts = []
            for r in range(world + 1):
                c = sharding.shard_range(n, world, r, 128)[0] if r < world else n
                while 0 < c < n and data[c] == data[c - 1]:
                    c += 1
                cuts.append(c)
            b0, ln = cuts[rank], cuts[rank + 1] - cuts[rank]
        shard = data[b0:b0 + ln]
        meta, payload = (flrl.fl_cpu_compress if kind == "fl" else flrl.rl_cpu_compress)(shard)
        my = torch.tensor([payload.size], dtype=torch.int64)
        sizes, offsets = torch.zeros(world, dtype=torch.int64), torch.zeros(world, dtype=torch.int64)
        sharding.exchange_sizes(dist, my, sizes, offsets)
        assert int(offsets[rank]) == int(sizes[:rank].sum())
        pay_t = torch.from_numpy(np.ascontiguousarray(payload))
        gather = torch.empty(int(sizes.sum()), dtype=torch.uint8) if rank == 0 else None
        sharding.concat_to_rank0(dist, pay_t, sizes, rank, gather)
        # the metadata (FL widths: offset known a priori; RL counts: same offsets as the values)
        metas = [None] * world
        dist.all_gather_object(metas, meta)
        if rank == 0:
            m_ref, p_ref = (flrl.fl_cpu_compress if kind == "fl" else flrl.rl_cpu_compress)(data)
            ok = np.array_equal(gather.numpy(), p_ref) and np.array_equal(np.concatenate(metas), m_ref)
            q.put(bool(ok))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("kind", ["fl", "rl"])
def test_two_rank_shards_concatenate_over_gloo(kind):
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + (os.getpid() % 300) + (0 if kind == "fl" else 1)
    n = (3 << 20) + 128 * 7 + 19
    procs = [ctx.Process(target=_worker, args=(r, 2, port, kind, n, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
        assert p.exitcode == 0
    assert q.get(timeout=5) is True


def test_shard_rule_matches_reference():
    sharding = importlib.import_module("fl-rl-compression-mpi_b200.sharding")
    n = 3275658  # the sample BMP
    for world in (1, 2, 4, 8):
        ranges = [sharding.shard_range(n, world, r) for r in range(world)]
        assert ranges[0][0] == 0 and sum(l for _, l in ranges) == n
        assert all(b % 128 == 0 for b, _ in ranges)
        assert all(l == (n // (128 * world)) * 128 for _, l in ranges[:-1])
