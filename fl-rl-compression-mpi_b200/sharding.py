"""Host-side logic of the multi-GPU path: contiguous shards, ONE exchange of the per-shard
compressed sizes, exclusive scan -> output offsets, concatenation on rank 0.  Backend-agnostic: bench.py runs it over
NCCL with device tensors, tests/test_sharding_gloo.py over gloo with CPU tensors (world_size 2, no GPU).

The in-process variant of the same logic (one process, all GPUs of the box) lives in csrc/flrl_host.cu."""
from __future__ import annotations


def shard_range(n: int, world: int, rank: int, unit: int = 128):
    """[begin, begin + length) of rank's shard: per = (n / (unit * world)) * unit, the last rank takes the rest
    (/root/reference/src/file_io.cu:46-51, in 64-bit arithmetic)."""
    per = (n // (unit * world)) * unit
    begin = rank * per
    return begin, (n - begin if rank == world - 1 else per)


def exchange_sizes(dist, my_size, sizes_out, offsets_out):
    """All-gather one integer per rank (the shard's compressed size) and turn the vector into exclusive offsets.
    my_size: 1-element int64 tensor; sizes_out / offsets_out: world-element int64 tensors on the same device."""
    import torch

    dist.all_gather_into_tensor(sizes_out, my_size) if sizes_out.is_cuda else dist.all_gather(list(sizes_out.split(1)), my_size)
    torch.cumsum(sizes_out, 0, out=offsets_out)
    offsets_out.sub_(sizes_out)
    return offsets_out


def concat_to_rank0(dist, payload, sizes, rank, gather_buf=None):
    """Every rank delivers payload[:sizes[rank]] to its offset in rank 0's gather_buf (the merge of the reference's
    gpuNCCLCompress, fl_gpu.cu:221-238, without moving payloads through an all-gather)."""
    s = [int(x) for x in sizes.tolist()]
    world = len(s)
    offs = [0]
    for x in s:
        offs.append(offs[-1] + x)
    if rank == 0:
        gather_buf[: s[0]].copy_(payload[: s[0]])
        ops = [dist.P2POp(dist.irecv, gather_buf[offs[r]:offs[r + 1]], r) for r in range(1, world) if s[r]]
    else:
        ops = [dist.P2POp(dist.isend, payload[: s[rank]], 0)] if s[rank] else []
    if ops:
        for req in dist.batch_isend_irecv(ops):
            req.wait()
    return offs
