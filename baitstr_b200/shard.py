"""Sharding of the select stage over several GPUs / ranks.

Reads are independent (/root/reference/src/select_STR_reads.pyx:294-319 looks at one read
at a time; the reference itself relies on it in pool.map, pyx:363), so the stage shards with
NO data-path collective: every rank scans a contiguous block of the input and the outputs
are concatenated in rank order, which is input order.  torch.distributed is used only for
rendezvous, barriers and scalar reductions of timings (bench.py)."""


def read_range(total, rank, world):
    """Contiguous block of `total` reads for `rank`: (first, count); blocks differ by <= 1."""
    base, extra = divmod(int(total), int(world))
    first = rank * base + min(rank, extra)
    return first, base + (1 if rank < extra else 0)


def file_block(paths, rank, world):
    """Contiguous block of the argv-ordered file list for `rank` (one shard file per GPU when
    len(paths) == world)."""
    first, count = read_range(len(paths), rank, world)
    return list(paths[first:first + count])


def reduce_scalar(dist, x, op="max", device=None):
    """max / sum of a python float over the ranks (identity without a process group)."""
    if dist is None:
        return x
    import torch
    if device is None:
        device = "cuda" if dist.get_backend() == "nccl" else "cpu"
    t = torch.tensor([x], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX if op == "max" else dist.ReduceOp.SUM)
    return float(t.item())


def gather_in_rank_order(dist, local_bytes):
    """Rank 0 gets the outputs of all ranks concatenated in rank (= input) order; others None."""
    if dist is None:
        return local_bytes
    world, rank = dist.get_world_size(), dist.get_rank()
    parts = [None] * world if rank == 0 else None
    dist.gather_object(local_bytes, parts, dst=0)
    return b"".join(parts) if rank == 0 else None
