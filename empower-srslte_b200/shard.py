"""Multi-GPU sharding of the decode path: code blocks are independent and HARQ state is per
(cell, HARQ process), so transport blocks are partitioned by CELL — every (re)transmission of a process lands on the
GPU that holds its soft buffer — and nothing is exchanged on the data path.  torch.distributed is used for the
rendezvous, the barrier around the timed region and the two scalar reductions below, nothing else."""


def owner(cell, world):
    """rank that owns every HARQ process of `cell` (soft buffers never move between GPUs)"""
    return int(cell) % int(world)


def partition(keys, rank, world):
    """indices into `keys` (iterable of (cell, harq_pid, ...) tuples) of the transport blocks this rank decodes"""
    return [i for i, k in enumerate(keys) if owner(k[0], world) == rank]


def reduce_job(dist, device, ms, **sums):
    """whole-job figures from per-rank ones: time = MAX over ranks (device-timed region), every keyword = SUM over ranks.
    `dist` is torch.distributed (or None for a single process). Returns (ms, dict)."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return float(ms), {k: float(v) for k, v in sums.items()}
    import torch
    t = torch.tensor([float(ms)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    names = sorted(sums)
    s = torch.tensor([float(sums[k]) for k in names], dtype=torch.float64, device=device)
    dist.all_reduce(s, op=dist.ReduceOp.SUM)
    return float(t.item()), {k: float(v) for k, v in zip(names, s.tolist())}
