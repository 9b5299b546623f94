"""torch.distributed plumbing for bench.py and the multi-process tests (gloo on CPU, nccl on GPUs): max / sum of one
scalar over the ranks, and the per-rank id counts that rebase the shards' offsets.  Not part of the product package
(hutoken_b200/ imports no PyTorch); no tensor of ids ever goes through a collective."""
from __future__ import annotations

import numpy as np


def all_reduce_scalar(x, op="max", device=None):
    """max / sum of one float over all ranks (the bench's max-over-ranks timing and whole-job totals)."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(x)
    t = torch.tensor([float(x)], dtype=torch.float64, device=device if device is not None else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX if op == "max" else dist.ReduceOp.SUM)
    return float(t.item())


def gather_counts(n_ids, device=None):
    """Every rank's id count, on every rank: the only data-dependent value shards have to exchange."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return [int(n_ids)]
    t = torch.tensor([int(n_ids)], dtype=torch.int64, device=device if device is not None else "cpu")
    outs = [torch.zeros_like(t) for _ in range(dist.get_world_size())]
    dist.all_gather(outs, t)
    return [int(o.item()) for o in outs]


def global_offsets(local_tok_offsets, n_ids, device=None):
    """This rank's token offsets rebased into the whole job's id space (+ the job's total id count)."""
    import torch.distributed as dist
    counts = gather_counts(n_ids, device)
    rank = dist.get_rank() if (dist.is_available() and dist.is_initialized()) else 0
    base = sum(counts[:rank])
    return np.asarray(local_tok_offsets, dtype=np.uint64) + np.uint64(base), sum(counts)
