"""Host-side logic of the multi-rank path (one process per GPU): which documents a rank encodes, and how the
per-shard results become one global result.

Replaces the reference's mutex-guarded "next document" counter shared by worker threads (taskqueue.c:16-35,
lib.c:779-794).  Shards are independent, so nothing crosses ranks on the data path; the only exchange is one
integer per rank (its id count) to rebase the per-shard offsets.  Nothing here imports PyTorch: the
torch.distributed plumbing of the bench and of the multi-process tests lives in tools/dist_torch.py.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import capi


def rank_env():
    """(rank, world, local_rank) from the torchrun environment; (0, 1, 0) outside it."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")),
            int(os.environ.get("LOCAL_RANK", "0")))


def plan_shards(doc_offsets, n_shards):
    """bounds[n_shards + 1]: shard s encodes documents [bounds[s], bounds[s+1]) -- contiguous, balanced by bytes.
    The same C function (hut_plan_shards) places the shards of a multi-device hut_encode_batch call."""
    lib = capi.load_library()
    offs = np.ascontiguousarray(doc_offsets, dtype=np.uint64)
    bounds = np.zeros(n_shards + 1, dtype=np.uint64)
    fn = lib.hut_plan_shards
    fn.argtypes = [C.c_void_p, C.c_uint64, C.c_int, C.c_void_p]
    rc = fn(offs.ctypes.data, C.c_uint64(len(offs) - 1), n_shards, bounds.ctypes.data)
    if rc != 0:
        raise ValueError("hut_plan_shards: bad argument")
    return bounds


def local_shard(data, doc_offsets, rank, world):
    """This rank's slice of a packed batch: (bytes view, rebased offsets, first document index)."""
    b = plan_shards(doc_offsets, world)
    lo, hi = int(b[rank]), int(b[rank + 1])
    offs = np.asarray(doc_offsets, dtype=np.uint64)
    base = offs[lo]
    return data[int(base):int(offs[hi])], offs[lo:hi + 1] - base, lo


def rebase_offsets(shard_tok_offsets, shard_n_ids):
    """Per-shard token offsets (each starting at 0, n_docs_s + 1 entries) -> the global n_docs + 1 offsets:
    one add per shard, no per-token work."""
    out, base = [], 0
    for tok, n in zip(shard_tok_offsets, shard_n_ids):
        tok = np.asarray(tok, dtype=np.uint64)
        out.append(tok[:-1] + np.uint64(base))
        base += int(n)
    out.append(np.array([base], dtype=np.uint64))
    return np.concatenate(out)
