"""N>1 path, host side, on CPU: world_size-2 gloo processes shard a batch by document (hut_plan_shards), each
"encodes" its shard with the ORACLE standing in for the device (the product path needs a GPU; this test checks the
sharding / rebasing / reduction logic around it), and the rebased per-shard results must equal the single-shard
result.  Mirrors the reference's threaded batch tests (tests/test_tokenizer.py:247-263: batch_encode with 1/3/4/8
threads equals the single-thread result)."""
import os
import socket
import sys

import numpy as np
import pytest

from common import oracle_for, pack
import synth


def test_plan_shards_properties():
    from hutoken_b200 import dist as D
    rng = np.random.default_rng(5)
    for n_docs in (0, 1, 2, 7, 1000):
        lens = rng.integers(0, 5000, size=n_docs)
        offs = np.zeros(n_docs + 1, dtype=np.uint64)
        offs[1:] = np.cumsum(lens)
        for n in (1, 2, 3, 8):
            b = D.plan_shards(offs, n)
            assert b[0] == 0 and b[-1] == n_docs and np.all(np.diff(b.astype(np.int64)) >= 0)
            if n_docs >= 100:  # balanced by bytes: no shard further than one document from its fair share
                per = np.diff(offs[b.astype(np.int64)].astype(np.int64))
                assert abs(per - int(offs[-1]) / n).max() <= lens.max() + 1
    # empty and skewed batches
    offs = np.array([0, 0, 0, 10, 10], dtype=np.uint64)
    b = D.plan_shards(offs, 2)
    assert b.tolist()[0] == 0 and b.tolist()[-1] == 4


def test_rebase_offsets():
    from hutoken_b200 import dist as D
    a = np.array([0, 3, 3], dtype=np.uint64)       # shard 0: 2 docs, 3 ids
    b = np.array([0, 5], dtype=np.uint64)          # shard 1: 1 doc, 5 ids
    c = np.array([0], dtype=np.uint64)             # shard 2: no docs
    out = D.rebase_offsets([a, b, c], [3, 5, 0])
    assert out.tolist() == [0, 3, 3, 8]


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, name, seed, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    import torch.distributed as dist
    from hutoken_b200 import dist as D
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        assert D.rank_env() == (rank, world, rank)
        data, offs = synth.corpus(seed, 1 << 20, special_every=250)
        orc = oracle_for(name)
        sdata, soffs, first_doc = D.local_shard(data, offs, rank, world)
        ids, tok = orc.encode_packed(np.concatenate([sdata, np.zeros(16, np.uint8)]), soffs)
        import dist_torch as DT
        gtok, total = DT.global_offsets(tok, len(ids))
        # the bench's reductions: max-over-ranks of a time, sum of the units
        tmax = DT.all_reduce_scalar(1.0 + rank, "max")
        bsum = DT.all_reduce_scalar(float(len(sdata)), "sum")
        q.put((rank, first_doc, ids, gtok, total, tmax, bsum))
    finally:
        dist.barrier()
        dist.destroy_process_group()


@pytest.mark.parametrize("name", ["gpt2_small", "llama_small"])
def test_two_rank_gloo_matches_single_shard(name):
    import torch.multiprocessing as mp
    world, seed = 2, 77
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, name, seed, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = sorted([q.get(timeout=300) for _ in range(world)], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    data, offs = synth.corpus(seed, 1 << 20, special_every=250)
    oids, otok = oracle_for(name).encode_packed(data, offs)
    ids = np.concatenate([g[2] for g in got])
    assert np.array_equal(ids, oids)
    # every rank's rebased offsets are its slice of the global offsets
    for rank, first_doc, _, gtok, total, tmax, bsum in got:
        assert total == len(oids)
        assert np.array_equal(gtok, otok[first_doc:first_doc + len(gtok)])
        assert tmax == float(world) and bsum == float(offs[-1])
