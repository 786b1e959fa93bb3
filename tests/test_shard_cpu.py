"""CPU, world_size 2 over gloo: the N>1 host logic (vt_b200/shard.py) -- contig/range sharding, per-rank processing,
counter sums and the ordered re-merge -- gives exactly the unsharded result.  The per-shard compute here is the
oracle (this is a test of the plumbing; the GPU path per rank is covered by the -m gpu tests)."""
import os
import sys

import numpy as np
import torch.distributed as dist
import torch.multiprocessing as mp

import helpers
from vt_b200 import shard
from vt_b200.api import Batch


def _inputs():
    rng = np.random.default_rng(77)
    names, seqs = helpers.synth_reference(rng, n_contig=5, length=6000)
    recs = helpers.synth_records(rng, seqs, 9000)
    return seqs, Batch.from_records(recs)


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    seqs, batch = _inputs()
    lo, hi = shard.plan_shards(batch.rid, world)[rank]
    sub = shard.slice_batch(batch, lo, hi)
    res = helpers.Oracle().normalize(seqs, sub)
    res.status, res.vtype = np.asarray(res.status[:sub.n_rec]), np.asarray(res.vtype[:sub.n_rec])
    parts = shard.gather_parts(shard.result_part(res, lo), dist)
    if rank == 0:
        q.put(shard.merge_results(parts))
    dist.barrier()
    dist.destroy_process_group()


def test_plan_covers_everything_and_prefers_contig_boundaries():
    rid = np.repeat(np.arange(6), [100, 300, 50, 250, 200, 100]).astype(np.int32)
    for n in (1, 2, 3, 4, 8):
        plan = shard.plan_shards(rid, n)
        assert plan[0][0] == 0 and plan[-1][1] == rid.shape[0]
        assert all(plan[k][1] == plan[k + 1][0] for k in range(n - 1))
        assert all(lo <= hi for lo, hi in plan)
    assert shard.plan_shards(rid, 2)[0][1] == 500  # nearest contig boundary (450) is 5% away: the contig is split
    rid2 = np.repeat(np.arange(2), [505, 495]).astype(np.int32)
    assert shard.plan_shards(rid2, 2)[0][1] == 505  # within 2%: cut on the contig boundary
    assert shard.plan_shards(np.zeros(0, np.int32), 4) == [(0, 0)] * 4


def test_two_rank_gloo_merge_equals_unsharded():
    world, port = 2, 29741
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    merged = q.get(timeout=300)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    seqs, batch = _inputs()
    exp = helpers.Oracle().normalize(seqs, batch)
    k = exp.n_changed
    assert merged["counters"] == list(exp.counters)
    assert np.array_equal(merged["status"], np.asarray(exp.status[:batch.n_rec]))
    assert np.array_equal(merged["vtype"], np.asarray(exp.vtype[:batch.n_rec]))
    for name in ("chg_rec", "chg_pos1", "chg_lt", "chg_le", "chg_rt"):
        assert np.array_equal(merged[name], np.asarray(getattr(exp, name)[:k])), name
    assert np.array_equal(merged["chg_off"], np.asarray(exp.chg_off[:k + 1]).astype(np.int64))
    assert np.array_equal(merged["chg_bytes"], np.asarray(exp.chg_bytes[:exp.n_changed_bytes]))
