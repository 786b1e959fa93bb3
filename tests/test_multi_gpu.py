"""GPU: several devices behind one call (vtn_multi_*, include/vtn.h).  The merged result of the sharded run must be bit-identical
to the single-device call on the whole batch -- on every GPU the box has, and (so that a one-GPU box tests the sharding too) with
the same device listed several times (independent contexts, streams and reference copies on one GPU)."""
import numpy as np
import pytest

import helpers
import test_bi_gpu as TB
from vt_b200 import synth
from vt_b200.api import Batch, MultiContext, shard_plan

pytestmark = pytest.mark.gpu


def _device_lists():
    try:
        import torch
        n = torch.cuda.device_count()
    except Exception:
        n = 0
    lists = [[0, 0], [0, 0, 0]]
    if n >= 2:
        lists.append(list(range(min(n, 8))))
    return lists


@pytest.mark.parametrize("devs", _device_lists())
def test_multi_normalize_bi_equals_single_device(ctx, oracle, devs):
    rng = np.random.default_rng(51)
    names, seqs = helpers.synth_reference(rng, n_contig=5, length=40000, lower=True, iupac=True)
    recs = [r for r in helpers.synth_records(rng, seqs, 90000, multi_frac=0.0) if len(r[2]) == 2 and r[0] >= 0]
    recs.sort(key=lambda r: (r[0], r[1]))
    b = Batch.from_records(recs)
    exp = oracle.normalize(seqs, b)
    with MultiContext(devs) as m:
        m.load_reference(names, seqs)
        bi = m.ctx0.bi_pack(b)
        try:
            TB._assert_bi_equal(m.normalize_bi(bi), exp, b.n_rec)
            TB._assert_bi_equal(m.normalize_bi(bi), exp, b.n_rec)  # staging buffers are reused
        finally:
            bi.close()


@pytest.mark.parametrize("devs", _device_lists())
def test_multi_decompose_normalize_equals_single_device(ctx, devs):
    """C3's generator (all records multiallelic): the sharded fused decompose -> normalize against the single-device call"""
    names, seqs = synth.make_reference(scale=0.001)
    b = synth.make_multiallelic(seqs, 60000, seed=synth.SEED_VCF_C3 + 3)
    ctx.load_reference(names, seqs)
    src, alt, dcnt, res = ctx.decompose(b, normalize=True)
    with MultiContext(devs) as m:
        m.load_reference(names, seqs)
        s2, a2, d2, r2 = m.decompose_normalize(b)
        s3, a3, d3, r3 = m.decompose_normalize(b)
    n = src.shape[0]
    for (s_, a_, d_, r_) in ((s2, a2, d2, r2), (s3, a3, d3, r3)):
        assert s_.shape[0] == n and d_ == dcnt
        np.testing.assert_array_equal(s_, src)
        np.testing.assert_array_equal(a_, alt)
        helpers.assert_norm_equal(r_, res, n)


def test_multi_more_shards_than_records(ctx, oracle):
    names, seqs = ["c1"], [np.frombuffer(b"ACGTACGTAAAAAAAAAACCCCCGGGGTTTTACGT" * 4, np.uint8)]
    b = Batch.from_records([(0, 12, [b"AA", b"A"]), (0, 30, [b"T", b"TT"])])
    exp = oracle.normalize(seqs, b)
    with MultiContext([0, 0, 0, 0]) as m:
        m.load_reference(names, seqs)
        bi = m.ctx0.bi_pack(b)
        try:
            TB._assert_bi_equal(m.normalize_bi(bi), exp, b.n_rec)
        finally:
            bi.close()
        mb = Batch.from_records([(0, 12, [b"AA", b"A", b"AAA"]), (0, 30, [b"T", b"TT", b"G"])])
        s2, a2, d2, r2 = m.decompose_normalize(mb)
    ctx.load_reference(names, seqs)
    src, alt, dcnt, res = ctx.decompose(mb, normalize=True)
    assert d2 == dcnt and np.array_equal(s2, src) and np.array_equal(a2, alt)
    helpers.assert_norm_equal(r2, res, src.shape[0])


def test_shard_plan_by_contig_then_range():
    rid = np.repeat(np.arange(6), [100, 300, 50, 250, 200, 100]).astype(np.int32)
    for n in (1, 2, 3, 4, 8):
        cut = shard_plan(rid, n)
        assert cut[0] == 0 and cut[-1] == rid.shape[0] and np.all(np.diff(cut.astype(np.int64)) >= 0)
    assert shard_plan(rid, 2)[1] == 500
    assert shard_plan(np.repeat(np.arange(2), [505, 495]).astype(np.int32), 2)[1] == 505
    assert list(shard_plan(rid, 4, align=128)) == [0, 256, 512, 768, 1000]
