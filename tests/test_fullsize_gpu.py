"""GPU, BASELINE.json's full batch size (10 M records, configuration C2; reference shrunk to 5% so that the test
generates its inputs in seconds): size-independent properties instead of an oracle run.
  * idempotence: normalizing the normalized output changes nothing (is_normalized holds for every output record);
  * conservation: no_variants + no_refs == records, the ten category counters sum to the changed count,
    new POS = POS - left_extended + left_trimmed, every new allele keeps >= 1 base;
  * a 1% sample of the records agrees bit-exactly with the CPU oracle;
  * the pipelined host-buffer call (1 M-record chunks) and the device-resident call give identical bytes."""
import numpy as np
import pytest

import helpers
from vt_b200 import synth
from vt_b200.api import Batch

pytestmark = pytest.mark.gpu
N = 10_000_000


def _apply(batch, res):
    """the batch vt would write: changed records get their new POS / alleles (numpy, vectorised)"""
    n = batch.n_rec
    k = res.n_changed
    lens = np.diff(batch.allele_off.astype(np.int64))
    delta = np.zeros(n, np.int64)
    rec = res.chg_rec[:k].astype(np.int64)
    delta[rec] = res.chg_le[:k].astype(np.int64) - res.chg_rt[:k].astype(np.int64) - res.chg_lt[:k].astype(np.int64)
    slot_rec = np.repeat(np.arange(n), np.diff(batch.allele_beg.astype(np.int64)))
    new_lens = lens + delta[slot_rec]
    assert new_lens.min() >= 1
    # source offset of every new allele in [old bytes | changed bytes]
    src_off = batch.allele_off[:-1].astype(np.int64).copy()
    changed = np.zeros(n, bool)
    changed[rec] = True
    first_slot = batch.allele_beg[:-1].astype(np.int64)
    # biallelic here: slot 0 at chg_off, slot 1 right after it
    base = batch.n_bytes
    co = res.chg_off[:k].astype(np.int64)
    src_off[first_slot[rec]] = base + co
    src_off[first_slot[rec] + 1] = base + co + new_lens[first_slot[rec]]
    both = np.concatenate([batch.data[:batch.n_bytes], res.chg_bytes[:res.n_changed_bytes]])
    seg, within = synth._ragged(new_lens)
    data = both[src_off[seg] + within]
    pos = batch.pos1.copy()
    pos[rec] = res.chg_pos1[:k]
    off = np.concatenate([[0], np.cumsum(new_lens)]).astype(np.uint32)
    return Batch(pos, batch.rid.copy(), batch.allele_beg.copy(), off, data)


def test_c2_full_size_properties(ctx, oracle):
    names, seqs = synth.make_reference(scale=0.05)
    batch = synth.make_biallelic(seqs, N, seed=synth.SEED_VCF_C2)
    ctx.load_reference(names, seqs)
    res = ctx.normalize(batch)                       # pipelined host-buffer call, 10 chunks
    k = res.n_changed
    cnt = res.counters
    assert cnt[10] + cnt[11] == N and sum(cnt[:10]) == k and k > 500_000
    rec = res.chg_rec[:k].astype(np.int64)
    assert np.all(np.diff(rec) > 0)                  # compacted in input order
    assert np.array_equal(res.chg_pos1[:k], batch.pos1[rec] - res.chg_le[:k].astype(np.int32) + res.chg_lt[:k].astype(np.int32))
    assert np.all((res.status[:N] & 0x10 != 0) == np.isin(np.arange(N), rec))
    # resident call == host-buffer call
    db = ctx.upload(batch)
    ctx.normalize_resident(db)
    res2 = ctx.fetch_result(db, batch)
    ctx.free(db)
    assert res2.n_changed == k and res2.counters == cnt
    assert np.array_equal(res2.status[:N], res.status[:N]) and np.array_equal(res2.chg_bytes[:res.n_changed_bytes], res.chg_bytes[:res.n_changed_bytes])
    # idempotence
    out = _apply(batch, res)
    again = ctx.normalize(out)
    assert again.n_changed == 0 and sum(again.counters[:10]) == 0
    assert not np.any(again.status[:N] & 0x13)       # still REF-consistent, nothing changed
    # a 1% contiguous sample against the oracle, bit-exact
    lo, hi = 4_000_000, 4_100_000
    from vt_b200 import shard
    sub = shard.slice_batch(batch, lo, hi)
    exp = oracle.normalize(seqs, sub)
    got = ctx.normalize(sub)
    helpers.assert_norm_equal(got, exp, sub.n_rec)


def test_c3_fused_equals_two_step(ctx, oracle):
    """decompose -> normalize fused on device == decompose plan, then normalize of the materialised biallelic batch
    (2 M multiallelic records); the plan equals the oracle's."""
    names, seqs = synth.make_reference(scale=0.02)
    batch = synth.make_multiallelic(seqs, 2_000_000, seed=synth.SEED_VCF_C3)
    ctx.load_reference(names, seqs)
    src, alt, dcnt, fres = ctx.decompose(batch, normalize=True)
    src2, alt2, dcnt2, _ = ctx.decompose(batch, normalize=False)
    assert np.array_equal(src, src2) and np.array_equal(alt, alt2) and dcnt == dcnt2
    esrc, ealt, ecnt = oracle.decompose_plan(batch)
    assert np.array_equal(src, esrc) and np.array_equal(alt, ealt) and dcnt == ecnt
    # materialise the biallelic batch on the host and normalize it separately
    n_out = src.shape[0]
    lens = np.diff(batch.allele_off.astype(np.int64))
    s0 = batch.allele_beg[:-1].astype(np.int64)[src.astype(np.int64)]
    s1 = s0 + alt.astype(np.int64)
    slot = np.empty(2 * n_out, np.int64)
    slot[0::2], slot[1::2] = s0, s1
    seg, within = synth._ragged(lens[slot])
    data = batch.data[batch.allele_off[:-1].astype(np.int64)[slot][seg] + within]
    off = np.concatenate([[0], np.cumsum(lens[slot])]).astype(np.uint32)
    bi = Batch(batch.pos1[src.astype(np.int64)].copy(), batch.rid[src.astype(np.int64)].copy(), (np.arange(n_out + 1) * 2).astype(np.uint32), off, data)
    sres = ctx.normalize(bi)
    assert sres.n_changed == fres.n_changed and sres.counters == fres.counters
    helpers.assert_norm_equal(fres, sres, n_out)
