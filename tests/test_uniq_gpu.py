"""GPU: vtn_uniq (vt uniq on the device) through the C ABI against the oracle's literal restatement of the
reference's synced reader, bit-exact on every output array, and against the golden vectors the unmodified reference
produced (tests/golden/uniq_*)."""
import os

import numpy as np
import pytest

import helpers
import uniq_helpers as U
from test_uniq_oracle import _batch, _read_in

pytestmark = pytest.mark.gpu


def _check(ctx, oracle, recs):
    b = _batch(recs)
    order, first, cnt, bad = ctx.uniq(b)
    eo, ef, ec, eb = oracle.uniq_plan(b)
    assert bad == eb == -1
    assert cnt == ec
    np.testing.assert_array_equal(order, eo)
    np.testing.assert_array_equal(first, ef)
    return order, first, cnt


@pytest.mark.parametrize("name", ["uniq_a", "uniq_b"])
def test_uniq_reference_golden(ctx, name):
    recs = _read_in(os.path.join(helpers.GOLDEN, name + "_in.vcf"))
    order, first, cnt, bad = ctx.uniq(_batch(recs))
    assert bad == -1
    assert U.expected_from_plan(recs, order, first) == U.read_kept(os.path.join(helpers.GOLDEN, name + "_out.vcf"))
    err = open(os.path.join(helpers.GOLDEN, name + "_out.stderr")).read()
    assert "Total number of observed variants   %d\n" % cnt[0] in err
    assert "Total number of unique variants     %d\n" % cnt[1] in err


@pytest.mark.parametrize("seed,n,mg,big", [(1, 5000, 4, 0), (2, 20000, 12, 0), (3, 30000, 9, 40), (4, 3000, 30, 7)])
def test_uniq_matches_oracle(ctx, oracle, seed, n, mg, big):
    recs = U.gen_uniq_records(np.random.default_rng(seed), n, max_group=mg, big_group_every=big)
    _check(ctx, oracle, recs)


def test_uniq_edge_cases(ctx, oracle):
    # empty batch
    order, first, cnt, bad = ctx.uniq(helpers.Batch.from_records([]))
    assert cnt == [0, 0] and bad == -1 and len(order) == 0
    # a single record; one group that is the whole file (no trailing record, first group)
    _check(ctx, oracle, [(0, 5, ["A", "C"], None)])
    _check(ctx, oracle, [(0, 5, ["A", "C"], None)] * 7)
    # all records distinct positions
    _check(ctx, oracle, [(0, 5 + i, ["A", "C"], None) for i in range(300)])
    # one group of 500 duplicates between singletons (global heap scratch), twice
    big = [(0, 3, ["G", "T"], None)] + [(0, 9, ["A", ["C", "CT", "C"][i % 3]], None) for i in range(500)] + [(0, 10, ["A", "C"], None)]
    _check(ctx, oracle, big + [(1, 9, al, ov) for (_, _, al, ov) in big[1:]])
    # prefix-ordered keys: "A,AT" < "AT,A" because ',' < 'T'; ALT "." (single allele); multiallelic ALT order ignored
    recs = [(0, 7, al, None) for al in (["AT", "A"], ["A", "AT"], ["A"], ["A", "AT", "C"], ["A", "C", "AT"], ["A", "AT"], ["AT", "A"], ["A"])]
    _check(ctx, oracle, recs)
    # contig change right after a group
    _check(ctx, oracle, [(0, 7, ["A", "C"], None)] * 3 + [(1, 7, ["A", "C"], None)] * 2 + [(2, 1, ["A", "C"], None)])


def test_uniq_not_in_order(ctx, oracle):
    recs = [(0, 10, ["A", "C"], None), (0, 10, ["A", "C"], None), (0, 12, ["A", "C"], None), (0, 11, ["A", "C"], None), (1, 13, ["A", "G"], None),
            (0, 50, ["A", "G"], None)]
    b = _batch(recs)
    _, _, _, bad = ctx.uniq(b)
    assert bad == 3 == oracle.uniq_plan(b)[3]


def test_uniq_fullsize_properties(ctx):
    """2 M records (size-independent properties): order is a permutation that keeps position groups in place, firsts = number of
    distinct (rid, pos, REF, sorted ALTs), and uniq of the survivors keeps everything (idempotence)."""
    rng = np.random.default_rng(7)
    n = 2_000_000
    pos = np.sort(rng.integers(1, 3_000_000, n)).astype(np.int32)
    rid = np.zeros(n, np.int32)
    bases = np.frombuffer(b"ACGT", np.uint8)
    ref = bases[rng.integers(0, 4, n)]
    alt = bases[(np.searchsorted(bases, ref) + rng.integers(1, 4, n)) % 4]
    data = np.stack([ref, alt], 1).reshape(-1)
    b = helpers.Batch(pos, rid, np.arange(0, 2 * n + 1, 2, dtype=np.uint32), np.arange(0, 2 * n + 1, dtype=np.uint32), data)
    order, first, cnt, bad = ctx.uniq(b)
    assert bad == -1 and cnt[0] == n
    assert np.array_equal(np.sort(order), np.arange(n, dtype=np.uint32))
    assert np.array_equal(pos[order], pos)  # groups stay in place
    key = pos.astype(np.int64) * 65536 + ref.astype(np.int64) * 256 + alt
    assert cnt[1] == len(np.unique(key)) == int(first.sum())
    ko = key[order]
    assert np.all(np.diff(ko) >= 0)  # sorted by (pos, REF, ALT) since single-base keys compare like their bytes
    assert np.array_equal(first[1:] == 1, np.diff(ko) != 0) and first[0] == 1
    keep = np.sort(order[first == 1])
    b2 = helpers.Batch(pos[keep], rid[keep], np.arange(0, 2 * len(keep) + 1, 2, dtype=np.uint32), np.arange(0, 2 * len(keep) + 1, dtype=np.uint32),
                       data.reshape(-1, 2)[keep].reshape(-1))
    _, first2, cnt2, _ = ctx.uniq(b2)
    assert cnt2 == [len(keep), len(keep)] and first2.all()
