"""GPU parity tests proper: CUDA path (through the C ABI, host buffers) vs the CPU oracle, bit-exact."""
import os

import numpy as np
import pytest

import helpers
from vt_b200.api import Batch

pytestmark = pytest.mark.gpu


def _run(ctx, oracle, names, seqs, recs):
    ctx.load_reference(names, seqs)
    b = Batch.from_records(recs)
    got = ctx.normalize(b)
    exp = oracle.normalize(seqs, b)
    helpers.assert_norm_equal(got, exp, b.n_rec)
    return got, exp


def test_ref_fetch_roundtrip(ctx):
    rng = np.random.default_rng(7)
    names, seqs = helpers.synth_reference(rng, n_contig=3, length=5000, lower=True, iupac=True)
    ctx.load_reference(names, seqs)
    for c, s in enumerate(seqs):
        L = s.shape[0]
        assert ctx.fetch(c, 0, L - 1) == bytes(s)
        assert ctx.fetch(c, 0, L - 1, upper=True) == bytes(s).upper()
        assert ctx.fetch(c, L - 3, L + 10) == bytes(s[L - 3:])      # clipped like faidx
        assert ctx.fetch(c, -5, 2) == bytes(s[0:3])
        assert ctx.fetch(c, L + 5, L + 9) == bytes(s[L - 1:])


def test_c1_golden_fixture(ctx, oracle):
    seq = helpers.chr20_reference()
    _, recs = helpers.read_vcf_sites(os.path.join(helpers.GOLDEN, "c1_normalize_in.vcf"), {"20": 0})
    got, _ = _run(ctx, oracle, ["20"], [seq], [(r[0], r[1], r[2]) for r in recs])
    import test_oracle_golden as T
    assert helpers.apply_normalize(recs, got, ["20"]) == T._golden_sites(os.path.join(helpers.GOLDEN, "c1_normalize_out.vcf"))
    assert got.counters[4] == 80 and got.counters[10] == 194 and sum(got.counters[:10]) == 80


@pytest.mark.parametrize("seed,lower,iupac", [(11, False, False), (12, True, False), (13, False, True), (14, True, True)])
def test_synthetic_mixed(ctx, oracle, seed, lower, iupac):
    rng = np.random.default_rng(seed)
    names, seqs = helpers.synth_reference(rng, n_contig=4, length=20000, lower=lower, iupac=iupac)
    recs = helpers.synth_records(rng, seqs, 60000)
    got, exp = _run(ctx, oracle, names, seqs, recs)
    assert exp.n_changed > 1000


def test_empty_and_tiny(ctx, oracle):
    names, seqs = ["c1"], [np.frombuffer(b"ACGTACGTAAAAAAAAAACCCCC", np.uint8)]
    _run(ctx, oracle, names, seqs, [])
    _run(ctx, oracle, names, seqs, [(0, 3, [b"G", b"T"])])
    _run(ctx, oracle, names, seqs, [(0, 12, [b"AA", b"A"]), (0, 1, [b"A", b"AA"]), (0, 3, [b"GT", b"G"]), (0, 23, [b"C", b"CC"])])


def test_contig_start_guard_and_single_allele(ctx, oracle):
    # c2 3 AA/A over AAAA... ends as c2 1 AA/A ; and the entered-but-unchanged case c2 1 aa/a
    names, seqs = ["c2"], [np.frombuffer(b"AAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAGT", np.uint8)]
    recs = [(0, 3, [b"AA", b"A"]), (0, 1, [b"aa", b"a"]), (0, 70, [b"A", b"AAAA"]), (0, 80, [b"AG", b"A"]), (0, 5, [b"A"]),
            (0, 40, [b"AAA", b"A", b"AA"]), (0, 40, [b"A", b"AA", b"AAA", b"AAAAAAA"])]
    got, exp = _run(ctx, oracle, names, seqs, recs)
    assert got.status[1] & 0x08 and not (got.status[1] & 0x10)
    assert int(got.chg_pos1[0]) == 1


def test_long_shift_and_long_alleles(ctx, oracle):
    rng = np.random.default_rng(5)
    unit = np.frombuffer(b"ACG", np.uint8)
    tract = np.resize(unit, 30000)
    flank = rng.choice(np.frombuffer(b"ACGT", np.uint8), size=500)
    seq = np.concatenate([flank, tract, flank])
    recs = []
    for k in range(200):
        p = 500 + 29000 + int(rng.integers(0, 900))
        L = 3 * int(rng.integers(1, 40))
        recs.append((0, p, [bytes(seq[p - 1:p - 1 + L + 1]), bytes(seq[p - 1:p])]))       # deletion of L bases
        recs.append((0, p, [bytes(seq[p - 1:p]), bytes(seq[p - 1:p]) + bytes(seq[p:p + L])]))  # insertion
    big = bytes(seq[100:100 + 5000])
    recs.append((0, 101, [big, big[:1]]))
    recs.append((0, 101, [big, big[:2500] + b"T" + big[2501:]]))
    recs.append((0, 600, [bytes(seq[599:599 + 9001]), bytes(seq[599:600])]))
    got, exp = _run(ctx, oracle, ["c1"], [seq], recs)
    assert int(got.chg_le.max()) > 20000


def test_tandem_repeat_tracts_warp_lookahead(ctx, oracle):
    """Short indels (fast path) at the right end of tandem-repeat tracts of 10..6000 bases: shifts beyond 64 bases go
    through the warp-cooperative look-ahead.  Tracts start at the contig start, contain N / lower-case / IUPAC bases and
    interruptions, so the look-ahead must hand back to the per-base walk at the right place."""
    rng = np.random.default_rng(77)
    acgt = np.frombuffer(b"ACGT", np.uint8)
    parts, tracts, at = [], [], 0

    def add(a):
        nonlocal at
        parts.append(a)
        at += a.shape[0]
    # a tract right at the contig start (cap reached), then flank / tract pairs
    for k in range(160):
        ulen = int(rng.integers(1, 7))
        unit = rng.choice(acgt, size=ulen)
        tlen = int(np.exp(rng.uniform(np.log(10), np.log(6000))))
        if k == 0:
            tlen = 2500
        else:
            add(rng.choice(acgt, size=int(rng.integers(40, 120))))
        tract = np.resize(unit, tlen).copy()
        kind = k % 8
        if kind == 3 and tlen > 200:   # an interruption somewhere inside
            tract[int(rng.integers(20, tlen - 20))] = ord("ACGT"[(int("ACGT".index(chr(tract[50]))) + 1) % 4])
        if kind == 4 and tlen > 300:   # N run inside the tract (compares equal at lag ulen only inside the run)
            q = int(rng.integers(100, tlen - 100))
            tract[q:q + int(rng.integers(1, 40))] = ord("N")
        if kind == 5 and tlen > 300:   # soft-masked stretch and an IUPAC code
            q = int(rng.integers(100, tlen - 100))
            tract[q:q + 30] = np.frombuffer(bytes(tract[q:q + 30]).lower(), np.uint8)
            tract[q + 40] = ord("R")
        tracts.append((at, tlen, ulen))
        add(tract)
    add(rng.choice(acgt, size=200))
    seq = np.concatenate(parts)
    recs = []
    for (t0, tlen, ulen) in tracts:
        for units in (1, 2, 4):
            L = ulen * units
            if L + 2 > tlen or L > 30:
                continue
            p = t0 + tlen - L  # 1-based position of the base before the last L tract bases
            if p < 2:
                continue
            anchor = bytes(seq[p - 1:p])
            recs.append((0, p, [bytes(seq[p - 1:p + L]), anchor]))              # deletion of the last copies
            recs.append((0, p + L, [bytes(seq[p + L - 1:p + L]), bytes(seq[p + L - 1:p + L]) + bytes(seq[p:p + L])]))  # insertion after the tract
            m = t0 + tlen // 2
            recs.append((0, m, [bytes(seq[m - 1:m + L]), bytes(seq[m - 1:m])]))   # mid-tract
    recs.sort(key=lambda r: r[1])
    got, exp = _run(ctx, oracle, ["c1"], [seq], recs)
    k = got.n_changed
    assert int(got.chg_le[:k].max()) > 2000 and int((got.chg_le[:k] > 64).sum()) > 100


def test_multi_tile_order_and_counts(ctx, oracle):
    rng = np.random.default_rng(21)
    names, seqs = helpers.synth_reference(rng, n_contig=2, length=200000, n_runs=False)
    recs = helpers.synth_records(rng, seqs, 300000, bad_ref_frac=0.0)
    got, exp = _run(ctx, oracle, names, seqs, recs)
    k = got.n_changed
    assert np.all(np.diff(got.chg_rec[:k].astype(np.int64)) > 0)  # compaction keeps input order


def test_chunked_host_call_matches_single_chunk(oracle):
    """vtn_normalize pipelines the batch in chunks; with a tiny chunk size the output must be identical."""
    import os
    from vt_b200 import Context
    rng = np.random.default_rng(77)
    names, seqs = helpers.synth_reference(rng, n_contig=2, length=30000)
    recs = helpers.synth_records(rng, seqs, 20000)
    b = Batch.from_records(recs)
    os.environ["VTN_CHUNK_REC"] = "1024"
    try:
        c2 = Context(0)
    finally:
        del os.environ["VTN_CHUNK_REC"]
    try:
        c2.load_reference(names, seqs)
        got = c2.normalize(b)
    finally:
        c2.close()
    exp = oracle.normalize(seqs, b)
    helpers.assert_norm_equal(got, exp, b.n_rec)
