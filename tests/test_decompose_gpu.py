"""GPU parity: decompose fan-out plan and fused decompose->normalize vs the CPU oracle, bit-exact."""
import os

import numpy as np
import pytest

import helpers
from vt_b200.api import Batch

pytestmark = pytest.mark.gpu


def _decomposed_records(recs, src, alt):
    out = []
    for s, a in zip(src, alt):
        rid, pos1, al = recs[int(s)][:3]
        out.append((rid, pos1, [al[0], al[int(a)]] if a else [al[0]]))
    return out


def test_decompose_golden_fixture(ctx, oracle):
    _, recs = helpers.read_vcf_sites(os.path.join(helpers.GOLDEN, "c1_decompose_in.vcf"), {})
    b = Batch.from_records([(0, r[1], r[2]) for r in recs])
    src, alt, cnt, _ = ctx.decompose(b)
    assert list(src) == [0, 0] and list(alt) == [1, 2] and cnt == [1, 0, 1, 1]


@pytest.mark.parametrize("seed", [31, 32])
def test_decompose_plan_matches_oracle(ctx, oracle, seed):
    rng = np.random.default_rng(seed)
    names, seqs = helpers.synth_reference(rng, n_contig=2, length=20000)
    recs = helpers.synth_records(rng, seqs, 70000, multi_frac=0.5)
    recs[5] = (0, 10, [b"A"])  # single-allele record
    b = Batch.from_records(recs)
    src, alt, cnt, _ = ctx.decompose(b)
    esrc, ealt, ecnt = oracle.decompose_plan(b)
    np.testing.assert_array_equal(src, esrc)
    np.testing.assert_array_equal(alt, ealt)
    assert cnt == ecnt


@pytest.mark.parametrize("seed", [41, 42])
def test_fused_decompose_normalize_matches_oracle(ctx, oracle, seed):
    rng = np.random.default_rng(seed)
    names, seqs = helpers.synth_reference(rng, n_contig=3, length=30000, lower=(seed == 42))
    recs = helpers.synth_records(rng, seqs, 50000, multi_frac=0.6)
    recs[7] = (0, 10, [b"A"])
    b = Batch.from_records(recs)
    ctx.load_reference(names, seqs)
    src, alt, cnt, got = ctx.decompose(b, normalize=True)
    esrc, ealt, ecnt = oracle.decompose_plan(b)
    np.testing.assert_array_equal(src, esrc)
    np.testing.assert_array_equal(alt, ealt)
    assert cnt == ecnt
    b2 = Batch.from_records(_decomposed_records(recs, esrc, ealt))
    exp = oracle.normalize(seqs, b2)
    helpers.assert_norm_equal(got, exp, b2.n_rec)


def test_decompose_empty(ctx):
    b = Batch.from_records([])
    src, alt, cnt, _ = ctx.decompose(b)
    assert len(src) == 0 and cnt == [0, 0, 0, 0]


def test_decompose_many_alts_per_record(ctx, oracle):
    """records with dozens of ALTs: more than 2048 outputs per 256-record tile, i.e. the emit kernel's binary-search
    path instead of its expansion table; fused result checked against the oracle as well"""
    rng = np.random.default_rng(51)
    names, seqs = helpers.synth_reference(rng, n_contig=1, length=30000, n_runs=False)
    s = seqs[0]
    recs = []
    for p in np.sort(rng.choice(np.arange(100, s.shape[0] - 100), size=3000, replace=False)):
        p = int(p)
        ref = bytes(s[p:p + 3])
        n_alt = int(rng.integers(1, 60))
        alts = []
        for q in range(n_alt):
            k = q % 3
            if k == 0:
                alts.append(ref[:1] + bytes(rng.choice(np.frombuffer(b"ACGT", np.uint8), size=1 + q // 3)) + ref[1:])
            elif k == 1:
                alts.append(ref[:1] if q < 3 else ref[:2] + bytes(rng.choice(np.frombuffer(b"ACGT", np.uint8), size=q // 3)))
            else:
                alts.append(bytes([b"ACGT"[(b"ACGT".index(ref[0]) + 1 + q // 3) % 4]]) + ref[1:])
        recs.append((0, p + 1, [ref] + alts))
    b = Batch.from_records(recs)
    ctx.load_reference(names, seqs)
    src, alt, cnt, got = ctx.decompose(b, normalize=True)
    esrc, ealt, ecnt = oracle.decompose_plan(b)
    np.testing.assert_array_equal(src, esrc)
    np.testing.assert_array_equal(alt, ealt)
    assert cnt == ecnt and len(src) > 60000
    b2 = Batch.from_records(_decomposed_records(recs, esrc, ealt))
    helpers.assert_norm_equal(got, oracle.normalize(seqs, b2), b2.n_rec)
