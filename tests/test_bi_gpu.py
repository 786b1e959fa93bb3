"""GPU parity of the compact biallelic layout (vtn_batch_bi / vtn_normalize_bi, include/vtn.h) against the CPU oracle:
the same records as the general batch, so status / vtype / changed records must come out bit-identical once sv, the
changed list and the notes are expanded back into vtn_norm_result's arrays."""
import os

import numpy as np
import pytest

import helpers
from vt_b200.api import Batch

pytestmark = pytest.mark.gpu


def _assert_bi_equal(res, exp, n_rec):
    status, vtype = res.expand()
    assert res.n_changed == exp.n_changed, (res.n_changed, exp.n_changed)
    assert res.n_changed_bytes == exp.n_changed_bytes
    np.testing.assert_array_equal(vtype[:n_rec], exp.vtype[:n_rec])
    np.testing.assert_array_equal(status[:n_rec], exp.status[:n_rec])
    k = exp.n_changed
    for name in ("chg_rec", "chg_pos1", "chg_lt", "chg_le", "chg_rt"):
        np.testing.assert_array_equal(getattr(res, name)[:k], getattr(exp, name)[:k], err_msg=name)
    np.testing.assert_array_equal(res.chg_off[:k + 1], exp.chg_off[:k + 1])
    np.testing.assert_array_equal(res.chg_bytes[:exp.n_changed_bytes], exp.chg_bytes[:exp.n_changed_bytes])
    assert list(res.counters) == list(exp.counters), (res.counters, exp.counters)
    # the notes are exactly the records whose status the sv byte and the changed list do not imply, ascending
    need = np.nonzero(((exp.status[:n_rec] & 0x04) != 0) | (((exp.status[:n_rec] & 0x08) != 0) & ((exp.status[:n_rec] & 0x10) == 0)))[0]
    np.testing.assert_array_equal(res.note_rec[:res.n_notes], need.astype(np.uint32))


def _run(ctx, oracle, names, seqs, recs, resident=True):
    recs = [r for r in recs if len(r[2]) == 2]
    ctx.load_reference(names, seqs)
    b = Batch.from_records(recs)
    exp = oracle.normalize(seqs, b)
    bi = ctx.bi_pack(b)
    try:
        res = ctx.normalize_bi(bi)
        _assert_bi_equal(res, exp, b.n_rec)
        if resident:
            db = ctx.upload_bi(bi)
            ctx.normalize_resident(db)
            ctx.normalize_resident(db)  # a second run over the same handle must give the same answer
            _assert_bi_equal(ctx.fetch_result_bi(db, bi), exp, b.n_rec)
            ctx.free(db)
    finally:
        bi.close()
    return res, exp, b


def test_bi_c1_golden_fixture(ctx, oracle):
    seq = helpers.chr20_reference()
    _, recs = helpers.read_vcf_sites(os.path.join(helpers.GOLDEN, "c1_normalize_in.vcf"), {"20": 0})
    res, exp, b = _run(ctx, oracle, ["20"], [seq], [(r[0], r[1], r[2]) for r in recs])
    assert res.counters[4] == 80 and res.counters[10] == 194 and sum(res.counters[:10]) == 80


@pytest.mark.parametrize("seed,lower,iupac", [(31, False, False), (32, True, False), (33, False, True), (34, True, True)])
def test_bi_synthetic_mixed(ctx, oracle, seed, lower, iupac):
    rng = np.random.default_rng(seed)
    names, seqs = helpers.synth_reference(rng, n_contig=4, length=20000, lower=lower, iupac=iupac)
    recs = helpers.synth_records(rng, seqs, 60000, multi_frac=0.0)
    res, exp, b = _run(ctx, oracle, names, seqs, recs)
    assert exp.n_changed > 1000 and res.n_notes > 0  # unknown contigs (rid -1) and contig-start records are in the mix


def test_bi_sorted_runs_large(ctx, oracle):
    """sorted by (contig, pos) like a real call set: one contig run per contig, tiles that straddle a run boundary"""
    rng = np.random.default_rng(41)
    names, seqs = helpers.synth_reference(rng, n_contig=5, length=150000, n_runs=False)
    recs = helpers.synth_records(rng, seqs, 300000, multi_frac=0.0, bad_ref_frac=0.0)
    recs = [r for r in recs if r[0] >= 0]
    recs.sort(key=lambda r: (r[0], r[1]))
    res, exp, b = _run(ctx, oracle, names, seqs, recs)
    k = res.n_changed
    assert np.all(np.diff(res.chg_rec[:k].astype(np.int64)) > 0)


def test_bi_empty_tiny_and_contig_start(ctx, oracle):
    names, seqs = ["c2"], [np.frombuffer(b"AAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAGT", np.uint8)]
    _run(ctx, oracle, names, seqs, [])
    _run(ctx, oracle, names, seqs, [(0, 3, [b"G", b"T"])])
    recs = [(0, 3, [b"AA", b"A"]), (0, 1, [b"aa", b"a"]), (0, 70, [b"A", b"AAAA"]), (0, 80, [b"AG", b"A"]), (-1, 5, [b"AT", b"A"]), (0, 81, [b"G", b"."]),
            (0, 81, [b"G", b"<DEL>"]), (0, 82, [b"T", b"T"])]
    res, exp, b = _run(ctx, oracle, names, seqs, recs)
    status, _ = res.expand()
    assert status[1] & 0x08 and not (status[1] & 0x10) and status[4] & 0x04
    assert int(res.chg_pos1[0]) == 1


def test_bi_long_alleles_and_shifts(ctx, oracle):
    """alleles of 255 bases and more travel through the ext table; 30 kb left shifts; 9 kb alleles"""
    rng = np.random.default_rng(5)
    unit = np.frombuffer(b"ACG", np.uint8)
    tract = np.resize(unit, 30000)
    flank = rng.choice(np.frombuffer(b"ACGT", np.uint8), size=500)
    seq = np.concatenate([flank, tract, flank])
    recs = []
    for k in range(200):
        p = 500 + 29000 + int(rng.integers(0, 900))
        L = 3 * int(rng.integers(1, 120))
        recs.append((0, p, [bytes(seq[p - 1:p - 1 + L + 1]), bytes(seq[p - 1:p])]))
        recs.append((0, p, [bytes(seq[p - 1:p]), bytes(seq[p - 1:p]) + bytes(seq[p:p + L])]))
    big = bytes(seq[100:100 + 5000])
    recs.append((0, 101, [big, big[:1]]))
    recs.append((0, 101, [big, big[:2500] + b"T" + big[2501:]]))
    recs.append((0, 600, [bytes(seq[599:599 + 9001]), bytes(seq[599:600])]))
    for L in (253, 254, 255, 256):  # around the escape value
        recs.append((0, 700, [bytes(seq[699:699 + L]), bytes(seq[699:700])]))
        recs.append((0, 700, [bytes(seq[699:700]), bytes(seq[699:699 + L])]))
    res, exp, b = _run(ctx, oracle, ["c1"], [seq], recs)
    assert int(res.chg_le.max()) > 20000


def test_bi_many_runs_unsorted(ctx, oracle):
    """contig ids that change from record to record: n_run ~ n_rec"""
    rng = np.random.default_rng(43)
    names, seqs = helpers.synth_reference(rng, n_contig=7, length=8000)
    recs = helpers.synth_records(rng, seqs, 20000, multi_frac=0.0)
    _run(ctx, oracle, names, seqs, recs)


def test_bi_chunked_host_call(oracle):
    from vt_b200 import Context
    rng = np.random.default_rng(77)
    names, seqs = helpers.synth_reference(rng, n_contig=2, length=30000)
    recs = [r for r in helpers.synth_records(rng, seqs, 20000, multi_frac=0.0) if len(r[2]) == 2]
    b = Batch.from_records(recs)
    os.environ["VTN_CHUNK_REC"] = "1024"
    try:
        c2 = Context(0)
    finally:
        del os.environ["VTN_CHUNK_REC"]
    try:
        c2.load_reference(names, seqs)
        bi = c2.bi_pack(b)
        res = c2.normalize_bi(bi)
        exp = oracle.normalize(seqs, b)
        _assert_bi_equal(res, exp, b.n_rec)
        bi.close()
    finally:
        c2.close()


def test_bi_rejects_multiallelic(ctx):
    from vt_b200.api import VtnError
    b = Batch.from_records([(0, 5, [b"A", b"C", b"G"])])
    with pytest.raises(VtnError):
        ctx.bi_pack(b)
