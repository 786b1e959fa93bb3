"""GPU tests of the lazily resident reference (vtn_ref_open_fasta / vtn_ref_require, include/vtn.h): contigs become resident the first
time a host-buffer call names them, and every result equals the one over the eagerly loaded FASTA and the CPU oracle."""
import os

import numpy as np
import pytest

import helpers
from vt_b200 import Context
from vt_b200.api import Batch

pytestmark = pytest.mark.gpu


def _fasta(tmp_path, seed=3, n_contig=5, length=70000, **kw):
    rng = np.random.default_rng(seed)
    names, seqs = helpers.synth_reference(rng, n_contig=n_contig, length=length, lower=True, iupac=True, **kw)
    p = str(tmp_path / "ref.fa")
    helpers.write_fasta(p, names, seqs)
    return rng, names, seqs, p


def test_lazy_equals_eager_and_oracle(oracle, tmp_path):
    rng, names, seqs, p = _fasta(tmp_path)
    recs = helpers.synth_records(rng, seqs, 40000)
    order = [3, 1, 4]  # contigs 0 and 2 are never touched
    c = Context(0)
    try:
        c.open_fasta(p)
        assert [c.lib.vtn_ref_contig_len(c.h, k) for k in range(len(names))] == [len(s) for s in seqs]
        for k in order:
            part = [r for r in recs if r[0] == k or r[0] < 0]
            b = Batch.from_records(part)
            helpers.assert_norm_equal(c.normalize(b), oracle.normalize(seqs, b), b.n_rec)
        # contig 0 through the fetch call, contig 2 through the compact layout
        assert c.fetch(0, 100, 5000) == bytes(seqs[0][100:5001])
        assert c.fetch(0, 100, 5000, upper=True) == bytes(seqs[0][100:5001]).upper()
        part = [r for r in recs if r[0] == 2 and len(r[2]) == 2]
        b = Batch.from_records(part)
        bi = c.bi_pack(b)
        res = c.normalize_bi(bi)
        exp = oracle.normalize(seqs, b)
        status, vtype = res.expand()
        np.testing.assert_array_equal(status[:b.n_rec], exp.status[:b.n_rec])
        assert res.n_changed == exp.n_changed
        bi.close()
        # everything resident now: one batch over all contigs, and the eager context gives the same bytes
        b = Batch.from_records(recs)
        got = c.normalize(b)
        helpers.assert_norm_equal(got, oracle.normalize(seqs, b), b.n_rec)
        c.require()
        e = Context(0)
        try:
            e.load_fasta(p)
            helpers.assert_norm_equal(e.normalize(b), got, b.n_rec)
            for k in range(len(names)):
                assert e.fetch(k, 0, len(seqs[k]) - 1) == c.fetch(k, 0, len(seqs[k]) - 1) == bytes(seqs[k])
        finally:
            e.close()
    finally:
        c.close()


def test_lazy_decompose_normalize_and_resident_handle(oracle, tmp_path):
    rng, names, seqs, p = _fasta(tmp_path, seed=4, n_contig=3, length=40000)
    recs = [r for r in helpers.synth_records(rng, seqs, 20000, multi_frac=0.5) if r[0] >= 0]
    b = Batch.from_records(recs)
    c = Context(0)
    try:
        c.open_fasta(p)
        db = c.upload(b)  # vtn_dbatch_create makes the contigs resident
        c.normalize_resident(db)
        helpers.assert_norm_equal(c.fetch_result(db, b), oracle.normalize(seqs, b), b.n_rec)
        c.free(db)
    finally:
        c.close()
    c = Context(0)
    e = Context(0)
    try:
        c.open_fasta(p)
        e.load_reference(names, seqs)
        gs, ga, gc, gr = c.decompose(b, normalize=True)
        es, ea, ec, er = e.decompose(b, normalize=True)
        np.testing.assert_array_equal(gs, es)
        np.testing.assert_array_equal(ga, ea)
        assert gc == ec
        helpers.assert_norm_equal(gr, er, len(gs))
    finally:
        c.close()
        e.close()


def test_lazy_index_built_when_fai_is_absent_and_short_file(tmp_path):
    from vt_b200.api import VtnError
    rng, names, seqs, p = _fasta(tmp_path, seed=5, n_contig=2, length=5000)
    os.remove(p + ".fai")
    c = Context(0)
    try:
        c.open_fasta(p)
        assert c.fetch(1, 0, 4999) == bytes(seqs[1])
        with open(p + ".fai", "w") as f:  # an index that promises more than the file holds
            f.write("c1\t%d\t4\t60\t61\nc2\t999999\t%d\t60\t61\n" % (len(seqs[0]), os.path.getsize(p) - 100))
        c.open_fasta(p)
        assert c.fetch(0, 0, 10) == bytes(seqs[0][:11])
        with pytest.raises(VtnError):
            c.fetch(1, 0, 10)
    finally:
        c.close()


def test_host_half_first_then_attach_and_prefetch(oracle, tmp_path):
    """vtn_fasta_open / vtn_fasta_prepack without a context, vtn_ref_attach, vtn_ref_prefetch of later contigs (also of one that is never
    used, of one that is already resident, and out of range): every result equals the oracle's"""
    rng, names, seqs, p = _fasta(tmp_path, seed=6, n_contig=5, length=60000)
    recs = [r for r in helpers.synth_records(rng, seqs, 30000) if r[0] >= 0]
    c = Context(0)
    try:
        c.attach_fasta(p, prepack=[names[1], names[1], "nope"])
        c.prefetch(2)
        c.prefetch(2)
        c.prefetch(4)   # never required: joined when the context is closed
        c.prefetch(99)
        for k in (1, 2, 3, 0):
            part = [r for r in recs if r[0] == k]
            b = Batch.from_records(part)
            helpers.assert_norm_equal(c.normalize(b), oracle.normalize(seqs, b), b.n_rec)
            c.prefetch(k)  # already resident: nothing to do
        assert c.fetch(1, 0, len(seqs[1]) - 1) == bytes(seqs[1])
        assert c.fetch(4, 10, 2000) == bytes(seqs[4][10:2001])
    finally:
        c.close()
    from vt_b200.api import VtnError
    c = Context(0)
    try:
        with pytest.raises(VtnError):
            c.attach_fasta(str(tmp_path / "missing.fa"))
    finally:
        c.close()
