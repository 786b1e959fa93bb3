"""GPU: vt annotate_indels through the C ABI (vtn_annotate_indels, vtn_ref_fetch_batch) and the command line, against the
UNMODIFIED reference's output for the same input (tests/golden/annot_*, made by oracle/_ref/vt_ref_trace): which records it keeps,
all 34 INFO values of every indel, FLANKSEQ, the exact list of (model, sequence) alignments FlankDetector issues with the
positions it reads back, the VCF bytes and the stderr text.  The region kernel (trim / right_align / left_align on the resident
reference) is also compared record by record with the literal CPU restatement in tests/annot_host_main.cpp."""
import os
import subprocess

import numpy as np
import pytest

import annot_helpers as A
import helpers
import test_annot_host_cpu as H
from test_cli_gpu import VTN, _run, _strip
from vt_b200.api import Batch

pytestmark = pytest.mark.gpu


def _golden_batch(ctx):
    ctx.load_fasta(A.REF_FA)
    recs = A.records(A.IN_VCF)
    return recs, Batch.from_records([(ctx.contig_id(r[0]), int(r[1]), [r[3]] + ([] if r[4] == "." else r[4].split(","))) for r in recs])


def test_annotate_abi_matches_reference(ctx, tmp_path):
    recs, b = _golden_batch(ctx)
    trace = str(tmp_path / "trace.txt")
    res = ctx.annotate_indels(b, trace=trace)
    ref = A.records(A.OUT_VCF)
    kept = [k for k in range(len(recs)) if res["flags"][k] & 1]
    assert len(kept) == len(ref) and len(kept) == len(recs) - 2
    n_indel = 0
    for k, o in zip(kept, ref):
        assert recs[k][:5] == o[:5]
        oi = A.parse_info(o[7])
        if res["flags"][k] & 2:
            v = res["vntr"][int(res["indel_idx"][k])]
            assert v["rec"] == k
            A.check_tags(v, oi)
            assert v["exact_ru_ambiguous"] == ("EXACT_RU_AMBIGUOUS" in oi)
            n_indel += 1
        else:
            assert not any(t.startswith("EX_") for t in oi)
    assert n_indel == 180 == len(res["vntr"])
    assert A.sorted_trace(trace) == A.golden_trace()
    assert res["n_alignments"] == len(A.golden_trace())


def test_region_kernel_matches_cpu_restatement(ctx):
    """flags, vtype and per ALT (left trim, right trim, right_align steps) of vta_region_kernel against the literal std::string loops"""
    H._build()
    out = subprocess.run([H.EXE, A.REF_FA, A.IN_VCF], check=True, capture_output=True, text=True).stdout
    recs, b = _golden_batch(ctx)
    res = ctx.annotate_indels(b)
    for k, l in enumerate(out.rstrip("\n").split("\n")):
        f = l.split("\t")
        assert int(f[1]) == int(res["flags"][k]) and int(f[2]) == int(res["vtype"][k])
        alts = [tuple(int(x) for x in t[4:].split(",")) for t in f if t.startswith("alt:")]
        s0 = int(b.allele_beg[k])
        got = [(int(res["trim_left"][s]), int(res["trim_right"][s]), int(res["shift"][s])) for s in range(s0 + 1, int(b.allele_beg[k + 1]))]
        if alts:
            assert alts == got, (k, alts, got)


def test_fetch_batch_matches_single_fetch(ctx):
    names, seqs = A.read_fasta(A.REF_FA)
    ctx.load_fasta(A.REF_FA)
    rng = np.random.default_rng(3)
    rid = rng.integers(0, len(seqs), 300)
    beg = np.array([int(rng.integers(-20, seqs[c].shape[0] + 10)) for c in rid])
    end = beg + rng.integers(-3, 1200, 300)
    got = ctx.fetch_batch(rid, beg, end, upper=True)
    for k in range(300):
        assert got[k] == ctx.fetch(int(rid[k]), int(beg[k]), int(end[k]), upper=True), k
    assert ctx.fetch_batch([], [], []) == []


def test_annotate_cli_byte_exact(tmp_path):
    out = str(tmp_path / "out.vcf")
    rc, err = _run(VTN, ["annotate_indels", A.IN_VCF, "-r", A.REF_FA, "-o", out])
    assert rc == 0, err
    assert open(out).read() == open(A.OUT_VCF).read()
    assert _strip(err) == _strip(open(A.OUT_ERR).read())


def test_annotate_cli_small_batches(tmp_path):
    """37-record device batches: the lock-step rounds and the ordered writer's carry-over must not change a byte"""
    out = str(tmp_path / "out.vcf")
    env = dict(os.environ, VTN_CLI_BATCH="37")
    p = subprocess.run([VTN, "annotate_indels", A.IN_VCF, "-r", A.REF_FA, "-o", out], capture_output=True, text=True, env=env)
    assert p.returncode == 0, p.stderr
    assert open(out).read() == open(A.OUT_VCF).read()
