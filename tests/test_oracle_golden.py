"""Pins the CPU oracle (oracle/vt_oracle.c) to the reference: its own golden fixtures (committed under
tests/golden/) and, when oracle/_ref/vt_ref (the unmodified reference binary) is present, differential runs."""
import os
import re
import subprocess

import numpy as np
import pytest

import helpers
from vt_b200.api import Batch


def _golden_sites(path):
    out = []
    for line in open(path):
        if line.startswith("#"):
            continue
        f = line.rstrip("\n").split("\t")
        m = re.search(r"OLD_VARIANT=([^;\t]+)", f[7])
        out.append((f[0], int(f[1]), f[3], f[4].split(","), m.group(1) if m else None))
    return out


def test_oracle_reproduces_normalize_golden(oracle):
    seq = helpers.chr20_reference()
    _, recs = helpers.read_vcf_sites(os.path.join(helpers.GOLDEN, "c1_normalize_in.vcf"), {"20": 0})
    b = Batch.from_records([(r[0], r[1], r[2]) for r in recs])
    res = oracle.normalize([seq], b)
    got = helpers.apply_normalize(recs, res, ["20"])
    exp = _golden_sites(os.path.join(helpers.GOLDEN, "c1_normalize_out.vcf"))
    assert got == exp
    # stats block of 01_OUT.stderr
    txt = open(os.path.join(helpers.GOLDEN, "c1_normalize_out.stderr")).read()
    nums = [int(x) for x in re.findall(r":\s+(\d+)\s*$", txt, re.M)]
    c = res.counters
    assert nums == [c[0], c[1], c[2], c[3], c[4], sum(c[0:5]), c[5], c[6], c[7], c[8], c[9], sum(c[5:10]), sum(c[0:10]), c[10], c[11]]
    assert c[4] == 80 and c[10] == 194


def test_oracle_reproduces_decompose_golden(oracle):
    _, recs = helpers.read_vcf_sites(os.path.join(helpers.GOLDEN, "c1_decompose_in.vcf"), {})
    b = Batch.from_records([(0, r[1], r[2]) for r in recs])
    src, alt, cnt = oracle.decompose_plan(b)
    exp = _golden_sites(os.path.join(helpers.GOLDEN, "c1_decompose_out.vcf"))
    got = [(recs[s][3][0], recs[s][1], recs[s][2][0], [recs[s][2][a]], None) for s, a in zip(src, alt)]
    assert got == exp
    txt = open(os.path.join(helpers.GOLDEN, "c1_decompose_out.stderr")).read()
    nums = [int(x) for x in re.findall(r":\s+(\d+)\s*$", txt, re.M)]
    assert nums == [cnt[0], cnt[1], cnt[2], cnt[3], cnt[3] + cnt[0]]


@pytest.mark.skipif(not os.path.exists(helpers.VT_REF), reason="reference binary oracle/_ref/vt_ref not built")
@pytest.mark.parametrize("seed,flags", [(1, ["-n"]), (2, ["-n"]), (3, ["-n"])])
def test_oracle_matches_reference_binary(oracle, tmp_path, seed, flags):
    """Differential: synthetic VCF through the UNMODIFIED reference `vt normalize -n` vs the restatement."""
    rng = np.random.default_rng(seed)
    names, seqs = helpers.synth_reference(rng, n_contig=3, length=6000, lower=(seed == 2), iupac=(seed == 3))
    recs = [r for r in helpers.synth_records(rng, seqs, 3000) if r[0] >= 0]
    recs.sort(key=lambda r: (r[0], r[1]))
    fa = str(tmp_path / "ref.fa")
    helpers.write_fasta(fa, names, seqs)
    vcf = str(tmp_path / "in.vcf")
    with open(vcf, "w") as f:
        f.write("##fileformat=VCFv4.2\n")
        for n, s in zip(names, seqs):
            f.write("##contig=<ID=%s,length=%d>\n" % (n, s.shape[0]))
        f.write("#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\n")
        for (c, p, al) in recs:
            f.write("%s\t%d\t.\t%s\t%s\t.\t.\t.\n" % (names[c], p, al[0].decode(), ",".join(a.decode() for a in al[1:])))
    out = str(tmp_path / "out.vcf")
    p = subprocess.run([helpers.VT_REF, "normalize", vcf, "-r", fa, "-o", out, "-w", "100000"] + flags, capture_output=True, text=True)
    assert p.returncode == 0, p.stderr[-2000:]
    exp = _golden_sites(out)
    b = Batch.from_records(recs)
    res = oracle.normalize(seqs, b)
    full = [(c, p, [a.decode() for a in al], [names[c]]) for (c, p, al) in recs]
    got = helpers.apply_normalize(full, res, names)
    assert len(got) == len(exp)
    assert got == exp
    nums = [int(x) for x in re.findall(r":\s+(\d+)\s*$", p.stderr[p.stderr.index("stats:"):], re.M)]
    c = res.counters
    assert nums[:5] == c[0:5] and nums[6:11] == c[5:10] and nums[13] == c[10] and nums[14] == c[11]
