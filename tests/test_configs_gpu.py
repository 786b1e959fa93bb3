"""GPU: the BASELINE.json configurations at their OWN generators (vt_b200/synth.py, the ones bench.py times), through the command
line against the UNMODIFIED reference binary on the same files, byte for byte (VCF output and stripped stderr):
  C4  synth.make_str_indels on a reference with 30 % tandem repeats (long left shifts), `normalize`
  C3  synth.make_multiallelic with the parity variant of SURVEY 8d: 3 samples GT:AD:DP:GQ:PL (AD Number=R, PL Number=G) and
      INFO AC / AF / AN / AD_R / GLX, `decompose -s` then `normalize`
  C2  synth.make_biallelic, `normalize`, and the compact layout (vtn_normalize_bi) on the same records against the oracle
Sizes: VTN_TEST_CONFIG_RECORDS records (default 200 000; SURVEY's parity variant is 1 M -- set the variable), reference scaled so
that the CPU side finishes in seconds."""
import os

import numpy as np
import pytest

import helpers
from test_cli_gpu import VTN, _run, _strip
from vt_b200 import synth

pytestmark = pytest.mark.gpu
needs_ref = pytest.mark.skipif(not os.path.exists(helpers.VT_REF), reason="reference binary only exists where oracle/_ref was built")
N = int(os.environ.get("VTN_TEST_CONFIG_RECORDS", 200000))


def _write_sites(path, names, seqs, batch, extra_header="", tail=lambda r, na: ".\t.\t."):
    data = batch.data.tobytes()
    with open(path, "w") as f:
        f.write("##fileformat=VCFv4.2\n")
        for n, s in zip(names, seqs):
            f.write("##contig=<ID=%s,length=%d>\n" % (n, s.shape[0]))
        f.write(extra_header)
        ab, ao = batch.allele_beg, batch.allele_off
        out = []
        for r in range(batch.n_rec):
            s0, s1 = int(ab[r]), int(ab[r + 1])
            al = [data[int(ao[s]):int(ao[s + 1])].decode() for s in range(s0, s1)]
            out.append("%s\t%d\t.\t%s\t%s\t%s\n" % (names[batch.rid[r]], batch.pos1[r], al[0], ",".join(al[1:]), tail(r, s1 - s0)))
        f.write("".join(out))


def _both(tmp_path, tool_args, stdin_vcf=None):
    """run vtn and vt_ref with the same arguments; -> (vtn bytes, ref bytes, vtn stderr, ref stderr)"""
    o1, o2 = str(tmp_path / "vtn.vcf"), str(tmp_path / "ref.vcf")
    rc1, e1 = _run(VTN, tool_args + ["-o", o1])
    rc2, e2 = _run(helpers.VT_REF, tool_args + ["-o", o2])
    assert rc1 == 0 and rc2 == 0, (e1[-2000:], e2[-2000:])
    return open(o1).read(), open(o2).read(), e1, e2


@needs_ref
def test_c4_generator_cli_matches_reference(tmp_path):
    tracts = []
    names, seqs = synth.make_reference(seed=synth.SEED_REF_C4, scale=0.002, str_frac=0.3, tracts_out=tracts)
    batch = synth.make_str_indels(seqs, tracts, N, seed=synth.SEED_VCF_C4)
    fa, vcf = str(tmp_path / "ref.fa"), str(tmp_path / "in.vcf")
    helpers.write_fasta(fa, names, seqs)
    _write_sites(vcf, names, seqs, batch, "#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\n")
    a, b, e1, e2 = _both(tmp_path, ["normalize", vcf, "-r", fa])
    assert a == b
    assert _strip(e1) == _strip(e2)
    shifted = sum(1 for l in a.split("\n") if "OLD_VARIANT=" in l)
    assert shifted > N // 2  # the generator puts 90 % of the indels at the right end of a tract


@needs_ref
def test_c2_generator_cli_matches_reference(tmp_path):
    names, seqs = synth.make_reference(scale=0.002)
    batch = synth.make_biallelic(seqs, N, seed=synth.SEED_VCF_C2)
    fa, vcf = str(tmp_path / "ref.fa"), str(tmp_path / "in.vcf")
    helpers.write_fasta(fa, names, seqs)
    _write_sites(vcf, names, seqs, batch, "#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\n")
    a, b, e1, e2 = _both(tmp_path, ["normalize", vcf, "-r", fa])
    assert a == b
    assert _strip(e1) == _strip(e2)


C3_HEADER = ('##INFO=<ID=AC,Number=A,Type=Integer,Description="x">\n##INFO=<ID=AF,Number=A,Type=Float,Description="x">\n'
             '##INFO=<ID=AN,Number=1,Type=Integer,Description="x">\n##INFO=<ID=AD_R,Number=R,Type=Integer,Description="x">\n'
             '##INFO=<ID=GLX,Number=G,Type=Float,Description="x">\n##FORMAT=<ID=GT,Number=1,Type=String,Description="x">\n'
             '##FORMAT=<ID=AD,Number=R,Type=Integer,Description="x">\n##FORMAT=<ID=DP,Number=1,Type=Integer,Description="x">\n'
             '##FORMAT=<ID=GQ,Number=1,Type=Integer,Description="x">\n##FORMAT=<ID=PL,Number=G,Type=Integer,Description="x">\n'
             "#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\tFORMAT\tS1\tS2\tS3\n")


@needs_ref
def test_c3_generator_decompose_s_then_normalize_matches_reference(tmp_path):
    """SURVEY 8d C3 parity variant: all records multiallelic, 3 samples GT:AD:DP:GQ:PL, decompose -s | normalize"""
    names, seqs = synth.make_reference(scale=0.002)
    n = max(N // 4, 1000)  # every record fans out to 2-6 and carries three samples: a quarter of the site count keeps the run short
    batch = synth.make_multiallelic(seqs, n, seed=synth.SEED_VCF_C3)
    rng = np.random.default_rng(1)
    ri = rng.integers(0, 99, size=(n, 40))
    rf = rng.random((n, 24))

    def tail(r, na):
        k = na - 1
        ng = na * (na + 1) // 2
        info = "AC=%s;AF=%s;AN=6;AD_R=%s;GLX=%s" % (",".join(str(x) for x in ri[r, :k]), ",".join("%.6f" % x for x in rf[r, :k]),
                                                  ",".join(str(x) for x in ri[r, 7:7 + na]), ",".join("%.3f" % (-20 * x) for x in rf[r, :min(ng, 24)]) if ng <= 24 else ".")
        smp = []
        for q in range(3):
            g0, g1 = int(ri[r, 20 + 2 * q]) % na, int(ri[r, 21 + 2 * q]) % na
            gt = "./." if ri[r, 30 + q] < 8 else "%d%s%d" % (min(g0, g1), "|" if ri[r, 33 + q] < 30 else "/", max(g0, g1))
            pl = ",".join(str(int(ri[r, (q + j) % 40]) * 2) for j in range(ng))
            smp.append("%s:%s:%d:%d:%s" % (gt, ",".join(str(x) for x in ri[r, 10 + q:10 + q + na]), int(ri[r, 36 + q]) + 1, int(ri[r, 37 + q]), pl))
        return "%s\tPASS\t%s\tGT:AD:DP:GQ:PL\t%s" % ("%.1f" % (30 * rf[r, 23]), info, "\t".join(smp))
    fa, vcf = str(tmp_path / "ref.fa"), str(tmp_path / "in.vcf")
    helpers.write_fasta(fa, names, seqs)
    _write_sites(vcf, names, seqs, batch, C3_HEADER, tail)
    d1, d2 = str(tmp_path / "d_vtn.vcf"), str(tmp_path / "d_ref.vcf")
    rc1, e1 = _run(VTN, ["decompose", "-s", vcf, "-o", d1])
    rc2, e2 = _run(helpers.VT_REF, ["decompose", "-s", vcf, "-o", d2])
    assert rc1 == 0 and rc2 == 0, (e1[-2000:], e2[-2000:])
    assert open(d1).read() == open(d2).read()
    assert _strip(e1) == _strip(e2)
    a, b, e1, e2 = _both(tmp_path, ["normalize", d1, "-r", fa])
    assert a == b
    assert _strip(e1) == _strip(e2)
    assert a.count("OLD_MULTIALLELIC=") > 2 * n and a.count("OLD_VARIANT=") > n // 20


def test_c4_generator_abi_matches_oracle(ctx, oracle):
    """the C4 generator through the C ABI on both layouts against the CPU oracle (every array bit-exact)"""
    import test_bi_gpu as TB
    tracts = []
    names, seqs = synth.make_reference(seed=synth.SEED_REF_C4, scale=0.001, str_frac=0.3, tracts_out=tracts)
    batch = synth.make_str_indels(seqs, tracts, min(N, 100000), seed=synth.SEED_VCF_C4 + 7)
    ctx.load_reference(names, seqs)
    exp = oracle.normalize(seqs, batch)
    helpers.assert_norm_equal(ctx.normalize(batch), exp, batch.n_rec)
    bi = ctx.bi_pack(batch)
    try:
        TB._assert_bi_equal(ctx.normalize_bi(bi), exp, batch.n_rec)
    finally:
        bi.close()
    k = exp.n_changed
    assert k > batch.n_rec // 2 and int(exp.chg_le[:k].max()) > 1000
