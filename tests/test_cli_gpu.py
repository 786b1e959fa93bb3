"""GPU: the command-line shim vt_b200/vtn (vt's `normalize` / `decompose` surface over the C ABI) against
 (a) the reference's own golden fixtures, byte for byte: VCF output and the stripped stderr (test/test.sh:4-7 strips
     `file.*` and `Time elapsed.*`), configuration C1;
 (b) the UNMODIFIED reference binary oracle/_ref/vt_ref (where it was built) on synthetic VCFs with INFO / FORMAT
     payloads, incl. `decompose -s`, `normalize -n` and window-limited ordering."""
import os
import re
import subprocess

import numpy as np
import pytest

import helpers

pytestmark = pytest.mark.gpu
VTN = os.path.join(helpers.ROOT, "vt_b200", "vtn")


def _strip(err):
    out = []
    for line in err.splitlines():
        line = re.sub(r"file.*", "file <stripped>", line)
        line = re.sub(r"Time elapsed.*", "Time elapsed <stripped>", line)
        out.append(line)
    return "\n".join(out).strip()


def _run(tool, args):
    p = subprocess.run([tool] + args, capture_output=True, text=True)
    return p.returncode, p.stderr


@pytest.fixture(scope="module")
def chr20_fasta(tmp_path_factory):
    d = tmp_path_factory.mktemp("chr20")
    fa = str(d / "20.fa")
    helpers.write_fasta(fa, ["20"], [helpers.chr20_reference()])
    return fa


def test_c1_normalize_golden_bytes(chr20_fasta, tmp_path):
    out = str(tmp_path / "out.vcf")
    rc, err = _run(VTN, ["normalize", os.path.join(helpers.GOLDEN, "c1_normalize_in.vcf"), "-r", chr20_fasta, "-o", out])
    assert rc == 0, err
    assert open(out).read() == open(os.path.join(helpers.GOLDEN, "c1_normalize_out.vcf")).read()
    assert _strip(err) == _strip(open(os.path.join(helpers.GOLDEN, "c1_normalize_out.stderr")).read())


def test_c1_decompose_golden_bytes(tmp_path):
    out = str(tmp_path / "out.vcf")
    rc, err = _run(VTN, ["decompose", os.path.join(helpers.GOLDEN, "c1_decompose_in.vcf"), "-o", out])
    assert rc == 0, err
    assert open(out).read() == open(os.path.join(helpers.GOLDEN, "c1_decompose_out.vcf")).read()
    assert _strip(err) == _strip(open(os.path.join(helpers.GOLDEN, "c1_decompose_out.stderr")).read())


HDR = """##fileformat=VCFv4.2
##FILTER=<ID=PASS,Description="All filters passed">
%s##INFO=<ID=AC,Number=A,Type=Integer,Description="x">
##INFO=<ID=AF,Number=A,Type=Float,Description="x">
##INFO=<ID=AN,Number=1,Type=Integer,Description="x">
##INFO=<ID=ADR,Number=R,Type=Integer,Description="x">
##INFO=<ID=GLX,Number=G,Type=Float,Description="x">
##INFO=<ID=DB,Number=0,Type=Flag,Description="x">
##INFO=<ID=NM,Number=.,Type=String,Description="x">
##FORMAT=<ID=GT,Number=1,Type=String,Description="x">
##FORMAT=<ID=AD,Number=R,Type=Integer,Description="x">
##FORMAT=<ID=DP,Number=1,Type=Integer,Description="x">
##FORMAT=<ID=GQ,Number=1,Type=Integer,Description="x">
##FORMAT=<ID=PL,Number=G,Type=Integer,Description="x">
##FORMAT=<ID=GL,Number=G,Type=Float,Description="x">
#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\tFORMAT\tS1\tS2\tS3
"""


def _synthetic_vcf(path, rng, names, seqs, n, multi):
    """records with REF from the reference (so that normalize's REF check passes), INFO/FORMAT payloads of every
    Number kind, floats that need re-serialisation"""
    with open(path, "w") as f:
        f.write(HDR % "".join("##contig=<ID=%s,length=%d>\n" % (nm, s.shape[0]) for nm, s in zip(names, seqs)))
        for c, s in enumerate(seqs):
            pos = np.sort(rng.choice(np.arange(50, s.shape[0] - 60), size=n, replace=False))
            for p in pos:
                p = int(p)
                n_alt = int(rng.integers(2, 5)) if multi else 1
                rl = int(rng.integers(1, 5))
                ref = bytes(s[p:p + rl]).decode()
                alts = []
                while len(alts) < n_alt:
                    k = int(rng.integers(0, 3))
                    if k == 0:
                        a = "ACGT"[int(rng.integers(0, 4))] + ref[1:]
                    elif k == 1:
                        a = ref[:1] + bytes(s[p + 1:p + 1 + int(rng.integers(1, 4))]).decode() + ref[1:]
                    else:
                        a = ref[:max(1, rl - int(rng.integers(1, 3)))]
                    if a != ref and a not in alts:
                        alts.append(a)
                na = n_alt + 1
                ng = na * (na + 1) // 2
                info = "AC=%s;AF=%s;AN=6;ADR=%s;GLX=%s;NM=a,b%s" % (
                    ",".join(str(int(x)) for x in rng.integers(0, 7, n_alt)),
                    ",".join("%.7f" % x for x in rng.random(n_alt)),
                    ",".join(str(int(x)) for x in rng.integers(0, 90, na)),
                    ",".join("%.4f" % (-x * 30) for x in rng.random(ng)),
                    ";DB" if rng.random() < 0.3 else "")
                samples = []
                for _ in range(3):
                    g = sorted(int(x) for x in rng.integers(0, na, 2))
                    sep = "|" if rng.random() < 0.3 else "/"
                    gt = "./." if rng.random() < 0.1 else "%d%s%d" % (g[0], sep, g[1])
                    samples.append("%s:%s:%d:%d:%s:%s" % (
                        gt, ",".join(str(int(x)) for x in rng.integers(0, 40, na)), int(rng.integers(1, 90)), int(rng.integers(0, 99)),
                        ",".join(str(int(x)) for x in rng.integers(0, 255, ng)), ",".join("%.3f" % (-x * 20) for x in rng.random(ng))))
                qual = "%.2f" % (rng.random() * 3000) if rng.random() < 0.8 else "."
                f.write("%s\t%d\t.\t%s\t%s\t%s\tPASS\t%s\tGT:AD:DP:GQ:PL:GL\t%s\n" % (names[c], p + 1, ref, ",".join(alts), qual, info, "\t".join(samples)))


needs_ref = pytest.mark.skipif(not os.path.exists(helpers.VT_REF), reason="reference binary only exists where oracle/_ref was built")


@needs_ref
@pytest.mark.parametrize("smart", [False, True])
def test_decompose_matches_reference_binary(tmp_path, smart):
    rng = np.random.default_rng(31 + smart)
    names, seqs = helpers.synth_reference(rng, n_contig=2, length=6000)
    vcf = str(tmp_path / "in.vcf")
    _synthetic_vcf(vcf, rng, names, seqs, 400, multi=True)
    flags = ["-s"] if smart else []
    o1, o2 = str(tmp_path / "a.vcf"), str(tmp_path / "b.vcf")
    rc1, e1 = _run(VTN, ["decompose", vcf, "-o", o1] + flags)
    rc2, e2 = _run(helpers.VT_REF, ["decompose", vcf, "-o", o2] + flags)
    assert rc1 == 0 and rc2 == 0, (e1, e2)
    assert open(o1).read() == open(o2).read()
    assert _strip(e1) == _strip(e2)


@needs_ref
@pytest.mark.parametrize("flags", [[], ["-w", "50"], ["-n"]])
def test_normalize_matches_reference_binary(tmp_path, flags):
    rng = np.random.default_rng(41 + len(flags))
    names, seqs = helpers.synth_reference(rng, n_contig=2, length=8000, n_runs=False)
    fa = str(tmp_path / "ref.fa")
    helpers.write_fasta(fa, names, seqs)
    vcf = str(tmp_path / "in.vcf")
    _synthetic_vcf(vcf, rng, names, seqs, 600, multi=False)
    o1, o2 = str(tmp_path / "a.vcf"), str(tmp_path / "b.vcf")
    rc1, e1 = _run(VTN, ["normalize", vcf, "-r", fa, "-o", o1] + flags)
    rc2, e2 = _run(helpers.VT_REF, ["normalize", vcf, "-r", fa, "-o", o2] + flags)
    assert rc1 == rc2 == 0, (e1[-2000:], e2[-2000:])
    assert open(o1).read() == open(o2).read()
    keep = lambda e: "\n".join(l for l in _strip(e).splitlines() if "Might not be sorted" not in l and not l.startswith("["))
    assert keep(e1) == keep(e2)


def _msgs(err):
    """bracketed messages with the source path of the reference build stripped: [/x/y/variant_manip.cpp:96 f] -> [variant_manip.cpp:96 f]"""
    out = []
    for l in err.splitlines():
        if l.startswith("[") or l.startswith("FAQ"):
            out.append(re.sub(r"^\[[^\] ]*/", "[", l))
    return out


@needs_ref
@pytest.mark.parametrize("flags", [[], ["-m"], ["-n"]])
def test_strictness_messages_and_exit_codes_match_reference(tmp_path, flags):
    """soft-masked and N reference bases under REF: same messages, same exit code, same output"""
    fa = str(tmp_path / "lc.fa")
    open(fa, "w").write(">c1\nACGTACGTacgtACGTACGTAAAAAAAAAACCCCCGGGGGTTTTTACGTNNNNACGTACGT\n")
    vcf = str(tmp_path / "lc.vcf")
    open(vcf, "w").write("##fileformat=VCFv4.2\n##contig=<ID=c1,length=60>\n#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\n"
                         "c1\t3\t.\tGT\tG\t.\t.\t.\nc1\t9\t.\tAC\tA\t.\t.\t.\nc1\t22\t.\taa\ta\t.\t.\t.\nc1\t48\t.\tTN\tT\t.\t.\t.\nc1\t55\t.\tCG\tC\t.\t.\t.\n")
    o1, o2 = str(tmp_path / "a.vcf"), str(tmp_path / "b.vcf")
    p1 = subprocess.run([VTN, "normalize", vcf, "-r", fa, "-o", o1] + flags, capture_output=True, text=True)
    p2 = subprocess.run([helpers.VT_REF, "normalize", vcf, "-r", fa, "-o", o2] + flags, capture_output=True, text=True)
    assert p1.returncode == p2.returncode
    assert _msgs(p1.stderr) == _msgs(p2.stderr)
    if p2.returncode == 0 and "-n" in flags:
        assert open(o1).read() == open(o2).read()


@needs_ref
def test_unknown_contig_message_matches_reference(tmp_path):
    fa = str(tmp_path / "lc.fa")
    open(fa, "w").write(">c1\nACGTACGTACGTACGTACGTAAAAAAAAAACCCCCGGGGGTTTTTACGT\n")
    vcf = str(tmp_path / "uc.vcf")
    open(vcf, "w").write("##fileformat=VCFv4.2\n##contig=<ID=zz,length=60>\n#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\nzz\t5\t.\tAC\tA\t.\t.\t.\n")
    p1 = subprocess.run([VTN, "normalize", vcf, "-r", fa, "-o", str(tmp_path / "a.vcf")], capture_output=True, text=True)
    p2 = subprocess.run([helpers.VT_REF, "normalize", vcf, "-r", fa, "-o", str(tmp_path / "b.vcf")], capture_output=True, text=True)
    assert p1.returncode == p2.returncode == 1
    assert _msgs(p1.stderr) == _msgs(p2.stderr)


@needs_ref
def test_header_pass_filter_placement_matches_reference(tmp_path):
    """htslib registers FILTER PASS right after the first header line and drops a later copy (vcf.c:809-829)"""
    vcf = str(tmp_path / "h.vcf")
    open(vcf, "w").write("##fileformat=VCFv4.2\n##contig=<ID=c1,length=60>\n##FILTER=<ID=q10,Description=\"x\">\n"
                         "##FILTER=<ID=PASS,Description=\"All filters passed\">\n#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\n"
                         "c1\t3\t.\tGT\tG,GA\t.\t.\t.\nc1\t9\t.\tA\tC\t.\tq10\t.\n")
    o1, o2 = str(tmp_path / "a.vcf"), str(tmp_path / "b.vcf")
    rc1, e1 = _run(VTN, ["decompose", vcf, "-o", o1])
    rc2, e2 = _run(helpers.VT_REF, ["decompose", vcf, "-o", o2])
    assert rc1 == rc2 == 0, (e1[-1000:], e2[-1000:])
    assert open(o1).read() == open(o2).read()


# ---- vt uniq ----
@pytest.mark.parametrize("name", ["uniq_a", "uniq_b"])
def test_uniq_cli_reference_golden_bytes(name, tmp_path):
    """vtn uniq against the output of the unmodified reference binary (tests/golden/make_uniq_golden.py), byte for byte"""
    out = str(tmp_path / "out.vcf")
    rc, err = _run(VTN, ["uniq", os.path.join(helpers.GOLDEN, name + "_in.vcf"), "-o", out])
    assert rc == 0, err
    assert open(out).read() == open(os.path.join(helpers.GOLDEN, name + "_out.vcf")).read()
    exp = open(os.path.join(helpers.GOLDEN, name + "_out.stderr")).read()
    assert _strip(err) == _strip(exp)


@pytest.mark.parametrize("name", ["uniq_a", "uniq_b"])
@pytest.mark.parametrize("batch", ["1", "7", "50"])
def test_uniq_cli_in_batches(name, batch, tmp_path):
    """the file in batches that end between position groups (a group larger than the batch stays whole): same bytes, same counts"""
    out = str(tmp_path / "out.vcf")
    p = subprocess.run([VTN, "uniq", os.path.join(helpers.GOLDEN, name + "_in.vcf"), "-o", out], capture_output=True, text=True,
                       env=dict(os.environ, VTN_CLI_BATCH=batch))
    assert p.returncode == 0, p.stderr
    assert open(out).read() == open(os.path.join(helpers.GOLDEN, name + "_out.vcf")).read()
    assert _strip(p.stderr) == _strip(open(os.path.join(helpers.GOLDEN, name + "_out.stderr")).read())


def test_uniq_cli_not_in_order_across_batches(tmp_path):
    import uniq_helpers as U
    vin = str(tmp_path / "in.vcf")
    U.write_uniq_vcf(vin, [(0, 10, ["A", "C"], None), (0, 10, ["A", "G"], None), (0, 12, ["A", "C"], None), (0, 11, ["A", "C"], None)])
    p = subprocess.run([VTN, "uniq", vin, "-o", str(tmp_path / "o.vcf")], capture_output=True, text=True, env=dict(os.environ, VTN_CLI_BATCH="3"))
    assert p.returncode == 1 and "VCF file not in order" in p.stderr


def test_uniq_cli_not_in_order(tmp_path):
    import uniq_helpers as U
    vin = str(tmp_path / "in.vcf")
    U.write_uniq_vcf(vin, [(0, 10, ["A", "C"], None), (0, 9, ["A", "C"], None)])
    rc, err = _run(VTN, ["uniq", vin, "-o", str(tmp_path / "o.vcf")])
    assert rc == 1 and "VCF file not in order" in err


@needs_ref
@pytest.mark.parametrize("window,batch", [(50, 64), (3, 37), (10, 1000000), (10000, 100)])
def test_normalize_device_writer_order_streaming(tmp_path, window, batch):
    """the writer's order comes from the device (vtn_writer_order), batch by batch with carried-over records: same bytes and the
    same `Might not be sorted` warnings, in the same order, as the reference's sequential BCFOrderedWriter"""
    rng = np.random.default_rng(77)
    names, seqs = helpers.synth_reference(rng, n_contig=2, length=8000, n_runs=False)
    fa = str(tmp_path / "ref.fa")
    helpers.write_fasta(fa, names, seqs)
    vcf = str(tmp_path / "in.vcf")
    _synthetic_vcf(vcf, rng, names, seqs, 900, multi=False)
    o1, o2 = str(tmp_path / "a.vcf"), str(tmp_path / "b.vcf")
    p1 = subprocess.run([VTN, "normalize", vcf, "-r", fa, "-o", o1, "-w", str(window)], capture_output=True, text=True,
                        env=dict(os.environ, VTN_CLI_BATCH=str(batch)))
    p2 = subprocess.run([helpers.VT_REF, "normalize", vcf, "-r", fa, "-o", o2, "-w", str(window)], capture_output=True, text=True)
    assert p1.returncode == p2.returncode == 0, (p1.stderr[-2000:], p2.stderr[-2000:])
    assert open(o1).read() == open(o2).read()
    assert _msgs(p1.stderr) == _msgs(p2.stderr)


# ---- vt decompose_blocksub ----
import blocksub_helpers as BS  # noqa: E402


@pytest.mark.parametrize("stem,flags,tag", BS.CASES + BS.ALIGN_CASES)
def test_blocksub_cli_reference_golden_bytes(stem, flags, tag, tmp_path):
    """vtn decompose_blocksub [-a] [-m] [-d N] [-p] against the reference's own fixtures (01, 02, 03) and vt_ref outputs, byte for byte"""
    vin, vout, verr = BS.paths(stem, tag)
    out = str(tmp_path / "out.vcf")
    rc, err = _run(VTN, ["decompose_blocksub"] + flags + [vin, "-o", out])
    assert rc == 0, err
    assert open(out).read() == open(vout).read()
    assert _strip(err) == _strip(open(verr).read())


def test_blocksub_cli_small_batches(tmp_path):
    vin, vout, _ = BS.paths("synth", "m")
    out = str(tmp_path / "out.vcf")
    p = subprocess.run([VTN, "decompose_blocksub", "-m", vin, "-o", out], capture_output=True, text=True, env=dict(os.environ, VTN_CLI_BATCH="23"))
    assert p.returncode == 0, p.stderr
    assert open(out).read() == open(vout).read()


@pytest.mark.parametrize("stem", ["unk", "renorm"])
def test_normalize_advice_goldens(tmp_path, stem):
    """made by the unmodified reference (tests/golden/make_advice_golden.py): two SNP-only contigs the FASTA lacks keep their order
    (the writer flushes on the header id, not the FASTA id); input that already carries OLD_VARIANT gets the value replaced in
    place and no second header line"""
    g = lambda x: os.path.join(helpers.GOLDEN, "advice_%s" % x)
    out = str(tmp_path / "out.vcf")
    rc, err = _run(VTN, ["normalize", g(stem + "_in.vcf"), "-r", g("ref.fa"), "-o", out])
    assert rc == 0, err
    assert open(out).read() == open(g(stem + "_out.vcf")).read()
    unpath = lambda e: "\n".join(re.sub(r"^\[[^\] ]*/", "[", l) for l in _strip(e).splitlines())  # the reference build's __FILE__ path
    assert unpath(err) == unpath(open(g(stem + "_out.stderr")).read())


# ---- -f: the filter expression in front of normalize / decompose (filter.cpp; goldens by tests/golden/make_filter_golden.py) ----
def _filter_cases():
    """every third `normalize` expression and all the `decompose` ones: tests/test_tools_cpu.py pins the engine itself on all of them
    (without a GPU); here the point is the plumbing -- VTYPE from the device, records leaving the block, counters, option lines"""
    p = os.path.join(helpers.GOLDEN, "filter_cases.tsv")
    rows = [tuple(l.rstrip("\n").split("\t")) for l in open(p)] if os.path.exists(p) else []
    return [r for r in rows if r[1] != "normalize" or int(r[0]) % 3 == 0 or r[2] in ("QUAL<=30", "FILTER.nope")]


@pytest.mark.parametrize("k,tool,exp", _filter_cases())
def test_filter_expression_goldens(k, tool, exp, tmp_path):
    """vtn normalize / decompose -f <exp> byte for byte against the unmodified reference, including its evaluation quirks (&& binds looser than
    ||, a <= between QUAL and an integer keeps its first answer, an undeclared FILTER id is true on every record that has a filter)"""
    vin, fa = os.path.join(helpers.GOLDEN, "filter_in.vcf"), os.path.join(helpers.GOLDEN, "advice_ref.fa")
    out = str(tmp_path / "out.vcf")
    args = tool.split() + ["-f", exp, vin, "-o", out] + (["-r", fa, "-n"] if tool == "normalize" else [])
    rc, err = _run(VTN, args)
    assert rc == 0, err
    assert open(out).read() == open(os.path.join(helpers.GOLDEN, "filter_%s_out.vcf" % k)).read()
    assert _strip(err) == _strip(open(os.path.join(helpers.GOLDEN, "filter_%s_out.stderr" % k)).read())


@pytest.mark.parametrize("exp", ["INFO.K>3", "VTYPE==REF", "QUAL+1>3", "VTYPE==(SNP|INDEL)", "DLEN>0"])
def test_filter_expression_outside_the_subset_fails_loudly(exp, tmp_path):
    vin, fa = os.path.join(helpers.GOLDEN, "filter_in.vcf"), os.path.join(helpers.GOLDEN, "advice_ref.fa")
    rc, err = _run(VTN, ["normalize", "-f", exp, vin, "-o", str(tmp_path / "o.vcf"), "-r", fa, "-n"])
    assert rc != 0 and "not handled" in err


# ---- compressed files either side: .vcf.gz in (gzip or BGZF) and out (BGZF), chosen like bcf_ordered_writer.cpp:35-83 ----
def _bgzf_blocks(raw):
    """walks the BGZF members of a file: [(total size, uncompressed size)]; raises if the structure is off"""
    out, p = [], 0
    while p < len(raw):
        assert raw[p:p + 4] == b"\x1f\x8b\x08\x04" and raw[p + 12:p + 16] == b"BC\x02\x00", p
        size = raw[p + 16] + 256 * raw[p + 17] + 1
        out.append((size, int.from_bytes(raw[p + size - 4:p + size], "little")))
        p += size
    assert p == len(raw)
    return out


def test_bgzf_output_and_compressed_input(chr20_fasta, tmp_path):
    import gzip
    golden = open(os.path.join(helpers.GOLDEN, "c1_normalize_out.vcf"), "rb").read()
    vin = os.path.join(helpers.GOLDEN, "c1_normalize_in.vcf")
    out = str(tmp_path / "out.vcf.gz")
    rc, err = _run(VTN, ["normalize", vin, "-r", chr20_fasta, "-o", out])
    assert rc == 0, err
    raw = open(out, "rb").read()
    assert gzip.decompress(raw) == golden
    blocks = _bgzf_blocks(raw)
    assert blocks[-1] == (28, 0) and all(u <= 0xff00 for _, u in blocks)  # the end-of-file member; htslib's block size
    # the same input as ordinary gzip and as BGZF (written by this tool: `vtn uniq` passes a sorted file without duplicates through)
    gz_in = str(tmp_path / "in.vcf.gz")
    with gzip.open(gz_in, "wb") as g:
        g.write(open(vin, "rb").read())
    out2 = str(tmp_path / "out2.vcf")
    rc, err = _run(VTN, ["normalize", gz_in, "-r", chr20_fasta, "-o", out2])
    assert rc == 0 and open(out2, "rb").read() == golden, err
    out3 = str(tmp_path / "out3.vcf")
    rc, err = _run(VTN, ["uniq", out, "-o", out3])  # BGZF in
    assert rc == 0, err
    plain = str(tmp_path / "plain.vcf")
    open(plain, "wb").write(golden)
    out4 = str(tmp_path / "out4.vcf")
    rc, err = _run(VTN, ["uniq", plain, "-o", out4])
    assert rc == 0 and open(out3, "rb").read() == open(out4, "rb").read()
    if os.path.exists(helpers.VT_REF):  # the unmodified reference reads what this tool wrote
        out5 = str(tmp_path / "view.vcf")
        p = subprocess.run([helpers.VT_REF, "view", out, "-o", out5], capture_output=True, text=True)
        assert p.returncode == 0, p.stderr
        assert [l for l in open(out5) if not l.startswith("##")] == [l for l in golden.decode().splitlines(True) if not l.startswith("##")]


def test_output_name_decides_the_format(tmp_path):
    vin = os.path.join(helpers.GOLDEN, "blocksub_01_in.vcf")
    rc, err = _run(VTN, ["decompose_blocksub", vin, "-o", str(tmp_path / "out.txt")])
    assert rc != 0 and "Not a VCF/BCF file" in err


def test_bgzf_large_multiblock_roundtrip(tmp_path):
    """a file of several hundred blocks: written compressed by one tool, read back block-parallel by the next, equal to the plain run"""
    import gzip
    lines = ["##fileformat=VCFv4.2", "##contig=<ID=1,length=100000000>", "##INFO=<ID=K,Number=1,Type=Integer,Description=\"k\">",
             "#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO"]
    rng = np.random.default_rng(5)
    for k in range(400000):
        ref = "ACGT"[k % 4] + "".join(rng.choice(list("ACGT"), 3))
        lines.append("1\t%d\trs%d\t%s\t%s\t.\t.\tK=%d" % (100 + 5 * k, k, ref, ref[::-1] if k % 3 else ref[:1], k))
    vin = str(tmp_path / "in.vcf")
    open(vin, "w").write("\n".join(lines) + "\n")
    a, b, c = str(tmp_path / "a.vcf.gz"), str(tmp_path / "b.vcf"), str(tmp_path / "c.vcf")
    assert _run(VTN, ["decompose_blocksub", vin, "-o", a])[0] == 0
    assert _run(VTN, ["decompose_blocksub", vin, "-o", b])[0] == 0
    assert gzip.decompress(open(a, "rb").read()) == open(b, "rb").read()
    assert len(_bgzf_blocks(open(a, "rb").read())) > 200
    assert _run(VTN, ["uniq", a, "-o", c])[0] == 0
    d = str(tmp_path / "d.vcf")
    assert _run(VTN, ["uniq", b, "-o", d])[0] == 0
    assert open(c, "rb").read() == open(d, "rb").read()


@pytest.mark.parametrize("ext", [".bcf", ".ubcf"])
def test_bcf_in_and_out(chr20_fasta, tmp_path, ext):
    """BCF2 either side: the reference's own BCF of the C1 input goes in, BCF comes out; decoded by this tool (uniq passes a sorted
    duplicate-free file through) and by the unmodified reference, the records equal the golden VCF"""
    golden = [l for l in open(os.path.join(helpers.GOLDEN, "c1_normalize_out.vcf")) if not l.startswith("##")]
    out = str(tmp_path / ("out" + ext))
    rc, err = _run(VTN, ["normalize", os.path.join(helpers.GOLDEN, "bcf_c1_normalize.bcf"), "-r", chr20_fasta, "-o", out])
    assert rc == 0, err
    assert _strip(err) == _strip(open(os.path.join(helpers.GOLDEN, "c1_normalize_out.stderr")).read())
    back = str(tmp_path / "back.vcf")
    rc, err = _run(VTN, ["uniq", out, "-o", back])
    assert rc == 0, err
    plain = str(tmp_path / "plain.vcf")
    assert _run(VTN, ["uniq", os.path.join(helpers.GOLDEN, "c1_normalize_out.vcf"), "-o", plain])[0] == 0
    assert [l for l in open(back) if not l.startswith("##")] == [l for l in open(plain) if not l.startswith("##")]
    if os.path.exists(helpers.VT_REF):
        view = str(tmp_path / "view.vcf")
        p = subprocess.run([helpers.VT_REF, "view", out, "-o", view], capture_output=True, text=True)
        assert p.returncode == 0, p.stderr
        assert [l for l in open(view) if not l.startswith("##")] == golden


def test_bcf_with_samples_through_decompose(tmp_path):
    vin = os.path.join(helpers.GOLDEN, "bcf_c1_decompose.bcf")
    a, b = str(tmp_path / "a.vcf"), str(tmp_path / "b.vcf")
    assert _run(VTN, ["decompose", "-s", vin, "-o", a])[0] == 0
    assert _run(VTN, ["decompose", "-s", os.path.join(helpers.GOLDEN, "c1_decompose_in.vcf"), "-o", b])[0] == 0
    assert [l for l in open(a) if not l.startswith("##")] == [l for l in open(b) if not l.startswith("##")]


# ---- -i / -I (goldens by tests/golden/make_interval_golden.py from the unmodified reference) ----
def _interval_cases():
    p = os.path.join(helpers.GOLDEN, "interval_cases.tsv")
    return [tuple(l.rstrip("\n").split("\t")) for l in open(p)] if os.path.exists(p) else []


@pytest.mark.parametrize("name,vin,extra", _interval_cases())
def test_intervals_goldens(name, vin, extra, tmp_path):
    """plain VCF / no index: the reference's message, everything read; with .tbi / .csi: the records that overlap, interval after interval
    (one linear pass per interval here instead of the index walk), repeated where intervals overlap"""
    out = str(tmp_path / "out.vcf")
    p = subprocess.run([VTN, "decompose"] + extra.split() + [vin, "-o", out], capture_output=True, text=True, cwd=helpers.GOLDEN)
    assert p.returncode == 0, p.stderr
    assert open(out).read() == open(os.path.join(helpers.GOLDEN, "interval_%s_out.vcf" % name)).read()
    assert _strip(p.stderr) == _strip(open(os.path.join(helpers.GOLDEN, "interval_%s_out.stderr" % name)).read())


# ---- the block reader of `vtn normalize`: line ends, block cuts, window 0 ----
def _small_normalize_input(tmp_path, n=700, big_info=0):
    rng = np.random.default_rng(81)
    names, seqs = helpers.synth_reference(rng, n_contig=2, length=9000, n_runs=False)
    fa = str(tmp_path / "ref.fa")
    helpers.write_fasta(fa, names, seqs)
    lines = ["##fileformat=VCFv4.2"] + ["##contig=<ID=%s,length=%d>" % (nm, s.shape[0]) for nm, s in zip(names, seqs)]
    lines += ['##INFO=<ID=S,Number=1,Type=String,Description="s">', '##INFO=<ID=AF,Number=A,Type=Float,Description="f">',
              "#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO"]
    for c, s in enumerate(seqs):
        for p in np.sort(rng.choice(np.arange(60, s.shape[0] - 60), size=n, replace=False)):
            p = int(p)
            ref = bytes(s[p:p + 3]).decode()
            k = int(rng.integers(0, 4))
            alt = ["ACGT"[(("ACGT".index(ref[0])) + 1) % 4] + ref[1:], ref[:1], ref + ref[-1], ref[:2]][k]
            info = [".", "AF=0.12345678", "S=x", "AF=0.5;S=" + "y" * big_info if big_info else "S=z"][int(rng.integers(0, 4))]
            lines.append("%s\t%d\t.\t%s\t%s\t%s\t.\t%s" % (names[c], p + 1, ref, alt, [".", "30", "12.5"][k % 3], info))
    return fa, lines


@pytest.mark.parametrize("variant", ["crlf", "no_final_newline", "line_longer_than_block", "tiny_batches"])
def test_normalize_block_reader_variants(tmp_path, variant):
    """the same records as LF text in one block, and as CRLF text / without the last line end / with a 1.3 MB line at 1 MB blocks /
    in 5-record blocks: identical output"""
    fa, lines = _small_normalize_input(tmp_path, big_info=1300000 if variant == "line_longer_than_block" else 0)
    plain, other = str(tmp_path / "plain.vcf"), str(tmp_path / "other.vcf")
    open(plain, "w").write("\n".join(lines) + "\n")
    env = dict(os.environ)
    if variant == "crlf":
        open(other, "wb").write(("\r\n".join(lines) + "\r\n").encode())
    elif variant == "no_final_newline":
        open(other, "w").write("\n".join(lines))
    elif variant == "line_longer_than_block":
        open(other, "w").write("\n".join(lines) + "\n")
        env["VTN_CLI_BLOCK_MB"] = "1"
    else:
        open(other, "w").write("\n".join(lines) + "\n")
        env["VTN_CLI_BATCH"] = "5"
    o1, o2 = str(tmp_path / "o1.vcf"), str(tmp_path / "o2.vcf")
    p1 = subprocess.run([VTN, "normalize", plain, "-r", fa, "-o", o1], capture_output=True, text=True)
    p2 = subprocess.run([VTN, "normalize", other, "-r", fa, "-o", o2], capture_output=True, text=True, env=env)
    assert p1.returncode == 0 and p2.returncode == 0, (p1.stderr[-1500:], p2.stderr[-1500:])
    assert open(o1, "rb").read() == open(o2, "rb").read()
    assert [l for l in _strip(p1.stderr).splitlines() if "file" not in l] == [l for l in _strip(p2.stderr).splitlines() if "file" not in l]


@needs_ref
@pytest.mark.parametrize("window", ["0", "1"])
def test_normalize_window_zero_and_one(tmp_path, window):
    fa, lines = _small_normalize_input(tmp_path, n=300)
    vin = str(tmp_path / "in.vcf")
    open(vin, "w").write("\n".join(lines) + "\n")
    o1, o2 = str(tmp_path / "o1.vcf"), str(tmp_path / "o2.vcf")
    p1 = subprocess.run([VTN, "normalize", vin, "-r", fa, "-o", o1, "-w", window], capture_output=True, text=True)
    p2 = subprocess.run([helpers.VT_REF, "normalize", vin, "-r", fa, "-o", o2, "-w", window], capture_output=True, text=True)
    assert p1.returncode == p2.returncode == 0, (p1.stderr[-1500:], p2.stderr[-1500:])
    assert open(o1).read() == open(o2).read()
    assert _msgs(p1.stderr) == _msgs(p2.stderr)


def test_tool_entry_points_in_process(chr20_fasta, tmp_path):
    """libvtn_tools.so used as a library (include/vtn_tools.h): vtn_tool_normalize called twice in one process, then vtn_tool_decompose;
    every resource comes back (no quick exit), the outputs equal the goldens"""
    import ctypes as C
    code = r'''
import ctypes as C, os, sys
lib = C.CDLL(sys.argv[1])
def call(fn, args):
    arr = (C.c_char_p * (len(args) + 1))(*[a.encode() for a in args], None)
    f = getattr(lib, fn)
    f.argtypes = [C.c_int, C.POINTER(C.c_char_p)]
    f.restype = C.c_int if fn == "vtn_tool_normalize" else None
    return f(len(args), arr)
for k in (1, 2):
    r = call("vtn_tool_normalize", ["normalize", sys.argv[2], "-r", sys.argv[3], "-o", sys.argv[4] + str(k) + ".vcf"])
    assert r == 1, r   # print == true: the caller prints the elapsed time
call("vtn_tool_decompose", ["decompose", sys.argv[5], "-o", sys.argv[4] + "d.vcf"])
print("done")
'''
    script = str(tmp_path / "drive.py")
    open(script, "w").write(code)
    import sys
    p = subprocess.run([sys.executable, script, os.path.join(helpers.ROOT, "vt_b200", "libvtn_tools.so"), os.path.join(helpers.GOLDEN, "c1_normalize_in.vcf"),
                        chr20_fasta, str(tmp_path / "out"), os.path.join(helpers.GOLDEN, "c1_decompose_in.vcf")], capture_output=True, text=True)
    assert p.returncode == 0 and "done" in p.stdout, p.stderr[-2000:]
    golden = open(os.path.join(helpers.GOLDEN, "c1_normalize_out.vcf")).read()
    assert open(str(tmp_path / "out1.vcf")).read() == golden and open(str(tmp_path / "out2.vcf")).read() == golden
    assert open(str(tmp_path / "outd.vcf")).read() == open(os.path.join(helpers.GOLDEN, "c1_decompose_out.vcf")).read()


def test_normalize_stdin_to_stdout(chr20_fasta):
    vin = open(os.path.join(helpers.GOLDEN, "c1_normalize_in.vcf"), "rb").read()
    p = subprocess.run([VTN, "normalize", "-", "-r", chr20_fasta, "-o", "-", "-q"], input=vin, capture_output=True)
    assert p.returncode == 0, p.stderr[-1000:]
    assert p.stdout == open(os.path.join(helpers.GOLDEN, "c1_normalize_out.vcf"), "rb").read()


# ---- the stress file (tests/golden/make_stress_golden.py): every tool on records that walk the corners of the format ----
@pytest.mark.parametrize("name,args", [("blocksub_a", ["decompose_blocksub", "-a"]), ("blocksub_p", ["decompose_blocksub", "-p"]), ("uniq", ["uniq"])])
def test_tools_on_the_stress_file(name, args, tmp_path):
    """(decompose is not among them: on this file's partial GL vectors and ploidy-3/4 genotypes the reference reads past its arrays and
    prints what it finds there, see make_stress_golden.py)"""
    out = str(tmp_path / "out.vcf")
    p = subprocess.run([VTN] + args + ["stress_in.vcf", "-o", out], capture_output=True, text=True, cwd=helpers.GOLDEN)
    assert p.returncode == 0, p.stderr[-2000:]
    exp = open(os.path.join(helpers.GOLDEN, "stress_%s_out.vcf" % name)).read().splitlines()
    got = open(out).read().splitlines()
    diff = [(g, e) for g, e in zip(got, exp) if g != e]
    assert len(got) == len(exp) and not diff, (len(got), len(exp), diff[:3])
    assert _strip(p.stderr) == _strip(open(os.path.join(helpers.GOLDEN, "stress_%s_out.stderr" % name)).read())


@pytest.mark.parametrize("name,args", [("decompose", ["decompose"]), ("decompose_s", ["decompose", "-s"])])
def test_decompose_on_the_well_formed_stress_file(name, args, tmp_path):
    """600 records with 1 to 6 ALTs, INFO / FORMAT fields of every Number kind and type, missing elements and fields, haploid and diploid
    genotypes (tests/golden/make_decompose_stress_golden.py): byte for byte the unmodified reference's output"""
    out = str(tmp_path / "out.vcf")
    p = subprocess.run([VTN] + args + ["dstress_in.vcf", "-o", out], capture_output=True, text=True, cwd=helpers.GOLDEN)
    assert p.returncode == 0, p.stderr[-2000:]
    exp = open(os.path.join(helpers.GOLDEN, "dstress_%s_out.vcf" % name)).read().splitlines()
    got = open(out).read().splitlines()
    diff = [(g, e) for g, e in zip(got, exp) if g != e]
    assert len(got) == len(exp) and not diff, (len(got), len(exp), len(diff), diff[:2])
    assert _strip(p.stderr) == _strip(open(os.path.join(helpers.GOLDEN, "dstress_%s_out.stderr" % name)).read())
