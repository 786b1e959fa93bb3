"""CPU tests of host-side pieces of the tool library that need no GPU: `vtn _floats` (the self-test hook of vtn_tools.cpp) prints, for
every input line, whether canon_float says htslib would print the value as written, and what put_float makes of it.  The pass-through
of untouched records relies on: canon_float(s) implies put_float(s) == s."""
import os
import subprocess

import numpy as np

import helpers

VTN = os.path.join(helpers.ROOT, "vt_b200", "vtn")


def test_canonical_floats_are_fixed_points_of_put_float():
    rng = np.random.default_rng(1)
    vals = []
    for _ in range(200000):
        k = int(rng.integers(0, 6))
        if k == 0:
            v = "%d" % rng.integers(0, 2000000)
        elif k == 1:
            v = "%.*f" % (int(rng.integers(0, 9)), rng.random() * 10 ** int(rng.integers(-5, 7)))
        elif k == 2:
            v = "%.*g" % (int(rng.integers(1, 9)), rng.random() * 10 ** int(rng.integers(-6, 8)))
        elif k == 3:
            v = "-" + "%.*f" % (int(rng.integers(0, 7)), rng.random() * 100)
        elif k == 4:
            v = str(rng.choice([".", "0", "-0", "0.0", "1.", "00", "01", "+1", "1e5", "1E-3", "nan", "inf", "0.0001", "0.00009", "999999", "1000000", "0.1000", "12.50",
                                "-0.5", "5.", "-", "0.25", "0.333333", "0.3333333"]))
        else:
            v = "0." + "0" * int(rng.integers(0, 5)) + "%d" % rng.integers(1, 999999)
        vals.append(v)
    p = subprocess.run([VTN, "_floats"], input="\n".join(vals) + "\n", capture_output=True, text=True)
    assert p.returncode == 0, p.stderr
    out = p.stdout.splitlines()
    assert len(out) == len(vals)
    n_canon = 0
    for v, l in zip(vals, out):
        c, _, t = l.partition("\t")
        if c == "1":
            n_canon += 1
            assert t == v, (v, t)
    assert n_canon > 50000
    # and the expected verdicts on the forms that matter
    exp = {".": 1, "0": 1, "-0": 1, "0.0": 0, "1.": 0, "00": 0, "01": 0, "+1": 0, "1e5": 0, "nan": 0, "0.0001": 1, "0.00009": 0, "999999": 1, "1000000": 0,
           "0.1000": 0, "12.50": 0, "-0.5": 1, "0.25": 1, "0.333333": 1, "0.3333333": 0, "30": 1, "12.5": 1}
    p = subprocess.run([VTN, "_floats"], input="\n".join(exp) + "\n", capture_output=True, text=True)
    got = {k: int(l.split("\t")[0]) for k, l in zip(exp, p.stdout.splitlines())}
    assert got == exp


def _filter_keep(exp, recs, meta):
    p = subprocess.run([VTN, "_filter", exp] + meta, input="".join("%d\t%d\t%s\t%s\n" % r for r in recs), capture_output=True, text=True)
    assert p.returncode == 0, p.stderr
    return [int(x) for x in p.stdout.split()]


def test_filter_expression_goldens_on_the_cpu():
    """the -f engine of the tools (vtn_tools.cpp: Filter, restating filter.cpp's parse and evaluate) without a GPU: the variant types come
    from the CPU oracle instead of the device, and the records it keeps must be the ones the unmodified reference wrote for the same
    expression (tests/golden/filter_*: `vt normalize -f`; the K INFO value identifies a record)"""
    import re
    vin = os.path.join(helpers.GOLDEN, "filter_in.vcf")
    meta = [l.rstrip("\n") for l in open(vin) if l.startswith("##FILTER")]
    fields = [l.rstrip("\n").split("\t") for l in open(vin) if not l.startswith("#")]
    seq = "".join(l.strip() for l in open(os.path.join(helpers.GOLDEN, "advice_ref.fa")) if not l.startswith(">"))
    recs = [(0, int(f[1]), [f[3].encode()] + ([] if f[4] == "." else [a.encode() for a in f[4].split(",")])) for f in fields]
    b = helpers.Batch.from_records(recs)
    vtype = helpers.Oracle().normalize([np.frombuffer(seq.encode(), np.uint8)], b).vtype
    rows = [(int(vtype[r]), len(recs[r][2]), fields[r][5], fields[r][6]) for r in range(len(recs))]
    cases = [l.rstrip("\n").split("\t") for l in open(os.path.join(helpers.GOLDEN, "filter_cases.tsv"))]
    n = 0
    for k, tool, exp in cases:
        if tool != "normalize":
            continue
        kept = {int(re.search(r"K=(\d+)", l).group(1)) for l in open(os.path.join(helpers.GOLDEN, "filter_%s_out.vcf" % k)) if not l.startswith("#")}
        got = _filter_keep(exp, rows, meta)
        assert {r for r in range(len(rows)) if got[r]} == kept, (k, exp)
        n += 1
    assert n >= 25


def test_filter_parse_rules():
    rows = [(1, 2, "50", "PASS"), (4, 3, ".", "q10"), (5, 2, "5", "."), (2, 2, "10.5", "q10;s50")]
    meta = ["##FILTER=<ID=q10,Description=\"q\">", "##FILTER=<ID=s50,Description=\"s\">"]
    # && is the root before || (filter.cpp:1563-1650: the operator with the lowest code splits first)
    assert _filter_keep("VTYPE==SNP||N_ALLELE>2&&QUAL>5", rows, meta) == [1, 0, 0, 0]
    assert _filter_keep("(VTYPE==SNP)||(N_ALLELE>2&&QUAL>5)", rows, meta) == [1, 0, 0, 0]
    # a missing QUAL: comparisons are false, ~ of them is true
    assert _filter_keep("QUAL<100", rows, meta) == [1, 0, 1, 1]
    assert _filter_keep("~(QUAL<100)", rows, meta) == [0, 1, 0, 0]
    # <= between QUAL and an integer keeps the answer of its first evaluation with a QUAL present (filter.cpp:307)
    assert _filter_keep("QUAL<=20", rows, meta) == [0, 0, 0, 0]
    assert _filter_keep("QUAL<=20.0", rows, meta) == [0, 0, 1, 1]
    # an undeclared FILTER id is true wherever a filter is set
    assert _filter_keep("FILTER.nope", rows, meta) == [1, 1, 0, 1]
    assert _filter_keep("FILTER.s50", rows, meta) == [0, 0, 0, 1]
    assert _filter_keep("VTYPE&INDEL", rows, meta) == [0, 1, 1, 0]
    for bad in ("INFO.AC>1", "QUAL+1>2", "VTYPE==REF"):
        p = subprocess.run([VTN, "_filter", bad], input="1\t2\t.\t.\n", capture_output=True, text=True)
        assert p.returncode != 0 and "not handled" in p.stderr


def test_tail_pass_through_rule(tmp_path):
    """tail_is_plain (which records' columns from QUAL on go out as they came in) against put_tail (what would be written otherwise):
    plain must imply identical text; the forms htslib rewrites must not be plain"""
    hdr = str(tmp_path / "h.vcf")
    open(hdr, "w").write('##fileformat=VCFv4.2\n##INFO=<ID=AF,Number=A,Type=Float,Description="f">\n##INFO=<ID=AC,Number=A,Type=Integer,Description="i">\n'
                         '##INFO=<ID=DB,Number=0,Type=Flag,Description="d">\n##FORMAT=<ID=GT,Number=1,Type=String,Description="g">\n'
                         '##FORMAT=<ID=GQ,Number=1,Type=Float,Description="q">\n##FORMAT=<ID=AD,Number=R,Type=Integer,Description="a">\n'
                         "#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\tFORMAT\tS1\tS2\n")
    cases = [
        (".\t.\t.", 1), ("30\tPASS\tAF=0.25;AC=3;DB", 1), ("12.5\tq10\tAF=0.5,0.125", 1), ("30.0\t.\t.", 0), ("1e3\t.\t.", 0), ("30\t.\tAF=0.250", 0),
        ("30\t.\tAF=1e-3", 0), ("30\t.\tAF=0.1234567", 0), ("30\t.\tAC=3;;DB", 0), ("30\t.\t;AC=3", 0), ("30\t.\tAC=3;", 0), ("30\t.\tAF=.", 1),
        ("30\t.\tXX=0.250", 1),  # an undeclared key is a string to htslib: as written
        (".\t.\t.\tGT:GQ:AD\t0/1:30:1,2\t1/1:12.5:3,4", 1), (".\t.\t.\tGT:GQ:AD\t0/1:30.0:1,2\t1/1:12.5:3,4", 0), (".\t.\t.\tGT:GQ:AD\t0/1:30\t1/1:12.5:3,4", 0),
        (".\t.\t.\tGT:GQ:AD\t0/1:.:1,2\t.:.:.", 1), (".\t.\tAF=0.5\tGT\t0/1\t1|1", 1), ("0\t.\t.", 1), ("00\t.\t.", 0), ("-1\t.\t.", 1),
    ]
    p = subprocess.run([VTN, "_plain", hdr], input="".join(c + "\n" for c, _ in cases), capture_output=True, text=True)
    assert p.returncode == 0, p.stderr
    out = p.stdout.splitlines()
    assert len(out) == len(cases)
    for (tail, exp), l in zip(cases, out):
        assert int(l[0]) == exp, (tail, l)
        if l[0] == "1":
            assert l[1:] == "\t" + tail, (tail, l)
    # a random sweep: wherever the rule says plain, put_tail agrees byte for byte
    rng = np.random.default_rng(3)
    fl = ["0.5", "0.25", "0.250", "1e-5", "3", "12.125", ".", "0.1234567", "100000", "0.00001", "-2.5", "7.0"]
    tails = []
    for _ in range(20000):
        info = ";".join(str(rng.choice(["AF=" + ",".join(rng.choice(fl, int(rng.integers(1, 3)))), "AC=%d" % rng.integers(0, 9), "DB", "", "ZZ=1.50"]))
                        for _ in range(int(rng.integers(1, 4)))) or "."
        smp = "" if rng.random() < 0.5 else "\tGT:GQ\t" + "\t".join("0/1:%s" % rng.choice(fl) if rng.random() < 0.9 else "0/1" for _ in range(2))
        tails.append("%s\t%s\t%s%s" % (rng.choice(fl), rng.choice([".", "PASS", "q10;s50"]), info, smp))
    p = subprocess.run([VTN, "_plain", hdr], input="".join(t + "\n" for t in tails), capture_output=True, text=True)
    out = p.stdout.splitlines()
    assert len(out) == len(tails)
    n_plain = 0
    for t, l in zip(tails, out):
        if l[0] == "1":
            n_plain += 1
            assert l[1:] == "\t" + t, (t, l)
    assert 500 < n_plain < 19000


def test_interval_parsing(tmp_path):
    """GenomeInterval::set / to_string and Program::parse_intervals (genome_interval.cpp:50-137, program.cpp:152-190)"""
    lst = str(tmp_path / "l.txt")
    open(lst, "w").write("2:100-400\n1:1000-1300\n2:100-400\nX\n")
    p = subprocess.run([VTN, "_intervals", "1:60-90,Y:5,2:100-400,chrUn_x", lst], capture_output=True, text=True)
    assert p.returncode == 0, p.stderr
    assert p.stdout.splitlines() == ["2:100-400\t100\t400", "1:1000-1300\t1000\t1300", "X\t1\t536870911", "1:60-90\t60\t90", "Y:5-5\t5\t5", "chrUn_x\t1\t536870911"]
    p = subprocess.run([VTN, "_intervals", "1:a-b"], capture_output=True, text=True)
    assert p.returncode != 0 and "Invalid genomic interval" in p.stderr


def test_put_tail_equals_the_reference_on_the_stress_file():
    """the tools' re-serialisation of QUAL / INFO / FORMAT (put_tail: htslib's float format, dropped trailing sample fields padded, ...)
    against what the unmodified reference printed for the same records (tests/golden/bcf_stress_view.vcf = `vt view` of stress_in.vcf)"""
    src = [l.rstrip("\n") for l in open(os.path.join(helpers.GOLDEN, "stress_in.vcf")) if not l.startswith("#")]
    ref = [l.rstrip("\n") for l in open(os.path.join(helpers.GOLDEN, "bcf_stress_view.vcf")) if not l.startswith("#")]
    tails = ["\t".join(l.split("\t")[5:]) for l in src]
    p = subprocess.run([VTN, "_plain", os.path.join(helpers.GOLDEN, "stress_in.vcf")], input="\n".join(tails) + "\n", capture_output=True, text=True)
    assert p.returncode == 0, p.stderr
    out = p.stdout.splitlines()
    assert len(out) == len(src) == len(ref) == 400
    for o, r in zip(out, ref):
        assert o[2:] == "\t".join(r.split("\t")[5:])
