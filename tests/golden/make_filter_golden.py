#!/usr/bin/env python3
"""Goldens for -f (the filter expression in front of normalize / decompose), made by the UNMODIFIED reference (oracle/_ref/vt_ref):
one input with every variant type, QUAL present / missing / fractional, FILTER PASS / q10 / q10;s50 / ., bi- and multiallelic records on
tests/golden/advice_ref.fa; one output per expression in tests/golden/filter_cases.tsv (tool, expression) -> filter_<k>_out.{vcf,stderr}.
Expressions the reference itself cannot evaluate are not goldens.  Run in the build container: python tests/golden/make_filter_golden.py"""
import os
import re
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers  # noqa: E402

VT = os.path.join(ROOT, "oracle", "_ref", "vt_ref")
fa = os.path.join(HERE, "advice_ref.fa")
seq = "".join(l.strip() for l in open(fa) if not l.startswith(">"))
rng = np.random.default_rng(17)
H = ("##fileformat=VCFv4.2\n##contig=<ID=1,length=3000>\n##FILTER=<ID=q10,Description=\"q\">\n##FILTER=<ID=s50,Description=\"s\">\n"
     "##INFO=<ID=K,Number=1,Type=Integer,Description=\"k\">\n#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\n")
quals = [".", "5", "10", "10.5", "30", "99.9", "0", "1e3"]
filts = [".", "PASS", "q10", "q10;s50", "s50"]
lines = []
pos = 20
for k in range(260):
    pos += int(rng.integers(3, 12))
    u = k % 13
    L = int(rng.integers(2, 6))
    ref = seq[pos - 1:pos - 1 + L]
    other = lambda c: "ACGT"[("ACGT".index(c) + 1 + int(rng.integers(0, 3))) % 4]
    if u < 4:
        ref = ref[:1]
        alt = other(ref)                                   # SNP
    elif u == 4:
        alt = ref[:1]                                      # deletion
    elif u == 5:
        alt = ref + seq[pos - 1 + L - 1] * 2               # insertion, padded
    elif u == 6:
        alt = "".join(other(c) for c in ref)               # MNP
    elif u == 7:
        alt = other(ref[0]) + ref[1:] + "T"                # SNP + indel
    elif u == 8:
        ref = ref[:1]
        alt = other(ref) + "," + ref + "AC"                # multiallelic SNP / insertion
    elif u == 9:
        alt = ref[:1] + "," + ref[:2] + "," + other(ref[0])  # three ALTs
    elif u == 10:
        ref = ref[:1]
        alt = "<DEL>"
    elif u == 11:
        ref = ref[:1]
        alt = "."
    else:
        alt = ref[:-1] + other(ref[-1])                    # SNP with padding on the left
    lines.append("1\t%d\t.\t%s\t%s\t%s\t%s\tK=%d" % (pos, ref, alt, quals[(k * 5 + k // 7) % len(quals)], filts[(k * 3 + k // 11) % len(filts)], k))
vin = os.path.join(HERE, "filter_in.vcf")
open(vin, "w").write(H + "\n".join(lines) + "\n")

EXPS = [("normalize", "VTYPE==SNP"), ("normalize", "VTYPE==INDEL&&N_ALLELE==2"), ("normalize", "VTYPE&INDEL"), ("normalize", "~(VTYPE==SNP)"),
        ("normalize", "QUAL>10"), ("normalize", "QUAL>=10.5"), ("normalize", "QUAL<30"), ("normalize", "QUAL<=30"), ("normalize", "QUAL==10"),
        ("normalize", "QUAL!=10"), ("normalize", "~(QUAL>10)"), ("normalize", "N_ALLELE>2||QUAL<10"), ("normalize", "N_ALLELE==2&&QUAL>5||VTYPE==MNP"),
        ("normalize", "VTYPE==MNP||N_ALLELE>2&&QUAL>5"), ("normalize", "(VTYPE==SNP||VTYPE==MNP)&&(QUAL>=30||N_ALLELE!=2)"), ("normalize", "PASS"),
        ("normalize", "~PASS"), ("normalize", "FILTER.q10"), ("normalize", "FILTER.q10&&FILTER.s50"), ("normalize", "FILTER.nope"), ("normalize", "N_FILTER==2"),
        ("normalize", "N_FILTER>=1&&PASS"), ("normalize", "VTYPE!=SNP&&VTYPE!=REF"), ("normalize", "(VTYPE&INDEL)==INDEL"), ("normalize", "VTYPE==SV"),
        ("normalize", "VTYPE==REF"), ("normalize", " VTYPE == SNP && QUAL > 5 "), ("normalize", "N_ALLELE>=2"), ("normalize", "N_ALLELE<2.5"),
        ("normalize", "VTYPE==5"), ("normalize", "VTYPE"), ("normalize", "QUAL<=30&&VTYPE==SNP"), ("normalize", "(VTYPE&SNP)&&~(VTYPE&INDEL)"),
        ("decompose", "N_ALLELE>2"), ("decompose", "VTYPE&INDEL&&QUAL>=10"), ("decompose", "~(N_ALLELE==2)"), ("decompose -s", "QUAL>5||PASS")]
cases = []
for k, (tool, exp) in enumerate(EXPS):
    out = os.path.join(HERE, "filter_%02d_out.vcf" % k)
    args = tool.split() + ["-f", exp, vin, "-o", out] + (["-r", fa, "-n"] if tool == "normalize" else [])
    p = subprocess.run([VT] + args, capture_output=True, text=True, cwd=HERE)
    if p.returncode != 0:
        print("reference cannot evaluate:", tool, repr(exp), p.stderr.strip().splitlines()[-1][:120])
        if os.path.exists(out):
            os.remove(out)
        continue
    err = re.sub(r"Time elapsed.*", "Time elapsed <stripped>", p.stderr).replace(HERE + "/", "")
    open(out[:-4] + ".stderr", "w").write(err)
    n = sum(1 for l in open(out) if not l.startswith("#"))
    cases.append((k, tool, exp, n))
    print("%02d %-12s %-55s %d records" % (k, tool, exp, n))
with open(os.path.join(HERE, "filter_cases.tsv"), "w") as f:
    for k, tool, exp, n in cases:
        f.write("%02d\t%s\t%s\n" % (k, tool, exp))
