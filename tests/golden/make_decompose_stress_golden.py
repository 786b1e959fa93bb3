#!/usr/bin/env python3
"""A well-formed multi-sample file for decompose / decompose -s: 2 to 6 ALTs, INFO and FORMAT fields of every Number kind (A, R, G, ., fixed)
and type, missing values and whole missing fields, haploid and diploid phased / unphased genotypes with missing alleles, dropped trailing
sample fields -- everything the reference indexes correctly: FORMAT vectors full-length or entirely missing, INFO vectors full-length with missing ELEMENTS
(an entirely missing Number=R Float INFO vector makes the reference's smart mode print values from beyond its one-element array).  Outputs of the UNMODIFIED reference
(oracle/_ref/vt_ref): dstress_decompose_out.vcf, dstress_decompose_s_out.vcf (+ .stderr).
Run in the build container: python tests/golden/make_decompose_stress_golden.py"""
import os
import re
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import helpers  # noqa: E402

rng = np.random.default_rng(31)
H = ['##fileformat=VCFv4.2', '##contig=<ID=1,length=100000000>', '##FILTER=<ID=q10,Description="q">',
     '##INFO=<ID=AC,Number=A,Type=Integer,Description="a">', '##INFO=<ID=AF,Number=A,Type=Float,Description="a">', '##INFO=<ID=AS,Number=A,Type=String,Description="a">',
     '##INFO=<ID=RD,Number=R,Type=Integer,Description="r">', '##INFO=<ID=RF,Number=R,Type=Float,Description="r">', '##INFO=<ID=GC,Number=G,Type=Integer,Description="g">',
     '##INFO=<ID=GF,Number=G,Type=Float,Description="g">', '##INFO=<ID=AN,Number=1,Type=Integer,Description="n">', '##INFO=<ID=VQ,Number=.,Type=Float,Description="v">',
     '##INFO=<ID=DB,Number=0,Type=Flag,Description="d">', '##INFO=<ID=TX,Number=1,Type=String,Description="t">',
     '##FORMAT=<ID=GT,Number=1,Type=String,Description="g">', '##FORMAT=<ID=AD,Number=R,Type=Integer,Description="a">', '##FORMAT=<ID=DP,Number=1,Type=Integer,Description="d">',
     '##FORMAT=<ID=GQ,Number=1,Type=Float,Description="q">', '##FORMAT=<ID=PL,Number=G,Type=Integer,Description="p">', '##FORMAT=<ID=GL,Number=G,Type=Float,Description="l">',
     '##FORMAT=<ID=AO,Number=A,Type=Integer,Description="o">', '##FORMAT=<ID=FT,Number=1,Type=String,Description="t">',
     '#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\tFORMAT\tS1\tS2\tS3\tS4']
L = []
pos = 1000


def vec(n, f, miss=0.1):
    return "." if rng.random() < miss else ",".join(f() for _ in range(n))


for k in range(600):
    pos += int(rng.integers(1, 200))
    n_alt = 1 if k % 5 == 0 else int(rng.integers(2, 7))
    ref = "ACGT"[k % 4] + "T" * (k % 3)
    alts = []
    while len(alts) < n_alt:
        a = str(rng.choice(["A", "C", "G", "T", "AT", "GTT", "ACGTA", "TTTTTT"])) + ("" if rng.random() < 0.7 else "C")
        if a != ref and a not in alts:
            alts.append(a)
    na = n_alt + 1
    ng2 = na * (na + 1) // 2
    I = lambda hi=99: (lambda: str(int(rng.integers(0, hi))))
    F = lambda: str(rng.choice(["0.5", "0.25", "1e-5", "0.123456789", "12", ".", "-3.75"]))
    info = []
    if rng.random() < 0.8: info.append("AC=" + vec(n_alt, I(), 0))
    if rng.random() < 0.8: info.append("AF=" + vec(n_alt, F, 0))
    if rng.random() < 0.4: info.append("AS=" + vec(n_alt, lambda: str(rng.choice(["x", "yy", "."])), 0))
    if rng.random() < 0.6: info.append("RD=" + vec(na, I(), 0))
    if rng.random() < 0.4: info.append("RF=" + vec(na, F, 0))
    if rng.random() < 0.5: info.append("GC=" + vec(ng2, I(), 0))
    if rng.random() < 0.4: info.append("GF=" + vec(ng2, F, 0))
    info.append("AN=%d" % rng.integers(2, 9))
    if rng.random() < 0.3: info.append("VQ=" + vec(int(rng.integers(1, 4)), F, 0))
    if rng.random() < 0.3: info.append("DB")
    if rng.random() < 0.3: info.append("TX=a;b".split(";")[0])
    keys = ["GT"] + [x for x in ["AD", "DP", "GQ", "PL", "GL", "AO", "FT"] if rng.random() < 0.6]
    smps = []
    hap = rng.random() < 0.2  # one ploidy per record: with haploid and diploid samples side by side the reference indexes PL / GL past their ends
    for s in range(4):
        v = []
        for key in keys:
            if key == "GT":
                al = lambda: "." if rng.random() < 0.1 else str(int(rng.integers(0, na)))
                v.append(al() if hap else al() + str(rng.choice(["/", "|"])) + al())
            elif key == "AD": v.append(vec(na, I(300)))
            elif key == "DP": v.append(vec(1, I(500)))
            elif key == "GQ": v.append(vec(1, F))
            elif key == "PL": v.append(vec(na if hap else ng2, I(999)))
            elif key == "GL": v.append(vec(na if hap else ng2, lambda: "%.2f" % (-rng.random() * 40)))
            elif key == "AO": v.append(vec(n_alt, I(200)))
            else: v.append(str(rng.choice([".", "PASS", "lowq"])))
        if rng.random() < 0.1:
            v = v[:max(1, len(v) - 1)]
        smps.append(":".join(v))
    L.append("1\t%d\t%s\t%s\t%s\t%s\t%s\t%s\t%s\t%s" % (pos, "." if k % 4 else "rs%d" % k, ref, ",".join(alts), rng.choice([".", "50", "33.3"]), rng.choice([".", "PASS", "q10"]),
                                                       ";".join(info), ":".join(keys), "\t".join(smps)))
open(os.path.join(HERE, "dstress_in.vcf"), "w").write("\n".join(H + L) + "\n")
for name, args in (("decompose", ["decompose"]), ("decompose_s", ["decompose", "-s"])):
    out = os.path.join(HERE, "dstress_%s_out.vcf" % name)
    p = subprocess.run([helpers.VT_REF] + args + ["dstress_in.vcf", "-o", out], capture_output=True, text=True, cwd=HERE)
    assert p.returncode == 0, p.stderr
    open(out[:-4] + ".stderr", "w").write(re.sub(r"Time elapsed.*", "Time elapsed <stripped>", p.stderr))
    txt = open(out).read()
    print(name, sum(1 for l in txt.splitlines() if not l.startswith("#")), "suspicious denormals:", len(re.findall(r"e-4[0-9]", txt)))
