#!/usr/bin/env python3
"""BCF2 fixtures made by the UNMODIFIED reference (oracle/_ref/vt_ref view): for a few VCF inputs of tests/golden, the BGZF-compressed BCF
(bcf_<stem>.bcf) and the VCF text the reference prints for it (bcf_<stem>_view.vcf).  tests/test_bcf_cpu.py decodes the former and
expects the latter; the encoder is checked through the decoder and, where the reference binary exists, through `vt view`.
Run in the build container: python tests/golden/make_bcf_golden.py"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import helpers  # noqa: E402

import numpy as np  # noqa: E402


def make_stress(path):
    """a file that walks the corners of the BCF2 typing rules: INFO vectors of up to 21 values (count >= 15 needs the long form), integers
    at every type boundary beside the reserved values, floats in exponent form / nan / inf, strings with commas and blanks, flags,
    INFO/END, a Character tag; FORMAT vectors of different lengths per sample, missing samples, dropped trailing fields, GT of ploidy
    1 to 4 with both separators and missing alleles, symbolic and '*' alleles, ALT '.', two contigs"""
    rng = np.random.default_rng(9)
    H = ['##fileformat=VCFv4.2', '##contig=<ID=1,length=100000000>', '##contig=<ID=chrX,length=100000000>', '##FILTER=<ID=q10,Description="q">',
         '##FILTER=<ID=s50,Description="s">', '##INFO=<ID=IV,Number=.,Type=Integer,Description="i">', '##INFO=<ID=FV,Number=.,Type=Float,Description="f">',
         '##INFO=<ID=S,Number=1,Type=String,Description="s">', '##INFO=<ID=FL,Number=0,Type=Flag,Description="fl">',
         '##INFO=<ID=END,Number=1,Type=Integer,Description="e">', '##INFO=<ID=CH,Number=1,Type=Character,Description="c">',
         '##FORMAT=<ID=GT,Number=1,Type=String,Description="g">', '##FORMAT=<ID=AD,Number=R,Type=Integer,Description="a">',
         '##FORMAT=<ID=GL,Number=G,Type=Float,Description="l">', '##FORMAT=<ID=FT,Number=1,Type=String,Description="t">',
         '##FORMAT=<ID=DP,Number=1,Type=Integer,Description="d">', '#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\tFORMAT\t' + '\t'.join('S%d' % i for i in range(5))]
    L = []
    pos = 100
    ints = [0, 1, -1, 127, -120, -121, 128, 32767, -32760, -32761, 32768, 2147483647, -2147483640, "."]
    fls = ["0.5", "1e-7", "3.25e10", "-0.0", "123456.789", ".", "7", "0.333333333", "nan", "inf", "-inf"]
    for k in range(400):
        pos += int(rng.integers(1, 500))
        chrom = "1" if k < 250 else "chrX"
        if k == 250:
            pos = 50
        ref = "ACGT"[k % 4] + ("G" * (k % 7))
        alts = []
        for j in range(int(rng.integers(1, 5))):
            alts.append(str(rng.choice(["A", "C", "T", "AT", "<DEL>", "<INS:ME>", "GTT", "*"])))
        alts = list(dict.fromkeys(alts))
        if k % 29 == 0:
            alts = ["."]
        info = []
        if k % 2 == 0:
            info.append("IV=" + ",".join(str(rng.choice(ints)) for _ in range(int(rng.integers(1, 22)))))
        if k % 3 == 0:
            info.append("FV=" + ",".join(str(rng.choice(fls)) for _ in range(int(rng.integers(1, 18)))))
        if k % 5 == 0:
            info.append("S=" + str(rng.choice(["abc", "a,b,c", "x y", "%3B", "."])))
        if k % 7 == 0:
            info.append("FL")
        if k % 11 == 0:
            info.append("END=%d" % (pos + int(rng.integers(0, 5000))))
        if k % 13 == 0:
            info.append("CH=Z")
        info = ";".join(info) or "."
        qual = str(rng.choice([".", "0", "30", "12.5", "1e-3", "999999", "1234567.5"]))
        filt = str(rng.choice([".", "PASS", "q10", "q10;s50"]))
        line = "%s\t%d\t%s\t%s\t%s\t%s\t%s\t%s" % (chrom, pos, "." if k % 3 else "rs%d" % k, ref, ",".join(alts), qual, filt, info)
        na = 1 + (0 if alts == ["."] else len(alts))
        keys = ["GT", "AD", "GL", "FT", "DP"][:int(rng.integers(1, 6))]
        smps = []
        for s in range(5):
            v = []
            for key in keys:
                if key == "GT":
                    v.append(str(rng.choice(["0/1", "1|0", "./.", ".", "0", "1/2/0", "0|0|1|1", ".|1", "%d/%d" % (na - 1, na - 1)])))
                elif key == "AD":
                    v.append(str(rng.choice([".", ",".join(str(rng.integers(0, 300)) for _ in range(na)), ",".join(str(rng.integers(0, 70000)) for _ in range(max(1, na - 1)))])))
                elif key == "GL":
                    v.append(str(rng.choice([".", ",".join("%.3f" % (-rng.random() * 50) for _ in range(na * (na + 1) // 2)), "-1.5,."])))
                elif key == "FT":
                    v.append(str(rng.choice([".", "PASS", "lowq;x", "averyveryverylongfiltername"])))
                else:
                    v.append(str(rng.choice([".", "0", "35", "70000"])))
            if rng.random() < 0.15:
                v = v[:max(1, len(v) - int(rng.integers(1, 3)))]
            smps.append(":".join(v))
        line += "\t" + ":".join(keys) + "\t" + "\t".join(smps)
        L.append(line)
    open(path, "w").write("\n".join(H + L) + "\n")


make_stress(os.path.join(HERE, "stress_in.vcf"))
STEMS = ["blocksub_02_in", "c1_decompose_in", "filter_in", "blocksub_uneven_in", "blocksub_03_in", "vcftext_in", "c1_normalize_in", "stress_in"]
for stem in STEMS:
    src = os.path.join(HERE, stem + ".vcf")
    short = stem[:-3]
    subprocess.run([helpers.VT_REF, "view", src, "-o", os.path.join(HERE, "bcf_%s.bcf" % short)], check=True, capture_output=True)
    subprocess.run([helpers.VT_REF, "view", os.path.join(HERE, "bcf_%s.bcf" % short), "-o", os.path.join(HERE, "bcf_%s_view.vcf" % short)], check=True,
                   capture_output=True)
    print(short, os.path.getsize(os.path.join(HERE, "bcf_%s.bcf" % short)))
