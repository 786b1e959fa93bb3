#!/usr/bin/env python3
"""Goldens for two `vt normalize` behaviours, made by the UNMODIFIED reference (oracle/_ref/vt_ref):
  unk_*    SNP-only records on two consecutive contigs the FASTA lacks: the ordered writer flushes at the contig change (it goes by
           the header dictionary id, bcf_ordered_writer.cpp:185-189), so U1 stays in front of U2
  renorm_* input that already declares and carries OLD_VARIANT: no second header line, the value is replaced in place
           (bcf_update_info_string), also for a record whose OLD_VARIANT is not the first key
run in the build container: python tests/golden/make_advice_golden.py"""
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers  # noqa: E402

VT = os.path.join(ROOT, "oracle", "_ref", "vt_ref")
rng = np.random.default_rng(3)
seq = rng.choice(np.frombuffer(b"ACGT", np.uint8), size=3000)
seq[9:13] = np.frombuffer(b"TTTT", np.uint8)
seq[8] = ord("A")
seq[599:604] = np.frombuffer(b"GCCCC", np.uint8)
fa = os.path.join(HERE, "advice_ref.fa")
helpers.write_fasta(fa, ["1"], [seq])
H = "##fileformat=VCFv4.2\n##contig=<ID=1,length=3000>\n"
C = "#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\n"
open(os.path.join(HERE, "advice_unk_in.vcf"), "w").write(
    H + "##contig=<ID=U1,length=100000>\n##contig=<ID=U2,length=100000>\n" + C + "1\t5\t.\t%s\tG\t.\t.\t.\n" % chr(seq[4]) +
    "U1\t50000\t.\tA\tC\t.\t.\t.\nU1\t50010\t.\tA\tC\t.\t.\t.\nU2\t45000\t.\tA\tC\t.\t.\t.\nU2\t100\t.\tA\tC\t.\t.\t.\n")
open(os.path.join(HERE, "advice_renorm_in.vcf"), "w").write(
    H + '##INFO=<ID=OLD_VARIANT,Number=.,Type=String,Description="Original chr:pos:ref:alt encoding">\n##INFO=<ID=AF,Number=A,Type=Float,Description="x">\n##INFO=<ID=AF2,Number=1,Type=Integer,Description="x">\n' + C +
    "1\t11\t.\tTT\tT\t.\t.\tOLD_VARIANT=1:99:XX/X;AF=0.5\n" + "1\t500\t.\t%s\tG\t.\t.\tAF=0.25;OLD_VARIANT=1:77:A/C\n" % chr(seq[499]) +
    "1\t602\t.\tCC\tC\t.\t.\tAF=0.125;OLD_VARIANT=1:88:G/GG;AF2=1\n")
for stem in ("unk", "renorm"):
    with open(os.path.join(HERE, "advice_%s_out.stderr" % stem), "w") as e:
        subprocess.check_call([VT, "normalize", os.path.join(HERE, "advice_%s_in.vcf" % stem), "-r", fa, "-o", os.path.join(HERE, "advice_%s_out.vcf" % stem)], stderr=e)
print(open(os.path.join(HERE, "advice_renorm_out.vcf")).read().split("#CHROM")[1])
