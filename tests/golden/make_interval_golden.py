#!/usr/bin/env python3
"""Goldens for -i / -I, made by the UNMODIFIED reference (oracle/_ref/vt_ref): one two-contig input as plain VCF, as BGZF VCF with and
without a .tbi, and as BCF with a .csi (written and indexed by the reference: `vt view`, `vt index`); `vt decompose` with interval
lists that overlap, repeat and come in another order than the file.  interval_cases.tsv: case, input, extra arguments.
Run in the build container: python tests/golden/make_interval_golden.py"""
import os
import re
import shutil
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import helpers  # noqa: E402

VT = helpers.VT_REF
rng = np.random.default_rng(23)
lines = ["##fileformat=VCFv4.2", "##contig=<ID=1,length=100000>", "##contig=<ID=2,length=100000>",
         "##INFO=<ID=END,Number=1,Type=Integer,Description=\"e\">", "##INFO=<ID=K,Number=1,Type=Integer,Description=\"k\">",
         "#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO"]
k = 0
for chrom in ("1", "2"):
    pos = 50
    for _ in range(150):
        pos += int(rng.integers(1, 30))
        L = int(rng.integers(1, 12))
        ref = "".join(rng.choice(list("ACGT"), L))
        u = k % 7
        alt = ("ACGT"["ACGT".index(ref[0]) ^ 1] + ref[1:]) if u < 3 else (ref[:1] if u == 3 else (ref + "A,C" if u == 4 else ("<DEL>" if u == 5 else ref[::-1] + ",G,T")))
        info = "K=%d" % k + (";END=%d" % (pos + 40) if u == 5 else "")
        lines.append("%s\t%d\t.\t%s\t%s\t.\t.\t%s" % (chrom, pos, ref, alt, info))
        k += 1
plain = os.path.join(HERE, "interval_in.vcf")
open(plain, "w").write("\n".join(lines) + "\n")
gz, bcf, noidx = os.path.join(HERE, "interval_in.vcf.gz"), os.path.join(HERE, "interval_in.bcf"), os.path.join(HERE, "interval_noidx.vcf.gz")
for out in (gz, bcf):
    subprocess.run([VT, "view", plain, "-o", out], check=True, capture_output=True)
    subprocess.run([VT, "index", out], check=True, capture_output=True)
shutil.copyfile(gz, noidx)
list_file = os.path.join(HERE, "interval_list.txt")
open(list_file, "w").write("2:100-400\n1:1000-1300\n2:100-400\n1\n")

CASES = [("plain", "interval_in.vcf", ["-i", "1:100-500"]), ("noidx", "interval_noidx.vcf.gz", ["-i", "1:100-500"]),
         ("gz_one", "interval_in.vcf.gz", ["-i", "1:100-500"]), ("gz_point", "interval_in.vcf.gz", ["-i", "2:700"]),
         ("gz_chrom", "interval_in.vcf.gz", ["-i", "2"]), ("gz_multi", "interval_in.vcf.gz", ["-i", "2:300-900,1:100-500,1:400-800"]),
         ("gz_none", "interval_in.vcf.gz", ["-i", "3:1-100,1:99000-99999"]), ("gz_list", "interval_in.vcf.gz", ["-I", "interval_list.txt", "-i", "1:60-90"]),
         ("bcf_multi", "interval_in.bcf", ["-i", "2:300-900,1:100-500,1:400-800"]), ("bcf_end", "interval_in.bcf", ["-i", "1:1-100000"])]
with open(os.path.join(HERE, "interval_cases.tsv"), "w") as tsv:
    for name, vin, extra in CASES:
        out = os.path.join(HERE, "interval_%s_out.vcf" % name)
        p = subprocess.run([VT, "decompose"] + extra + [vin, "-o", out], capture_output=True, text=True, cwd=HERE)
        assert p.returncode == 0, p.stderr
        open(out[:-4] + ".stderr", "w").write(re.sub(r"Time elapsed.*", "Time elapsed <stripped>", p.stderr))
        n = sum(1 for l in open(out) if not l.startswith("#"))
        tsv.write("%s\t%s\t%s\n" % (name, vin, " ".join(extra)))
        print(name, n, [l for l in p.stderr.splitlines() if l.startswith("[E")])
