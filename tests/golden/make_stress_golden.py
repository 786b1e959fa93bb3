#!/usr/bin/env python3
"""The tools on the stress file (tests/golden/stress_in.vcf, made by make_bcf_golden.py: long vectors, type boundaries, nan / inf, padded
FORMAT vectors, ploidy 1-4, dropped trailing sample fields, symbolic and '*' alleles), run through the UNMODIFIED reference
(oracle/_ref/vt_ref): decompose_blocksub -a / -p, uniq -> stress_<tool>_out.{vcf,stderr}.
Run in the build container: python tests/golden/make_stress_golden.py"""
import os
import re
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import helpers  # noqa: E402

# decompose is left out on purpose: on this file's partial GL vectors and ploidy-3/4 genotypes the reference indexes past its arrays
# and prints whatever it finds there (values like 5.60519e-45 that change from build to build), so there is nothing to pin
RUNS = [("blocksub_a", ["decompose_blocksub", "-a"]), ("blocksub_p", ["decompose_blocksub", "-p"]), ("uniq", ["uniq"])]
for name, args in RUNS:
    out = os.path.join(HERE, "stress_%s_out.vcf" % name)
    p = subprocess.run([helpers.VT_REF] + args + ["stress_in.vcf", "-o", out], capture_output=True, text=True, cwd=HERE)
    assert p.returncode == 0, p.stderr
    open(out[:-4] + ".stderr", "w").write(re.sub(r"Time elapsed.*", "Time elapsed <stripped>", p.stderr))
    print(name, sum(1 for l in open(out) if not l.startswith("#")))
