#!/usr/bin/env python3
"""Generates tests/golden/vcftext_{in,out}.vcf: synthetic record lines of every ALT shape (tests/vcf_text_helpers.py) parsed and
re-emitted by the UNMODIFIED reference (`vt view`, oracle/_ref/vt_ref -> htslib vcf_parse1).  Build container only."""
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import helpers  # noqa: E402
import vcf_text_helpers as T  # noqa: E402

lines = T.gen_lines(np.random.default_rng(31), 700, long_frac=0.005)
vin, vout = os.path.join(HERE, "vcftext_in.vcf"), os.path.join(HERE, "vcftext_out.vcf")
open(vin, "w").write(T.HDR + T.COLS + "\n" + "\n".join(lines) + "\n")
p = subprocess.run([helpers.VT_REF, "view", vin, "-o", vout], capture_output=True, text=True, cwd=HERE)
assert p.returncode == 0, p.stderr
print(len(lines), "records")
