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

STEMS = ["blocksub_02_in", "c1_decompose_in", "filter_in", "blocksub_uneven_in", "blocksub_03_in", "vcftext_in", "c1_normalize_in"]
for stem in STEMS:
    src = os.path.join(HERE, stem + ".vcf")
    short = stem[:-3]
    subprocess.run([helpers.VT_REF, "view", src, "-o", os.path.join(HERE, "bcf_%s.bcf" % short)], check=True, capture_output=True)
    subprocess.run([helpers.VT_REF, "view", os.path.join(HERE, "bcf_%s.bcf" % short), "-o", os.path.join(HERE, "bcf_%s_view.vcf" % short)], check=True,
                   capture_output=True)
    print(short, os.path.getsize(os.path.join(HERE, "bcf_%s.bcf" % short)))
