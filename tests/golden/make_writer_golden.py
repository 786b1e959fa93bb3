#!/usr/bin/env python3
"""Generates tests/golden/writer_{a,b}_{in.vcf,out.vcf,out.stderr}: jittered synthetic inputs pushed through the UNMODIFIED
reference's BCFOrderedWriter by `vt view -w W` (oracle/_ref/vt_ref).  Build container only; outputs are committed."""
import os
import re
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import helpers  # noqa: E402
import test_writer_oracle as W  # noqa: E402

for (name, window), seed, jitter in zip(W.GOLD, (21, 22), (12, 9)):
    rid, pos = W.gen_jittered(np.random.default_rng(seed), 1200, jitter, p_back=0.02)
    vin, vout = os.path.join(HERE, name + "_in.vcf"), os.path.join(HERE, name + "_out.vcf")
    W.write_vcf(vin, rid, pos)
    p = subprocess.run([helpers.VT_REF, "view", "-w", str(window), vin, "-o", vout], capture_output=True, text=True, cwd=HERE)
    assert p.returncode == 0, p.stderr
    err = re.sub(r"Time elapsed.*", "Time elapsed <stripped>", p.stderr).replace(HERE + "/", "")
    open(os.path.join(HERE, name + "_out.stderr"), "w").write(err)
    print(name, len(rid), "records,", err.count("Might not be sorted"), "warnings")
