#!/usr/bin/env python3
"""Generates tests/golden/uniq_{a,b}_{in.vcf,out.vcf,out.stderr}: synthetic inputs (tests/uniq_helpers.py) run through the
UNMODIFIED reference `vt uniq` (oracle/_ref/vt_ref, built by oracle/Makefile from /root/reference).  Run in the build
container only; the outputs are committed."""
import os
import re
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import helpers  # noqa: E402
import uniq_helpers as U  # noqa: E402

for name, seed, n, mg, big in (("uniq_a", 11, 600, 8, 0), ("uniq_b", 12, 1500, 12, 25)):
    recs = U.gen_uniq_records(np.random.default_rng(seed), n, max_group=mg, big_group_every=big)
    vin = os.path.join(HERE, name + "_in.vcf")
    vout = os.path.join(HERE, name + "_out.vcf")
    U.write_uniq_vcf(vin, recs)
    p = subprocess.run([helpers.VT_REF, "uniq", vin, "-o", vout], capture_output=True, text=True, cwd=HERE)
    assert p.returncode == 0, p.stderr
    err = re.sub(r"Time elapsed.*", "Time elapsed <stripped>", p.stderr)
    err = err.replace(HERE + "/", "")
    open(os.path.join(HERE, name + "_out.stderr"), "w").write(err)
    print(name, len(recs), "records")
