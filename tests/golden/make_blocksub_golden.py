#!/usr/bin/env python3
"""tests/golden/blocksub_*: (a) the reference's own fixtures test/decompose_blocksub/01_* (default) and 03_* (-p), copied; (b) the
four inputs of test/test_mnv.sh (test/decompose_blocksub/mnv/*.vcf) and one synthetic input, run through the UNMODIFIED reference
`vt decompose_blocksub` / `-m` / `-m -d 1` (oracle/_ref/vt_ref).  Build container only; outputs are committed."""
import os
import re
import shutil
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import helpers  # noqa: E402

REF = "/root/reference/test/decompose_blocksub"
for src, dst in (("01_IN_even_length.vcf", "blocksub_01_in.vcf"), ("01_OUT_even_length.vcf", "blocksub_01_out.vcf"),
                 ("01_OUT_even_length.stderr", "blocksub_01_out.stderr"), ("03_IN_phased_even_length.vcf", "blocksub_03_in.vcf"),
                 ("03_OUT_phased_even_length.vcf", "blocksub_03_out.vcf"), ("03_OUT_phased_even_length.stderr", "blocksub_03_out.stderr"),
                 ("02_IN_uneven_length.vcf", "blocksub_02_in.vcf"), ("02_OUT_uneven_length.vcf", "blocksub_02_out.vcf"),
                 ("02_OUT_uneven_length.stderr", "blocksub_02_out.stderr")):
    shutil.copyfile(os.path.join(REF, src), os.path.join(HERE, dst))
for name in ("simple_mnv", "1bp_dist_mnv", "2bp_dist_mnv", "complex_mnv"):
    shutil.copyfile(os.path.join(REF, "mnv", name + ".vcf"), os.path.join(HERE, "blocksub_%s_in.vcf" % name))

# a synthetic sites-only input: block substitutions of every shape between SNVs, indels and multiallelic records
rng = np.random.default_rng(51)
lines = ["##fileformat=VCFv4.2", "##contig=<ID=1,length=100000000>", "##INFO=<ID=K,Number=1,Type=Integer,Description=\"k\">",
         "#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO"]
pos = 100
for k in range(400):
    pos += int(rng.integers(1, 40))
    u = rng.random()
    L = int(rng.integers(2, 14))
    ref = "".join(rng.choice(list("ACGT"), L))
    if u < 0.6:
        alt = "".join(c if rng.random() < 0.55 else rng.choice(list("ACGT")) for c in ref)
    elif u < 0.7:
        alt = ref[:1]
    elif u < 0.8:
        alt = ref + "T"
    elif u < 0.9:
        alt = ref[::-1] + "," + ref[:1]
    else:
        ref, alt = ref[:1], "ACGT"[(("ACGT".index(ref[0])) + 1) % 4]
    lines.append("1\t%d\t.\t%s\t%s\t.\t.\tK=%d" % (pos, ref, alt, k))
open(os.path.join(HERE, "blocksub_synth_in.vcf"), "w").write("\n".join(lines) + "\n")

# uneven block substitutions for the alignment mode (-a): REF / ALT made from one another by substitutions, insertions and deletions,
# a few with samples (phased output -p rewrites GT and adds PS), one multiallelic record (only ALT 1 is aligned)
rng = np.random.default_rng(52)
lines = ["##fileformat=VCFv4.2", "##contig=<ID=1,length=100000000>", "##INFO=<ID=K,Number=1,Type=Integer,Description=\"k\">",
         "##FORMAT=<ID=GT,Number=1,Type=String,Description=\"g\">", "#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\tFORMAT\tS1\tS2"]
pos = 100
for k in range(300):
    pos += int(rng.integers(1, 60))
    L = int(rng.integers(1, 25))
    ref = "".join(rng.choice(list("ACGT"), L))
    alt = []
    for c in ref:
        u = rng.random()
        if u < 0.70:
            alt.append(c)
        elif u < 0.82:
            alt.append(str(rng.choice(list("ACGT"))))
        elif u < 0.91:
            alt.append(c + "".join(rng.choice(list("ACGT"), int(rng.integers(1, 4)))))
    alt = "".join(alt) or "A"
    if k % 37 == 5:
        alt = alt + "," + ref[:1] + "TT"
    if k % 41 == 7:
        ref, alt = ref.lower(), alt.lower() if k % 2 else alt
    gts = ["0/1", "1|1", "./.", "0|1", "1/1"]
    lines.append("1\t%d\t.\t%s\t%s\t.\t.\tK=%d\tGT\t%s\t%s" % (pos, ref, alt, k, gts[k % 5], gts[(k // 5) % 5]))
open(os.path.join(HERE, "blocksub_uneven_in.vcf"), "w").write("\n".join(lines) + "\n")

runs = [("02", ["-a"]), ("uneven", ["-a"]), ("uneven", ["-a", "-p"]), ("uneven", ["-a", "-m"]), ("uneven", []), ("synth", ["-a"]),
        ("simple_mnv", []), ("simple_mnv", ["-m"]), ("1bp_dist_mnv", ["-m"]), ("2bp_dist_mnv", ["-m"]), ("complex_mnv", ["-m"]),
        ("synth", []), ("synth", ["-m"]), ("synth", ["-m", "-d", "1"]), ("synth", ["-m", "-d", "4"])]
for name, flags in runs:
    tag = "".join(f.replace("-", "") for f in flags) or "default"
    vin = os.path.join(HERE, "blocksub_%s_in.vcf" % name)
    vout = os.path.join(HERE, "blocksub_%s_%s_out.vcf" % (name, tag))
    p = subprocess.run([helpers.VT_REF, "decompose_blocksub"] + flags + [vin, "-o", vout], capture_output=True, text=True, cwd=HERE)
    assert p.returncode == 0, p.stderr
    err = re.sub(r"Time elapsed.*", "Time elapsed <stripped>", p.stderr).replace(HERE + "/", "")
    open(vout[:-4] + ".stderr", "w").write(err)
    print(name, flags, "ok")

# the freshly built reference must reproduce its own fixture for -a; the duplicate is not kept
for ext in (".vcf", ".stderr"):
    a, b = os.path.join(HERE, "blocksub_02_a_out" + ext), os.path.join(HERE, "blocksub_02_out" + ext)
    if ext == ".vcf":
        assert open(a).read() == open(b).read(), "vt_ref does not reproduce test/decompose_blocksub/02_OUT_uneven_length.vcf"
    os.remove(a)
