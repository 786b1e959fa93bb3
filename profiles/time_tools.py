#!/usr/bin/env python3
"""Wall clock of the line-based tools (decompose, decompose -s, uniq, decompose_blocksub) against the unmodified reference on one
synthetic multiallelic file with 3 samples (the C3 parity variant, 300 k records) on tmpfs; outputs compared."""
import json, os, subprocess, sys, tempfile, time, filecmp
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
VTN, VT = os.path.join(ROOT, "vt_b200", "vtn"), os.path.join(ROOT, "oracle", "_ref", "vt_ref")
tmp = tempfile.mkdtemp(dir="/dev/shm")
rng = np.random.default_rng(5)
n = int(os.environ.get("N", 300000))
lines = ["##fileformat=VCFv4.2", "##contig=<ID=1,length=250000000>", '##INFO=<ID=AC,Number=A,Type=Integer,Description="x">',
         '##INFO=<ID=AF,Number=A,Type=Float,Description="x">', '##FORMAT=<ID=GT,Number=1,Type=String,Description="x">',
         '##FORMAT=<ID=AD,Number=R,Type=Integer,Description="x">', '##FORMAT=<ID=DP,Number=1,Type=Integer,Description="x">',
         '##FORMAT=<ID=PL,Number=G,Type=Integer,Description="x">', "#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\tFORMAT\tS1\tS2\tS3"]
pos = 1000
B = "ACGT"
for k in range(n):
    pos += int(rng.integers(1, 300))
    na = 1 if k % 3 else int(rng.integers(2, 4))
    ref = B[k % 4]
    alts = [B[(k + 1 + j) % 4] + ("T" * j) for j in range(na)]
    ng = (na + 1) * (na + 2) // 2
    smp = ["0/1:%s:30:%s" % (",".join("7" for _ in range(na + 1)), ",".join(str(10 * j) for j in range(ng))) for _ in range(3)]
    lines.append("1\t%d\t.\t%s\t%s\t50\tPASS\tAC=%s;AF=%s\tGT:AD:DP:PL\t%s" % (pos, ref, ",".join(alts), ",".join("1" for _ in alts), ",".join("0.25" for _ in alts), "\t".join(smp)))
vin = os.path.join(tmp, "in.vcf")
open(vin, "w").write("\n".join(lines) + "\n")
out = {}
for name, args in (("decompose", ["decompose"]), ("decompose -s", ["decompose", "-s"]), ("uniq", ["uniq"]), ("decompose_blocksub", ["decompose_blocksub"])):
    row = {}
    for tool, exe in (("vtn", VTN), ("vt", VT)):
        o = os.path.join(tmp, "%s_%s.vcf" % (tool, name.replace(" ", "")))
        best = 1e9
        for _ in range(2):
            t0 = time.perf_counter()
            p = subprocess.run([exe] + args + [vin, "-o", o], capture_output=True, env=dict(os.environ, VTN_CLI_TIMING="1"))
            best = min(best, time.perf_counter() - t0)
        if tool == "vtn":
            row["phases"] = {l.split("]")[1].rsplit(None, 2)[0].strip(): float(l.split()[-2]) for l in p.stderr.decode().splitlines() if l.startswith("[vtn timing]")}
        row[tool + "_s"] = round(best, 3)
        row[tool + "_rc"] = p.returncode
    row["identical"] = filecmp.cmp(os.path.join(tmp, "vtn_%s.vcf" % name.replace(" ", "")), os.path.join(tmp, "vt_%s.vcf" % name.replace(" ", "")), shallow=False)
    row["speedup"] = round(row["vt_s"] / row["vtn_s"], 2)
    out[name] = row
print(json.dumps({"records": n, "bytes": os.path.getsize(vin), "tools": out}, indent=1))
