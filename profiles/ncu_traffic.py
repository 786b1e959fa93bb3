#!/usr/bin/env python3
"""Records the DRAM traffic of the dominant kernel from an `ncu --set full` report in profiles/traffic.json, keyed by the hash of
the kernel sources (bench.py: kernel_source_hash), so that bench.py's roofline.traffic always belongs to the kernels it times.
usage: ncu_traffic.py report.ncu-rep <workload> <layout> <kernel substring> <records> <ref_scale> <source note>"""
import csv, json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
rep, workload, layout, kname, records, ref_scale, note = sys.argv[1:8]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[0]
units = rows[1]
ik, ir, iw = hdr.index("Kernel Name"), hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")


def to_bytes(v, u):
    return float(v.replace(",", "")) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]


vals = [to_bytes(r[ir], units[ir]) + to_bytes(r[iw], units[iw]) for r in rows[2:] if kname in r[ik]]
assert vals, "no launch of %s in %s" % (kname, rep)
import importlib.util
spec = importlib.util.spec_from_file_location("bench_hash", os.path.join(ROOT, "bench.py"))
src = open(os.path.join(ROOT, "bench.py")).read()
import hashlib
h = hashlib.sha1()
for f in ("vtn_normalize.cu", "vtn_fast.cuh", "vtn_ref_dev.cuh", "vtn_internal.h", "vtn_scan.cuh"):
    h.update(open(os.path.join(ROOT, "vt_b200", "csrc", f), "rb").read())
key = "%s:%s:%s" % (h.hexdigest()[:16], workload, layout)
path = os.path.join(ROOT, "profiles", "traffic.json")
t = json.load(open(path)) if os.path.exists(path) else {}
t[key] = {"dram_bytes": sum(vals) / len(vals), "launches": len(vals), "kernel": kname, "records": int(records), "ref_scale": float(ref_scale), "source": note}
json.dump(t, open(path, "w"), indent=1, sort_keys=True)
print(key, t[key])
