#!/usr/bin/env python3
"""Correlates the per-SASS-instruction counters of an ncu report with CUDA source lines (via nvdisasm -g on the
kernel's cubin) and prints the hottest lines.  usage: ncu_lines.py report.ncu-rep cubin kernel_substr [topN]"""
import csv, re, subprocess, sys, collections
rep, cubin, kname = sys.argv[1], sys.argv[2], sys.argv[3]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
out = open(rep).read() if rep.endswith(".csv") else subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
# a report with several kernels has one table per kernel ("Kernel Name" row, header row, instructions): keep kname's
starts = [i for i, r in enumerate(rows) if r and r[0] == 'Kernel Name']
if starts:
    pick = next((i for i in starts if kname in rows[i][1]), starts[0])
    end = next((i for i in starts if i > pick), len(rows))
    rows = rows[pick:end]
h = rows[1]
iA, iS, iN, iI, iT = h.index('Address'), h.index('Source'), h.index('# Samples'), h.index('Instructions Executed'), h.index('Thread Instructions Executed')
ilsb = h.index('stall_long_sb'); ibr = h.index('stall_branch_resolving')
ins = [(r[iS].strip(), int(r[iN] or 0), int(r[iI] or 0), int(r[iT] or 0), int(r[ilsb] or 0), int(r[ibr] or 0)) for r in rows[2:] if len(r) > iT]
dis = subprocess.run(['nvdisasm', '-g', '-c', cubin], capture_output=True, text=True).stdout.splitlines()
# walk the kernel's function body, tracking current line
lines = []
cur = None; infn = False
for l in dis:
    if l.startswith('.text.') or re.match(r'\s*\.section\s+\.text', l):
        infn = kname in l
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m: cur = (m.group(1).split('/')[-1], int(m.group(2)))
    if infn and re.match(r'\s+/\*[0-9a-f]{4}\*/', l):
        lines.append(cur)
print('sass instrs: ncu %d, nvdisasm %d' % (len(ins), len(lines)))
agg = collections.defaultdict(lambda: [0, 0, 0, 0, 0])
for k, (src, ns, ni, nt, lsb, br) in enumerate(ins):
    key = lines[k] if k < len(lines) else None
    a = agg[key]; a[0] += ns; a[1] += ni; a[2] += nt; a[3] += lsb; a[4] += br
tot_s = sum(a[0] for a in agg.values()); tot_i = sum(a[1] for a in agg.values())
print('total samples %d, warp instr %d' % (tot_s, tot_i))
print('%-28s %8s %6s %12s %6s %6s %8s %8s' % ('line', 'samples', '%', 'warp_instr', '%', 'thr/i', 'long_sb', 'branch'))
for key, a in sorted(agg.items(), key=lambda kv: -kv[1][int(sys.argv[5]) if len(sys.argv) > 5 else 0])[:top]:
    print('%-28s %8d %6.1f %12d %6.1f %6.1f %8d %8d' % ('%s:%s' % key if key else '?', a[0], 100.0 * a[0] / max(tot_s, 1), a[1], 100.0 * a[1] / max(tot_i, 1), a[2] / max(a[1], 1), a[3], a[4]))
