"""Shared by the decompose_blocksub tests: golden cases and plan -> site columns."""
import os

import helpers

# (input stem, flags, output tag)
CASES = [("01", [], None), ("03", ["-p"], None), ("simple_mnv", [], "default"), ("simple_mnv", ["-m"], "m"), ("1bp_dist_mnv", ["-m"], "m"),
         ("2bp_dist_mnv", ["-m"], "m"), ("complex_mnv", ["-m"], "m"), ("synth", [], "default"), ("synth", ["-m"], "m"),
         ("synth", ["-m", "-d", "1"], "md1"), ("synth", ["-m", "-d", "4"], "md4"), ("uneven", [], "default")]
# the alignment mode (-a): the reference's own fixture 02 and vt_ref outputs
ALIGN_CASES = [("02", ["-a"], None), ("uneven", ["-a"], "a"), ("uneven", ["-a", "-p"], "ap"), ("uneven", ["-a", "-m"], "am"), ("synth", ["-a"], "a")]


def paths(stem, tag):
    vin = os.path.join(helpers.GOLDEN, "blocksub_%s_in.vcf" % stem)
    out = os.path.join(helpers.GOLDEN, "blocksub_%s_out" % stem if tag is None else "blocksub_%s_%s_out" % (stem, tag))
    return vin, out + ".vcf", out + ".stderr"


def mode(flags):
    keep = "-m" in flags
    dist = int(flags[flags.index("-d") + 1]) if "-d" in flags else 2
    return keep, dist


def read_sites(path):
    """[(chrom, pos, ref, alt)] of a VCF, in file order"""
    out = []
    for line in open(path):
        if line.startswith("#"):
            continue
        f = line.rstrip("\n").split("\t")
        out.append((f[0], int(f[1]), f[3], f[4]))
    return out


def batch_of(sites):
    chroms = sorted(set(s[0] for s in sites), key=[s[0] for s in sites].index)
    return helpers.Batch.from_records([(chroms.index(c), p, [r] + ([] if a == "." else a.split(","))) for (c, p, r, a) in sites])


def apply_plan(sites, src, off, ln):
    """site columns of the planned outputs in plan order (the writer re-sorts them within its window afterwards)"""
    out = []
    for s, o, l in zip(src, off, ln):
        c, p, r, a = sites[int(s)]
        out.append((c, p, r, a) if l == 0 else (c, p + int(o), r[int(o):int(o) + int(l)], a[int(o):int(o) + int(l)]))
    return out


def apply_plan_align(sites, src, off, ln, align):
    """the same with the alignment mode: a record with chunks (align[r] non-empty) is replaced by them (only ALT 1 is cut)"""
    out = []
    for s, o, l in zip(src, off, ln):
        c, p, r, a = sites[int(s)]
        if align[int(s)]:
            a1 = a.split(",")[0]
            out += [(c, p + pr, r[pr:pr + lr], a1[pa:pa + la]) for (pr, pa, lr, la) in align[int(s)]]
        else:
            out.append((c, p, r, a) if l == 0 else (c, p + int(o), r[int(o):int(o) + int(l)], a[int(o):int(o) + int(l)]))
    return out


def align_counters(cnt, align):
    """the reference's four counters in alignment mode from the plain plan's (decompose_blocksub.cpp:288-299)"""
    k = [len(a) for a in align if a]
    return [cnt[0], cnt[1] + len(k), cnt[2] + sum(k), cnt[3] + sum(k) - len(k)]


def stable_by_pos(cols):
    return sorted(cols, key=lambda x: x[1])


def stats_of(err):
    import re
    return [int(x) for x in re.findall(r":\s+(\d+)\n", err[err.index("stats:"):])]
