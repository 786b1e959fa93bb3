"""Generators and VCF plumbing for the `vt uniq` row: position groups with duplicate, distinct and
multiallelic records, each record tagged K=<input index> so that WHICH duplicate survived can be read off the output."""
import numpy as np

UNIQ_HDR = """##fileformat=VCFv4.2
##FILTER=<ID=PASS,Description="All filters passed">
##contig=<ID=1,length=100000000>
##contig=<ID=2,length=100000000>
##contig=<ID=3,length=100000000>
##INFO=<ID=OLD_VARIANT,Number=.,Type=String,Description="Original chr:pos:ref:alt encoding">
##INFO=<ID=K,Number=1,Type=Integer,Description="input index">
#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO
"""
CONTIGS = ["1", "2", "3"]


def _allele(rng, maxlen):
    return "".join(rng.choice(list("ACGT"), size=int(rng.integers(1, maxlen + 1))))


def gen_uniq_records(rng, n, max_group=9, big_group_every=0, old_variant_frac=0.3):
    """-> list of (rid, pos1, [alleles], old_variant or None), sorted by (rid, pos1), many duplicates per position."""
    recs = []
    rid, pos = 0, 10
    g = 0
    while len(recs) < n:
        g += 1
        m = int(rng.integers(1, max_group + 1))
        if rng.random() < 0.5:
            m = 1 + int(rng.random() < 0.3)
        if big_group_every and g % big_group_every == 0:
            m = int(rng.integers(40, 120))
        pool = []
        for _ in range(max(1, int(rng.integers(1, 4)))):
            ref = _allele(rng, 3)
            na = 1 if rng.random() < 0.8 else int(rng.integers(2, 5))
            alts = []
            while len(alts) < na:
                a = _allele(rng, 3)
                if a != ref and a not in alts:
                    alts.append(a)
            pool.append([ref] + alts)
        if rng.random() < 0.03:
            pool.append([_allele(rng, 2)])  # ALT "." : a single allele
        for _ in range(m):
            al = list(pool[int(rng.integers(0, len(pool)))])
            if len(al) > 2 and rng.random() < 0.5:
                alts = al[1:]
                rng.shuffle(alts)  # same set of ALTs in another order compares equal (sorted allele string)
                al = [al[0]] + list(alts)
            ov = None
            if rng.random() < old_variant_frac:
                ov = "%s:%d:%s/%s" % (CONTIGS[rid], pos + int(rng.integers(0, 3)), al[0] + "A", (al[1] if len(al) > 1 else "C") + "A")
                if rng.random() < 0.2:
                    ov += "," + ov.replace(":", ":1", 1)
            recs.append((rid, pos, al, ov))
        pos += int(rng.integers(1, 4))
        if rng.random() < 0.01 and rid < 2:
            rid += 1
            pos = 5
    return recs[:n]


def write_uniq_vcf(path, recs):
    with open(path, "w") as f:
        f.write(UNIQ_HDR)
        for k, (rid, pos, al, ov) in enumerate(recs):
            info = "K=%d" % k + (";OLD_VARIANT=" + ov if ov else "")
            f.write("%s\t%d\t.\t%s\t%s\t.\t.\t%s\n" % (CONTIGS[rid], pos, al[0], ",".join(al[1:]) if len(al) > 1 else ".", info))


def read_kept(path):
    """-> [(K, OLD_VARIANT string or None)] of the records in an output VCF, in file order"""
    out = []
    for line in open(path):
        if line.startswith("#"):
            continue
        info = dict((kv.split("=", 1) + [None])[:2] for kv in line.rstrip("\n").split("\t")[7].split(";"))
        out.append((int(info["K"]), info.get("OLD_VARIANT")))
    return out


def expected_from_plan(recs, order, first):
    """what uniq writes given a plan: crecs[0] of every group with the OLD_VARIANT strings of the group merged in pop
    order, each distinct string once (uniq.cpp:114-143; the map key is the whole comma-joined INFO value)"""
    out = []
    k = 0
    n = len(order)
    while k < n:
        assert first[k]
        e = k + 1
        while e < n and not first[e]:
            e += 1
        head = int(order[k])
        ov = recs[head][3]
        if e - k > 1:
            seen, merged = set(), []
            for j in range(k, e):
                v = recs[int(order[j])][3]
                if v is not None and v not in seen:
                    seen.add(v)
                    merged.append(v)
            if merged:
                ov = ",".join(merged)
        out.append((head, ov))
        k = e
    return out
