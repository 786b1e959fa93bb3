#!/usr/bin/env python3
"""Regenerates tests/golden/ from the reference's own test fixtures (run in the build container only;
/root/reference does not exist on the GPU box).

Fixtures:
  /root/reference/test/normalize/01_IN.vcf, 01_OUT.vcf, 01_OUT.stderr     -> c1_normalize_*.{vcf,stderr}
  /root/reference/test/decompose/01_IN_multi.vcf, 01_OUT_multi.*          -> c1_decompose_*.{vcf,stderr}
  /root/reference/test/ref/20.fa.gz is a MISSING blob (.MISSING_LARGE_BLOBS); the bases of chr20 that
  the fixture pins are reconstructed by replaying right_trim_or_left_extend (variant_manip.cpp:513-583)
  from every OLD_VARIANT to its OUT record and written as chr20_pinned.tsv (pos1 <TAB> base).
  tests/helpers.py materialises a 63,025,520-bp FASTA ('N' + pinned bases, 60-col lines) from it.
"""
import os
import shutil
import sys

REF = "/root/reference/test"
HERE = os.path.dirname(os.path.abspath(__file__))


def read_vcf(path):
    recs = []
    for line in open(path):
        if line.startswith("#"):
            continue
        f = line.rstrip("\n").split("\t")
        recs.append((f[0], int(f[1]), f[3], f[4].split(","), f[7]))
    return recs


def main():
    pairs = [
        ("normalize/01_IN.vcf", "c1_normalize_in.vcf"),
        ("normalize/01_OUT.vcf", "c1_normalize_out.vcf"),
        ("normalize/01_OUT.stderr", "c1_normalize_out.stderr"),
        ("decompose/01_IN_multi.vcf", "c1_decompose_in.vcf"),
        ("decompose/01_OUT_multi.vcf", "c1_decompose_out.vcf"),
        ("decompose/01_OUT_multi.stderr", "c1_decompose_out.stderr"),
        ("ref/20.fa.gz.fai", "chr20.fai"),
    ]
    for src, dst in pairs:
        shutil.copyfile(os.path.join(REF, src), os.path.join(HERE, dst))

    ref = {}

    def put(p, b):
        if p in ref and ref[p] != b:
            raise SystemExit("conflict at %d: %s vs %s" % (p, ref[p], b))
        ref[p] = b

    IN = read_vcf(os.path.join(REF, "normalize/01_IN.vcf"))
    OUT = read_vcf(os.path.join(REF, "normalize/01_OUT.vcf"))
    for recs in (IN, OUT):
        for (_, pos, r, _, _) in recs:
            for i, c in enumerate(r.upper()):
                put(pos + i, c)
    n_replayed = 0
    for (chrom, pos, r, alts, info) in OUT:
        ov = [x for x in info.split(";") if x.startswith("OLD_VARIANT=")]
        if not ov:
            continue
        c, opos, al = ov[0][len("OLD_VARIANT="):].split(":")
        A = [a.upper() for a in al.split("/")]
        pos1 = int(opos)
        while True:
            if any(len(a) == 0 for a in A):
                pos1 -= 1
                if pos1 not in ref:
                    put(pos1, max(A, key=len)[-1])  # only value for which the shift continues
                A = [ref[pos1] + a for a in A]
            elif all(a[-1] == A[0][-1] for a in A) and not (pos1 == 1 and any(len(a) == 1 for a in A)):
                A = [a[:-1] for a in A]
            else:
                break
        while all(len(a) > 1 and a[0] == A[0][0] for a in A):
            A = [a[1:] for a in A]
            pos1 += 1
        assert (pos1, A) == (pos, [r] + alts), (pos1, A, pos, r, alts)
        n_replayed += 1
    with open(os.path.join(HERE, "chr20_pinned.tsv"), "w") as f:
        for p in sorted(ref):
            f.write("%d\t%s\n" % (p, ref[p]))
    print("pinned %d bases, replayed %d OLD_VARIANT records" % (len(ref), n_replayed), file=sys.stderr)


if __name__ == "__main__":
    main()
