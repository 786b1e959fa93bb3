#!/usr/bin/env python3
"""Golden vectors for annotate_indels, made by the UNMODIFIED reference: oracle/_ref/vt_ref_trace is vt_ref with
link-time taps on the AHMM / LFHMM / RFHMM calls (oracle/fd_trace.cpp).  Run in the build container (needs /root/reference built):
    python tests/golden/make_annot_golden.py
writes tests/golden/annot_ref.fa(.fai), annot_in.vcf, annot_out.vcf, annot_out.stderr, annot_trace.txt.gz"""
import gzip
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers  # noqa: E402
from vt_b200 import synth  # noqa: E402

VT = os.path.join(ROOT, "oracle", "_ref", "vt_ref")
VT_TRACE = os.path.join(ROOT, "oracle", "_ref", "vt_ref_trace")


def main():
    tmp = "/tmp/annot_golden"
    os.makedirs(tmp, exist_ok=True)
    tr = []
    names, seqs = synth.make_reference(seed=77, scale=0.00012, str_frac=0.3, tracts_out=tr, contigs=synth.GRCH38[:3])
    # soft-masked stretch and a few IUPAC codes away from the tracts' ends: fetch_seq upper-cases, the census sees non-ACGT bytes
    seqs[1][5000:5200] = np.frombuffer(bytes(seqs[1][5000:5200]).lower(), np.uint8)
    fa = os.path.join(HERE, "annot_ref.fa")
    helpers.write_fasta(fa, names, seqs)
    rng = np.random.default_rng(5)
    b = synth.make_str_indels(seqs, tr, 150, seed=9)
    b2 = synth.make_biallelic(seqs, 60, seed=10)
    lines = []
    for bb in (b, b2):
        data = bb.data.tobytes()
        ao = bb.allele_off
        for r in range(bb.n_rec):
            o0, o1, o2 = int(ao[2 * r]), int(ao[2 * r + 1]), int(ao[2 * r + 2])
            lines.append((int(bb.rid[r]), int(bb.pos1[r]), data[o0:o1].decode(), data[o1:o2].decode()))
    # multiallelic indels at the left end of a tract: deletions of one and two units; a deletion and an insertion; an indel + a SNP
    for c, (starts, tls, uls) in enumerate(tr):
        s = seqs[c]
        picked = 0
        for st, tl, ul in zip(starts, tls, uls):
            st, tl, ul = int(st), int(tl), int(ul)
            if tl < 4 * ul + 4 or st < 200 or s[st - 1] == ord("N") or picked >= 8:
                continue
            a = st - 1
            anchor = bytes(s[a:a + 1]).decode()
            u1 = bytes(s[a + 1:a + 1 + ul]).decode()
            u2 = bytes(s[a + 1:a + 1 + 2 * ul]).decode()
            if "N" in anchor + u2:
                continue
            kind = picked % 4
            if kind == 0:
                lines.append((c, a + 1, anchor + u2, anchor + u1 + "," + anchor))
            elif kind == 1:
                lines.append((c, a + 1, anchor + u1, anchor + "," + anchor + u1 + u1))
            elif kind == 2:
                snp = "ACGT"[("ACGT".index(anchor) + 1) % 4]
                lines.append((c, a + 1, anchor + u1, anchor + "," + snp + u1))
            else:
                lines.append((c, a + 1, anchor, anchor + u1 + "," + anchor + u2 + "," + anchor + u1[:1]))
            picked += 1
    # records the reference drops: an unnormalized deletion, an allele with N; and MNP / SNP records that only get FLANKSEQ
    s = seqs[0]
    p = 20000
    while s[p] == ord("N") or s[p + 1] != s[p + 2]:
        p += 1
    lines.append((0, p + 1, bytes(s[p:p + 3]).decode(), bytes(s[p:p + 2]).decode()))      # last bases equal: not normalized
    lines.append((0, p + 40, "N", "NA"))
    lines.append((0, p + 60, bytes(s[p + 59:p + 61]).decode(), "TT" if bytes(s[p + 59:p + 61]) != b"TT" else "GG"))  # MNP
    lines.sort(key=lambda x: (x[0], x[1]))
    raw = os.path.join(tmp, "raw.vcf")
    with open(raw, "w") as f:
        f.write("##fileformat=VCFv4.2\n")
        for n, q in zip(names, seqs):
            f.write("##contig=<ID=%s,length=%d>\n" % (n, q.shape[0]))
        f.write('##INFO=<ID=AC,Number=A,Type=Integer,Description="ac">\n')
        f.write("#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\n")
        for k, (c, pos, ref, alt) in enumerate(lines):
            f.write("%s\t%d\t.\t%s\t%s\t.\tPASS\t%s\n" % (names[c], pos, ref, alt, "." if k % 3 else "AC=" + ",".join("1" for _ in alt.split(","))))
    # biallelic STR indels come right-aligned from the generator: normalize + uniq them with the reference first; the multiallelic and
    # special records above are written as they are (annotate_indels itself skips what is not normalized)
    norm = os.path.join(tmp, "norm.vcf")
    subprocess.check_call([VT, "normalize", raw, "-r", fa, "-o", norm, "-n", "-q"], stderr=subprocess.DEVNULL)
    uq = os.path.join(HERE, "annot_in.vcf")
    subprocess.check_call([VT, "uniq", norm, "-o", uq], stderr=subprocess.DEVNULL)
    # put the unnormalized record back (normalize fixed it): append a copy of the raw line in position order is not needed -- the
    # reference's filter is exercised by the N record and by re-adding the raw unnormalized line at its own position
    text = open(uq).read().rstrip("\n").split("\n")
    hdr = [l for l in text if l.startswith("#")]
    body = [l for l in text if not l.startswith("#")]
    body.append("%s\t%d\t.\t%s\t%s\t.\tPASS\t." % (names[0], p + 1, bytes(s[p:p + 3]).decode(), bytes(s[p:p + 2]).decode()))
    body.sort(key=lambda l: (names.index(l.split("\t")[0]), int(l.split("\t")[1])))
    open(uq, "w").write("\n".join(hdr + body) + "\n")
    out = os.path.join(HERE, "annot_out.vcf")
    trace = os.path.join(tmp, "trace.txt")
    env = dict(os.environ, VT_HMM_TRACE=trace)
    with open(os.path.join(HERE, "annot_out.stderr"), "w") as ferr:
        subprocess.check_call([VT_TRACE, "annotate_indels", uq, "-r", fa, "-o", out], stderr=ferr, env=env)
    with open(trace, "rb") as f, gzip.GzipFile(os.path.join(HERE, "annot_trace.txt.gz"), "wb", mtime=0) as g:
        g.write(f.read())
    n = sum(1 for l in open(out) if not l.startswith("#"))
    print("annot golden: %d records in, %d out, %d alignments" % (len(body), n, sum(1 for _ in open(trace))))


if __name__ == "__main__":
    main()
