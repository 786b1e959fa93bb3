"""Generators for the VCF-text tokenizer row."""
import numpy as np

HDR = ("##fileformat=VCFv4.2\n##FILTER=<ID=PASS,Description=\"All filters passed\">\n##contig=<ID=1,length=100000000>\n"
       "##contig=<ID=chrX_alt,length=100000000>\n##contig=<ID=HLA-A*01:01,length=100000000>\n"
       "##ALT=<ID=DEL,Description=\"d\">\n##INFO=<ID=K,Number=1,Type=Integer,Description=\"k\">\n"
       "##INFO=<ID=SVTYPE,Number=1,Type=String,Description=\"t\">\n##FORMAT=<ID=GT,Number=1,Type=String,Description=\"g\">\n")
COLS = "#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO"
CHROMS = ["1", "chrX_alt", "HLA-A*01:01"]


def _al(rng, maxlen, alphabet="ACGT"):
    return "".join(rng.choice(list(alphabet), size=int(rng.integers(1, maxlen + 1))))


def gen_lines(rng, n, samples=0, long_frac=0.01):
    """record lines (no line ends) with every ALT shape the path meets: SNV, indels, multiallelic, '.', '*', symbolic, lower case, IUPAC"""
    out = []
    pos = 1
    for k in range(n):
        pos += int(rng.integers(0, 300))
        c = CHROMS[min(2, k * 3 // max(n, 1))]
        ref = _al(rng, 1 if rng.random() < 0.7 else 6, "ACGTacgtN" if rng.random() < 0.1 else "ACGT")
        if rng.random() < long_frac:
            ref = _al(rng, 400) + "A" * int(rng.integers(0, 300))
        u = rng.random()
        if u < 0.05:
            alt = "."
        elif u < 0.08:
            alt = "<DEL>"
        elif u < 0.10:
            alt = "*," + _al(rng, 3)
        elif u < 0.12:
            alt = ref[0] + "[" + c + ":" + str(pos + 50) + "["
        else:
            alt = ",".join(_al(rng, 1 if rng.random() < 0.7 else 8) for _ in range(1 if rng.random() < 0.8 else int(rng.integers(2, 7))))
        qual = "." if rng.random() < 0.5 else "%d" % int(rng.integers(1, 99))
        filt = "." if rng.random() < 0.5 else "PASS"
        info = "K=%d" % k + (";SVTYPE=DEL" if alt == "<DEL>" else "")
        line = "\t".join([c, str(pos), "." if rng.random() < 0.7 else "rs%d" % k, ref, alt, qual, filt, info])
        if samples:
            line += "\tGT" + "".join("\t" + rng.choice(["0/0", "0/1", "1/1", "./."]) for _ in range(samples))
        out.append(line)
    return out


def site_columns(text, f):
    """CHROM, POS, REF, ALT strings rebuilt from the tokenizer's arrays, one tuple per record"""
    out = []
    by = bytes(f["bytes"])
    for r in range(f["n_rec"]):
        o = int(f["line_off"][r])
        chrom = text[o:o + int(f["chrom_len"][r])].decode()
        s0, s1 = int(f["allele_beg"][r]), int(f["allele_beg"][r + 1])
        al = [by[int(f["allele_off"][s]):int(f["allele_off"][s + 1])].decode() for s in range(s0, s1)]
        out.append((chrom, int(f["pos1"][r]), al[0] if al else None, ",".join(al[1:]) if len(al) > 1 else "."))
    return out
