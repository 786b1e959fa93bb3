"""Shared by the annotate_indels tests: parsing of the reference's annotated VCF and of the alignment traces
(oracle/fd_trace.cpp format).  Goldens are made by tests/golden/make_annot_golden.py with the unmodified reference."""
import gzip
import os

import numpy as np

import helpers

FLOAT_TAGS = ("ENTROPY", "ENTROPY2", "KL_DIVERGENCE", "KL_DIVERGENCE2", "REF", "SCORE")
G = helpers.GOLDEN
REF_FA, IN_VCF, OUT_VCF, OUT_ERR, TRACE_GZ = (os.path.join(G, x) for x in ("annot_ref.fa", "annot_in.vcf", "annot_out.vcf", "annot_out.stderr", "annot_trace.txt.gz"))


def parse_info(s):
    d = {}
    for kv in s.split(";"):
        if "=" in kv:
            k, v = kv.split("=", 1)
            d[k] = v
        elif kv and kv != ".":
            d[kv] = True
    return d


def records(path):
    return [l.rstrip("\n").split("\t") for l in open(path) if not l.startswith("#")]


def htslib_float(x):
    """how htslib prints a float32 INFO value (6 significant digits, %g-like)"""
    return "%g" % np.float32(x)


def read_fasta(path):
    names, seqs = [], []
    for l in open(path):
        if l.startswith(">"):
            names.append(l[1:].split()[0])
            seqs.append([])
        else:
            seqs[-1].append(l.strip())
    return names, [np.frombuffer("".join(s).encode(), np.uint8) for s in seqs]


def golden_trace():
    return [l.split() for l in gzip.open(TRACE_GZ, "rt")]


def sorted_trace(path):
    """the driver's trace ("<indel> <stage> ...", rounds batched over all indels) in the reference's per-indel order"""
    t = [l.split() for l in open(path)]
    t.sort(key=lambda f: (int(f[0]), int(f[1])))
    return [f[2:] for f in t]


def check_tags(mine, ref_info):
    """mine: dict tag -> str / number / tuple (one annotated indel), ref_info: parsed INFO of the reference's record"""
    for k, v in mine.items():
        if not (k.startswith("EX_") or k.startswith("FZ_")):
            continue
        assert k in ref_info, k
        if k.split("_", 1)[1] in FLOAT_TAGS:
            assert htslib_float(v) == ref_info[k], (k, v, ref_info[k])
        elif isinstance(v, tuple):
            assert ",".join(str(x) for x in v) == ref_info[k], (k, v, ref_info[k])
        else:
            assert str(v) == ref_info[k], (k, v, ref_info[k])
    n_ref = sum(1 for k in ref_info if k.startswith("EX_") or k.startswith("FZ_"))
    assert n_ref == sum(1 for k in mine if k.startswith("EX_") or k.startswith("FZ_"))
