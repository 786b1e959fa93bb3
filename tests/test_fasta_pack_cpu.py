"""CPU test of the host half of the resident reference (vt_b200/csrc/vtn_fasta_pack.h): the parallel, geometry-driven packer
against a byte-by-byte restatement of the isgraph stream faidx reads (tests/fasta_pack_main.cpp), on regular, CRLF,
irregular, index-less, gzip and truncated FASTA files."""
import gzip
import os
import subprocess

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def packer(tmp_path_factory):
    out = tmp_path_factory.mktemp("fp") / "fasta_pack_main"
    subprocess.run(["g++", "-O2", "-std=c++17", "-pthread", os.path.join(HERE, "fasta_pack_main.cpp"), "-lz", "-o", str(out)], check=True)
    return str(out)


def _contig(rng, n, lower=True, iupac=True):
    s = rng.choice(np.frombuffer(b"ACGT", np.uint8), size=n)
    if n > 100:
        if lower:
            a = int(rng.integers(0, n - 50))
            s[a:a + int(rng.integers(1, 50))] |= 0x20
        if iupac:
            for _ in range(4):
                a = int(rng.integers(0, n - 40))
                s[a:a + int(rng.integers(1, 40))] = ord("N")
            for ch in b"RYKMnry*-":
                s[int(rng.integers(0, n))] = ch
    return s.tobytes()


def _write(path, contigs, width=60, eol=b"\n", fai=True, widths=None):
    """widths: per-line widths cycle for an irregular file"""
    idx = []
    with open(path, "wb") as f:
        for name, seq in contigs:
            f.write(b">" + name.encode() + b" some description" + eol)
            off = f.tell()
            k = 0
            i = 0
            while k < len(seq):
                w = widths[i % len(widths)] if widths else width
                f.write(seq[k:k + w] + eol)
                k += w
                i += 1
            idx.append((name, len(seq), off, width, width + len(eol)))
    if fai:
        with open(path + ".fai", "w") as f:
            for e in idx:
                f.write("%s\t%d\t%d\t%d\t%d\n" % e)


def _run(packer, path, threads=4):
    return subprocess.run([packer, path, str(threads)], capture_output=True, text=True, check=True).stdout.strip()


def _contigs(rng):
    return [("1", _contig(rng, 3 * (1 << 20) + 12345)), ("2", _contig(rng, 600)), ("empty", b""), ("3", _contig(rng, 60 * 1000)),
            ("4", _contig(rng, 31)), ("5", _contig(rng, 1 << 20)), ("6", _contig(rng, (1 << 20) + 1))]


@pytest.mark.parametrize("width,eol", [(60, b"\n"), (70, b"\r\n"), (1, b"\n"), (1 << 21, b"\n")])
def test_regular_geometry(packer, tmp_path, width, eol):
    rng = np.random.default_rng(width)
    cs = _contigs(rng)
    p = str(tmp_path / "r.fa")
    _write(p, cs, width=width, eol=eol)
    total = sum(len(s) for _, s in cs)
    for thr in (1, 5):
        out = _run(packer, p, thr).split()
        assert out[0] == "ok" and int(out[1]) == len(cs) and int(out[2]) == total and int(out[3]) > 10, out


def test_built_index_matches_fai(packer, tmp_path):
    rng = np.random.default_rng(9)
    cs = [c for c in _contigs(rng) if c[1]]
    a, b = str(tmp_path / "a.fa"), str(tmp_path / "b.fa")
    _write(a, cs, fai=True)
    _write(b, cs, fai=False)
    assert _run(packer, a) == _run(packer, b)


def test_irregular_lines_fall_back_to_the_stream(packer, tmp_path):
    """lines of varying width: the geometry check fails and the contig is packed from the isgraph stream, like faidx reads it"""
    rng = np.random.default_rng(10)
    cs = [("1", _contig(rng, 2 * (1 << 20) + 77)), ("2", _contig(rng, 5000))]
    p = str(tmp_path / "i.fa")
    _write(p, cs, width=60, widths=[60, 60, 60, 45, 80, 60], fai=True)
    out = _run(packer, p).split()
    assert out[0] == "ok" and int(out[2]) == sum(len(s) for _, s in cs)
    # blank lines and trailing blanks inside a contig
    with open(str(tmp_path / "j.fa"), "wb") as f:
        f.write(b">x\nACGTNN\n\nACGTAC \nAC\n>y\nacgt\n")
    with open(str(tmp_path / "j.fa.fai"), "w") as f:
        f.write("x\t14\t3\t6\t7\ny\t4\t25\t4\t5\n")
    assert _run(packer, str(tmp_path / "j.fa")).split()[:3] == ["ok", "2", "18"]


def test_gzip_and_truncated(packer, tmp_path):
    rng = np.random.default_rng(11)
    cs = [("1", _contig(rng, 100000)), ("2", _contig(rng, 777))]
    p = str(tmp_path / "g.fa")
    _write(p, cs)
    with open(p, "rb") as f, gzip.open(p + ".gz", "wb") as g:
        g.write(f.read())
    os.rename(p + ".fai", p + ".gz.fai")
    assert _run(packer, p + ".gz").split()[0] == "ok"
    # an index that promises more bases than the file holds
    with open(p + ".fai", "w") as f:
        f.write("1\t100000\t20\t60\t61\n2\t999\t101706\t60\t61\n")
    assert _run(packer, p).split() == ["short", "2"]
