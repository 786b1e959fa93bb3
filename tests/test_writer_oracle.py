"""CPU: the oracle's literal emulation of BCFOrderedWriter (oracle/vt_oracle.c vto_writer_order; bcf_ordered_writer.cpp:141-300)
pinned against the UNMODIFIED reference: `vt view -w W` runs every record through that writer.  Golden vectors
(tests/golden/writer_*) were produced by oracle/_ref/vt_ref with tests/golden/make_writer_golden.py; where the binary exists
the comparison is also run live."""
import os
import subprocess

import numpy as np
import pytest

import helpers

HDR = """##fileformat=VCFv4.2
##FILTER=<ID=PASS,Description="All filters passed">
##contig=<ID=1,length=100000000>
##contig=<ID=2,length=100000000>
##contig=<ID=3,length=100000000>
##INFO=<ID=K,Number=1,Type=Integer,Description="arrival index">
#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO
"""


def gen_jittered(rng, n, jitter, p_back=0.01, start=1):
    """positions that mostly grow, with local disorder up to `jitter` and occasional long jumps back (window overflow);
    contig runs may repeat (a rid change flushes the writer)"""
    rid = np.zeros(n, np.int32)
    pos = np.zeros(n, np.int32)
    cur, c = start, 0
    for i in range(n):
        cur += int(rng.integers(0, 4))
        p = cur - int(rng.integers(0, jitter + 1))
        if rng.random() < p_back:
            p = cur - int(rng.integers(jitter, 6 * jitter + 2))
        if rng.random() < 0.004:
            c = int(rng.integers(0, 3))
            cur = start + int(rng.integers(0, 50))
            p = cur
        rid[i] = c
        pos[i] = max(p, 1 if rng.random() < 0.999 else 0)
    return rid, pos


def write_vcf(path, rid, pos):
    with open(path, "w") as f:
        f.write(HDR)
        for k in range(len(rid)):
            f.write("%d\t%d\t.\tA\tC\t.\t.\tK=%d\n" % (rid[k] + 1, pos[k], k))


def read_vcf(path):
    rid, pos, ks = [], [], []
    for line in open(path):
        if line[0] == "#":
            continue
        f = line.split("\t")
        rid.append(int(f[0]) - 1)
        pos.append(int(f[1]))
        ks.append(int(f[7].strip().split("=")[1]))
    return np.array(rid, np.int32), np.array(pos, np.int32), np.array(ks, np.uint32)


GOLD = [("writer_a", 30), ("writer_b", 7)]


@pytest.mark.parametrize("name,window", GOLD)
def test_writer_oracle_matches_reference_golden(name, window):
    rid, pos, _ = read_vcf(os.path.join(helpers.GOLDEN, name + "_in.vcf"))
    _, _, ks = read_vcf(os.path.join(helpers.GOLDEN, name + "_out.vcf"))
    order, late = helpers.Oracle().writer_order(rid, pos, window)
    np.testing.assert_array_equal(order, ks)
    n_warn = open(os.path.join(helpers.GOLDEN, name + "_out.stderr")).read().count("Might not be sorted")
    assert int((late != np.iinfo(np.int32).min).sum()) == n_warn


@pytest.mark.skipif(not os.path.exists(helpers.VT_REF), reason="reference binary only exists in the build container")
@pytest.mark.parametrize("seed,window,jitter", [(1, 50, 10), (2, 5, 10), (3, 1, 3), (4, 200, 40)])
def test_writer_oracle_matches_reference_live(seed, window, jitter, tmp_path):
    rid, pos = gen_jittered(np.random.default_rng(seed), 3000, jitter)
    vin, vout = str(tmp_path / "in.vcf"), str(tmp_path / "out.vcf")
    write_vcf(vin, rid, pos)
    p = subprocess.run([helpers.VT_REF, "view", "-w", str(window), vin, "-o", vout], capture_output=True, text=True)
    assert p.returncode == 0, p.stderr
    _, _, ks = read_vcf(vout)
    order, late = helpers.Oracle().writer_order(rid, pos, window)
    np.testing.assert_array_equal(order, ks)
    assert int((late != np.iinfo(np.int32).min).sum()) == p.stderr.count("Might not be sorted")
