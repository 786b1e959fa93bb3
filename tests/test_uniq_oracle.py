"""CPU: the oracle's restatement of `vt uniq` (oracle/vt_oracle.c vto_uniq_plan: BCFSyncedReader's buffer + libstdc++
priority queue, uniq.cpp:110-156) pinned against the UNMODIFIED reference: committed golden vectors produced by
oracle/_ref/vt_ref (tests/golden/make_uniq_golden.py) and, where the binary exists, live differential runs."""
import os
import subprocess

import numpy as np
import pytest

import helpers
import uniq_helpers as U


def _batch(recs):
    return helpers.Batch.from_records([(r, p, al) for (r, p, al, _) in recs])


def _read_in(path):
    recs = []
    for line in open(path):
        if line.startswith("#"):
            continue
        f = line.rstrip("\n").split("\t")
        info = dict((kv.split("=", 1) + [None])[:2] for kv in f[7].split(";"))
        recs.append((U.CONTIGS.index(f[0]), int(f[1]), [f[3]] + ([] if f[4] == "." else f[4].split(",")), info.get("OLD_VARIANT")))
    return recs


@pytest.mark.parametrize("name", ["uniq_a", "uniq_b"])
def test_uniq_oracle_matches_reference_golden(name):
    recs = _read_in(os.path.join(helpers.GOLDEN, name + "_in.vcf"))
    order, first, cnt, bad = helpers.Oracle().uniq_plan(_batch(recs))
    assert bad == -1
    exp = U.read_kept(os.path.join(helpers.GOLDEN, name + "_out.vcf"))
    assert U.expected_from_plan(recs, order, first) == exp
    err = open(os.path.join(helpers.GOLDEN, name + "_out.stderr")).read()
    assert "Total number of observed variants   %d\n" % cnt[0] in err
    assert "Total number of unique variants     %d\n" % cnt[1] in err


def test_uniq_oracle_out_of_order():
    recs = [(0, 10, ["A", "C"], None), (0, 10, ["A", "C"], None), (0, 12, ["A", "C"], None), (0, 11, ["A", "C"], None), (0, 13, ["A", "G"], None)]
    _, _, _, bad = helpers.Oracle().uniq_plan(_batch(recs))
    assert bad == 3


@pytest.mark.skipif(not os.path.exists(helpers.VT_REF), reason="reference binary only exists in the build container")
@pytest.mark.parametrize("seed", [1, 2, 3])
def test_uniq_oracle_matches_reference_live(seed, tmp_path):
    rng = np.random.default_rng(seed)
    recs = U.gen_uniq_records(rng, 4000, max_group=6 + 3 * seed, big_group_every=50 if seed == 3 else 0)
    vin, vout = str(tmp_path / "in.vcf"), str(tmp_path / "out.vcf")
    U.write_uniq_vcf(vin, recs)
    p = subprocess.run([helpers.VT_REF, "uniq", vin, "-o", vout], capture_output=True, text=True)
    assert p.returncode == 0, p.stderr
    order, first, cnt, bad = helpers.Oracle().uniq_plan(_batch(recs))
    assert bad == -1 and cnt[0] == len(recs)
    assert U.expected_from_plan(recs, order, first) == U.read_kept(vout)
    assert "Total number of unique variants     %d\n" % cnt[1] in p.stderr
