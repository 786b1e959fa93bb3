"""The VariantManip-shaped facade (include/vtn_variant_manip.hpp): compiled here, then the C1 fixture and a synthetic mix are replayed
through it RECORD BY RECORD with the control flow of normalize.cpp:188-250 (tests/variant_manip_main.cpp); every line must equal what the
CPU oracle says about the record (type, REF code, is_normalized, contains_N, new POS / alleles, left_trimmed / left_extended /
right_trimmed)."""
import os
import subprocess

import numpy as np
import pytest

import helpers
from vt_b200.api import Batch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def harness():
    out = os.path.join(helpers.ROOT, "tests", "_build")
    os.makedirs(out, exist_ok=True)
    exe = os.path.join(out, "variant_manip_main")
    lib = os.path.join(helpers.ROOT, "vt_b200")
    subprocess.run(["g++", "-std=c++17", "-O1", "-Wall", os.path.join(helpers.ROOT, "tests", "variant_manip_main.cpp"), "-L" + lib, "-lvtn",
                    "-Wl,-rpath," + lib, "-o", exe], check=True)
    return exe


def _expected(oracle, names, seqs, recs):
    b = Batch.from_records(recs)
    exp = oracle.normalize(seqs, b)
    chg = {int(r): k for k, r in enumerate(exp.chg_rec[:exp.n_changed])}
    lines = []
    for r, (rid, pos1, alleles) in enumerate(recs):
        st, vt = int(exp.status[r]), int(exp.vtype[r])
        code = st & 3 if vt != 1 else 0  # normalize.cpp:204: SNPs are not checked
        has_n = any(b"N" in a and not (i and a.startswith(b"<")) for i, a in enumerate(alleles))
        if code == 0 and not (vt & 48):
            norm = not (st & 0x08)
        else:
            norm = None  # not reached by the normalize loop: checked through vtn_variant_props' own test below
        if r in chg:
            k = chg[r]
            o, d, new = int(exp.chg_off[k]), int(exp.chg_le[k]) - int(exp.chg_rt[k]) - int(exp.chg_lt[k]), []
            for a in alleles:
                new.append(bytes(exp.chg_bytes[o:o + len(a) + d]))
                o += len(a) + d
            tail = "%d %s %s %d %d %d" % (exp.chg_pos1[k], new[0].decode(), b",".join(new[1:]).decode(), exp.chg_lt[k], exp.chg_le[k], exp.chg_rt[k])
        else:
            al = [a.upper() if (st & 0x08) else a for a in alleles]
            tail = "%d %s %s 0 0 0" % (pos1, al[0].decode(), b",".join(al[1:]).decode() if len(al) > 1 else ".")
        lines.append((names[rid], pos1, alleles, vt, code, norm, has_n, tail))
    return lines


def _replay(harness, oracle, tmp_path, names, seqs, recs):
    fa = str(tmp_path / "ref.fa")
    helpers.write_fasta(fa, names, seqs)
    sites = str(tmp_path / "sites.tsv")
    with open(sites, "w") as f:
        for rid, pos1, alleles in recs:
            f.write("%s\t%d\t%s\t%s\n" % (names[rid], pos1, alleles[0].decode(), b",".join(alleles[1:]).decode() if len(alleles) > 1 else "."))
    p = subprocess.run([harness, fa, sites], capture_output=True, text=True)
    assert p.returncode == 0, p.stderr[-2000:]
    got = p.stdout.splitlines()
    exp = _expected(oracle, names, seqs, recs)
    assert len(got) == len(exp)
    n_changed = 0
    for g, (chrom, pos1, alleles, vt, code, norm, has_n, tail) in zip(got, exp):
        head, mid, out = [x.strip() for x in g.split("|")]
        t, c, nm, hn = [int(x) for x in mid.split()]
        assert (t, c, bool(hn)) == (vt, code, has_n), (g, vt, code, has_n)
        if norm is not None:
            assert bool(nm) == norm, g
        assert out == tail, (g, tail)
        n_changed += not tail.endswith(" 0 0 0")
    return n_changed, p.stderr


def test_facade_replays_the_c1_fixture(harness, oracle, tmp_path):
    seq = helpers.chr20_reference()
    _, recs = helpers.read_vcf_sites(os.path.join(helpers.GOLDEN, "c1_normalize_in.vcf"), {"20": 0})
    recs = [(r[0], r[1], [a.encode() for a in r[2]]) for r in recs]
    n_changed, _ = _replay(harness, oracle, tmp_path, ["20"], [seq], recs)
    assert n_changed == 80  # the 80 records 01_OUT.vcf carries an OLD_VARIANT for


def test_facade_synthetic_mix_with_masked_and_inconsistent_reference(harness, oracle, tmp_path):
    rng = np.random.default_rng(91)
    names, seqs = helpers.synth_reference(rng, n_contig=3, length=30000, lower=True, iupac=True)
    recs = [r for r in helpers.synth_records(rng, seqs, 1500) if r[0] >= 0]
    n_changed, err = _replay(harness, oracle, tmp_path, names, seqs, recs)
    assert n_changed > 50 and "reference bases not consistent" in err


def test_variant_props_batch(ctx, oracle):
    """vtn_variant_props against the oracle's vtype and a direct restatement of is_normalized / contains_N (variant_manip.cpp:109-222)"""
    rng = np.random.default_rng(92)
    names, seqs = helpers.synth_reference(rng, n_contig=2, length=20000, lower=True)
    recs = helpers.synth_records(rng, seqs, 30000)
    recs += [(0, 5, [b"A"]), (0, 5, [b"a", b"A"]), (0, 5, [b"AC", b"ac"]), (0, 7, [b"N", b"<DEL>"]), (0, 7, [b"A", b"<N>"]), (0, 7, [b"AT", b"A", b"ATT", b"at"])]
    b = Batch.from_records(recs)
    vtype, props = ctx.variant_props(b)
    ctx.load_reference(names, seqs)
    exp = oracle.normalize(seqs, b)
    np.testing.assert_array_equal(vtype, exp.vtype[:b.n_rec])

    def is_norm(al):
        al = [a.upper() for a in al]
        if len(al) < 2:
            return True
        if len(al) == 2 and len(al[0]) == 1 and len(al[1]) == 1:
            return True
        if all(a == al[0] for a in al[1:]):
            return True
        one = any(len(a) == 1 for a in al)
        first = all(a[0] == al[0][0] for a in al)
        last = all(a[-1] == al[0][-1] for a in al)
        return not (last or (not one and first))

    for r, (_, _, al) in enumerate(recs):
        assert bool(props[r] & 1) == is_norm(al), (r, al)
        assert bool(props[r] & 2) == any(b"N" in a and not (i and a.startswith(b"<")) for i, a in enumerate(al)), (r, al)
