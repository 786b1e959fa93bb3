"""GPU: vtn_vcf_tokenize (VCF record text -> SoA batch on the device) through the C ABI against the oracle's per-line
restatement of htslib's parse, every output array bit-exact; and against the reference's golden pair (tests/golden/vcftext_*)."""
import os

import numpy as np
import pytest

import helpers
import vcf_text_helpers as T
from test_vcf_oracle import _body, _ref_columns

pytestmark = pytest.mark.gpu
KEYS = ("line_off", "chrom_len", "pos1", "tail_off", "status", "allele_beg", "allele_off", "bytes")


def _check(ctx, oracle, text):
    got, exp = ctx.vcf_tokenize(text), oracle.vcf_tokenize(text)
    assert (got["n_rec"], got["n_allele"], got["n_bytes"]) == (exp["n_rec"], exp["n_allele"], exp["n_bytes"])
    for k in KEYS:
        np.testing.assert_array_equal(got[k], exp[k], err_msg=k)
    return got


def test_vcf_reference_golden(ctx):
    text = _body(os.path.join(helpers.GOLDEN, "vcftext_in.vcf"))
    f = ctx.vcf_tokenize(text)
    assert T.site_columns(text, f) == _ref_columns(os.path.join(helpers.GOLDEN, "vcftext_out.vcf"))


@pytest.mark.parametrize("seed,n,samples", [(1, 3000, 0), (2, 20000, 0), (3, 4000, 5), (4, 1500, 300)])
def test_vcf_matches_oracle(ctx, oracle, seed, n, samples):
    lines = T.gen_lines(np.random.default_rng(seed), n, samples=samples, long_frac=0.02)
    _check(ctx, oracle, ("\n".join(lines) + "\n").encode())


def test_vcf_edge_cases(ctx, oracle):
    f = ctx.vcf_tokenize(b"")
    assert f["n_rec"] == 0
    _check(ctx, oracle, b"1\t5\t.\tA\tC\t.\t.\t.")                     # one line, no line end
    _check(ctx, oracle, b"1\t5\t.\tA\tC\t.\t.\t.\n")
    _check(ctx, oracle, b"1\t5\t.\tA\tC\t.\t.\t.\r\n1\t6\tx\tAC\t.\t.\t.\tK=1\nbroken\tline\n1\t-3\t.\tacg\tA,,T\t.\t.\t.")
    _check(ctx, oracle, b"\n\n1\t5\t.\tA\tC\t.\t.\t.\n\n")              # empty lines are malformed records (status 1)
    _check(ctx, oracle, b"1\t 77\t.\tA\t\t.\t.\t.\n1\t99999999999999999999\t.\tA\tC\t.\t.\t.\n1\t+8x\t.\t.\t.\t.\t.\t.\n")
    # a 70 kb allele across many tiles, lines that end exactly on tile boundaries
    big = b"1\t10\t.\t" + b"ACGT" * 17500 + b"\t" + b"A," * 300 + b"T\t.\t.\t.\n"
    pad = b"1\t11\t.\tA\tC\t.\t.\t" + b"X" * (8192 - 17 - len(big) % 8192) + b"\n"
    _check(ctx, oracle, big + pad + big)
    text = b"".join(b"1\t%d\t.\tA\tG\t.\t.\tK=%07d\n" % (i, i) for i in range(4096))
    assert len(text) % 32 != 0 or True
    _check(ctx, oracle, text)


def test_vcf_capacity_errors(ctx):
    from vt_b200.api import VtnError
    text = b"1\t5\t.\tA\tC,G\t.\t.\t.\n" * 10
    with pytest.raises(VtnError):
        ctx.vcf_tokenize(text, cap_rec=5)
    with pytest.raises(VtnError):
        ctx.vcf_tokenize(text, cap_allele=7)


def test_vcf_fullsize_properties(ctx):
    """2 M sites-only lines (72 MB): positions, allele counts and bytes against numpy's own parse of the generator's columns"""
    rng = np.random.default_rng(9)
    n = 2_000_000
    pos = np.sort(rng.integers(1, 250_000_000, n))
    bases = np.array(list("ACGT"))
    ref = bases[rng.integers(0, 4, n)]
    alt = bases[rng.integers(0, 4, n)]
    multi = rng.random(n) < 0.1
    lines = ["20\t%d\t.\t%s\t%s%s\t.\tPASS\tK=%d" % (pos[i], ref[i], alt[i], ",TT" if multi[i] else "", i) for i in range(n)]
    text = ("\n".join(lines) + "\n").encode()
    f = ctx.vcf_tokenize(text)
    assert f["n_rec"] == n and not f["status"].any()
    np.testing.assert_array_equal(f["pos1"], pos.astype(np.int32))
    np.testing.assert_array_equal(np.diff(f["allele_beg"].astype(np.int64)), 2 + multi.astype(np.int64))
    assert f["n_bytes"] == 2 * n + 2 * int(multi.sum())
    assert (f["chrom_len"] == 2).all()
    first = f["bytes"][f["allele_off"][f["allele_beg"][:-1]]]
    np.testing.assert_array_equal(first, np.frombuffer("".join(ref).encode(), np.uint8))
    starts = np.frombuffer(text, np.uint8)[f["line_off"][:-1].astype(np.int64)]
    assert (starts == ord("2")).all() and int(f["line_off"][-1]) == len(text)
