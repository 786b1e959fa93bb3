"""CPU: the oracle's restatement of the site-column parse (oracle/vt_oracle.c vto_vcf_tokenize; htslib vcf_parse1 +
bcf_unpack(BCF_UN_STR)) pinned against the UNMODIFIED reference: `vt view` parses every line with htslib and prints the record
back, so CHROM / POS / REF / ALT of its output are what the parse produced.  Golden pair tests/golden/vcftext_* (made by
make_vcftext_golden.py with oracle/_ref/vt_ref); live where the binary exists."""
import os
import subprocess

import numpy as np
import pytest

import helpers
import vcf_text_helpers as T


def _body(path):
    return b"".join(l for l in open(path, "rb") if not l.startswith(b"#"))


def _ref_columns(path):
    return [(f[0], int(f[1]), f[3], f[4]) for f in (l.rstrip("\n").split("\t") for l in open(path) if not l.startswith("#"))]


def test_vcf_oracle_matches_reference_golden():
    text = _body(os.path.join(helpers.GOLDEN, "vcftext_in.vcf"))
    f = helpers.Oracle().vcf_tokenize(text)
    assert not f["status"].any()
    assert T.site_columns(text, f) == _ref_columns(os.path.join(helpers.GOLDEN, "vcftext_out.vcf"))
    # the pass-through tail starts at the QUAL column
    for r in (0, 1, f["n_rec"] - 1):
        o = int(f["line_off"][r])
        line = text[o:int(f["line_off"][r + 1])].rstrip(b"\n").split(b"\t")
        assert text[o + int(f["tail_off"][r]):].startswith(line[5] + b"\t" + line[6])


def test_vcf_oracle_line_ends_and_malformed():
    text = b"1\t5\t.\tA\tC\t.\t.\t.\r\n1\t6\tx\tAC\t.\t.\t.\tK=1\nbroken\tline\n1\t-3\t.\tacg\tA,,T\t.\t.\t."
    f = helpers.Oracle().vcf_tokenize(text)
    assert f["n_rec"] == 4 and list(f["status"]) == [0, 0, 1, 0]
    cols = T.site_columns(text, f)
    assert cols[0] == ("1", 5, "A", "C") and cols[1] == ("1", 6, "AC", ".") and cols[3] == ("1", -3, "acg", "A,,T")
    assert list(f["allele_beg"]) == [0, 2, 3, 3, 7]


@pytest.mark.skipif(not os.path.exists(helpers.VT_REF), reason="reference binary only exists in the build container")
@pytest.mark.parametrize("seed,samples", [(1, 0), (2, 3)])
def test_vcf_oracle_matches_reference_live(seed, samples, tmp_path):
    lines = T.gen_lines(np.random.default_rng(seed), 1500, samples=samples)
    vin, vout = str(tmp_path / "in.vcf"), str(tmp_path / "out.vcf")
    cols = T.COLS + ("\tFORMAT" + "".join("\tS%d" % i for i in range(samples)) if samples else "")
    open(vin, "w").write(T.HDR + cols + "\n" + "\n".join(lines) + "\n")
    p = subprocess.run([helpers.VT_REF, "view", vin, "-o", vout], capture_output=True, text=True)
    assert p.returncode == 0, p.stderr
    text = _body(vin)
    f = helpers.Oracle().vcf_tokenize(text)
    assert T.site_columns(text, f) == _ref_columns(vout)
