"""CPU test of vt_b200/csrc/vtn_bcf.h (tests/bcf_main.cpp): BCF2 files written by the unmodified reference (tests/golden/bcf_*.bcf, made by
make_bcf_golden.py) decode to the VCF text the reference prints for them; VCF text encoded to BCF2 (BGZF and raw) decodes back to the
same records, and, where the reference binary exists, `vt view` prints the same for our file as for its own."""
import os
import subprocess

import pytest

import helpers

HERE = os.path.dirname(os.path.abspath(__file__))
STEMS = ["blocksub_02", "c1_decompose", "filter", "blocksub_uneven", "blocksub_03", "vcftext", "c1_normalize", "stress"]


@pytest.fixture(scope="module")
def exe(tmp_path_factory):
    out = tmp_path_factory.mktemp("bcf") / "bcf_main"
    subprocess.run(["g++", "-O1", "-std=c++17", "-pthread", os.path.join(HERE, "bcf_main.cpp"), "-lz", "-o", str(out)], check=True)
    return str(out)


def _records(path):
    return [l for l in open(path) if not l.startswith("##")]


@pytest.mark.parametrize("stem", STEMS)
def test_decode_reference_bcf(exe, tmp_path, stem):
    out = str(tmp_path / "d.vcf")
    subprocess.check_call([exe, "d", os.path.join(helpers.GOLDEN, "bcf_%s.bcf" % stem), out])
    assert _records(out) == _records(os.path.join(helpers.GOLDEN, "bcf_%s_view.vcf" % stem))
    # the header comes out in its VCF form: no IDX attributes
    assert not any(",IDX=" in l for l in open(out) if l.startswith("##"))


@pytest.mark.parametrize("ext", [".bcf", ".ubcf"])
@pytest.mark.parametrize("stem", STEMS)
def test_encode_then_decode(exe, tmp_path, stem, ext):
    src = os.path.join(helpers.GOLDEN, stem + "_in.vcf")
    enc, out = str(tmp_path / ("e" + ext)), str(tmp_path / "d.vcf")
    subprocess.check_call([exe, "c", src, enc])
    head = open(enc, "rb").read(5)
    assert head == (b"BCF\x02\x02" if ext == ".ubcf" else b"\x1f\x8b\x08\x04\x00")
    subprocess.check_call([exe, "d", enc, out])
    assert _records(out) == _records(os.path.join(helpers.GOLDEN, "bcf_%s_view.vcf" % stem))
    if os.path.exists(helpers.VT_REF):
        view = str(tmp_path / "v.vcf")
        subprocess.run([helpers.VT_REF, "view", enc, "-o", view], check=True, capture_output=True)
        assert _records(view) == _records(os.path.join(helpers.GOLDEN, "bcf_%s_view.vcf" % stem))


def test_undefined_tags_fail_loudly(exe, tmp_path):
    src = str(tmp_path / "x.vcf")
    open(src, "w").write("##fileformat=VCFv4.2\n##contig=<ID=1,length=1000>\n#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\n1\t5\t.\tA\tC\t.\t.\tZZ=1\n")
    p = subprocess.run([exe, "c", src, str(tmp_path / "x.bcf")], capture_output=True, text=True)
    assert p.returncode != 0 and "not defined in the header" in p.stderr
