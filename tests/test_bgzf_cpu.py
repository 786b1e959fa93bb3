"""CPU test of vt_b200/csrc/vtn_bgzf.h (tests/bgzf_main.cpp): the BGZF writer's output is a valid multi-member gzip with BC subfields and
the end-of-file member; the reader gives back the bytes of BGZF, ordinary gzip (several members) and plain files."""
import gzip
import os
import subprocess

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def exe(tmp_path_factory):
    out = tmp_path_factory.mktemp("bz") / "bgzf_main"
    subprocess.run(["g++", "-O2", "-std=c++17", "-pthread", os.path.join(HERE, "bgzf_main.cpp"), "-lz", "-o", str(out)], check=True)
    return str(out)


@pytest.mark.parametrize("level", ["-1", "0", "1"])
@pytest.mark.parametrize("size", [0, 1, 0xff00, 0xff01, 5_000_000])
def test_bgzf_roundtrip(exe, tmp_path, level, size):
    rng = np.random.default_rng(size + 1)
    data = (b"1\t12345\t.\tACGT\tA\t.\t.\tK=1\n" * (size // 27 + 1))[:size // 2] + rng.integers(0, 256, size - size // 2, dtype=np.uint8).tobytes()
    src, gz, back = str(tmp_path / "in"), str(tmp_path / "x.gz"), str(tmp_path / "out")
    open(src, "wb").write(data)
    subprocess.check_call([exe, "c", src, gz, level])
    raw = open(gz, "rb").read()
    assert gzip.decompress(raw) == data
    p, n = 0, 0
    while p < len(raw):
        assert raw[p:p + 4] == b"\x1f\x8b\x08\x04" and raw[p + 12:p + 16] == b"BC\x02\x00"
        p += raw[p + 16] + 256 * raw[p + 17] + 1
        n += 1
    assert p == len(raw) and n == (size + 0xff00 - 1) // 0xff00 + 1
    assert raw[-28:] == bytes([0x1f, 0x8b, 8, 4, 0, 0, 0, 0, 0, 0xff, 6, 0, 66, 67, 2, 0, 0x1b, 0, 3, 0, 0, 0, 0, 0, 0, 0, 0, 0])
    subprocess.check_call([exe, "d", gz, back])
    assert open(back, "rb").read() == data


def test_reader_takes_gzip_and_plain_and_rejects_damage(exe, tmp_path):
    data = os.urandom(300000) + b"\n" + b"abc\n" * 1000
    src, back = str(tmp_path / "in"), str(tmp_path / "out")
    open(src, "wb").write(data)
    subprocess.check_call([exe, "d", src, back])
    assert open(back, "rb").read() == data
    gz = str(tmp_path / "m.gz")
    open(gz, "wb").write(gzip.compress(data[:1000]) + gzip.compress(data[1000:]))
    subprocess.check_call([exe, "d", gz, back])
    assert open(back, "rb").read() == data
    bz = str(tmp_path / "b.gz")
    subprocess.check_call([exe, "c", src, bz])
    raw = bytearray(open(bz, "rb").read())
    raw[len(raw) // 2] ^= 0x55
    open(bz, "wb").write(bytes(raw))
    assert subprocess.run([exe, "d", bz, back]).returncode != 0
