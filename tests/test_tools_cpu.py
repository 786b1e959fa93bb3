"""CPU tests of host-side pieces of the tool library that need no GPU: `vtn _floats` (the self-test hook of vtn_tools.cpp) prints, for
every input line, whether canon_float says htslib would print the value as written, and what put_float makes of it.  The pass-through
of untouched records relies on: canon_float(s) implies put_float(s) == s."""
import os
import subprocess

import numpy as np

import helpers

VTN = os.path.join(helpers.ROOT, "vt_b200", "vtn")


def test_canonical_floats_are_fixed_points_of_put_float():
    rng = np.random.default_rng(1)
    vals = []
    for _ in range(200000):
        k = int(rng.integers(0, 6))
        if k == 0:
            v = "%d" % rng.integers(0, 2000000)
        elif k == 1:
            v = "%.*f" % (int(rng.integers(0, 9)), rng.random() * 10 ** int(rng.integers(-5, 7)))
        elif k == 2:
            v = "%.*g" % (int(rng.integers(1, 9)), rng.random() * 10 ** int(rng.integers(-6, 8)))
        elif k == 3:
            v = "-" + "%.*f" % (int(rng.integers(0, 7)), rng.random() * 100)
        elif k == 4:
            v = str(rng.choice([".", "0", "-0", "0.0", "1.", "00", "01", "+1", "1e5", "1E-3", "nan", "inf", "0.0001", "0.00009", "999999", "1000000", "0.1000", "12.50",
                                "-0.5", "5.", "-", "0.25", "0.333333", "0.3333333"]))
        else:
            v = "0." + "0" * int(rng.integers(0, 5)) + "%d" % rng.integers(1, 999999)
        vals.append(v)
    p = subprocess.run([VTN, "_floats"], input="\n".join(vals) + "\n", capture_output=True, text=True)
    assert p.returncode == 0, p.stderr
    out = p.stdout.splitlines()
    assert len(out) == len(vals)
    n_canon = 0
    for v, l in zip(vals, out):
        c, _, t = l.partition("\t")
        if c == "1":
            n_canon += 1
            assert t == v, (v, t)
    assert n_canon > 50000
    # and the expected verdicts on the forms that matter
    exp = {".": 1, "0": 1, "-0": 1, "0.0": 0, "1.": 0, "00": 0, "01": 0, "+1": 0, "1e5": 0, "nan": 0, "0.0001": 1, "0.00009": 0, "999999": 1, "1000000": 0,
           "0.1000": 0, "12.50": 0, "-0.5": 1, "0.25": 1, "0.333333": 1, "0.3333333": 0, "30": 1, "12.5": 1}
    p = subprocess.run([VTN, "_floats"], input="\n".join(exp) + "\n", capture_output=True, text=True)
    got = {k: int(l.split("\t")[0]) for k, l in zip(exp, p.stdout.splitlines())}
    assert got == exp
