"""GPU parity of the HMM kernels (vt_b200/csrc/vtn_hmm.cu) through the C ABI (vtn_hmm_align): bit-exact scores,
exact paths / positions / counts against (a) the known answers the UNMODIFIED reference classes produced
(tests/golden/hmm_*.txt) and (b) the CPU oracle on seeded random FlankDetector-shaped inputs, including
multi-band (read > 32), maximum-size (1023) and rejected inputs."""
import os

import numpy as np
import pytest

import hmm_helpers as H
from vt_b200 import api as A


def _res_dict(row, path):
    d = {"optimal_score": int(row["optimal_score"].view(np.uint32)), "optimal_state": int(row["optimal_state"]),
         "optimal_probe_len": int(row["optimal_probe_len"]), "path_len": int(row["path_len"]),
         "motif_count": int(row["motif_count"]), "exact_motif_count": int(row["exact_motif_count"]),
         "motif_concordance": int(row["motif_concordance"].view(np.uint32)), "trf_score": int(row["trf_score"]),
         "frac_no_repeats": int(row["frac_no_repeats"].view(np.uint32)), "last_motif_pos": int(row["last_motif_pos"]),
         "path": [int(x) for x in path[:int(row["path_len"])]]}
    for k in range(12):
        d["pos%d" % k] = int(row["pos"][k])
    return d


def _run(ctx, cases, stride=2100):
    out, path = ctx.hmm_align(A.HmmBatch(cases), path_stride=stride)
    return out, path


def test_hmm_transitions_cpu_side():
    # host code of libvtn (no device): same tables as the reference classes
    T, _ = H.parse_answers(open(os.path.join(H.GOLDEN, "hmm_answers.txt")).read())
    for kind in (A.AHMM, A.LFHMM, A.RFHMM):
        assert (A.hmm_transitions(kind).view(np.uint32) == T[kind].view(np.uint32)).all()


@pytest.mark.gpu
def test_hmm_golden_vectors(ctx):
    cases = H.parse_cases(open(os.path.join(H.GOLDEN, "hmm_cases.txt")).read())
    _, ans = H.parse_answers(open(os.path.join(H.GOLDEN, "hmm_answers.txt")).read())
    out, path = _run(ctx, cases)
    for idx, (c, (kind, exp)) in enumerate(zip(cases, ans)):
        assert out[idx]["status"] == 0, "case %d status %d" % (idx, out[idx]["status"])
        H.compare(kind, _res_dict(out[idx], path[idx]), exp, "golden case %d %r" % (idx, c[:3]))


@pytest.mark.gpu
def test_hmm_random_vs_oracle(ctx):
    rng = np.random.default_rng(20260107)
    cases = H.gen_cases(rng, 1500, max_read=300) + H.gen_cases(rng, 40, max_read=1000)
    out, path = _run(ctx, cases)
    o = H.HmmOracle()
    n_checked = 0
    for idx, c in enumerate(cases):
        exp = o.align(*c)
        if exp is None:
            assert out[idx]["status"] != 0
            continue
        assert out[idx]["status"] == 0, "case %d status %d %r" % (idx, out[idx]["status"], c[:3])
        H.compare(c[0], _res_dict(out[idx], path[idx]), exp, "random case %d %r len %d" % (idx, c[:3], len(c[3])))
        n_checked += 1
    assert n_checked > 1500


@pytest.mark.gpu
def test_hmm_limits_and_rejections(ctx):
    cases = [
        (A.AHMM, "AT", "", ("AT" * 600)[:1100]),            # truncated to 1023 (ahmm.cpp:476-480)
        (A.LFHMM, "CA", "ACGTA", "ACGTA" + "CA" * 509),     # rlen 1023: plen 1028 leaves the 1024 rows
        (A.LFHMM, "CA", "ACGTA", "ACGTA" + "CA" * 505),     # largest that fits
        (A.RFHMM, "CA", "TTGAC", "CA" * 505 + "TTGAC"),
        (A.LFHMM, "CA", "ACGTA", "A" * 1024),               # lfhmm.cpp:498-502
        (A.AHMM, "", "", "ACGT"),                           # empty motif
        (A.AHMM, "A", "", "A"),                             # 1x1
    ]
    out, path = _run(ctx, cases)
    o = H.HmmOracle()
    assert out[0]["status"] == 0 and out[0]["optimal_probe_len"] <= 1023
    H.compare(A.AHMM, _res_dict(out[0], path[0]), o.align(*cases[0]), "truncated AHMM")
    assert out[1]["status"] == 1 and out[4]["status"] == 1
    for k in (2, 3, 6):
        assert out[k]["status"] == 0
        H.compare(cases[k][0], _res_dict(out[k], path[k]), o.align(*cases[k]), "limit case %d" % k)
    assert out[5]["status"] == 2


@pytest.mark.gpu
def test_hmm_resident_rerun_is_idempotent(ctx):
    rng = np.random.default_rng(3)
    cases = H.gen_cases(rng, 300, max_read=200)
    b = A.HmmBatch(cases)
    db = ctx.hmm_upload(b, path_stride=0)
    ctx.hmm_align_resident(db)
    r1, _ = ctx.hmm_fetch(db, b.n)
    ctx.hmm_align_resident(db)
    r2, _ = ctx.hmm_fetch(db, b.n)
    cells, pc = ctx.hmm_work(db)
    ctx.hmm_free(db)
    assert r1.tobytes() == r2.tobytes()
    assert cells > 0 and pc > cells
    # paths not requested: everything else must still agree with the host-buffer call
    r3, _ = ctx.hmm_align(b)
    assert r1.tobytes() == r3.tobytes()
