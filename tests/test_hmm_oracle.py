"""CPU: the HMM restatement (oracle/hmm_oracle.c) against the committed known-answer vectors that the UNMODIFIED
reference classes produced (tests/golden/make_hmm_golden.py), and -- only where oracle/_ref/hmm_ref exists,
i.e. in the build container -- a live differential run on fresh random cases."""
import os

import numpy as np
import pytest

import hmm_helpers as H


def _golden():
    cases = H.parse_cases(open(os.path.join(H.GOLDEN, "hmm_cases.txt")).read())
    T, ans = H.parse_answers(open(os.path.join(H.GOLDEN, "hmm_answers.txt")).read())
    assert len(cases) == len(ans) and len(cases) > 100
    return cases, T, ans


def test_transition_tables_match_reference():
    _, T, _ = _golden()
    o = H.HmmOracle()
    for kind in (H.AHMM, H.LFHMM, H.RFHMM):
        assert (o.transitions(kind).view(np.uint32) == T[kind].view(np.uint32)).all()
    # quirk that must survive: RFHMM T[I][M] stays -inf (rfhmm.cpp:197)
    assert np.isneginf(o.transitions(H.RFHMM)[4][2])


def test_oracle_matches_golden_vectors():
    cases, _, ans = _golden()
    o = H.HmmOracle()
    for idx, (c, (kind, exp)) in enumerate(zip(cases, ans)):
        assert kind == c[0]
        got = o.align(*c)
        assert got is not None, "case %d" % idx
        H.compare(kind, got, exp, "golden case %d" % idx)


def test_oracle_rejects_what_the_reference_cannot_align():
    o = H.HmmOracle()
    assert o.align(H.LFHMM, "CA", "ACGTA", "A" * 1024) is None      # lfhmm.cpp:498-502 exit(1)
    assert o.align(H.AHMM, "CA", "", "CA" * 600)["optimal_probe_len"] <= 1023  # ahmm.cpp:476-480 truncation


@pytest.mark.skipif(not os.path.exists(H.HMM_REF), reason="reference harness only exists in the build container")
def test_oracle_matches_live_reference():
    rng = np.random.default_rng(7)
    cases = H.gen_cases(rng, 250, max_read=260) + H.gen_cases(rng, 4, max_read=1000)
    _, ans = H.run_ref(cases)
    o = H.HmmOracle()
    for idx, (c, (kind, exp)) in enumerate(zip(cases, ans)):
        H.compare(kind, o.align(*c), exp, "live case %d" % idx)
