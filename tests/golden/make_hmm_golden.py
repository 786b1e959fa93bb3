"""Generates tests/golden/hmm_cases.txt + hmm_answers.txt: known-answer vectors for AHMM / LFHMM / RFHMM
produced by the UNMODIFIED reference classes (oracle/_ref/hmm_ref, built from /root/reference by
oracle/Makefile; harness oracle/hmm_ref_main.cpp) with FlankDetector's parameters
(flank_detector.cpp:44-48).  Run in the build container only; the vectors travel, the reference does not."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hmm_helpers as H

H.build_oracle()
rng = np.random.default_rng(20260106)
cases = H.gen_cases(rng, 90, max_read=160)
cases += H.gen_cases(rng, 6, max_read=1000)
# hand-made corners: exact tracts, homopolymers, flank-only reads, a read >= 1024 (AHMM truncates, ahmm.cpp:476-480)
cases += [
    (H.AHMM, "CAG", "", "CAG" * 7),
    (H.AHMM, "A", "", "A" * 40),
    (H.AHMM, "AC", "", "ACACACATACACACCACAC"),
    (H.AHMM, "ACGT", "", "TTTTTTTT"),
    (H.AHMM, "AT", "", ("AT" * 520)[:1030]),
    (H.LFHMM, "CAG", "ACGTA", "ACGTA" + "CAG" * 12 + "TTTTGGGGAC"),
    (H.LFHMM, "T", "GACCA", "GACCA" + "T" * 30 + "GGCA" * 20),
    (H.LFHMM, "CA", "GGGTG", "GGGTG" + "AAAAAAAAAAAAAA"),
    (H.LFHMM, "CA", "GGGTG", "GGGTG"),
    (H.RFHMM, "CAG", "TTGAC", "ACGTACGGTT" + "CAG" * 12 + "TTGAC"),
    (H.RFHMM, "A", "CGGTC", "GCGCGTGCA" + "A" * 25 + "CGGTC"),
    (H.RFHMM, "AAG", "CGGTC", "AAG" * 20 + "CGGTC"),
    (H.RFHMM, "TG", "CGGTC", "ACACACACACACACAC" + "CGGTC"),
]
T, ans = H.run_ref(cases)
assert len(ans) == len(cases)
here = os.path.dirname(os.path.abspath(__file__))
open(os.path.join(here, "hmm_cases.txt"), "w").write(H.cases_text(cases))
out = H.subprocess.run([H.HMM_REF] + [repr(float(p)) for p in H.FD_PARAMS[:5]], input=H.cases_text(cases).encode(),
                       stdout=H.subprocess.PIPE, stderr=H.subprocess.DEVNULL, check=True).stdout.decode()
open(os.path.join(here, "hmm_answers.txt"), "w").write(out)
print("wrote %d cases" % len(cases))
