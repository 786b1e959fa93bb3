"""CPU-only: libvtn's annotate_indels control logic (vt_b200/csrc/vtn_annot_host.h -- motif census and picking, anchor clipping,
phase choice, slen growth, purity, composition / entropy) driven by the CPU oracles instead of the device
(tests/annot_host_main.cpp), against what the UNMODIFIED reference wrote for the same input: all 34 INFO values of every indel,
which records are dropped, and the exact list of (model, sequence) alignments FlankDetector issues with the positions it reads
back (oracle/_ref/vt_ref_trace; goldens under tests/golden/annot_*).  This is the pin of SURVEY 8 a18 and 8f-3."""
import os
import subprocess

import annot_helpers as A
import helpers

BUILD = os.path.join(helpers.ROOT, "tests", "_build")
EXE = os.path.join(BUILD, "annot_host_main")


def _build():
    os.makedirs(BUILD, exist_ok=True)
    srcs = [os.path.join(helpers.ROOT, "tests", "annot_host_main.cpp"), os.path.join(helpers.ROOT, "vt_b200", "csrc", "vtn_annot_host.h"),
            os.path.join(helpers.ORACLE_DIR, "hmm_oracle.c"), os.path.join(helpers.ORACLE_DIR, "vt_oracle.c")]
    if os.path.exists(EXE) and all(os.path.getmtime(EXE) >= os.path.getmtime(s) for s in srcs):
        return
    for c in ("hmm_oracle", "vt_oracle"):
        subprocess.check_call(["gcc", "-O2", "-w", "-c", "-o", os.path.join(BUILD, c + ".o"), os.path.join(helpers.ORACLE_DIR, c + ".c")])
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-o", EXE, srcs[0], os.path.join(BUILD, "hmm_oracle.o"), os.path.join(BUILD, "vt_oracle.o"), "-lm"])


def test_annotate_control_logic_matches_reference(tmp_path):
    _build()
    trace = str(tmp_path / "trace.txt")
    out = subprocess.run([EXE, A.REF_FA, A.IN_VCF, trace], check=True, capture_output=True, text=True).stdout
    mine = [l.split("\t") for l in out.rstrip("\n").split("\n")]
    recs, ref = A.records(A.IN_VCF), A.records(A.OUT_VCF)
    assert len(mine) == len(recs)
    kept = [(r, m) for r, m in zip(recs, mine) if int(m[1]) & 1]
    assert len(kept) == len(ref) and len(kept) < len(recs)  # the unnormalized record and the N record are dropped
    n_indel = 0
    for (r, m), o in zip(kept, ref):
        assert r[:5] == o[:5]
        oi = A.parse_info(o[7])
        if int(m[1]) & 2:
            n_indel += 1
            mi = A.parse_info(m[3])
            for k, v in mi.items():
                if k.split("_", 1)[-1] in A.FLOAT_TAGS and k != "EX_REPEAT_TRACT":
                    assert A.htslib_float(float(v)) == oi[k], (r[:2], k, v, oi[k])
                else:
                    assert v == oi[k], (r[:2], k, v, oi.get(k))
            assert sum(1 for k in oi if k[:3] in ("EX_", "FZ_") or k == "EXACT_RU_AMBIGUOUS") == len(mi)
        else:
            assert not any(k.startswith("EX_") for k in oi)
    assert n_indel == 180
    assert A.sorted_trace(trace) == A.golden_trace()
