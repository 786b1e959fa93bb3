"""CPU: the reduced-form recurrences + traceback the CUDA kernel is built from (vt_b200/csrc/vtn_hmm_core.cuh,
host+device inline code) driven by row-major loops (tests/hmm_core_host.cpp) against the reference's known answers
and the oracle.  This checks the header's logic only; the product path is the CUDA kernel (tests/test_hmm_gpu.py)."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

import hmm_helpers as H

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "_build", "libhmm_core_host.so")


class VtnHmmResult(ctypes.Structure):
    _fields_ = [("optimal_score", ctypes.c_float), ("optimal_state", ctypes.c_int32),
                ("optimal_probe_len", ctypes.c_int32), ("path_len", ctypes.c_int32),
                ("pos", ctypes.c_int32 * 12), ("motif_count", ctypes.c_int32),
                ("exact_motif_count", ctypes.c_int32), ("motif_concordance", ctypes.c_float),
                ("trf_score", ctypes.c_int32), ("frac_no_repeats", ctypes.c_float),
                ("last_motif_pos", ctypes.c_int32), ("status", ctypes.c_int32), ("reserved", ctypes.c_int32)]


@pytest.fixture(scope="module")
def core():
    os.makedirs(os.path.dirname(SO), exist_ok=True)
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-w", "-fPIC", "-shared", "-o", SO,
                           os.path.join(HERE, "hmm_core_host.cpp")])
    lib = ctypes.CDLL(SO)
    par = H.HmmoParams(*H.FD_PARAMS)
    path = (ctypes.c_int32 * 4096)()

    def align(kind, motif, flank, read):
        res = VtnHmmResult()
        m, f, r = motif.encode(), (flank or "").encode(), read.encode()
        lib.hmmcore_align(kind, ctypes.byref(par), m, len(m), f, len(f), r, len(r), ctypes.byref(res), path, 4096)
        if res.status != 0:
            return None
        return H.result_dict(res, list(path[:res.path_len]))

    def transitions(kind):
        T = (ctypes.c_float * 49)()
        lib.hmmcore_transitions(kind, ctypes.byref(par), T)
        return np.frombuffer(T, dtype=np.float32).copy().reshape(7, 7)

    align.transitions = transitions
    return align


def test_core_tables_match_reference(core):
    T, _ = H.parse_answers(open(os.path.join(H.GOLDEN, "hmm_answers.txt")).read())
    for kind in (H.AHMM, H.LFHMM, H.RFHMM):
        assert (core.transitions(kind).view(np.uint32) == T[kind].view(np.uint32)).all()


def test_core_matches_golden_vectors(core):
    cases = H.parse_cases(open(os.path.join(H.GOLDEN, "hmm_cases.txt")).read())
    _, ans = H.parse_answers(open(os.path.join(H.GOLDEN, "hmm_answers.txt")).read())
    for idx, (c, (kind, exp)) in enumerate(zip(cases, ans)):
        got = core(*c)
        assert got is not None, "case %d" % idx
        H.compare(kind, got, exp, "golden case %d %r" % (idx, c[:3]))


def test_core_matches_oracle_random(core):
    rng = np.random.default_rng(11)
    o = H.HmmOracle()
    cases = H.gen_cases(rng, 400, max_read=200) + H.gen_cases(rng, 3, max_read=900)
    for idx, c in enumerate(cases):
        exp = o.align(*c)
        got = core(*c)
        if exp is None:
            assert got is None
            continue
        assert got is not None, "case %d %r" % (idx, c)
        H.compare(c[0], got, exp, "random case %d %r" % (idx, c[:3]))
