"""Test helpers for the HMM family (AHMM / LFHMM / RFHMM): case generator shaped like FlankDetector's calls
(flank_detector.cpp:171-281, 646-712), ctypes wrapper of the CPU restatement oracle/hmm_oracle.c, the
runner/parser for the unmodified reference harness oracle/_ref/hmm_ref.  TEST INFRASTRUCTURE ONLY."""
import ctypes
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
HMM_REF = os.path.join(ORACLE_DIR, "_ref", "hmm_ref")
HMM_ORACLE_SO = os.path.join(ORACLE_DIR, "_ref", "libhmm_oracle.so")
GOLDEN = os.path.join(ROOT, "tests", "golden")

AHMM, LFHMM, RFHMM = 0, 1, 2
KIND_CH = "ALR"
# FlankDetector's fixed parameters (flank_detector.cpp:44-48) and qual 'K' (:74)
FD_PARAMS = (1e-7, 1e-7, 0.01, 0.01, 5.0, 42)

RESULT_FIELDS = ["optimal_score", "optimal_state", "optimal_probe_len", "path_len"] + \
    ["pos%d" % k for k in range(12)] + \
    ["motif_count", "exact_motif_count", "motif_concordance", "trf_score", "frac_no_repeats", "last_motif_pos"]


class HmmoParams(ctypes.Structure):
    _fields_ = [("delta", ctypes.c_float), ("epsilon", ctypes.c_float), ("tau", ctypes.c_float),
                ("eta", ctypes.c_float), ("mismatch_penalty", ctypes.c_float), ("pl", ctypes.c_int32)]


class HmmoResult(ctypes.Structure):
    _fields_ = [("optimal_score", ctypes.c_float), ("optimal_state", ctypes.c_int32),
                ("optimal_probe_len", ctypes.c_int32), ("path_len", ctypes.c_int32),
                ("pos", ctypes.c_int32 * 12), ("motif_count", ctypes.c_int32),
                ("exact_motif_count", ctypes.c_int32), ("motif_concordance", ctypes.c_float),
                ("trf_score", ctypes.c_int32), ("frac_no_repeats", ctypes.c_float),
                ("last_motif_pos", ctypes.c_int32)]


def f2bits(x):
    return int(np.float32(x).view(np.uint32))


def bits2f(b):
    return float(np.uint32(b).view(np.float32))


def build_oracle():
    subprocess.check_call(["make", "-s", "-C", ORACLE_DIR, "_ref/libhmm_oracle.so"])
    if os.path.isdir("/root/reference"):
        subprocess.check_call(["make", "-s", "-j8", "-C", ORACLE_DIR, "_ref/hmm_ref"])


class HmmOracle:
    """One reused 'object' per model kind, like FlankDetector's ahmm/lfhmm/rfhmm members: stale fields carry
    over between calls exactly as in the reference (and in hmm_ref, which starts them from the same values)."""

    def __init__(self, params=FD_PARAMS):
        if not os.path.exists(HMM_ORACLE_SO):
            build_oracle()
        self.lib = ctypes.CDLL(HMM_ORACLE_SO)
        self.lib.hmmo_align.restype = ctypes.c_int
        self.par = HmmoParams(*params)
        self.state = [HmmoResult() for _ in range(3)]
        for s in self.state:
            s.last_motif_pos = -1
        self.path = (ctypes.c_int32 * 4096)()

    def transitions(self, kind):
        T = (ctypes.c_float * 49)()
        self.lib.hmmo_transitions(kind, ctypes.byref(self.par), T)
        return np.frombuffer(T, dtype=np.float32).copy().reshape(7, 7)

    def align(self, kind, motif, flank, read):
        """-> dict of RESULT_FIELDS (+ 'path': list of int tracks), or None where the reference would die."""
        res = self.state[kind]
        rc = self.lib.hmmo_align(kind, ctypes.byref(self.par), motif.encode(), (flank or "").encode(),
                                 read.encode(), ctypes.byref(res), self.path)
        if rc != 0:
            return None
        return result_dict(res, list(self.path[:res.path_len]))


def result_dict(res, path):
    d = {"optimal_score": f2bits(res.optimal_score), "optimal_state": res.optimal_state,
         "optimal_probe_len": res.optimal_probe_len, "path_len": res.path_len,
         "motif_count": res.motif_count, "exact_motif_count": res.exact_motif_count,
         "motif_concordance": f2bits(res.motif_concordance), "trf_score": res.trf_score,
         "frac_no_repeats": f2bits(res.frac_no_repeats), "last_motif_pos": res.last_motif_pos, "path": path}
    for k in range(12):
        d["pos%d" % k] = res.pos[k]
    return d


def parse_answer(line):
    """One hmm_ref answer line -> (kind, dict); floats stay as bit patterns."""
    f = line.split()
    kind = KIND_CH.index(f[0])
    d = {"optimal_score": int(f[1], 16), "optimal_state": int(f[2]), "optimal_probe_len": int(f[3]),
         "path_len": int(f[4])}
    for k in range(12):
        d["pos%d" % k] = int(f[5 + k])
    d["motif_count"] = int(f[17])
    d["exact_motif_count"] = int(f[18])
    d["motif_concordance"] = int(f[19], 16)
    d["trf_score"] = int(f[20])
    d["frac_no_repeats"] = int(f[21], 16)
    d["last_motif_pos"] = int(f[22])
    d["path"] = [int(x, 16) - (1 << 32) if int(x, 16) >= (1 << 31) else int(x, 16) for x in f[23:]]
    return kind, d


def parse_answers(text):
    T, ans = {}, []
    for line in text.splitlines():
        if not line.strip():
            continue
        if line.startswith("T "):
            f = line.split()
            n = 6 if f[1] == "A" else 7
            m = np.full((7, 7), -np.inf, dtype=np.float32)
            m[:n, :n] = np.array([int(x, 16) for x in f[2:]], dtype=np.uint32).view(np.float32).reshape(n, n)
            T[KIND_CH.index(f[1])] = m
        else:
            ans.append(parse_answer(line))
    return T, ans


def cases_text(cases):
    return "".join("%s %s %s %s\n" % (KIND_CH[k], m, f if f else "-", r) for k, m, f, r in cases)


def parse_cases(text):
    out = []
    for line in text.splitlines():
        f = line.split()
        if len(f) == 4:
            out.append((KIND_CH.index(f[0]), f[1], "" if f[2] == "-" else f[2], f[3]))
    return out


def run_ref(cases, params=FD_PARAMS):
    """Run the unmodified reference classes on the cases (only where oracle/_ref/hmm_ref was built)."""
    out = subprocess.run([HMM_REF] + [repr(float(p)) for p in params[:5]], input=cases_text(cases).encode(),
                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, check=True).stdout.decode()
    return parse_answers(out)


# which fields the reference defines per model (the rest are stale / not members; see oracle/hmm_ref_main.cpp)
def defined_fields(kind, d):
    base = ["optimal_score", "optimal_state", "optimal_probe_len", "path_len", "path"] + ["pos%d" % k for k in range(12)]
    if kind == AHMM:
        return base + ["motif_count", "exact_motif_count", "motif_concordance", "trf_score", "frac_no_repeats",
                       "last_motif_pos"]
    if kind == LFHMM:
        out = base + ["motif_count", "exact_motif_count", "motif_concordance", "trf_score"]
        # last_motif_pos (and frac_no_repeats, which reads it) is set only when the path leaves Z for M/D
        # (lfhmm.cpp:790-811); otherwise it is whatever the previous call left behind
        return out
    # RFHMM: motif_count only when the path goes MR -> M/D (rfhmm.cpp:745-781); checked via pos6 (motif_end model)
    return base + (["motif_count"] if d["pos6"] != -1 else [])


BASES = "ACGT"


def rand_seq(rng, n):
    return "".join(BASES[i] for i in rng.integers(0, 4, n))


def mutate(rng, s, rate):
    out = []
    for ch in s:
        r = rng.random()
        if r < rate / 3:
            continue
        if r < 2 * rate / 3:
            out.append(BASES[rng.integers(0, 4)])
            continue
        out.append(ch)
        if r < rate:
            out.append(BASES[rng.integers(0, 4)])
    return "".join(out)


def gen_cases(rng, n, max_read=400, kinds=(AHMM, LFHMM, RFHMM)):
    """(kind, motif, flank, read) tuples shaped like FlankDetector's: 5 bp flanks, repeat tracts with a few
    interruptions embedded in random sequence; read lengths up to max_read (< 1019 for the flanked models)."""
    cases = []
    for _ in range(n):
        kind = kinds[rng.integers(0, len(kinds))]
        mlen = int(rng.integers(1, 7))
        motif = rand_seq(rng, mlen)
        total = int(rng.integers(8, max_read + 1))
        tract_len = int(rng.integers(1, total + 1))
        tract = (motif * (tract_len // mlen + 2))[int(rng.integers(0, mlen)):][:tract_len]
        tract = mutate(rng, tract, [0.0, 0.01, 0.03, 0.1][rng.integers(0, 4)])
        pad = max(0, total - len(tract))
        if kind == AHMM:
            read = tract if rng.random() < 0.7 else tract + rand_seq(rng, min(pad, 6))
            flank = ""
            if len(read) < 2:
                read = motif + motif
        elif kind == LFHMM:
            flank = rand_seq(rng, 5)
            read = flank + tract + rand_seq(rng, pad)
            if rng.random() < 0.2:
                read = mutate(rng, read[:5], 0.3) + read[5:]
        else:
            flank = rand_seq(rng, 5)
            read = rand_seq(rng, pad) + tract + flank
        cases.append((kind, motif, flank, read[:max_read + 10]))
    return cases


def compare(kind, got, exp, what=""):
    for k in defined_fields(kind, exp):
        if got[k] != exp[k]:
            raise AssertionError("%s: field %s differs: got %r expected %r" % (what, k, got[k] if k != "path" else "path", exp[k] if k != "path" else "path"))
