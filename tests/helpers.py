"""Test helpers: ctypes binding of the CPU oracle (oracle/vt_oracle.c), VCF site parsing, synthetic
reference / record generators.  The oracle is the checker only; nothing here is on the product path."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")
ORACLE_DIR = os.path.join(ROOT, "oracle")
ORACLE_SO = os.path.join(ORACLE_DIR, "_ref", "libvt_oracle.so")
VT_REF = os.path.join(ORACLE_DIR, "_ref", "vt_ref")

from vt_b200.api import Batch  # noqa: E402

u8p = C.POINTER(C.c_uint8)
u16p = C.POINTER(C.c_uint16)
u32p = C.POINTER(C.c_uint32)
i32p = C.POINTER(C.c_int32)
u64p = C.POINTER(C.c_uint64)
i64p = C.POINTER(C.c_int64)


class VtoRef(C.Structure):
    _fields_ = [("n_contig", C.c_int32), ("len", i64p), ("seq", C.POINTER(C.c_void_p))]


class VtoBatch(C.Structure):
    _fields_ = [("n_rec", C.c_uint32), ("pos1", i32p), ("rid", i32p), ("allele_beg", u32p), ("allele_off", u32p), ("bytes", u8p)]


class VtoNormOut(C.Structure):
    _fields_ = [
        ("status", u8p), ("vtype", u8p), ("n_changed", C.c_uint32), ("n_changed_bytes", C.c_uint64),
        ("chg_rec", u32p), ("chg_pos1", i32p), ("chg_lt", u32p), ("chg_le", u32p), ("chg_rt", u32p),
        ("chg_off", u64p), ("chg_bytes", u8p), ("counters", C.c_uint32 * 12),
    ]


def build_oracle():
    if not os.path.exists(ORACLE_SO) or os.path.getmtime(ORACLE_SO) < os.path.getmtime(os.path.join(ORACLE_DIR, "vt_oracle.c")):
        subprocess.check_call(["make", "-s", "-C", ORACLE_DIR, "_ref/libvt_oracle.so"])


class OracleResult:
    pass


class Oracle:
    def __init__(self):
        build_oracle()
        self.lib = C.CDLL(ORACLE_SO)
        self.lib.vto_normalize_batch.restype = C.c_int32
        self.lib.vto_normalize_batch.argtypes = [C.POINTER(VtoRef), C.POINTER(VtoBatch), C.POINTER(VtoNormOut)]
        self.lib.vto_decompose_plan.restype = C.c_int32
        self.lib.vto_decompose_plan.argtypes = [C.POINTER(VtoBatch), u32p, u16p, u64p, u32p]
        self.lib.vto_blocksub_plan.restype = C.c_int32
        self.lib.vto_blocksub_plan.argtypes = [C.POINTER(VtoBatch), C.c_int32, C.c_int32, u32p, u32p, u32p, u64p, u32p]
        self.lib.vto_vcf_tokenize.restype = C.c_int64
        self.lib.vto_vcf_tokenize.argtypes = [C.c_char_p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64, u64p, u32p, i32p, u32p, u8p, u32p, u32p, u8p]
        self.lib.vto_writer_order.restype = C.c_int32
        self.lib.vto_writer_order.argtypes = [C.c_uint32, i32p, i32p, C.c_int32, u32p, i32p]
        self.lib.vto_uniq_plan.restype = C.c_int32
        self.lib.vto_uniq_plan.argtypes = [C.POINTER(VtoBatch), u32p, u8p, u32p, i64p]

    @staticmethod
    def _ref(seqs):
        arrs = [np.frombuffer(s, np.uint8) if isinstance(s, (bytes, bytearray)) else np.ascontiguousarray(s, np.uint8) for s in seqs]
        lens = np.array([a.shape[0] for a in arrs], np.int64)
        ptrs = (C.c_void_p * max(len(arrs), 1))(*[a.ctypes.data if a.size else None for a in arrs])
        r = VtoRef(len(arrs), lens.ctypes.data_as(i64p), ptrs)
        r._keep = (arrs, lens, ptrs)
        return r

    @staticmethod
    def _batch(b):
        data = b.data if b.data.size else np.zeros(1, np.uint8)
        vb = VtoBatch(b.n_rec, b.pos1.ctypes.data_as(i32p), b.rid.ctypes.data_as(i32p), b.allele_beg.ctypes.data_as(u32p),
                      b.allele_off.ctypes.data_as(u32p), data.ctypes.data_as(u8p))
        vb._keep = data
        return vb

    def normalize(self, seqs, b):
        n, nb = max(b.n_rec, 1), max(b.n_bytes, 1)
        o = OracleResult()
        o.status = np.zeros(n, np.uint8)
        o.vtype = np.zeros(n, np.uint8)
        o.chg_rec = np.zeros(n, np.uint32)
        o.chg_pos1 = np.zeros(n, np.int32)
        o.chg_lt = np.zeros(n, np.uint32)
        o.chg_le = np.zeros(n, np.uint32)
        o.chg_rt = np.zeros(n, np.uint32)
        o.chg_off = np.zeros(n + 1, np.uint64)
        o.chg_bytes = np.zeros(nb, np.uint8)
        out = VtoNormOut(o.status.ctypes.data_as(u8p), o.vtype.ctypes.data_as(u8p), 0, 0, o.chg_rec.ctypes.data_as(u32p),
                         o.chg_pos1.ctypes.data_as(i32p), o.chg_lt.ctypes.data_as(u32p), o.chg_le.ctypes.data_as(u32p),
                         o.chg_rt.ctypes.data_as(u32p), o.chg_off.ctypes.data_as(u64p), o.chg_bytes.ctypes.data_as(u8p))
        ref = self._ref(seqs)
        vb = self._batch(b)
        rc = self.lib.vto_normalize_batch(C.byref(ref), C.byref(vb), C.byref(out))
        assert rc == 0
        o.n_changed = int(out.n_changed)
        o.n_changed_bytes = int(out.n_changed_bytes)
        o.counters = list(out.counters)
        return o

    def decompose_plan(self, b):
        bound = int(b.n_allele) + int(b.n_rec) + 1
        src = np.zeros(bound, np.uint32)
        alt = np.zeros(bound, np.uint16)
        n_out = C.c_uint64(0)
        cnt = (C.c_uint32 * 4)()
        vb = self._batch(b)
        rc = self.lib.vto_decompose_plan(C.byref(vb), src.ctypes.data_as(u32p), alt.ctypes.data_as(u16p), C.byref(n_out), cnt)
        assert rc == 0
        k = int(n_out.value)
        return src[:k], alt[:k], list(cnt)


    def blocksub_plan(self, b, keep_mnv=False, max_mnv_dist=2):
        """-> (src_rec, off, len, [no_variants, no_biallelic_blocksub, no_additional_snps, new_no_variants])"""
        bound = int(b.n_bytes) + int(b.n_rec) + 1
        src, off, ln = np.zeros(bound, np.uint32), np.zeros(bound, np.uint32), np.zeros(bound, np.uint32)
        n_out = C.c_uint64(0)
        cnt = (C.c_uint32 * 4)()
        vb = self._batch(b)
        rc = self.lib.vto_blocksub_plan(C.byref(vb), 1 if keep_mnv else 0, max_mnv_dist, src.ctypes.data_as(u32p), off.ctypes.data_as(u32p),
                                        ln.ctypes.data_as(u32p), C.byref(n_out), cnt)
        assert rc == 0, rc
        k = int(n_out.value)
        return src[:k], off[:k], ln[:k], list(cnt)

    def blocksub_align(self, b):
        """decompose_blocksub -a -> per record a list of (pos_ref, pos_alt, len_ref, len_alt); [] where the mode does not apply"""
        cap = int(b.n_bytes) + 1
        beg, cnt = np.zeros(b.n_rec + 1, np.uint32), np.zeros(max(b.n_rec, 1), np.uint32)
        arr = [np.zeros(cap, np.uint32) for _ in range(4)]
        vb = self._batch(b)
        self.lib.vto_blocksub_align.restype = C.c_int32
        self.lib.vto_blocksub_align.argtypes = [C.POINTER(VtoBatch)] + [u32p] * 6
        rc = self.lib.vto_blocksub_align(C.byref(vb), beg.ctypes.data_as(u32p), cnt.ctypes.data_as(u32p), *[a.ctypes.data_as(u32p) for a in arr])
        assert rc == 0, rc
        return [[tuple(int(a[k]) for a in arr) for k in range(int(beg[r]), int(beg[r]) + int(cnt[r]))] for r in range(b.n_rec)]

    def vcf_tokenize(self, text):
        n = len(text)
        cap_rec = text.count(b"\n") + 1
        cap_allele = text.count(b",") + 2 * cap_rec
        a = {"line_off": np.zeros(cap_rec + 1, np.uint64), "chrom_len": np.zeros(cap_rec, np.uint32), "pos1": np.zeros(cap_rec, np.int32),
             "tail_off": np.zeros(cap_rec, np.uint32), "status": np.zeros(cap_rec, np.uint8), "allele_beg": np.zeros(cap_rec + 1, np.uint32),
             "allele_off": np.zeros(cap_allele + 1, np.uint32), "bytes": np.zeros(max(n, 1), np.uint8)}
        r = self.lib.vto_vcf_tokenize(text, n, cap_rec, cap_allele, n, a["line_off"].ctypes.data_as(u64p), a["chrom_len"].ctypes.data_as(u32p),
                                      a["pos1"].ctypes.data_as(i32p), a["tail_off"].ctypes.data_as(u32p), a["status"].ctypes.data_as(u8p),
                                      a["allele_beg"].ctypes.data_as(u32p), a["allele_off"].ctypes.data_as(u32p), a["bytes"].ctypes.data_as(u8p))
        assert r >= 0
        r = int(r)
        na = int(a["allele_beg"][r])
        nb = int(a["allele_off"][na])
        return {"n_rec": r, "n_allele": na, "n_bytes": nb, "line_off": a["line_off"][:r + 1], "chrom_len": a["chrom_len"][:r], "pos1": a["pos1"][:r],
                "tail_off": a["tail_off"][:r], "status": a["status"][:r], "allele_beg": a["allele_beg"][:r + 1], "allele_off": a["allele_off"][:na + 1],
                "bytes": a["bytes"][:nb]}

    def writer_order(self, rid, pos1, window):
        rid = np.ascontiguousarray(rid, np.int32)
        pos1 = np.ascontiguousarray(pos1, np.int32)
        n = len(rid)
        order = np.zeros(max(n, 1), np.uint32)
        late = np.zeros(max(n, 1), np.int32)
        rc = self.lib.vto_writer_order(n, rid.ctypes.data_as(i32p), pos1.ctypes.data_as(i32p), window, order.ctypes.data_as(u32p), late.ctypes.data_as(i32p))
        assert rc == 0
        return order[:n], late[:n]

    def uniq_plan(self, b):
        """-> (order, first, [no_total_variants, no_unique_variants], not_in_order)"""
        order = np.zeros(max(b.n_rec, 1), np.uint32)
        first = np.zeros(max(b.n_rec, 1), np.uint8)
        cnt = (C.c_uint32 * 2)()
        bad = C.c_int64(-1)
        vb = self._batch(b)
        rc = self.lib.vto_uniq_plan(C.byref(vb), order.ctypes.data_as(u32p), first.ctypes.data_as(u8p), cnt, C.byref(bad))
        assert rc == 0
        k = int(cnt[0])
        return order[:k], first[:k], list(cnt), int(bad.value)


def assert_norm_equal(got, exp, n_rec):
    """got: vt_b200.NormResult (CUDA), exp: OracleResult.  Bit-exact on every array."""
    assert got.n_changed == exp.n_changed, (got.n_changed, exp.n_changed)
    assert got.n_changed_bytes == exp.n_changed_bytes
    np.testing.assert_array_equal(got.status[:n_rec], exp.status[:n_rec])
    np.testing.assert_array_equal(got.vtype[:n_rec], exp.vtype[:n_rec])
    k = exp.n_changed
    for name in ("chg_rec", "chg_pos1", "chg_lt", "chg_le", "chg_rt"):
        np.testing.assert_array_equal(getattr(got, name)[:k], getattr(exp, name)[:k], err_msg=name)
    np.testing.assert_array_equal(got.chg_off[:k + 1], exp.chg_off[:k + 1])
    np.testing.assert_array_equal(got.chg_bytes[:exp.n_changed_bytes], exp.chg_bytes[:exp.n_changed_bytes])
    assert list(got.counters) == list(exp.counters), (got.counters, exp.counters)


# ---------------------------------------------------------------------------------------------
def read_vcf_sites(path, contig_ids):
    """-> (header lines, records) with records = (rid, pos1, [alleles], full fields list)"""
    hdr, recs = [], []
    for line in open(path):
        if line.startswith("#"):
            hdr.append(line)
            continue
        f = line.rstrip("\n").split("\t")
        recs.append((contig_ids.get(f[0], -1), int(f[1]), [f[3]] + f[4].split(","), f))
    return hdr, recs


def chr20_reference():
    """The reconstructed chr20 of the normalize fixture: 63,025,520 x 'N' with the 779 pinned bases."""
    n = int(open(os.path.join(GOLDEN, "chr20.fai")).read().split("\t")[1])
    seq = np.full(n, ord("N"), np.uint8)
    for line in open(os.path.join(GOLDEN, "chr20_pinned.tsv")):
        p, b = line.split()
        seq[int(p) - 1] = ord(b)
    return seq


def write_fasta(path, names, seqs, width=60):
    with open(path, "wb") as f, open(path + ".fai", "w") as fi:
        off = 0
        for name, s in zip(names, seqs):
            s = np.frombuffer(s, np.uint8) if isinstance(s, (bytes, bytearray)) else s
            h = (">%s\n" % name).encode()
            f.write(h)
            off += len(h)
            n = s.shape[0]
            full = n // width
            body = np.empty(n + (n + width - 1) // width, np.uint8)
            if full:
                m = body[: full * (width + 1)].reshape(full, width + 1)
                m[:, :width] = s[: full * width].reshape(full, width)
                m[:, width] = 10
            if n % width:
                body[full * (width + 1): -1] = s[full * width:]
                body[-1] = 10
            f.write(body.tobytes())
            fi.write("%s\t%d\t%d\t%d\t%d\n" % (name, n, off, width, width + 1))
            off += body.shape[0]


def apply_normalize(recs, res, names):
    """Re-create vt normalize's output site list from a result: [(chrom, pos1, ref, alts, old_variant|None)],
    in stable windowed order (bcf_ordered_writer.cpp:141-210 == stable sort by pos within a contig run)."""
    out = []
    k = 0
    for r, (rid, pos1, alleles, f) in enumerate(recs):
        st = int(res.status[r])
        if st & 0x10:
            assert int(res.chg_rec[k]) == r
            new = res.changed_alleles(k, [len(a) for a in alleles]) if hasattr(res, "changed_alleles") else None
            if new is None:
                o = int(res.chg_off[k]); d = int(res.chg_le[k]) - int(res.chg_rt[k]) - int(res.chg_lt[k]); new = []
                for a in alleles:
                    new.append(bytes(res.chg_bytes[o:o + len(a) + d])); o += len(a) + d
            old = "%s:%d:%s" % (f[0], pos1, "/".join(a.upper() for a in alleles))
            out.append((f[0], int(res.chg_pos1[k]), new[0].decode(), [x.decode() for x in new[1:]], old))
            k += 1
        else:
            al = [a.upper() for a in alleles] if st & 0x08 else alleles
            out.append((f[0], pos1, al[0], al[1:], None))
    # stable sort by pos inside runs of the same contig
    res_sorted, i = [], 0
    while i < len(out):
        j = i
        while j < len(out) and out[j][0] == out[i][0]:
            j += 1
        res_sorted += sorted(out[i:j], key=lambda x: x[1])
        i = j
    return res_sorted


# ---------------------------------------------------------------------------------------------
def synth_reference(rng, n_contig=3, length=20000, repeats=True, lower=False, n_runs=True, iupac=False):
    """Small repeat-rich contigs with optional N runs / lower-case / IUPAC bases."""
    seqs = []
    for c in range(n_contig):
        L = int(length * (0.5 + rng.random()))
        s = rng.choice(np.frombuffer(b"ACGT", np.uint8), size=L, p=[0.3, 0.2, 0.2, 0.3])
        if repeats:
            p = 0
            while p < L - 300:
                p += int(rng.integers(20, 400))
                if p >= L:
                    break
                unit = rng.choice(np.frombuffer(b"ACGT", np.uint8), size=int(rng.integers(1, 7)))
                tl = int(rng.integers(4, 200))
                tract = np.resize(unit, tl)
                e = min(L, p + tl)
                s[p:e] = tract[: e - p]
                p = e
        if n_runs:
            for _ in range(3):
                a = int(rng.integers(0, L - 50))
                s[a:a + int(rng.integers(1, 40))] = ord("N")
        if iupac:
            idx = rng.integers(0, L, size=max(L // 500, 1))
            s[idx] = rng.choice(np.frombuffer(b"RYKMSWnrBDHV", np.uint8), size=idx.shape[0])
        if lower:
            for _ in range(5):
                a = int(rng.integers(0, L - 100))
                e = a + int(rng.integers(1, 100))
                s[a:e] = np.where((s[a:e] >= 65) & (s[a:e] <= 90), s[a:e] + 32, s[a:e])
        seqs.append(np.ascontiguousarray(s))
    return ["c%d" % (i + 1) for i in range(n_contig)], seqs


def synth_records(rng, seqs, n, multi_frac=0.2, bad_ref_frac=0.01, symbolic_frac=0.01, lower_frac=0.02, near_start_frac=0.02,
                  max_len=12):
    """Random records: SNVs, MNPs, padded / right-aligned indels in and outside repeats, multiallelics,
    symbolic / breakend ALTs, '.', lower-case alleles, REF mismatches, records at contig ends."""
    A = np.frombuffer(b"ACGT", np.uint8)
    recs = []
    for _ in range(n):
        c = int(rng.integers(0, len(seqs)))
        s = seqs[c]
        L = s.shape[0]
        u = rng.random()
        if u < near_start_frac:
            pos1 = int(rng.integers(1, 6))
        elif u < 2 * near_start_frac:
            pos1 = int(rng.integers(max(L - 5, 1), L + 3))
        else:
            pos1 = int(rng.integers(1, L + 1))
        rl = int(rng.integers(1, max_len)) if rng.random() < 0.6 else 1
        ref = bytes(s[pos1 - 1: pos1 - 1 + rl]) or b"A"
        if len(ref) < rl:
            ref = ref + bytes(rng.choice(A, size=rl - len(ref)))
        n_alt = 1 if rng.random() > multi_frac else int(rng.integers(2, 5))
        alts = []
        for _a in range(n_alt):
            k = rng.random()
            if k < 0.30:  # SNV/MNP-ish: mutate one base
                a = bytearray(ref)
                j = int(rng.integers(0, len(a)))
                a[j] = int(rng.choice(A))
                alts.append(bytes(a))
            elif k < 0.55:  # deletion keeping a prefix of ref (often unnormalized / shiftable)
                keep = int(rng.integers(1, len(ref) + 1))
                alts.append(ref[:keep])
            elif k < 0.80:  # insertion of a copy of following/preceding reference bases
                il = int(rng.integers(1, 8))
                src = bytes(s[pos1 - 1 + len(ref): pos1 - 1 + len(ref) + il]) if rng.random() < 0.5 else bytes(rng.choice(A, size=il))
                at = int(rng.integers(1, len(ref) + 1))
                alts.append(ref[:at] + src + ref[at:])
            elif k < 0.90:  # suffix-sharing form
                il = int(rng.integers(1, 6))
                alts.append(bytes(rng.choice(A, size=il)) + ref[-1:])
            elif k < 0.90 + symbolic_frac * 5:
                alts.append(rng.choice([b"<DEL>", b"<VNTR>", b"<STR>", b"<ST12>", b"<VN3>", b"<VNx>", b"<ST1.5>", b"<STx>", b"A[c1:5[",
                                        b"]c1:7]T", b".", b"*", b"<X>", b"N", b"<INS:ME>"]).item())
            else:
                alts.append(bytes(rng.choice(A, size=int(rng.integers(1, max_len)))))
        if rng.random() < bad_ref_frac:
            r = bytearray(ref)
            r[int(rng.integers(0, len(r)))] = int(rng.choice(np.frombuffer(b"ACGTN", np.uint8)))
            ref = bytes(r)
        if rng.random() < lower_frac:
            ref = ref.lower()
            alts = [a.lower() if rng.random() < 0.7 else a for a in alts]
        rid = c if rng.random() > 0.003 else -1
        recs.append((rid, pos1, [ref] + alts))
    return recs
