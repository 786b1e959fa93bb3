"""Thin numpy-facing wrappers over the C ABI (include/vtn.h) for tests, bench and tooling.
All arithmetic happens in libvtn.so on the GPU; these classes only hold buffers."""
import ctypes as C

import numpy as np

from . import _lib as L


class VtnError(RuntimeError):
    pass


def _p(a, ty):
    return a.ctypes.data_as(ty)


class Batch:
    """SoA record batch (see vtn_batch in include/vtn.h)."""

    def __init__(self, pos1, rid, allele_beg, allele_off, data):
        self.pos1 = np.ascontiguousarray(pos1, dtype=np.int32)
        self.rid = np.ascontiguousarray(rid, dtype=np.int32)
        self.allele_beg = np.ascontiguousarray(allele_beg, dtype=np.uint32)
        self.allele_off = np.ascontiguousarray(allele_off, dtype=np.uint32)
        self.data = np.ascontiguousarray(data, dtype=np.uint8)
        self.n_rec = int(self.pos1.shape[0])
        assert self.rid.shape[0] == self.n_rec and self.allele_beg.shape[0] == self.n_rec + 1
        self.n_allele = int(self.allele_beg[-1]) if self.n_rec else 0
        assert self.allele_off.shape[0] == self.n_allele + 1
        self.n_bytes = int(self.allele_off[-1]) if self.n_allele else 0

    @staticmethod
    def from_records(records):
        """records: iterable of (rid, pos1, [allele bytes/str, ...])."""
        pos1, rid, abeg, aoff, chunks = [], [], [0], [0], []
        for (c, p, alleles) in records:
            rid.append(c)
            pos1.append(p)
            for a in alleles:
                b = a.encode() if isinstance(a, str) else bytes(a)
                chunks.append(b)
                aoff.append(aoff[-1] + len(b))
            abeg.append(abeg[-1] + len(alleles))
        data = np.frombuffer(b"".join(chunks), dtype=np.uint8) if chunks else np.zeros(0, np.uint8)
        return Batch(np.array(pos1, np.int32), np.array(rid, np.int32), np.array(abeg, np.uint32), np.array(aoff, np.uint32), data)

    def alleles(self, r):
        b, e = int(self.allele_beg[r]), int(self.allele_beg[r + 1])
        return [bytes(self.data[int(self.allele_off[s]):int(self.allele_off[s + 1])]) for s in range(b, e)]

    def as_c(self):
        data = self.data if self.data.size else np.zeros(1, np.uint8)
        self._keep = data
        return L.VtnBatch(
            self.n_rec, self.n_allele, self.n_bytes, _p(self.pos1, L.c_i32p), _p(self.rid, L.c_i32p),
            _p(self.allele_beg, L.c_u32p), _p(self.allele_off, L.c_u32p), _p(data, L.c_u8p))


class NormResult:
    """Host arrays for vtn_norm_result, sized for the worst case of a batch."""

    def __init__(self, n_rec, n_bytes, ctx=None):
        """ctx given: the arrays live in pinned host memory from vtn_host_alloc (async D2H overlaps)."""
        z = (lambda n, dt: ctx.pinned_empty(n, dt)) if ctx is not None else (lambda n, dt: np.zeros(n, dt))
        self.status = z(max(n_rec, 1), np.uint8)
        self.vtype = z(max(n_rec, 1), np.uint8)
        self.chg_rec = z(max(n_rec, 1), np.uint32)
        self.chg_pos1 = z(max(n_rec, 1), np.int32)
        self.chg_lt = z(max(n_rec, 1), np.uint32)
        self.chg_le = z(max(n_rec, 1), np.uint32)
        self.chg_rt = z(max(n_rec, 1), np.uint32)
        self.chg_off = z(n_rec + 1, np.uint64)
        self.chg_bytes = z(max(n_bytes, 1), np.uint8)
        self.c = L.VtnNormResult()
        self.c.status = _p(self.status, L.c_u8p)
        self.c.vtype = _p(self.vtype, L.c_u8p)
        self.c.cap_changed = n_rec
        self.c.cap_changed_bytes = n_bytes
        self.c.chg_rec = _p(self.chg_rec, L.c_u32p)
        self.c.chg_pos1 = _p(self.chg_pos1, L.c_i32p)
        self.c.chg_lt = _p(self.chg_lt, L.c_u32p)
        self.c.chg_le = _p(self.chg_le, L.c_u32p)
        self.c.chg_rt = _p(self.chg_rt, L.c_u32p)
        self.c.chg_off = _p(self.chg_off, L.c_u64p)
        self.c.chg_bytes = _p(self.chg_bytes, L.c_u8p)
        self.n_rec = n_rec

    @property
    def n_changed(self):
        return int(self.c.n_changed)

    @property
    def n_changed_bytes(self):
        return int(self.c.n_changed_bytes)

    @property
    def counters(self):
        """12 counters in normalize.cpp:65-78 order: bi lt rt lt_rt rt_la la | multi ... | no_variants no_refs"""
        k = self.c.counters
        return list(k.bi) + list(k.multi) + [k.no_variants, k.no_refs]

    def changed_alleles(self, k, old_lens):
        """new alleles of compacted entry k given the record's old allele lengths"""
        o = int(self.chg_off[k])
        d = int(self.chg_le[k]) - int(self.chg_rt[k]) - int(self.chg_lt[k])
        out = []
        for l in old_lens:
            nl = l + d
            out.append(bytes(self.chg_bytes[o:o + nl]))
            o += nl
        assert o == int(self.chg_off[k + 1])
        return out


class NormResultBi:
    """Host arrays for vtn_norm_result_bi (compact biallelic layout), sized for the worst case of a batch."""

    def __init__(self, n_rec, n_bytes, ctx=None, want_sv=True):
        z = (lambda n, dt: ctx.pinned_empty(n, dt)) if ctx is not None else (lambda n, dt: np.zeros(n, dt))
        self.sv = z(max(n_rec, 1), np.uint8) if want_sv else None
        self.chg_rec = z(max(n_rec, 1), np.uint32)
        self.chg_pos1 = z(max(n_rec, 1), np.int32)
        self.chg_lt = z(max(n_rec, 1), np.uint32)
        self.chg_le = z(max(n_rec, 1), np.uint32)
        self.chg_rt = z(max(n_rec, 1), np.uint32)
        self.chg_off = z(n_rec + 1, np.uint64)
        self.chg_bytes = z(max(n_bytes, 1), np.uint8)
        self.note_rec = z(max(n_rec, 1), np.uint32)
        self.note_status = z(max(n_rec, 1), np.uint8)
        self.c = L.VtnNormResultBi()
        self.c.sv = _p(self.sv, L.c_u8p) if want_sv else None
        self.c.cap_changed = n_rec
        self.c.cap_changed_bytes = n_bytes
        self.c.cap_notes = n_rec
        self.c.chg_rec = _p(self.chg_rec, L.c_u32p)
        self.c.chg_pos1 = _p(self.chg_pos1, L.c_i32p)
        self.c.chg_lt = _p(self.chg_lt, L.c_u32p)
        self.c.chg_le = _p(self.chg_le, L.c_u32p)
        self.c.chg_rt = _p(self.chg_rt, L.c_u32p)
        self.c.chg_off = _p(self.chg_off, L.c_u64p)
        self.c.chg_bytes = _p(self.chg_bytes, L.c_u8p)
        self.c.note_rec = _p(self.note_rec, L.c_u32p)
        self.c.note_status = _p(self.note_status, L.c_u8p)
        self.n_rec = n_rec

    n_changed = property(lambda self: int(self.c.n_changed))
    n_changed_bytes = property(lambda self: int(self.c.n_changed_bytes))
    n_notes = property(lambda self: int(self.c.n_notes))
    counters = NormResult.counters
    changed_alleles = NormResult.changed_alleles

    def expand(self):
        """(status[n_rec], vtype[n_rec]) as vtn_norm_result would hold them: sv split into its two fields, the status of a
        changed record rebuilt from (lt, le, rt) with normalize.cpp:254-311's category rule, the notes laid over."""
        n = self.n_rec
        sv = self.sv[:n]
        vtype = (sv & 0x3f).astype(np.uint8)
        status = (sv >> 6).astype(np.uint8)
        k = self.n_changed
        lt, le, rt = (x[:k].astype(np.int64) for x in (self.chg_lt, self.chg_le, self.chg_rt))
        cat = np.where(le > 0, np.where(rt > le, 4, 5), np.where((lt > 0) & (rt > le), 3, np.where(lt > 0, 1, np.where(rt > le, 2, 0))))
        status[self.chg_rec[:k]] |= (0x08 | 0x10 | (cat << 5)).astype(np.uint8)
        m = self.n_notes
        status[self.note_rec[:m]] = self.note_status[:m]
        return status, vtype


class BiBatch:
    """A biallelic Batch re-packed into the compact layout (vtn_batch_bi) by the library's host helper; pinned memory."""

    def __init__(self, ctx, batch):
        self.ctx = ctx
        self.n_rec, self.n_bytes = batch.n_rec, batch.n_bytes
        self.h = C.c_void_p()
        cb = batch.as_c()
        ctx._chk(ctx.lib.vtn_bi_host_create(ctx.h, C.byref(cb), C.byref(self.h)), "vtn_bi_host_create")
        self.view = ctx.lib.vtn_bi_host_view(self.h)
        v = self.view.contents
        self.n_run, self.n_ext = int(v.n_run), int(v.n_ext)
        n_tiles = (self.n_rec + 127) // 128
        self.h2d_bytes = 4 * self.n_rec + 2 * self.n_rec + self.n_bytes + 4 * (n_tiles + 1) + 8 * self.n_run + 4 + 12 * self.n_ext

    def close(self):
        if self.h:
            self.ctx.lib.vtn_bi_host_destroy(self.ctx.h, self.h)
            self.h = None


class Context:
    """vtn_ctx: one GPU, its stream pair and the resident reference."""

    def __init__(self, device=0):
        self.lib = L.load()
        h = C.c_void_p()
        rc = self.lib.vtn_ctx_create(C.byref(h), device)
        if rc:
            raise VtnError("vtn_ctx_create: %s" % self.lib.vtn_last_error(None).decode())
        self.h = h

    def close(self):
        if self.h:
            self.lib.vtn_ctx_destroy(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _chk(self, rc, what):
        if rc < 0:
            raise VtnError("%s failed (%d): %s" % (what, rc, self.lib.vtn_last_error(self.h).decode()))
        return rc

    def pinned_empty(self, n, dtype):
        """numpy array over pinned host memory owned by the library (never freed before the ctx)"""
        dt = np.dtype(dtype)
        p = C.c_void_p()
        nbytes = max(int(n) * dt.itemsize, 1)
        self._chk(self.lib.vtn_host_alloc(self.h, nbytes, C.byref(p)), "vtn_host_alloc")
        buf = (C.c_uint8 * nbytes).from_address(p.value)
        a = np.frombuffer(buf, dtype=dt, count=int(n))
        a[...] = 0
        return a

    # ---- reference ----
    def load_fasta(self, path):
        self._chk(self.lib.vtn_ref_load_fasta(self.h, path.encode()), "vtn_ref_load_fasta")

    def open_fasta(self, path):
        """index only; a contig becomes resident the first time a batch names it (vtn_ref_open_fasta)"""
        self._chk(self.lib.vtn_ref_open_fasta(self.h, path.encode()), "vtn_ref_open_fasta")

    def attach_fasta(self, path, prepack=()):
        """vtn_fasta_open + vtn_fasta_prepack (host only) + vtn_ref_attach: what a tool does while its context is still being created"""
        h = C.c_void_p()
        rc = self.lib.vtn_fasta_open(path.encode(), C.byref(h))
        if rc:
            raise VtnError("vtn_fasta_open: %s" % self.lib.vtn_last_error(None).decode())
        for name in prepack:
            self.lib.vtn_fasta_prepack(h, self.lib.vtn_fasta_contig_id(h, name.encode()))
        self._chk(self.lib.vtn_ref_attach(self.h, h), "vtn_ref_attach")

    def prefetch(self, rid):
        self._chk(self.lib.vtn_ref_prefetch(self.h, int(rid)), "vtn_ref_prefetch")

    def require(self, rid=None):
        if rid is None:
            self._chk(self.lib.vtn_ref_require_all(self.h), "vtn_ref_require_all")
        else:
            self._chk(self.lib.vtn_ref_require(self.h, int(rid)), "vtn_ref_require")

    def load_reference(self, names, seqs):
        """seqs: list of bytes / uint8 arrays with the raw FASTA bases of each contig"""
        arrs = [np.frombuffer(s, dtype=np.uint8) if isinstance(s, (bytes, bytearray)) else np.ascontiguousarray(s, np.uint8) for s in seqs]
        n = len(arrs)
        cn = (C.c_char_p * max(n, 1))(*[x.encode() for x in names])
        ln = np.array([a.shape[0] for a in arrs], np.int64)
        ptr = (C.c_void_p * max(n, 1))(*[a.ctypes.data if a.size else None for a in arrs])
        self._chk(self.lib.vtn_ref_load_memory(self.h, n, cn, _p(ln, L.c_i64p), ptr), "vtn_ref_load_memory")

    def contig_id(self, name):
        return self.lib.vtn_ref_contig_id(self.h, name.encode())

    def fetch(self, rid, beg0, end0, upper=False):
        cap = max(end0 - beg0 + 1, 1)
        out = np.zeros(cap, np.uint8)
        n = self._chk(self.lib.vtn_ref_fetch(self.h, rid, beg0, end0, 1 if upper else 0, _p(out, L.c_u8p), cap), "vtn_ref_fetch")
        return bytes(out[:n])

    # ---- normalize / decompose through HOST buffers ----
    def normalize(self, batch, result=None):
        res = result or NormResult(batch.n_rec, batch.n_bytes)
        cb = batch.as_c()
        self._chk(self.lib.vtn_normalize(self.h, C.byref(cb), C.byref(res.c)), "vtn_normalize")
        return res

    def bi_pack(self, batch):
        return BiBatch(self, batch)

    def normalize_bi(self, bi, result=None):
        """vtn_normalize_bi through HOST buffers (compact biallelic layout)"""
        res = result or NormResultBi(bi.n_rec, bi.n_bytes)
        self._chk(self.lib.vtn_normalize_bi(self.h, bi.view, C.byref(res.c)), "vtn_normalize_bi")
        return res

    def upload_bi(self, bi):
        h = C.c_void_p()
        self._chk(self.lib.vtn_dbatch_create_bi(self.h, bi.view, C.byref(h)), "vtn_dbatch_create_bi")
        return h

    def fetch_result_bi(self, db, bi):
        res = NormResultBi(bi.n_rec, bi.n_bytes)
        self._chk(self.lib.vtn_dbatch_fetch_result_bi(self.h, db, C.byref(res.c)), "vtn_dbatch_fetch_result_bi")
        return res

    def decompose(self, batch, normalize=False):
        bound = int(self.lib.vtn_decompose_out_bound(C.byref(batch.as_c())))
        src = np.zeros(max(bound, 1), np.uint32)
        alt = np.zeros(max(bound, 1), np.uint16)
        n_out = C.c_uint64(0)
        dc = L.VtnDecompCounters()
        cb = batch.as_c()
        if not normalize:
            self._chk(self.lib.vtn_decompose(self.h, C.byref(cb), _p(src, L.c_u32p), _p(alt, L.c_u16p), C.byref(n_out), C.byref(dc)), "vtn_decompose")
            res = None
        else:
            # output bytes bound: every output record carries REF + one ALT
            nb = 0
            if batch.n_rec:
                lens = np.diff(batch.allele_off.astype(np.int64))
                na = np.diff(batch.allele_beg.astype(np.int64))
                ref_len = lens[batch.allele_beg[:-1].astype(np.int64)]
                nb = int((np.maximum(na - 1, 1) * ref_len).sum() + lens.sum())
            res = NormResult(bound, nb)
            self._chk(self.lib.vtn_decompose_normalize(self.h, C.byref(cb), _p(src, L.c_u32p), _p(alt, L.c_u16p), C.byref(n_out), C.byref(dc), C.byref(res.c)), "vtn_decompose_normalize")
        k = int(n_out.value)
        counters = [dc.no_variants, dc.no_biallelic, dc.no_multiallelic, dc.no_additional_biallelic]
        return src[:k], alt[:k], counters, res

    def uniq(self, batch):
        """vt uniq plan (include/vtn.h vtn_uniq): -> (order, first, [no_total_variants, no_unique_variants], not_in_order)"""
        order = np.zeros(max(batch.n_rec, 1), np.uint32)
        first = np.zeros(max(batch.n_rec, 1), np.uint8)
        uc = L.VtnUniqCounters()
        bad = C.c_int64(-1)
        cb = batch.as_c()
        self._chk(self.lib.vtn_uniq(self.h, C.byref(cb), _p(order, L.c_u32p), _p(first, L.c_u8p), C.byref(uc), C.byref(bad)), "vtn_uniq")
        return order[:batch.n_rec], first[:batch.n_rec], [uc.no_total_variants, uc.no_unique_variants], int(bad.value)

    def writer_order(self, rid, pos1, window, want_late=True, want_pending=False):
        """BCFOrderedWriter's output order (include/vtn.h vtn_writer_order) -> (order, late_front or None[, n_pending])"""
        rid = np.ascontiguousarray(rid, np.int32)
        pos1 = np.ascontiguousarray(pos1, np.int32)
        n = int(rid.shape[0])
        order = np.zeros(max(n, 1), np.uint32)
        late = np.zeros(max(n, 1), np.int32) if want_late else None
        pend = C.c_uint32(0)
        self._chk(self.lib.vtn_writer_order(self.h, n, _p(rid, L.c_i32p), _p(pos1, L.c_i32p), int(window), _p(order, L.c_u32p),
                                            _p(late, L.c_i32p) if want_late else None, C.byref(pend)), "vtn_writer_order")
        if want_pending:
            return order[:n], (late[:n] if want_late else None), int(pend.value)
        return order[:n], (late[:n] if want_late else None)

    def writer_order_resident(self, db, window):
        self._chk(self.lib.vtn_writer_order_resident(self.h, db, int(window)), "vtn_writer_order_resident")

    def fetch_writer_order(self, db, n_rec):
        order = np.zeros(max(n_rec, 1), np.uint32)
        late = np.zeros(max(n_rec, 1), np.int32)
        pend = C.c_uint32(0)
        self._chk(self.lib.vtn_dbatch_fetch_writer_order(self.h, db, _p(order, L.c_u32p), _p(late, L.c_i32p), C.byref(pend)), "vtn_dbatch_fetch_writer_order")
        return order[:n_rec], late[:n_rec], int(pend.value)

    @staticmethod
    def _vcf_fields(cap_rec, cap_allele, cap_bytes):
        f = L.VtnVcfFields()
        f.cap_rec, f.cap_allele, f.cap_bytes = cap_rec, cap_allele, cap_bytes
        arr = {"line_off": np.zeros(cap_rec + 1, np.uint64), "chrom_len": np.zeros(max(cap_rec, 1), np.uint32), "pos1": np.zeros(max(cap_rec, 1), np.int32),
               "tail_off": np.zeros(max(cap_rec, 1), np.uint32), "status": np.zeros(max(cap_rec, 1), np.uint8), "allele_beg": np.zeros(cap_rec + 1, np.uint32),
               "allele_off": np.zeros(cap_allele + 1, np.uint32), "bytes": np.zeros(max(cap_bytes, 1), np.uint8)}
        for k, a in arr.items():
            setattr(f, k, a.ctypes.data_as(type(getattr(f, k))))
        return f, arr

    @staticmethod
    def _vcf_result(f, arr):
        r, na, nb = int(f.n_rec), int(f.n_allele), int(f.n_bytes)
        return {"n_rec": r, "n_allele": na, "n_bytes": nb, "line_off": arr["line_off"][:r + 1], "chrom_len": arr["chrom_len"][:r], "pos1": arr["pos1"][:r],
                "tail_off": arr["tail_off"][:r], "status": arr["status"][:r], "allele_beg": arr["allele_beg"][:r + 1],
                "allele_off": arr["allele_off"][:na + 1], "bytes": arr["bytes"][:nb]}

    @staticmethod
    def _vcf_caps(text, cap_rec, cap_allele, cap_bytes):
        cap_rec = text.count(b"\n") + 1 if cap_rec is None else cap_rec
        cap_allele = text.count(b",") + 2 * cap_rec if cap_allele is None else cap_allele
        return cap_rec, cap_allele, (len(text) if cap_bytes is None else cap_bytes)

    def vcf_tokenize(self, text, cap_rec=None, cap_allele=None, cap_bytes=None):
        """VCF record text (bytes) -> dict of the vtn_vcf_fields arrays (include/vtn.h vtn_vcf_tokenize)"""
        f, arr = self._vcf_fields(*self._vcf_caps(text, cap_rec, cap_allele, cap_bytes))
        self._chk(self.lib.vtn_vcf_tokenize(self.h, text, len(text), C.byref(f)), "vtn_vcf_tokenize")
        return self._vcf_result(f, arr)

    def vcf_upload(self, text, cap_rec=None, cap_allele=None, cap_bytes=None):
        caps = self._vcf_caps(text, cap_rec, cap_allele, cap_bytes)
        h = C.c_void_p()
        self._chk(self.lib.vtn_vcf_dtext_create(self.h, text, len(text), caps[0], caps[1], caps[2], C.byref(h)), "vtn_vcf_dtext_create")
        return h, caps

    def vcf_tokenize_resident(self, h):
        self._chk(self.lib.vtn_vcf_tokenize_resident(self.h, h), "vtn_vcf_tokenize_resident")

    def vcf_fetch(self, h, caps):
        f, arr = self._vcf_fields(*caps)
        self._chk(self.lib.vtn_vcf_dtext_fetch(self.h, h, C.byref(f)), "vtn_vcf_dtext_fetch")
        return self._vcf_result(f, arr)

    def vcf_free(self, h):
        self.lib.vtn_vcf_dtext_destroy(self.h, h)

    def blocksub(self, batch, keep_mnv=False, max_mnv_dist=2):
        """decompose_blocksub plan (include/vtn.h vtn_blocksub) -> (src_rec, off, len, [4 counters])"""
        cb = batch.as_c()
        bound = int(self.lib.vtn_blocksub_out_bound(C.byref(cb)))
        src, off, ln = np.zeros(bound, np.uint32), np.zeros(bound, np.uint32), np.zeros(bound, np.uint32)
        n_out = C.c_uint64(0)
        bc = L.VtnBlocksubCounters()
        self._chk(self.lib.vtn_blocksub(self.h, C.byref(cb), 1 if keep_mnv else 0, int(max_mnv_dist), _p(src, L.c_u32p), _p(off, L.c_u32p),
                                        _p(ln, L.c_u32p), bound, C.byref(n_out), C.byref(bc)), "vtn_blocksub")
        k = int(n_out.value)
        return src[:k], off[:k], ln[:k], [bc.no_variants, bc.no_biallelic_blocksub, bc.no_additional_snps, bc.new_no_variants]

    def blocksub_align(self, batch):
        """decompose_blocksub -a (vtn_blocksub_align) -> per record a list of (pos_ref, pos_alt, len_ref, len_alt); [] where the mode does not apply"""
        cb = batch.as_c()
        cap = int(self.lib.vtn_blocksub_align_bound(C.byref(cb)))
        beg = np.zeros(batch.n_rec + 1, np.uint32)
        cnt = np.zeros(max(batch.n_rec, 1), np.uint32)
        arr = [np.zeros(max(cap, 1), np.uint32) for _ in range(4)]
        plan = L.VtnBlocksubAlignPlan(cap, 0, _p(beg, L.c_u32p), _p(cnt, L.c_u32p), *[_p(a, L.c_u32p) for a in arr])
        self._chk(self.lib.vtn_blocksub_align(self.h, C.byref(cb), C.byref(plan)), "vtn_blocksub_align")
        return [[tuple(int(a[k]) for a in arr) for k in range(int(beg[r]), int(beg[r]) + int(cnt[r]))] for r in range(batch.n_rec)]

    def blocksub_resident(self, db, keep_mnv=False, max_mnv_dist=2):
        self._chk(self.lib.vtn_blocksub_resident(self.h, db, 1 if keep_mnv else 0, int(max_mnv_dist)), "vtn_blocksub_resident")

    def uniq_resident(self, db):
        self._chk(self.lib.vtn_uniq_resident(self.h, db), "vtn_uniq_resident")

    def fetch_uniq(self, db, n_rec):
        order = np.zeros(max(n_rec, 1), np.uint32)
        first = np.zeros(max(n_rec, 1), np.uint8)
        uc = L.VtnUniqCounters()
        bad = C.c_int64(-1)
        self._chk(self.lib.vtn_dbatch_fetch_uniq(self.h, db, _p(order, L.c_u32p), _p(first, L.c_u8p), C.byref(uc), C.byref(bad)), "vtn_dbatch_fetch_uniq")
        return order[:n_rec], first[:n_rec], [uc.no_total_variants, uc.no_unique_variants], int(bad.value)

    # ---- device-resident batches ----
    def upload(self, batch):
        h = C.c_void_p()
        cb = batch.as_c()
        self._chk(self.lib.vtn_dbatch_create(self.h, C.byref(cb), C.byref(h)), "vtn_dbatch_create")
        return h

    def free(self, db):
        self.lib.vtn_dbatch_destroy(self.h, db)

    def normalize_resident(self, db):
        self._chk(self.lib.vtn_normalize_resident(self.h, db), "vtn_normalize_resident")

    def decompose_resident(self, db, fuse_normalize):
        self._chk(self.lib.vtn_decompose_resident(self.h, db, 1 if fuse_normalize else 0), "vtn_decompose_resident")

    def fetch_result(self, db, batch):
        res = NormResult(batch.n_rec, batch.n_bytes)
        self._chk(self.lib.vtn_dbatch_fetch_result(self.h, db, C.byref(res.c)), "vtn_dbatch_fetch_result")
        return res

    def variant_props(self, batch):
        """VariantManip's allele-only predicates for a batch -> (vtype, props); props bit 0 is_normalized, bit 1 contains_N"""
        vt = np.zeros(max(batch.n_rec, 1), np.uint8)
        pr = np.zeros(max(batch.n_rec, 1), np.uint8)
        cb = batch.as_c()
        self._chk(self.lib.vtn_variant_props(self.h, C.byref(cb), _p(vt, L.c_u8p), _p(pr, L.c_u8p)), "vtn_variant_props")
        return vt[:batch.n_rec], pr[:batch.n_rec]

    def annotate_indels(self, batch, trace=None, raw=False):
        """vt annotate_indels for a batch (include/vtn.h vtn_annotate_indels) -> dict: flags, vtype, indel_idx, trim_left / trim_right /
        shift per allele slot, vntr (list of dicts with the 34 INFO values per indel), n_alignments."""
        n = batch.n_rec
        flags, vtype = np.zeros(max(n, 1), np.uint8), np.zeros(max(n, 1), np.uint8)
        idx = np.zeros(max(n, 1), np.uint32)
        tl, tr, sh = (np.zeros(batch.n_allele + 1, np.uint32) for _ in range(3))
        vn = (L.VtnVntrAnnot * max(n, 1))()
        cap_str = 4 * batch.n_bytes + 64 * n + 4096
        strs = C.create_string_buffer(cap_str)
        r = L.VtnAnnotResult(_p(flags, L.c_u8p), _p(vtype, L.c_u8p), _p(idx, L.c_u32p), _p(tl, L.c_u32p), _p(tr, L.c_u32p), _p(sh, L.c_u32p),
                             n, cap_str, 0, 0, 0, vn, C.cast(strs, C.c_char_p))
        self._chk(self.lib.vtn_annot_set_trace(self.h, trace.encode() if trace else None), "vtn_annot_set_trace")
        cb = batch.as_c()
        try:
            self._chk(self.lib.vtn_annotate_indels(self.h, C.byref(cb), C.byref(r)), "vtn_annotate_indels")
        finally:
            self.lib.vtn_annot_set_trace(self.h, None)
        if raw:  # counts only (bench): no per-indel dicts
            return {"flags": flags[:n], "vtype": vtype[:n], "indel_idx": idx[:n], "n_indel": int(r.n_indel), "n_alignments": int(r.n_alignments)}
        raw = strs.raw
        out = []
        for k in range(int(r.n_indel)):
            a = vn[k]
            d = {"rec": int(a.rec), "exact_ru_ambiguous": bool(a.exact_ru_ambiguous), "is_large_repeat_tract": bool(a.is_large_repeat_tract)}
            for pre, sd in (("EX", a.exact), ("FZ", a.fuzzy)):
                st = [raw[sd.str_off[i]:sd.str_off[i] + sd.str_len[i]].decode() for i in range(3)]
                d.update({pre + "_MOTIF": st[0], pre + "_RU": st[1], pre + "_BASIS": st[2], pre + "_MLEN": int(sd.mlen), pre + "_BLEN": int(sd.blen),
                          pre + "_REPEAT_TRACT": (int(sd.beg1), int(sd.end1)), pre + "_COMP": tuple(int(x) for x in sd.comp),
                          pre + "_ENTROPY": np.float32(sd.entropy), pre + "_ENTROPY2": np.float32(sd.entropy2), pre + "_KL_DIVERGENCE": np.float32(sd.kl_divergence),
                          pre + "_KL_DIVERGENCE2": np.float32(sd.kl_divergence2), pre + "_REF": np.float32(sd.ref), pre + "_RL": int(sd.rl), pre + "_LL": int(sd.ll),
                          pre + "_RU_COUNTS": (int(sd.no_perfect_ru), int(sd.no_ru)), pre + "_SCORE": np.float32(sd.score), pre + "_TRF_SCORE": int(sd.trf_score)})
            out.append(d)
        return {"flags": flags[:n], "vtype": vtype[:n], "indel_idx": idx[:n], "trim_left": tl[:batch.n_allele], "trim_right": tr[:batch.n_allele],
                "shift": sh[:batch.n_allele], "vntr": out, "n_alignments": int(r.n_alignments)}

    def fetch_batch(self, rid, beg0, end0, upper=True):
        """list of reference windows in one kernel (include/vtn.h vtn_ref_fetch_batch) -> list of bytes"""
        rid = np.ascontiguousarray(rid, np.int32)
        beg0 = np.ascontiguousarray(beg0, np.int64)
        end0 = np.ascontiguousarray(end0, np.int64)
        n = int(rid.shape[0])
        cap = int(np.maximum(end0 - beg0 + 1, 1).sum()) + 1
        off = np.zeros(n + 1, np.uint64)
        out = np.zeros(cap, np.uint8)
        self._chk(self.lib.vtn_ref_fetch_batch(self.h, n, _p(rid, L.c_i32p), _p(beg0, L.c_i64p), _p(end0, L.c_i64p), 1 if upper else 0,
                                               _p(off, L.c_u64p), _p(out, L.c_u8p), cap), "vtn_ref_fetch_batch")
        return [bytes(out[int(off[k]):int(off[k + 1])]) for k in range(n)]

    def sync(self):
        self._chk(self.lib.vtn_sync(self.h), "vtn_sync")

    def timer_start(self):
        self._chk(self.lib.vtn_timer_start(self.h), "vtn_timer_start")

    def timer_stop(self):
        ms = C.c_float(0)
        self._chk(self.lib.vtn_timer_stop(self.h, C.byref(ms)), "vtn_timer_stop")
        return float(ms.value)

    def launch_count(self):
        return int(self.lib.vtn_launch_count(self.h))


def shard_plan(rid, n_shard, align=1):
    """vtn_shard_plan: record cuts of one contiguous range per shard (host code, no device needed)"""
    rid = np.ascontiguousarray(rid, np.int32)
    cut = np.zeros(n_shard + 1, np.uint32)
    rc = L.load().vtn_shard_plan(_p(rid, L.c_i32p), int(rid.shape[0]), int(n_shard), int(align), _p(cut, L.c_u32p))
    if rc:
        raise VtnError("vtn_shard_plan failed (%d)" % rc)
    return cut


class MultiContext:
    """vtn_multi: several devices behind one call (one vtn_ctx per entry of device_ids; ids may repeat)."""

    def __init__(self, device_ids):
        self.lib = L.load()
        ids = np.ascontiguousarray(device_ids, np.int32)
        h = C.c_void_p()
        rc = self.lib.vtn_multi_create(C.byref(h), _p(ids, L.c_i32p), int(ids.shape[0]))
        if rc:
            raise VtnError("vtn_multi_create: %s" % self.lib.vtn_multi_last_error(None).decode())
        self.h = h
        self.n_dev = int(ids.shape[0])
        # slot 0's context through the single-device wrapper: pinned buffers, bi packing, contig ids
        self.ctx0 = Context.__new__(Context)
        self.ctx0.lib = self.lib
        self.ctx0.h = C.c_void_p(self.lib.vtn_multi_ctx(self.h, 0))

    def close(self):
        if self.h:
            self.lib.vtn_multi_destroy(self.h)
            self.h = None
            self.ctx0.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _chk(self, rc, what):
        if rc < 0:
            raise VtnError("%s failed (%d): %s" % (what, rc, self.lib.vtn_multi_last_error(self.h).decode()))
        return rc

    def load_reference(self, names, seqs):
        arrs = [np.frombuffer(s, dtype=np.uint8) if isinstance(s, (bytes, bytearray)) else np.ascontiguousarray(s, np.uint8) for s in seqs]
        n = len(arrs)
        cn = (C.c_char_p * max(n, 1))(*[x.encode() for x in names])
        ln = np.array([a.shape[0] for a in arrs], np.int64)
        ptr = (C.c_void_p * max(n, 1))(*[a.ctypes.data if a.size else None for a in arrs])
        self._chk(self.lib.vtn_multi_ref_load_memory(self.h, n, cn, _p(ln, L.c_i64p), ptr), "vtn_multi_ref_load_memory")

    def normalize_bi(self, bi, result=None):
        res = result or NormResultBi(bi.n_rec, bi.n_bytes)
        self._chk(self.lib.vtn_multi_normalize_bi(self.h, bi.view, C.byref(res.c)), "vtn_multi_normalize_bi")
        return res

    def decompose_normalize(self, batch, bufs=None):
        """-> (src_rec, alt_idx, decompose counters, NormResult indexed by output record); bufs = (src, alt, NormResult) to reuse"""
        if bufs is None:
            bound = int(self.lib.vtn_decompose_out_bound(C.byref(batch.as_c())))
            lens = np.diff(batch.allele_off.astype(np.int64))
            na = np.diff(batch.allele_beg.astype(np.int64))
            ref_len = lens[batch.allele_beg[:-1].astype(np.int64)] if batch.n_rec else np.zeros(0, np.int64)
            nb = int((np.maximum(na - 1, 1) * ref_len).sum() + lens.sum()) if batch.n_rec else 0
            bufs = (np.zeros(max(bound, 1), np.uint32), np.zeros(max(bound, 1), np.uint16), NormResult(bound, nb))
        src, alt, res = bufs
        n_out = C.c_uint64(0)
        dc = L.VtnDecompCounters()
        cb = batch.as_c()
        self._chk(self.lib.vtn_multi_decompose_normalize(self.h, C.byref(cb), _p(src, L.c_u32p), _p(alt, L.c_u16p), C.byref(n_out), C.byref(dc), C.byref(res.c)),
                  "vtn_multi_decompose_normalize")
        k = int(n_out.value)
        return src[:k], alt[:k], [dc.no_variants, dc.no_biallelic, dc.no_multiallelic, dc.no_additional_biallelic], res


# ---------------------------------------------------------------------------------------------------------
# HMM family (AHMM / LFHMM / RFHMM): host-side mirror of set_model + align + the result members
# ---------------------------------------------------------------------------------------------------------
AHMM, LFHMM, RFHMM = 0, 1, 2
# FlankDetector's parameters (flank_detector.cpp:44-48) and qual 'K' (:74)
FLANK_DETECTOR_PARAMS = (1e-7, 1e-7, 0.01, 0.01, 5.0, 42)

HMM_RESULT_DTYPE = np.dtype([
    ("optimal_score", np.float32), ("optimal_state", np.int32), ("optimal_probe_len", np.int32), ("path_len", np.int32),
    ("pos", np.int32, (12,)), ("motif_count", np.int32), ("exact_motif_count", np.int32),
    ("motif_concordance", np.float32), ("trf_score", np.int32), ("frac_no_repeats", np.float32),
    ("last_motif_pos", np.int32), ("status", np.int32), ("reserved", np.int32)])
assert HMM_RESULT_DTYPE.itemsize == C.sizeof(L.VtnHmmResult) == 96


class HmmBatch:
    """Alignments as (kind, motif, flank, read) with flank = '' for AHMM; packed as vtn_hmm_batch."""

    def __init__(self, cases):
        self.n = len(cases)
        self.kind = np.array([c[0] for c in cases], dtype=np.uint8)
        parts, off, o = [], [0], 0
        for _, motif, flank, read in cases:
            for s in (motif, flank or "", read):
                b = s.encode() if isinstance(s, str) else bytes(s)
                parts.append(b)
                o += len(b)
                off.append(o)
        self.seq_off = np.array(off, dtype=np.uint32)
        self.data = np.frombuffer(b"".join(parts) + b"\0", dtype=np.uint8).copy()
        self.n_bytes = o

    def as_c(self):
        return L.VtnHmmBatch(self.n, _p(self.kind, L.c_u8p), _p(self.seq_off, L.c_u32p), _p(self.data, L.c_u8p), self.n_bytes)


def hmm_transitions(kind, params=FLANK_DETECTOR_PARAMS):
    """initialize_T's table (7x7, reference state numbering) as libvtn computes it (host code, no device needed)."""
    T = (C.c_float * 49)()
    par = L.VtnHmmParams(*params)
    rc = L.load().vtn_hmm_transitions(kind, C.byref(par), T)
    if rc:
        raise VtnError("vtn_hmm_transitions failed (%d)" % rc)
    return np.frombuffer(T, dtype=np.float32).copy().reshape(7, 7)


def _hmm_align(self, batch, params=FLANK_DETECTOR_PARAMS, path_stride=0):
    """Host-buffer call -> (structured result array [n], path array [n, path_stride] or None)."""
    par = L.VtnHmmParams(*params)
    out = np.zeros(batch.n, dtype=HMM_RESULT_DTYPE)
    path = np.zeros((batch.n, path_stride), dtype=np.int32) if path_stride else None
    cb = batch.as_c()
    self._chk(self.lib.vtn_hmm_align(self.h, C.byref(par), C.byref(cb), out.ctypes.data_as(C.c_void_p),
                                     _p(path, L.c_i32p) if path_stride else None, path_stride), "vtn_hmm_align")
    return out, path


def _hmm_upload(self, batch, params=FLANK_DETECTOR_PARAMS, path_stride=0):
    par = L.VtnHmmParams(*params)
    cb = batch.as_c()
    db = C.c_void_p()
    self._chk(self.lib.vtn_hmm_dbatch_create(self.h, C.byref(par), C.byref(cb), path_stride, C.byref(db)), "vtn_hmm_dbatch_create")
    return db


def _hmm_align_resident(self, db):
    self._chk(self.lib.vtn_hmm_align_resident(self.h, db), "vtn_hmm_align_resident")


def _hmm_fetch(self, db, n, path_stride=0):
    out = np.zeros(n, dtype=HMM_RESULT_DTYPE)
    path = np.zeros((n, path_stride), dtype=np.int32) if path_stride else None
    self._chk(self.lib.vtn_hmm_dbatch_fetch(self.h, db, out.ctypes.data_as(C.c_void_p),
                                            _p(path, L.c_i32p) if path_stride else None), "vtn_hmm_dbatch_fetch")
    return out, path


def _hmm_work(self, db):
    cells, pc = C.c_uint64(), C.c_uint64()
    self._chk(self.lib.vtn_hmm_dbatch_work(self.h, db, C.byref(cells), C.byref(pc)), "vtn_hmm_dbatch_work")
    return cells.value, pc.value


def _hmm_free(self, db):
    self.lib.vtn_hmm_dbatch_destroy(self.h, db)


Context.hmm_align = _hmm_align
Context.hmm_upload = _hmm_upload
Context.hmm_align_resident = _hmm_align_resident
Context.hmm_fetch = _hmm_fetch
Context.hmm_work = _hmm_work
Context.hmm_free = _hmm_free
