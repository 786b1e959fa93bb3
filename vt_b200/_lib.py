"""ctypes binding of libvtn.so (include/vtn.h).  Plumbing only: every call goes to the CUDA library;
there is no Python or CPU implementation behind these wrappers and loading fails loudly if the
library is missing."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("VTN_LIB_PATH") or os.path.join(_HERE, "libvtn.so")  # override only for kernel experiments

c_u8p = C.POINTER(C.c_uint8)
c_u16p = C.POINTER(C.c_uint16)
c_u32p = C.POINTER(C.c_uint32)
c_i32p = C.POINTER(C.c_int32)
c_u64p = C.POINTER(C.c_uint64)
c_i64p = C.POINTER(C.c_int64)


class VtnBatch(C.Structure):
    _fields_ = [
        ("n_rec", C.c_uint32),
        ("n_allele", C.c_uint32),
        ("n_bytes", C.c_uint64),
        ("pos1", c_i32p),
        ("rid", c_i32p),
        ("allele_beg", c_u32p),
        ("allele_off", c_u32p),
        ("bytes", c_u8p),
    ]


class VtnNormCounters(C.Structure):
    _fields_ = [("bi", C.c_uint32 * 5), ("multi", C.c_uint32 * 5), ("no_variants", C.c_uint32), ("no_refs", C.c_uint32)]


class VtnNormResult(C.Structure):
    _fields_ = [
        ("status", c_u8p),
        ("vtype", c_u8p),
        ("cap_changed", C.c_uint32),
        ("cap_changed_bytes", C.c_uint64),
        ("n_changed", C.c_uint32),
        ("n_changed_bytes", C.c_uint64),
        ("chg_rec", c_u32p),
        ("chg_pos1", c_i32p),
        ("chg_lt", c_u32p),
        ("chg_le", c_u32p),
        ("chg_rt", c_u32p),
        ("chg_off", c_u64p),
        ("chg_bytes", c_u8p),
        ("counters", VtnNormCounters),
    ]


class VtnBlocksubAlignPlan(C.Structure):
    _fields_ = [("cap_chunk", C.c_uint64), ("n_chunks", C.c_uint64), ("chunk_beg", c_u32p), ("n_chunk", c_u32p), ("pos_ref", c_u32p), ("pos_alt", c_u32p),
                ("len_ref", c_u32p), ("len_alt", c_u32p)]


class VtnBatchBi(C.Structure):
    _fields_ = [("n_rec", C.c_uint32), ("n_run", C.c_uint32), ("n_ext", C.c_uint32), ("n_bytes", C.c_uint64), ("pos1", c_i32p),
                ("lens", c_u16p), ("bytes", c_u8p), ("tile_off", c_u32p), ("run_start", c_u32p), ("run_rid", c_i32p),
                ("ext_rec", c_u32p), ("ext_len", c_u32p)]


class VtnNormResultBi(C.Structure):
    _fields_ = [("sv", c_u8p), ("cap_changed", C.c_uint32), ("cap_changed_bytes", C.c_uint64), ("n_changed", C.c_uint32),
                ("n_changed_bytes", C.c_uint64), ("chg_rec", c_u32p), ("chg_pos1", c_i32p), ("chg_lt", c_u32p), ("chg_le", c_u32p),
                ("chg_rt", c_u32p), ("chg_off", c_u64p), ("chg_bytes", c_u8p), ("cap_notes", C.c_uint32), ("n_notes", C.c_uint32),
                ("note_rec", c_u32p), ("note_status", c_u8p), ("counters", VtnNormCounters)]


class VtnVcfFields(C.Structure):
    _fields_ = [("cap_rec", C.c_uint32), ("cap_allele", C.c_uint32), ("cap_bytes", C.c_uint64), ("n_rec", C.c_uint32), ("n_allele", C.c_uint32),
                ("n_bytes", C.c_uint64), ("line_off", C.POINTER(C.c_uint64)), ("chrom_len", C.POINTER(C.c_uint32)), ("pos1", C.POINTER(C.c_int32)),
                ("tail_off", C.POINTER(C.c_uint32)), ("status", C.POINTER(C.c_uint8)), ("allele_beg", C.POINTER(C.c_uint32)),
                ("allele_off", C.POINTER(C.c_uint32)), ("bytes", C.POINTER(C.c_uint8))]


class VtnBlocksubCounters(C.Structure):
    _fields_ = [("no_variants", C.c_uint32), ("no_biallelic_blocksub", C.c_uint32), ("no_additional_snps", C.c_uint32), ("new_no_variants", C.c_uint32)]


class VtnUniqCounters(C.Structure):
    _fields_ = [("no_total_variants", C.c_uint32), ("no_unique_variants", C.c_uint32)]


class VtnDecompCounters(C.Structure):
    _fields_ = [
        ("no_variants", C.c_uint32),
        ("no_biallelic", C.c_uint32),
        ("no_multiallelic", C.c_uint32),
        ("no_additional_biallelic", C.c_uint32),
    ]


class VtnHmmParams(C.Structure):
    _fields_ = [("delta", C.c_float), ("epsilon", C.c_float), ("tau", C.c_float), ("eta", C.c_float),
                ("mismatch_penalty", C.c_float), ("pl", C.c_int32)]


class VtnHmmBatch(C.Structure):
    _fields_ = [("n", C.c_uint32), ("kind", c_u8p), ("seq_off", c_u32p), ("bytes", c_u8p), ("n_bytes", C.c_uint64)]


class VtnHmmResult(C.Structure):
    _fields_ = [("optimal_score", C.c_float), ("optimal_state", C.c_int32), ("optimal_probe_len", C.c_int32),
                ("path_len", C.c_int32), ("pos", C.c_int32 * 12), ("motif_count", C.c_int32),
                ("exact_motif_count", C.c_int32), ("motif_concordance", C.c_float), ("trf_score", C.c_int32),
                ("frac_no_repeats", C.c_float), ("last_motif_pos", C.c_int32), ("status", C.c_int32),
                ("reserved", C.c_int32)]


class VtnVntrSide(C.Structure):
    _fields_ = [("beg1", C.c_int32), ("end1", C.c_int32), ("mlen", C.c_int32), ("blen", C.c_int32), ("comp", C.c_int32 * 4), ("rl", C.c_int32),
                ("ll", C.c_int32), ("no_perfect_ru", C.c_int32), ("no_ru", C.c_int32), ("trf_score", C.c_int32), ("entropy", C.c_float),
                ("entropy2", C.c_float), ("kl_divergence", C.c_float), ("kl_divergence2", C.c_float), ("ref", C.c_float), ("score", C.c_float),
                ("str_off", C.c_uint32 * 3), ("str_len", C.c_uint32 * 3)]


class VtnVntrAnnot(C.Structure):
    _fields_ = [("rec", C.c_uint32), ("exact_ru_ambiguous", C.c_uint8), ("is_large_repeat_tract", C.c_uint8), ("reserved", C.c_uint8 * 2),
                ("exact", VtnVntrSide), ("fuzzy", VtnVntrSide)]


class VtnAnnotResult(C.Structure):
    _fields_ = [("flags", c_u8p), ("vtype", c_u8p), ("indel_idx", c_u32p), ("trim_left", c_u32p), ("trim_right", c_u32p), ("shift", c_u32p),
                ("cap_indel", C.c_uint32), ("cap_str", C.c_uint64), ("n_indel", C.c_uint32), ("n_str", C.c_uint64), ("n_alignments", C.c_uint64),
                ("vntr", C.POINTER(VtnVntrAnnot)), ("strings", C.c_char_p)]


# name -> (restype, argtypes); also the list the CPU test checks against include/vtn.h
SIGNATURES = {
    "vtn_ctx_create": (C.c_int, [C.POINTER(C.c_void_p), C.c_int]),
    "vtn_ctx_destroy": (C.c_int, [C.c_void_p]),
    "vtn_last_error": (C.c_char_p, [C.c_void_p]),
    "vtn_sync": (C.c_int, [C.c_void_p]),
    "vtn_host_alloc": (C.c_int, [C.c_void_p, C.c_size_t, C.POINTER(C.c_void_p)]),
    "vtn_host_free": (C.c_int, [C.c_void_p, C.c_void_p]),
    "vtn_ref_load_fasta": (C.c_int, [C.c_void_p, C.c_char_p]),
    "vtn_blocksub_align_bound": (C.c_uint64, [C.POINTER(VtnBatch)]),
    "vtn_blocksub_align": (C.c_int, [C.c_void_p, C.POINTER(VtnBatch), C.POINTER(VtnBlocksubAlignPlan)]),
    "vtn_variant_props": (C.c_int, [C.c_void_p, C.POINTER(VtnBatch), c_u8p, c_u8p]),
    "vtn_ref_open_fasta": (C.c_int, [C.c_void_p, C.c_char_p]),
    "vtn_ref_require": (C.c_int, [C.c_void_p, C.c_int]),
    "vtn_ref_require_all": (C.c_int, [C.c_void_p]),
    "vtn_fasta_open": (C.c_int, [C.c_char_p, C.POINTER(C.c_void_p)]),
    "vtn_fasta_n_contig": (C.c_int, [C.c_void_p]),
    "vtn_fasta_contig_id": (C.c_int, [C.c_void_p, C.c_char_p]),
    "vtn_fasta_prepack": (C.c_int, [C.c_void_p, C.c_int]),
    "vtn_fasta_close": (None, [C.c_void_p]),
    "vtn_ref_attach": (C.c_int, [C.c_void_p, C.c_void_p]),
    "vtn_ref_prefetch": (C.c_int, [C.c_void_p, C.c_int]),
    "vtn_ref_load_memory": (C.c_int, [C.c_void_p, C.c_int32, C.POINTER(C.c_char_p), c_i64p, C.POINTER(C.c_void_p)]),
    "vtn_ref_n_contig": (C.c_int, [C.c_void_p]),
    "vtn_ref_contig_id": (C.c_int, [C.c_void_p, C.c_char_p]),
    "vtn_ref_contig_len": (C.c_int64, [C.c_void_p, C.c_int]),
    "vtn_ref_contig_name": (C.c_char_p, [C.c_void_p, C.c_int]),
    "vtn_ref_resident_bytes": (C.c_int64, [C.c_void_p]),
    "vtn_ref_fetch": (C.c_int64, [C.c_void_p, C.c_int, C.c_int64, C.c_int64, C.c_int, c_u8p, C.c_int64]),
    "vtn_normalize": (C.c_int, [C.c_void_p, C.POINTER(VtnBatch), C.POINTER(VtnNormResult)]),
    "vtn_dbatch_create": (C.c_int, [C.c_void_p, C.POINTER(VtnBatch), C.POINTER(C.c_void_p)]),
    "vtn_dbatch_destroy": (C.c_int, [C.c_void_p, C.c_void_p]),
    "vtn_normalize_resident": (C.c_int, [C.c_void_p, C.c_void_p]),
    "vtn_dbatch_fetch_result": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(VtnNormResult)]),
    "vtn_bi_host_create": (C.c_int, [C.c_void_p, C.POINTER(VtnBatch), C.POINTER(C.c_void_p)]),
    "vtn_bi_host_view": (C.POINTER(VtnBatchBi), [C.c_void_p]),
    "vtn_bi_host_destroy": (C.c_int, [C.c_void_p, C.c_void_p]),
    "vtn_normalize_bi": (C.c_int, [C.c_void_p, C.POINTER(VtnBatchBi), C.POINTER(VtnNormResultBi)]),
    "vtn_dbatch_create_bi": (C.c_int, [C.c_void_p, C.POINTER(VtnBatchBi), C.POINTER(C.c_void_p)]),
    "vtn_dbatch_fetch_result_bi": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(VtnNormResultBi)]),
    "vtn_decompose_out_bound": (C.c_uint64, [C.POINTER(VtnBatch)]),
    "vtn_decompose": (C.c_int, [C.c_void_p, C.POINTER(VtnBatch), c_u32p, c_u16p, c_u64p, C.POINTER(VtnDecompCounters)]),
    "vtn_decompose_normalize": (
        C.c_int,
        [C.c_void_p, C.POINTER(VtnBatch), c_u32p, c_u16p, c_u64p, C.POINTER(VtnDecompCounters), C.POINTER(VtnNormResult)],
    ),
    "vtn_decompose_resident": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int]),
    "vtn_uniq": (C.c_int, [C.c_void_p, C.POINTER(VtnBatch), c_u32p, c_u8p, C.POINTER(VtnUniqCounters), C.POINTER(C.c_int64)]),
    "vtn_writer_order": (C.c_int, [C.c_void_p, C.c_uint32, c_i32p, c_i32p, C.c_int32, c_u32p, c_i32p, c_u32p]),
    "vtn_writer_order_resident": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32]),
    "vtn_dbatch_fetch_writer_order": (C.c_int, [C.c_void_p, C.c_void_p, c_u32p, c_i32p, c_u32p]),
    "vtn_vcf_tokenize": (C.c_int, [C.c_void_p, C.c_char_p, C.c_uint64, C.POINTER(VtnVcfFields)]),
    "vtn_vcf_dtext_create": (C.c_int, [C.c_void_p, C.c_char_p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint64, C.POINTER(C.c_void_p)]),
    "vtn_vcf_dtext_destroy": (C.c_int, [C.c_void_p, C.c_void_p]),
    "vtn_vcf_tokenize_resident": (C.c_int, [C.c_void_p, C.c_void_p]),
    "vtn_vcf_dtext_fetch": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(VtnVcfFields)]),
    "vtn_blocksub_out_bound": (C.c_uint64, [C.POINTER(VtnBatch)]),
    "vtn_blocksub": (C.c_int, [C.c_void_p, C.POINTER(VtnBatch), C.c_int, C.c_int, c_u32p, c_u32p, c_u32p, C.c_uint64, c_u64p, C.POINTER(VtnBlocksubCounters)]),
    "vtn_blocksub_noalloc": (C.c_int, [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                      C.c_uint64, C.c_void_p, C.c_void_p]),
    "vtn_blocksub_resident": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int]),
    "vtn_dbatch_fetch_blocksub": (C.c_int, [C.c_void_p, C.c_void_p, c_u32p, c_u32p, c_u32p, C.c_uint64, c_u64p, C.POINTER(VtnBlocksubCounters)]),
    "vtn_uniq_resident": (C.c_int, [C.c_void_p, C.c_void_p]),
    "vtn_dbatch_fetch_uniq": (C.c_int, [C.c_void_p, C.c_void_p, c_u32p, c_u8p, C.POINTER(VtnUniqCounters), C.POINTER(C.c_int64)]),
    "vtn_dbatch_fetch_decompose": (C.c_int, [C.c_void_p, C.c_void_p, c_u32p, c_u16p, c_u64p, C.POINTER(VtnDecompCounters)]),
    "vtn_hmm_align": (C.c_int, [C.c_void_p, C.POINTER(VtnHmmParams), C.POINTER(VtnHmmBatch), C.c_void_p, c_i32p, C.c_uint32]),
    "vtn_hmm_dbatch_create": (C.c_int, [C.c_void_p, C.POINTER(VtnHmmParams), C.POINTER(VtnHmmBatch), C.c_uint32, C.POINTER(C.c_void_p)]),
    "vtn_hmm_dbatch_destroy": (C.c_int, [C.c_void_p, C.c_void_p]),
    "vtn_hmm_align_resident": (C.c_int, [C.c_void_p, C.c_void_p]),
    "vtn_hmm_dbatch_fetch": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, c_i32p]),
    "vtn_hmm_transitions": (C.c_int, [C.c_int, C.POINTER(VtnHmmParams), C.POINTER(C.c_float)]),
    "vtn_hmm_dbatch_work": (C.c_int, [C.c_void_p, C.c_void_p, c_u64p, c_u64p]),
    "vtn_annotate_indels": (C.c_int, [C.c_void_p, C.POINTER(VtnBatch), C.POINTER(VtnAnnotResult)]),
    "vtn_annot_set_trace": (C.c_int, [C.c_void_p, C.c_char_p]),
    "vtn_ref_fetch_batch": (C.c_int, [C.c_void_p, C.c_uint32, c_i32p, c_i64p, c_i64p, C.c_int, c_u64p, c_u8p, C.c_uint64]),
    "vtn_multi_create": (C.c_int, [C.POINTER(C.c_void_p), c_i32p, C.c_int]),
    "vtn_multi_destroy": (C.c_int, [C.c_void_p]),
    "vtn_multi_n_dev": (C.c_int, [C.c_void_p]),
    "vtn_multi_ctx": (C.c_void_p, [C.c_void_p, C.c_int]),
    "vtn_multi_last_error": (C.c_char_p, [C.c_void_p]),
    "vtn_multi_ref_load_fasta": (C.c_int, [C.c_void_p, C.c_char_p]),
    "vtn_multi_ref_load_memory": (C.c_int, [C.c_void_p, C.c_int32, C.POINTER(C.c_char_p), c_i64p, C.POINTER(C.c_void_p)]),
    "vtn_shard_plan": (C.c_int, [c_i32p, C.c_uint32, C.c_uint32, C.c_uint32, c_u32p]),
    "vtn_multi_normalize_bi": (C.c_int, [C.c_void_p, C.POINTER(VtnBatchBi), C.POINTER(VtnNormResultBi)]),
    "vtn_multi_decompose_normalize": (
        C.c_int,
        [C.c_void_p, C.POINTER(VtnBatch), c_u32p, c_u16p, c_u64p, C.POINTER(VtnDecompCounters), C.POINTER(VtnNormResult)],
    ),
    "vtn_timer_start": (C.c_int, [C.c_void_p]),
    "vtn_timer_stop": (C.c_int, [C.c_void_p, C.POINTER(C.c_float)]),
    "vtn_launch_count": (C.c_uint64, [C.c_void_p]),
}

_lib = None


def load():
    """dlopen libvtn.so (built in-tree by vt_b200/csrc/Makefile).  Raises if it is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("libvtn.so not built (%s): run `python -c 'import __graft_entry__ as g; g.build()'`" % LIB_PATH)
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib
