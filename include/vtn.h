/* include/vtn.h -- C ABI of libvtn (vt_b200): the B200-native drop-in for vt's per-record
 * normalization hot path.  Plain pointers and sizes only; no C++/torch types.
 *
 * What each entry point replaces in the reference (paths relative to atks/vt):
 *   vtn_ref_*            VariantManip::VariantManip(ref_fasta)  variant_manip.cpp:31-43  (fai_load) and every
 *                        faidx_fetch_seq / faidx_fetch_uc_seq call under it (variant_manip.cpp:69, :565;
 *                        hts_utils.cpp:41-51; lib/htslib/faidx.c:832-881)
 *   vtn_normalize*       the body of Igor::normalize()'s record loop, normalize.cpp:188-322:
 *                        classify_variant (variant_manip.cpp:227-508), is_not_ref_consistent (:61-103),
 *                        is_normalized (:109-197), right_trim_or_left_extend (:513-583), left_trim (:588-615),
 *                        and the ten category counters + no_variants/no_refs (normalize.cpp:254-322)
 *   vtn_decompose*       the allele fan-out of Igor::decompose(), decompose.cpp:210-392 / :393-957, and its
 *                        four counters (:139-143)
 *   vtn_hmm_*            AHMM/LFHMM/RFHMM::align (ahmm.cpp:468-587, lfhmm.cpp:489-654, rfhmm.cpp:516+)
 *
 * The library never exits or prints: every call returns 0 on success or a negative vtn_err code, and
 * vtn_last_error() gives the text.  Per-record problems (REF inconsistent, unknown contig) are reported in
 * the status byte so that the caller can reproduce the reference's messages / exit(1) in input order.
 * There is no CPU fallback: without a CUDA device vtn_ctx_create fails.
 */
#ifndef VTN_H
#define VTN_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct vtn_ctx vtn_ctx;       /* one per process per GPU; owns a stream pair and the resident reference */
typedef struct vtn_dbatch vtn_dbatch; /* a record batch resident in HBM (inputs + result buffers) */

enum vtn_err
{
    VTN_OK = 0,
    VTN_E_CUDA = -1,     /* CUDA runtime error (text in vtn_last_error) */
    VTN_E_ARG = -2,      /* bad argument */
    VTN_E_IO = -3,       /* FASTA / .fai could not be read */
    VTN_E_NOMEM = -4,
    VTN_E_CAPACITY = -5, /* caller-provided output capacity too small */
    VTN_E_STATE = -6     /* e.g. normalize before a reference is loaded */
};

/* VT_* allele type bits, identical to allele.h:30-37 */
#define VTN_VT_REF 0
#define VTN_VT_SNP 1
#define VTN_VT_MNP 2
#define VTN_VT_INDEL 4
#define VTN_VT_CLUMPED 8
#define VTN_VT_VNTR 16
#define VTN_VT_SV 32

/* status byte of one record */
#define VTN_ST_REFCODE_MASK 0x03 /* is_not_ref_consistent: 0 consistent, 1 N-mismatch only, 2 mismatch */
#define VTN_ST_FETCH_FAIL 0x04   /* contig unknown to the reference (faidx_fetch_seq would return NULL) */
#define VTN_ST_ENTERED 0x08      /* entered the trim path: alleles are upper-cased in place (normalize.cpp:223-233) */
#define VTN_ST_CHANGED 0x10      /* left_trimmed || left_extended || right_trimmed (normalize.cpp:239) */
#define VTN_ST_CAT_SHIFT 5       /* bits 5-7: 0 none, 1 lt, 2 rt, 3 lt_rt, 4 rt_la, 5 la (normalize.cpp:254-311) */

/* A batch of records, struct-of-arrays.  Record r has alleles allele_beg[r] .. allele_beg[r+1]-1 (slot
 * indices; slot 0 of a record is REF); allele slot s is bytes[allele_off[s] .. allele_off[s+1]).  Every
 * allele has length >= 1 (VCF's "." counts as one byte). */
typedef struct vtn_batch
{
    uint32_t n_rec;
    uint32_t n_allele;          /* total allele slots = allele_beg[n_rec] */
    uint64_t n_bytes;           /* total allele bytes = allele_off[n_allele], < 2^32 */
    const int32_t* pos1;        /* [n_rec] 1-based POS */
    const int32_t* rid;         /* [n_rec] contig id from vtn_ref_contig_id, or -1 */
    const uint32_t* allele_beg; /* [n_rec+1] */
    const uint32_t* allele_off; /* [n_allele+1] */
    const uint8_t* bytes;       /* [n_bytes] */
} vtn_batch;

typedef struct vtn_norm_counters
{
    uint32_t bi[5];    /* no_lt, no_rt, no_lt_rt, no_rt_la, no_la            normalize.cpp:68-72 */
    uint32_t multi[5]; /* no_multi_lt .. no_multi_la                         normalize.cpp:74-78 */
    uint32_t no_variants; /* normalize.cpp:65 */
    uint32_t no_refs;     /* normalize.cpp:66 */
} vtn_norm_counters;

/* Result of normalize, caller-allocated host arrays.  status/vtype have one entry per input record; the
 * chg_* arrays hold only the changed records, compacted IN INPUT ORDER (warp-scan compaction on device). */
typedef struct vtn_norm_result
{
    uint8_t* status;            /* [n_rec] */
    uint8_t* vtype;             /* [n_rec] VT_* bitmask of classify_variant */
    uint32_t cap_changed;       /* in: capacity of chg_rec/pos1/lt/le/rt (and chg_off has cap_changed+1) */
    uint64_t cap_changed_bytes; /* in: capacity of chg_bytes */
    uint32_t n_changed;         /* out */
    uint64_t n_changed_bytes;   /* out */
    uint32_t* chg_rec;          /* record index */
    int32_t* chg_pos1;          /* new 1-based POS */
    uint32_t* chg_lt;           /* left_trimmed */
    uint32_t* chg_le;           /* left_extended */
    uint32_t* chg_rt;           /* right_trimmed */
    uint64_t* chg_off;          /* [n_changed+1] offset of the record's new alleles in chg_bytes */
    uint8_t* chg_bytes;         /* new alleles (upper-case) concatenated in allele order; new length of
                                   allele i = old length - rt + le - lt */
    vtn_norm_counters counters; /* out: totals of THIS call */
} vtn_norm_result;

typedef struct vtn_decomp_counters
{
    uint32_t no_variants, no_biallelic, no_multiallelic, no_additional_biallelic; /* decompose.cpp:139-143 */
} vtn_decomp_counters;

/* ---- context ---- */
int vtn_ctx_create(vtn_ctx** ctx, int device);
int vtn_ctx_destroy(vtn_ctx* ctx);
const char* vtn_last_error(const vtn_ctx* ctx); /* ctx may be NULL: last error of vtn_ctx_create */
int vtn_sync(vtn_ctx* ctx);

/* ---- pinned host memory for batches (optional; any host pointer is accepted, pinned is faster) ---- */
int vtn_host_alloc(vtn_ctx* ctx, size_t bytes, void** p);
int vtn_host_free(vtn_ctx* ctx, void* p);

/* ---- reference: replaces fai_load + faidx_fetch_* ---- */
/* Load FASTA (+ .fai, built in memory when absent; plain or gzip/bgzip) and make it resident in HBM. */
int vtn_ref_load_fasta(vtn_ctx* ctx, const char* fasta_path);
/* Same from memory: contig c has raw FASTA bases seq[c][0..len[c]) (case preserved, any isgraph byte). */
int vtn_ref_load_memory(vtn_ctx* ctx, int32_t n_contig, const char* const* names, const int64_t* len,
                        const uint8_t* const* seq);
int vtn_ref_n_contig(const vtn_ctx* ctx);
int vtn_ref_contig_id(const vtn_ctx* ctx, const char* name); /* -1 when unknown */
int64_t vtn_ref_contig_len(const vtn_ctx* ctx, int rid);
const char* vtn_ref_contig_name(const vtn_ctx* ctx, int rid);
int64_t vtn_ref_resident_bytes(const vtn_ctx* ctx);
/* faidx_fetch_seq-style read-back from the DEVICE copy (0-based inclusive, clipped like faidx.c:832-860);
 * upper!=0 gives faidx_fetch_uc_seq.  Returns bases written or <0. */
int64_t vtn_ref_fetch(vtn_ctx* ctx, int rid, int64_t beg0, int64_t end0, int upper, uint8_t* out, int64_t cap);

/* ---- normalize ---- */
/* Host-buffer call (the reference-facing one): H2D, kernels, D2H of the results; returns after completion. */
int vtn_normalize(vtn_ctx* ctx, const vtn_batch* in, vtn_norm_result* out);

/* Device-resident variant: upload once, run many times (bench / pipelines). */
int vtn_dbatch_create(vtn_ctx* ctx, const vtn_batch* in, vtn_dbatch** db);
int vtn_dbatch_destroy(vtn_ctx* ctx, vtn_dbatch* db);
int vtn_normalize_resident(vtn_ctx* ctx, vtn_dbatch* db);                       /* async on the ctx stream */
int vtn_dbatch_fetch_result(vtn_ctx* ctx, vtn_dbatch* db, vtn_norm_result* out); /* D2H of the last run */

/* ---- decompose ---- */
/* Fan-out plan: out record k comes from input record src_rec[k] and keeps alleles {0, alt_idx[k]}
 * (alt_idx 0 = record had a single allele).  src_rec/alt_idx need capacity vtn_decompose_out_bound(in). */
uint64_t vtn_decompose_out_bound(const vtn_batch* in);
int vtn_decompose(vtn_ctx* ctx, const vtn_batch* in, uint32_t* src_rec, uint16_t* alt_idx, uint64_t* n_out,
                  vtn_decomp_counters* counters);
/* Fused decompose -> normalize on device: result arrays are indexed by OUTPUT (biallelic) record. */
int vtn_decompose_normalize(vtn_ctx* ctx, const vtn_batch* in, uint32_t* src_rec, uint16_t* alt_idx,
                            uint64_t* n_out, vtn_decomp_counters* dcounters, vtn_norm_result* out);
int vtn_decompose_resident(vtn_ctx* ctx, vtn_dbatch* db, int fuse_normalize); /* async; for bench */
int vtn_dbatch_fetch_decompose(vtn_ctx* ctx, vtn_dbatch* db, uint32_t* src_rec, uint16_t* alt_idx,
                               uint64_t* n_out, vtn_decomp_counters* counters);

/* ---- timing on the ctx stream (CUDA events) ---- */
int vtn_timer_start(vtn_ctx* ctx);
int vtn_timer_stop(vtn_ctx* ctx, float* ms); /* records, synchronises, returns elapsed ms */
/* number of kernels libvtn launched on this ctx since creation */
uint64_t vtn_launch_count(const vtn_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif
