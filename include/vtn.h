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
 *   vtn_normalize_bi     the same loop over the compact biallelic layout (vtn_batch_bi)
 *   vtn_decompose*       the allele fan-out of Igor::decompose(), decompose.cpp:210-392 / :393-957, and its
 *                        four counters (:139-143)
 *   vtn_uniq*            the record loop of Igor::uniq(), uniq.cpp:110-156, over BCFSyncedReader::read_next_position
 *                        (bcf_synced_reader.cpp:388-433, :594-637; comparator bcf_synced_reader.h:86-122)
 *   vtn_writer_order     BCFOrderedWriter::write / flush, bcf_ordered_writer.cpp:141-210, :252-300
 *   vtn_vcf_tokenize     vcf_parse1 + bcf_unpack(BCF_UN_STR) for CHROM/POS/REF/ALT (lib/htslib/vcf.c) behind
 *                        BCFOrderedReader::read, bcf_ordered_reader.cpp:238-298
 *   vtn_blocksub*        the record loop of Igor::decompose_blocksub(), decompose_blocksub.cpp:185-421 (without -a)
 *   vtn_annotate_indels  the record loop of Igor::annotate_indels(), annotate_indels.cpp:293-410, and VNTRAnnotator::annotate under it
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
/* The way the reference itself uses a FASTA (fai_load reads only the index, faidx_fetch_seq seeks into the file, variant_manip.cpp:35,
 * 70): open the file and its index, reserve the HBM for the whole genome, and make a contig resident (packed by all host threads
 * straight from the mapped file) the first time it is needed.  Every entry point that takes HOST contig ids (vtn_normalize,
 * vtn_normalize_bi, vtn_decompose_normalize, vtn_dbatch_create*, vtn_annotate_indels, vtn_ref_fetch*) does that itself; a caller
 * that builds device batches another way calls vtn_ref_require / vtn_ref_require_all first.  Contig ids, names and lengths are
 * valid right after the open.  vtn_ref_load_fasta == vtn_ref_open_fasta + vtn_ref_require_all. */
int vtn_ref_open_fasta(vtn_ctx* ctx, const char* fasta_path);
int vtn_ref_require(vtn_ctx* ctx, int rid); /* no-op for a resident or unknown contig */
int vtn_ref_require_all(vtn_ctx* ctx);
/* The host half on its own.  Creating a CUDA context takes 0.25-2 s; a tool that starts it on another thread can meanwhile open the
 * FASTA (vtn_fasta_open: mapping + index, no device), read the first records of its input and pack their contig on the host
 * (vtn_fasta_prepack), then hand the handle to the context (vtn_ref_attach == what vtn_ref_open_fasta does after its own open; the
 * context owns the handle from then on, also when the call fails).  vtn_ref_prefetch packs one more contig on a background thread:
 * a sorted call set walks the contigs in order, so asking for contig c + 1 when c first appears leaves vtn_ref_require only the
 * upload.  vtn_last_error(NULL) has the text of a failed vtn_fasta_open. */
typedef struct vtn_fasta vtn_fasta;
int vtn_fasta_open(const char* fasta_path, vtn_fasta** out);
int vtn_fasta_n_contig(const vtn_fasta* f);
int vtn_fasta_contig_id(const vtn_fasta* f, const char* name); /* -1 when unknown */
int vtn_fasta_prepack(vtn_fasta* f, int rid);
void vtn_fasta_close(vtn_fasta* f); /* only for a handle that was never attached */
int vtn_ref_attach(vtn_ctx* ctx, vtn_fasta* f);
int vtn_ref_prefetch(vtn_ctx* ctx, int rid);
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

/* ---- compact biallelic batch (the common case of `vt normalize` input: a sorted, decomposed call set) ----
 * The same records as a vtn_batch whose records all have two alleles, in a layout that costs an SNV 8 bytes instead of 22
 * (pos1 4 + lens 2 + its two bases), on PCIe and in HBM alike:
 *   lens[r]     = REF length | ALT length << 8; VTN_BI_LEN_ESC (255) in either byte = an allele of 255 bases or more, whose
 *                 lengths are in ext_len[2k], ext_len[2k+1] with ext_rec[k] == r (ext_rec ascending)
 *   bytes       = REF then ALT of record 0, of record 1, ... (what vtn_batch::bytes holds for biallelic records)
 *   tile_off[t] = offset in bytes of the first record of tile t (VTN_BI_TILE records); tile_off[n_tiles] = n_bytes.  The host
 *                 knows it while it appends records; the device needs no batch-wide scan.
 *   contig ids as RUNS: records [run_start[k], run_start[k+1]) lie on contig run_rid[k] (vtn_ref_contig_id, or -1); a sorted
 *                 VCF has one run per contig.  run_start[0] = 0, run_start[n_run] = n_rec.
 * Result (vtn_norm_result_bi): the changed records exactly as in vtn_norm_result; per record one byte sv[r] = VT_* bits |
 * is_not_ref_consistent code << 6 (optional: NULL = not copied back); the rare records whose status carries more than that
 * and is not implied by the changed list (VTN_ST_FETCH_FAIL, or VTN_ST_ENTERED without VTN_ST_CHANGED) are listed in
 * note_rec / note_status (ascending).  status of a changed record = VTN_ST_ENTERED | VTN_ST_CHANGED | category of (lt, le, rt). */
#define VTN_BI_TILE 128
#define VTN_BI_LEN_ESC 255
typedef struct vtn_batch_bi
{
    uint32_t n_rec;
    uint32_t n_run;
    uint32_t n_ext;
    uint64_t n_bytes;          /* < 2^32 */
    const int32_t* pos1;       /* [n_rec] */
    const uint16_t* lens;      /* [n_rec] */
    const uint8_t* bytes;      /* [n_bytes] */
    const uint32_t* tile_off;  /* [ceil(n_rec / VTN_BI_TILE) + 1] */
    const uint32_t* run_start; /* [n_run + 1] */
    const int32_t* run_rid;    /* [n_run] */
    const uint32_t* ext_rec;   /* [n_ext] */
    const uint32_t* ext_len;   /* [2 * n_ext] */
} vtn_batch_bi;

typedef struct vtn_norm_result_bi
{
    uint8_t* sv;                /* [n_rec] or NULL */
    uint32_t cap_changed;       /* in */
    uint64_t cap_changed_bytes; /* in */
    uint32_t n_changed;         /* out */
    uint64_t n_changed_bytes;   /* out */
    uint32_t* chg_rec;
    int32_t* chg_pos1;
    uint32_t* chg_lt;
    uint32_t* chg_le;
    uint32_t* chg_rt;
    uint64_t* chg_off;          /* [n_changed+1] */
    uint8_t* chg_bytes;
    uint32_t cap_notes;         /* in */
    uint32_t n_notes;           /* out */
    uint32_t* note_rec;
    uint8_t* note_status;       /* full VTN_ST_* byte of that record */
    vtn_norm_counters counters;
} vtn_norm_result_bi;

/* Host helper: re-pack a vtn_batch whose records are all biallelic (VTN_E_ARG otherwise) into pinned host arrays owned by
 * the returned object (what a reader that fills the compact layout directly would hold). */
typedef struct vtn_bi_host vtn_bi_host;
int vtn_bi_host_create(vtn_ctx* ctx, const vtn_batch* in, vtn_bi_host** h);
const vtn_batch_bi* vtn_bi_host_view(const vtn_bi_host* h);
int vtn_bi_host_destroy(vtn_ctx* ctx, vtn_bi_host* h);
/* Host-buffer call, pipelined in chunks like vtn_normalize (H2D, kernels, D2H overlap); returns after completion. */
int vtn_normalize_bi(vtn_ctx* ctx, const vtn_batch_bi* in, vtn_norm_result_bi* out);
/* Device-resident variant: vtn_normalize_resident / vtn_dbatch_destroy work on the handle. */
int vtn_dbatch_create_bi(vtn_ctx* ctx, const vtn_batch_bi* in, vtn_dbatch** db);
int vtn_dbatch_fetch_result_bi(vtn_ctx* ctx, vtn_dbatch* db, vtn_norm_result_bi* out);

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

/* ---- uniq: drop duplicate variants (vt uniq) ----
 * The input must be sorted like vt requires (bcf_is_in_order, hts_utils.cpp:1042-1050: rid ascending, pos ascending
 * within a rid; rid >= 0).  order[k] = input index of the k-th record the reference's synced reader hands out: records
 * are grouped by (rid, pos1) in input order, and inside a position group they come out sorted by strcmp of
 * "REF,ALT" (ALTs of a multiallelic record sorted first, hts_utils.cpp:1136-1167), equal records in the order
 * libstdc++'s priority queue surfaces them (which is NOT file order; reproduced exactly).  first[k] = 1 iff order[k] is
 * crecs[0] of a read_next_position call = the record `vt uniq` writes; the entries with first[k] == 0 that follow are its
 * duplicates, whose OLD_VARIANT values uniq merges into it in this order (uniq.cpp:114-143).
 * *not_in_order = -1, or the index of the first record that is smaller than its predecessor (the reference prints
 * "VCF file not in order" and exits; order/first are unspecified then). */
typedef struct vtn_uniq_counters
{
    uint32_t no_total_variants, no_unique_variants; /* uniq.cpp:48-49, :147-148 */
} vtn_uniq_counters;
int vtn_uniq(vtn_ctx* ctx, const vtn_batch* in, uint32_t* order, uint8_t* first, vtn_uniq_counters* counters,
             int64_t* not_in_order);
int vtn_uniq_resident(vtn_ctx* ctx, vtn_dbatch* db); /* async on the ctx stream; for bench / pipelines */
int vtn_dbatch_fetch_uniq(vtn_ctx* ctx, vtn_dbatch* db, uint32_t* order, uint8_t* first, vtn_uniq_counters* counters,
                          int64_t* not_in_order);

/* ---- ordered writer: the order in which BCFOrderedWriter (bcf_ordered_writer.cpp:141-210, :252-300) writes a stream ----
 * rid[i] / pos1[i] = contig id and (new) 1-based position of the i-th record handed to the writer, window = its sorting
 * window in bp (normalize: 10000, normalize.cpp:101,139; 0 = write straight through).  order[k] = index of the k-th record
 * written: the list-insertion / windowed-flush rules are reproduced exactly, including records that arrive more than
 * `window` behind (they are written late, not sorted in) and the flush at every rid change.  late_front (optional, may
 * be NULL): for a record that triggers the writer's "Might not be sorted for window size" warning (:163-179) the position
 * of the buffer's front at that moment (the text's "[last record]"), INT32_MIN otherwise.
 * Streaming: n_pending (optional) receives how many records at the END of order[] are still inside the writer's buffer
 * after the last record (they would only leave with the closing flush).  A caller with more input writes
 * order[0 .. n - n_pending) now and hands the pending records, in arrival order, to the next call in front of the new
 * records: replaying them rebuilds the same buffer without writing anything, so the concatenated output equals one call
 * over the whole stream (tests/test_writer_gpu.py). */
int vtn_writer_order(vtn_ctx* ctx, uint32_t n, const int32_t* rid, const int32_t* pos1, int32_t window, uint32_t* order,
                     int32_t* late_front, uint32_t* n_pending);
/* Device-resident variant over the batch's own rid / pos1 (bench / pipelines; window > 0): async on the ctx stream. */
int vtn_writer_order_resident(vtn_ctx* ctx, vtn_dbatch* db, int32_t window);
int vtn_dbatch_fetch_writer_order(vtn_ctx* ctx, vtn_dbatch* db, uint32_t* order, int32_t* late_front, uint32_t* n_pending);

/* ---- VCF text -> batch: the five site columns of every record line, parsed on the device ----
 * Replaces htslib's per-line text parse for what the path reads: vcf_parse1 (lib/htslib/vcf.c; tab split, CHROM, POS,
 * "REF,ALT" -> d.allele[], ALT "." = no alternate allele) + bcf_unpack(BCF_UN_STR)/bcf_get_allele, i.e. what
 * BCFOrderedReader::read (bcf_ordered_reader.cpp:238-298) hands to normalize.cpp:196 / decompose.cpp:205.
 * text = record lines only (no '#' header), each ended by '\n' (a final line without one is accepted, "\r\n" too).
 * Record r is text[line_off[r] .. line_off[r+1]) (with its line end); CHROM is its first chrom_len[r] bytes; the columns
 * from QUAL on start at line_off[r] + tail_off[r] and are not parsed (pass them through).  pos1 / allele_beg /
 * allele_off / bytes are exactly the arrays of vtn_batch; rid is the caller's (header dictionary) lookup of CHROM.
 * status[r] = 1 for a line with fewer than 8 columns (no alleles are emitted for it).
 * All arrays are caller-allocated with the capacities given; on VTN_E_CAPACITY the n_* fields hold what is needed. */
typedef struct vtn_vcf_fields
{
    uint32_t cap_rec, cap_allele; /* in */
    uint64_t cap_bytes;           /* in */
    uint32_t n_rec, n_allele;     /* out */
    uint64_t n_bytes;             /* out */
    uint64_t* line_off;           /* [cap_rec+1] */
    uint32_t* chrom_len;          /* [cap_rec] */
    int32_t* pos1;                /* [cap_rec] */
    uint32_t* tail_off;           /* [cap_rec] */
    uint8_t* status;              /* [cap_rec] */
    uint32_t* allele_beg;         /* [cap_rec+1] */
    uint32_t* allele_off;         /* [cap_allele+1] */
    uint8_t* bytes;               /* [cap_bytes] */
} vtn_vcf_fields;
int vtn_vcf_tokenize(vtn_ctx* ctx, const char* text, uint64_t n_text, vtn_vcf_fields* out);
/* Device-resident variant (bench / pipelines): upload once, tokenize many times, fetch the arrays of the last run. */
typedef struct vtn_vcf_dtext vtn_vcf_dtext;
int vtn_vcf_dtext_create(vtn_ctx* ctx, const char* text, uint64_t n_text, uint32_t cap_rec, uint32_t cap_allele,
                         uint64_t cap_bytes, vtn_vcf_dtext** h);
int vtn_vcf_dtext_destroy(vtn_ctx* ctx, vtn_vcf_dtext* h);
int vtn_vcf_tokenize_resident(vtn_ctx* ctx, vtn_vcf_dtext* h);
int vtn_vcf_dtext_fetch(vtn_ctx* ctx, vtn_vcf_dtext* h, vtn_vcf_fields* out);

/* ---- decompose_blocksub: block substitutions -> SNPs (or MNVs) ----
 * Replaces the record loop of Igor::decompose_blocksub() (decompose_blocksub.cpp:185-421); its alignment mode (-a) is vtn_blocksub_align below.
 * Output k comes from input record src_rec[k]: len[k] == 0 -> the record itself, unchanged; otherwise a new record at
 * POS = pos1 + off[k] with alleles REF[off, off+len) / ALT[off, off+len) (one base per differing position; with keep_mnv
 * (-m) one run of differences at most max_mnv_dist (-d, default 2) apart).  Outputs are in input order; the reference then
 * writes them through a 1000-bp ordered writer (decompose_blocksub.cpp:133; vtn_writer_order).  A block substitution without
 * any difference emits nothing, like the reference.  Capacity: vtn_blocksub_out_bound(in). */
typedef struct vtn_blocksub_counters
{
    uint32_t no_variants, no_biallelic_blocksub, no_additional_snps, new_no_variants; /* decompose_blocksub.cpp:157-160 */
} vtn_blocksub_counters;
uint64_t vtn_blocksub_out_bound(const vtn_batch* in);
int vtn_blocksub(vtn_ctx* ctx, const vtn_batch* in, int keep_mnv, int max_mnv_dist, uint32_t* src_rec, uint32_t* off,
                 uint32_t* len, uint64_t cap_out, uint64_t* n_out, vtn_blocksub_counters* counters);
int vtn_blocksub_resident(vtn_ctx* ctx, vtn_dbatch* db, int keep_mnv, int max_mnv_dist); /* async; bench / pipelines */
int vtn_dbatch_fetch_blocksub(vtn_ctx* ctx, vtn_dbatch* db, uint32_t* src_rec, uint32_t* off, uint32_t* len, uint64_t cap_out,
                              uint64_t* n_out, vtn_blocksub_counters* counters);

/* ---- decompose_blocksub -a: alignment mode ----
 * Replaces the aggressive branch of the same loop (decompose_blocksub.cpp:190-303) and NeedlemanWunsch::align / trace_path
 * (needle.cpp:36-133; match 0, mismatch -1, gap -1, needle.h:37): a record whose REF and first ALT differ in length, with REF longer
 * than one base and ALT longer than one base, is cut along the global alignment of REF[1:] and ALT[1:] (first bases force-aligned)
 * into chunks = one aligned column plus the gap columns after it, kept when they hold a mismatch or a gap.  Chunk c of record r
 * (chunk_beg[r] <= c < chunk_beg[r] + n_chunk[r]) is the new record POS = pos1 + pos_ref[c], REF = REF[pos_ref, +len_ref),
 * ALT = ALT[pos_alt, +len_alt); n_chunk[r] == 0 means the mode does not apply to record r (use vtn_blocksub's plan for it).  The
 * reference counts ++no_biallelic_blocksub per such record and ++no_additional_snps, ++new_no_variants per chunk (:288-299). */
typedef struct vtn_blocksub_align_plan
{
    uint64_t cap_chunk; /* in: capacity of the four chunk arrays; vtn_blocksub_align_bound(in) always suffices */
    uint64_t n_chunks;  /* out: slots laid out (chunk_beg[n_rec]) */
    uint32_t* chunk_beg; /* [n_rec + 1] */
    uint32_t* n_chunk;   /* [n_rec] */
    uint32_t *pos_ref, *pos_alt, *len_ref, *len_alt; /* [cap_chunk] */
} vtn_blocksub_align_plan;
uint64_t vtn_blocksub_align_bound(const vtn_batch* in);
int vtn_blocksub_align(vtn_ctx* ctx, const vtn_batch* in, vtn_blocksub_align_plan* plan);

/* ---- decompose_blocksub without any allocation ----
 * No counterpart in the reference: added for callers that cannot call cudaMalloc (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid it).  The same plan as vtn_blocksub, but inputs and
 * outputs are DEVICE pointers the caller owns, the work is queued on the ctx stream and the call returns without waiting
 * (vtn_sync), and the scratch the kernels need (4 bytes per record + the per-tile words) is a fixed device array inside the
 * library: at most VTN_BLOCKSUB_NOALLOC_MAX_REC records per call (4 MiB + 8 KiB held by every GPU that loads the library until
 * exit); more are rejected with VTN_E_ARG and a text that names the limit.  The call allocates and frees nothing; the first call
 * on a device looks the array's device address up (cudaGetSymbolAddress), so make one call per device before a capture begins.
 * Every call shares that array: calls that overlap on one GPU (two contexts / streams) race -- run them one after the other.
 * d_bytes must be 4-byte aligned and readable for 16 bytes past the last allele byte (the kernels load whole words);
 * d_src_rec / d_off / d_len have capacity cap_out (outputs beyond it are not written); *d_n_out receives the number of outputs
 * (compare it with cap_out), d_counters the four words of vtn_blocksub_counters. */
#define VTN_BLOCKSUB_NOALLOC_MAX_REC (1u << 20)
int vtn_blocksub_noalloc(vtn_ctx* ctx, uint32_t n_rec, const uint32_t* d_allele_beg, const uint32_t* d_allele_off, const uint8_t* d_bytes, int keep_mnv,
                         int max_mnv_dist, uint32_t* d_src_rec, uint32_t* d_off, uint32_t* d_len, uint64_t cap_out, uint64_t* d_n_out,
                         uint32_t* d_counters);

/* ---- HMM family: AHMM / LFHMM / RFHMM Viterbi realignment (annotate_indels' FlankDetector) ----
 * Replaces <model>::set_model + <model>::align + the result members/getters:
 *   AHMM  ahmm.h:94-557, ahmm.cpp:240-247 (set_model), :468-587 (align), :592-730 (trace_path, collect_statistics)
 *   LFHMM lfhmm.h, lfhmm.cpp:281-292, :489-654, :659-861
 *   RFHMM rfhmm.h, rfhmm.cpp:306-318, :516-648, :650-840
 * and the set_delta/epsilon/tau/eta/mismatch_penalty + initialize_T sequence (flank_detector.cpp:50-72).
 * One alignment per warp on the device; scores are computed with the reference's float expression order and
 * come out bit-identical, paths and integer results exactly (tolerance stated in north_star: 1e-4 abs on the
 * score; observed difference 0). */
#define VTN_HMM_AHMM 0
#define VTN_HMM_LFHMM 1
#define VTN_HMM_RFHMM 2

typedef struct vtn_hmm_params
{
    float delta, epsilon, tau, eta, mismatch_penalty; /* ahmm.h:74-92; FlankDetector: 1e-7,1e-7,.01,.01,5 */
    int32_t pl; /* base quality - 33 of every read base; FlankDetector passes 'K' = 42 (flank_detector.cpp:74) */
} vtn_hmm_params;

/* Alignment a: kind[a]; motif = bytes[seq_off[3a] .. seq_off[3a+1]), flank = [seq_off[3a+1] .. seq_off[3a+2])
 * (left flank for LFHMM, right flank for RFHMM, empty for AHMM), read = [seq_off[3a+2] .. seq_off[3a+3]).
 * Limits: read <= 1023 (AHMM truncates longer reads to 1023 like ahmm.cpp:476-480; the flanked models report
 * VTN_HMM_ST_TOO_LONG where the reference exits), motif 1..255, flank 1..31. */
typedef struct vtn_hmm_batch
{
    uint32_t n;
    const uint8_t* kind;     /* [n] */
    const uint32_t* seq_off; /* [3n+1] */
    const uint8_t* bytes;
    uint64_t n_bytes;
} vtn_hmm_batch;

#define VTN_HMM_ST_OK 0
#define VTN_HMM_ST_TOO_LONG 1 /* lfhmm.cpp:498-502 / rfhmm.cpp:524-528: the reference prints and exit(1)s */
#define VTN_HMM_ST_BAD_MODEL 2 /* empty motif/read, flank or motif longer than the limits above */
#define VTN_HMM_ST_NOPATH 3    /* no finite end cell: the reference would index outside its arrays */

/* What align() leaves in the object (ahmm.h:109-132); 96 bytes per alignment.
 * pos[] = lflank_start[MODEL,READ], lflank_end[..], motif_start, motif_end, rflank_start, rflank_end.
 * Members the reference does not (re)compute in a call are reported as: last_motif_pos -1 (LFHMM unless the path
 * leaves Z), motif_count 0 (RFHMM unless the path goes MR->M/D); exact_motif_count / motif_concordance /
 * trf_score / frac_no_repeats are 0 for RFHMM (not members of its align). */
typedef struct vtn_hmm_result
{
    float optimal_score;
    int32_t optimal_state, optimal_probe_len, path_len;
    int32_t pos[12];
    int32_t motif_count, exact_motif_count;
    float motif_concordance;
    int32_t trf_score;
    float frac_no_repeats;
    int32_t last_motif_pos;
    int32_t status; /* VTN_HMM_ST_* */
    int32_t reserved;
} vtn_hmm_result;

/* Host-buffer call: H2D, one kernel, D2H; returns after completion.  path (optional, may be NULL) receives the
 * optimal path of alignment a as reference track words [u:5|d:3|c:12|p:12] (ahmm.h:57-67) in forward order at
 * path[a*path_stride ...]; entries beyond path_stride are dropped (path_len still reports the full length). */
int vtn_hmm_align(vtn_ctx* ctx, const vtn_hmm_params* par, const vtn_hmm_batch* in, vtn_hmm_result* out,
                  int32_t* path, uint32_t path_stride);
/* Device-resident variant for bench / pipelines. */
typedef struct vtn_hmm_dbatch vtn_hmm_dbatch;
int vtn_hmm_dbatch_create(vtn_ctx* ctx, const vtn_hmm_params* par, const vtn_hmm_batch* in, uint32_t path_stride,
                          vtn_hmm_dbatch** db);
int vtn_hmm_dbatch_destroy(vtn_ctx* ctx, vtn_hmm_dbatch* db);
int vtn_hmm_align_resident(vtn_ctx* ctx, vtn_hmm_dbatch* db); /* async on the ctx stream */
int vtn_hmm_dbatch_fetch(vtn_ctx* ctx, vtn_hmm_dbatch* db, vtn_hmm_result* out, int32_t* path);
/* the transition table initialize_T would build (7x7 row-major, reference state numbering of that model) */
int vtn_hmm_transitions(int kind, const vtn_hmm_params* par, float* T49);
/* DP cells / predecessor evaluations (SURVEY 8d "W") of the last resident run, for roofline accounting */
int vtn_hmm_dbatch_work(vtn_ctx* ctx, vtn_hmm_dbatch* db, uint64_t* cells, uint64_t* proc_comps);

/* ---- annotate_indels: VNTR annotation of indels ----
 * Replaces the record loop of Igor::annotate_indels() (annotate_indels.cpp:293-410) and everything under
 * VNTRAnnotator::annotate (vntr_annotator.cpp:62-149) for a batch of records:
 *   classify_variant, is_normalized, contains_N (variant_manip.cpp:227-508, :109-197, :202-222)             device
 *   CandidateRegionExtractor::pick_candidate_region(EXACT_LEFT_RIGHT_ALIGNMENT) (candidate_region_extractor.cpp:162-267)  device
 *   CandidateMotifPicker::update_exact_repeat_unit / generate_candidate_motifs / next_motif (candidate_motif_picker.cpp)   host
 *   FlankDetector::detect_flanks(EXACT|FUZZY), compute_purity_score, compute_composition_and_entropy (flank_detector.cpp:103-887):
 *   every RFHMM / LFHMM / AHMM alignment on the device (vtn_hmm_*), one batched launch per slen round over all indels still open;
 *   the reference fetches (ReferenceSequence::fetch_seq / fetch_base) from the resident reference by vtn_ref_fetch_batch's kernel.
 * flags[r]: VTN_ANNOT_KEEP = the reference writes the record (it drops records that are not normalized or contain N,
 * annotate_indels.cpp:306-323); VTN_ANNOT_INDEL = it is annotated, vntr[indel_idx[r]] holds the 34 INFO values (:342-387).
 * Per allele slot of an annotated record (optional, all three or none): what trim() took from the left / right of REF and that
 * ALT and how many bases right_align moved them -- enough to reprint the two stderr lines per ALT the reference writes
 * unconditionally (candidate_region_extractor.cpp:195, :199). */
#define VTN_ANNOT_KEEP 1
#define VTN_ANNOT_INDEL 2
typedef struct vtn_vntr_side /* the exact_* or fuzzy_* members of VNTR (vntr.h) */
{
    int32_t beg1, end1;           /* *_REPEAT_TRACT */
    int32_t mlen, blen;           /* *_MLEN, *_BLEN */
    int32_t comp[4];              /* *_COMP */
    int32_t rl, ll;               /* *_RL, *_LL */
    int32_t no_perfect_ru, no_ru; /* *_RU_COUNTS */
    int32_t trf_score;            /* *_TRF_SCORE */
    float entropy, entropy2, kl_divergence, kl_divergence2, ref, score;
    uint32_t str_off[3], str_len[3]; /* motif, ru, basis inside vtn_annot_result::strings */
} vtn_vntr_side;
typedef struct vtn_vntr_annot
{
    uint32_t rec;                 /* input record */
    uint8_t exact_ru_ambiguous;   /* EXACT_RU_AMBIGUOUS flag */
    uint8_t is_large_repeat_tract;
    uint8_t reserved[2];
    vtn_vntr_side exact, fuzzy;
} vtn_vntr_annot;
typedef struct vtn_annot_result
{
    uint8_t* flags;        /* [n_rec] */
    uint8_t* vtype;        /* [n_rec] */
    uint32_t* indel_idx;   /* [n_rec] index into vntr[], UINT32_MAX otherwise */
    uint32_t* trim_left;   /* [n_allele] or NULL */
    uint32_t* trim_right;  /* [n_allele] or NULL */
    uint32_t* shift;       /* [n_allele] or NULL */
    uint32_t cap_indel;    /* in */
    uint64_t cap_str;      /* in */
    uint32_t n_indel;      /* out */
    uint64_t n_str;        /* out */
    uint64_t n_alignments; /* out: HMM alignments run on the device */
    vtn_vntr_annot* vntr;  /* [cap_indel] */
    char* strings;         /* [cap_str] */
} vtn_annot_result;
int vtn_annotate_indels(vtn_ctx* ctx, const vtn_batch* in, vtn_annot_result* out);
/* Test / diagnostic tap: every alignment the driver issues is written to `path` as "<indel> <stage> <kind> <motif> <flank> <read>
 * <what the driver reads back>" (stage 0 exact purity, 1 RFHMM rounds, 2 LFHMM rounds, 3 fuzzy purity) -- the format of
 * oracle/fd_trace.cpp, which taps the unmodified reference classes the same way.  NULL or "" switches it off. */
int vtn_annot_set_trace(vtn_ctx* ctx, const char* path);

/* ---- VariantManip's allele-only predicates for a batch ----
 * Replaces, record by record: VariantManip::classify_variant's type mask (variant_manip.cpp:337-490; vtype[r], VTN_VT_* bits, may be
 * NULL), VariantManip::is_normalized (:109-197; VTN_PROP_NORMALIZED) and VariantManip::contains_N (:202-222; VTN_PROP_CONTAINS_N)
 * -- what annotate_indels.cpp:295-316 and normalize.cpp:218 ask of a record.  Needs no reference; in->pos1 / in->rid are not read. */
#define VTN_PROP_NORMALIZED 1
#define VTN_PROP_CONTAINS_N 2
int vtn_variant_props(vtn_ctx* ctx, const vtn_batch* in, uint8_t* vtype, uint8_t* props);
/* ReferenceSequence::fetch_seq for a list of windows (0-based inclusive, clipped like faidx.c:832-860), one kernel:
 * out[off[k] .. off[k+1]) = window k; upper != 0 = faidx_fetch_uc_seq.  An unknown contig yields an empty window. */
int vtn_ref_fetch_batch(vtn_ctx* ctx, uint32_t n, const int32_t* rid, const int64_t* beg0, const int64_t* end0, int upper, uint64_t* off,
                        uint8_t* out, uint64_t cap);

/* ---- several GPUs of one box behind one call ----
 * The path shards trivially (records independent, reference read-only): no collective, no NCCL.  vtn_multi holds one vtn_ctx per
 * entry of device_ids (an id may repeat: two contexts on one device, which is how a single-GPU box tests the sharding); the
 * reference is packed and uploaded once and replicated to the other devices by peer copies (NVLink / NVSwitch where the devices
 * reach each other).  A call cuts the batch into one contiguous record range per device (vtn_shard_plan: by contig, then by
 * record range -- vt's own way to go parallel is one process per interval, program.cpp:152-190), runs every range on its device
 * from its own host thread, and puts the results back together in input order: what comes back is bit-identical to the
 * single-device call on the whole batch. */
typedef struct vtn_multi vtn_multi;
int vtn_multi_create(vtn_multi** m, const int* device_ids, int n_dev);
int vtn_multi_destroy(vtn_multi* m);
int vtn_multi_n_dev(const vtn_multi* m);
vtn_ctx* vtn_multi_ctx(vtn_multi* m, int slot); /* the context of one slot, for the vtn_ref_* queries and single-device calls */
const char* vtn_multi_last_error(const vtn_multi* m);
int vtn_multi_ref_load_fasta(vtn_multi* m, const char* fasta_path);
int vtn_multi_ref_load_memory(vtn_multi* m, int32_t n_contig, const char* const* names, const int64_t* len, const uint8_t* const* seq);
/* cut[0..n_shard]: shard s = records [cut[s], cut[s+1]); equal record counts, cuts on multiples of align, a cut within a sixteenth
 * of a shard of a contig change moves onto it */
int vtn_shard_plan(const int32_t* rid, uint32_t n_rec, uint32_t n_shard, uint32_t align, uint32_t* cut);
int vtn_multi_normalize_bi(vtn_multi* m, const vtn_batch_bi* in, vtn_norm_result_bi* out);
int vtn_multi_decompose_normalize(vtn_multi* m, const vtn_batch* in, uint32_t* src_rec, uint16_t* alt_idx, uint64_t* n_out,
                                  vtn_decomp_counters* dcounters, vtn_norm_result* out);

/* ---- timing on the ctx stream (CUDA events) ---- */
int vtn_timer_start(vtn_ctx* ctx);
int vtn_timer_stop(vtn_ctx* ctx, float* ms); /* records, synchronises, returns elapsed ms */
/* number of kernels libvtn launched on this ctx since creation */
uint64_t vtn_launch_count(const vtn_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif
