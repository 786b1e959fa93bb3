/* oracle/vt_oracle.h -- TEST INFRASTRUCTURE ONLY (never linked or called by the product path).
 *
 * Plain-C CPU restatement of the reference's per-record normalization path, written as a literal
 * loop-for-loop restatement (NOT the closed form the CUDA kernels use) so that it is an independent
 * check.  Each function cites the reference file:line it follows (paths relative to /root/reference).
 *
 * Parity pinning: tests/test_oracle_golden.py checks this restatement against the reference's own
 * golden fixture (test/normalize/01_IN.vcf -> 01_OUT.vcf, test/decompose/01_IN_multi.vcf ->
 * 01_OUT_multi.vcf; committed under tests/golden/) and, in the build container, against the
 * unmodified reference binary oracle/_ref/vt_ref on synthetic inputs.
 *
 * Batch layout = the SoA layout of include/vtn.h (same arrays, same meaning) so tests compare
 * array-for-array.
 */
#ifndef VT_ORACLE_H
#define VT_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* in-memory reference: per contig the raw FASTA bases (isgraph bytes, case preserved) */
typedef struct vto_ref
{
    int32_t n_contig;
    const int64_t* len;         /* [n_contig] */
    const uint8_t* const* seq;  /* [n_contig] */
} vto_ref;

typedef struct vto_batch
{
    uint32_t n_rec;
    const int32_t* pos1;        /* [n_rec] 1-based POS */
    const int32_t* rid;         /* [n_rec] contig index into vto_ref, <0 = unknown */
    const uint32_t* allele_beg; /* [n_rec+1] first allele slot of record r */
    const uint32_t* allele_off; /* [n_allele+1] byte offset of allele slot */
    const uint8_t* bytes;
} vto_batch;

/* same bit meaning as include/vtn.h */
#define VTO_ST_REFCODE_MASK 0x03
#define VTO_ST_FETCH_FAIL 0x04
#define VTO_ST_ENTERED 0x08
#define VTO_ST_CHANGED 0x10
#define VTO_ST_CAT_SHIFT 5

typedef struct vto_norm_out
{
    uint8_t* status;      /* [n_rec] */
    uint8_t* vtype;       /* [n_rec] */
    uint32_t n_changed;   /* out */
    uint64_t n_changed_bytes; /* out */
    uint32_t* chg_rec;    /* [cap] record index */
    int32_t* chg_pos1;    /* [cap] new POS */
    uint32_t* chg_lt;     /* [cap] */
    uint32_t* chg_le;     /* [cap] */
    uint32_t* chg_rt;     /* [cap] */
    uint64_t* chg_off;    /* [cap+1] byte offset of the record's first new allele in chg_bytes */
    uint8_t* chg_bytes;   /* new alleles, concatenated in allele order, lengths = old - rt + le - lt */
    uint32_t counters[12]; /* lt rt lt_rt rt_la la | multi lt rt lt_rt rt_la la | no_variants no_refs */
} vto_norm_out;

/* variant_manip.cpp:227-508 -> VT_* bitmask (allele.h:30-37) */
int32_t vto_classify(int32_t n_allele, const uint8_t* const* al, const uint32_t* len);
/* variant_manip.cpp:61-103 (+ faidx.c:832-860 clipping); returns 0/1/2, -1 when the contig is unknown */
int32_t vto_is_not_ref_consistent(const vto_ref* ref, int32_t rid, int32_t pos1, const uint8_t* a0, uint32_t len0);
/* variant_manip.cpp:109-197 */
int32_t vto_is_normalized(int32_t n_allele, const uint8_t* const* al, const uint32_t* len);
/* normalize.cpp:174-330 over a whole batch; returns 0, or -1 on capacity/alloc failure */
int32_t vto_normalize_batch(const vto_ref* ref, const vto_batch* in, vto_norm_out* out);

/* decompose.cpp:210-392 / 393-957 allele fan-out: for every record with n_allele>2 emits n_allele-1
 * (src_rec, alt_idx) pairs; records with <=2 alleles emit (src_rec, 1) (or alt_idx 0 when n_allele==1).
 * counters: no_variants, no_biallelic, no_multiallelic, no_additional_biallelic (decompose.cpp:139-143) */
int32_t vto_decompose_plan(const vto_batch* in, uint32_t* src_rec, uint16_t* alt_idx, uint64_t* n_out, uint32_t counters[4]);

/* vt uniq (uniq.cpp:110-156 over BCFSyncedReader SYNC_BY_VAR, one file, sequential reading): pop order of the
 * reader's priority queue (libstdc++ heap restated), crecs[0] flags, counters {no_total_variants,
 * no_unique_variants}; *not_in_order = index of the record where the reference exits with "VCF file not in
 * order", else -1.  Pinned against oracle/_ref/vt_ref uniq (tests/test_uniq_oracle.py, tests/golden/uniq_*). */
int32_t vto_uniq_plan(const vto_batch* in, uint32_t* order, uint8_t* first, uint32_t counters[2], int64_t* not_in_order);

/* BCFOrderedWriter::write / flush (bcf_ordered_writer.cpp:141-210, :252-300) as a literal list emulation: order[k] = arrival
 * index of the k-th record written, late_front[v] = front position when v triggered the "Might not be sorted" warning
 * (INT32_MIN otherwise).  Pinned through the command-line tests against oracle/_ref/vt_ref (window-limited ordering). */
int32_t vto_writer_order(uint32_t n, const int32_t* rid, const int32_t* pos1, int32_t window, uint32_t* order, int32_t* late_front);

/* VCF record text -> CHROM length, POS, REF / ALT alleles (SoA of include/vtn.h) and the offset of the pass-through tail, as
 * htslib's vcf_parse1 + bcf_unpack(BCF_UN_STR) hand them to the path; see vt_oracle.c.  Returns n_rec or -1 (capacity). */
int64_t vto_vcf_tokenize(const char* text, uint64_t n_text, uint64_t cap_rec, uint64_t cap_allele, uint64_t cap_bytes,
                         uint64_t* line_off, uint32_t* chrom_len, int32_t* pos1, uint32_t* tail_off, uint8_t* status,
                         uint32_t* allele_beg, uint32_t* allele_off, uint8_t* bytes);

/* vt decompose_blocksub without -a (decompose_blocksub.cpp:185-421): output k = (src_rec, off, len), len 0 = the record itself;
 * counters no_variants, no_biallelic_blocksub, no_additional_snps, new_no_variants.  Pinned to the reference's own fixtures
 * (test/decompose_blocksub 01, 03 and the mnv directory) and to oracle/_ref/vt_ref (tests/test_blocksub_oracle.py). */
int32_t vto_blocksub_plan(const vto_batch* in, int32_t keep_mnv, int32_t max_mnv_dist, uint32_t* src_rec, uint32_t* off, uint32_t* len,
                          uint64_t* n_out, uint32_t counters[4]);

/* vt decompose_blocksub -a (decompose_blocksub.cpp:190-303, needle.cpp:36-133): per record the chunks of the alignment of REF[1:] and
 * ALT[1:]; chunk_beg[n_rec + 1] is laid out with min(ref_len, alt_len) slots per record the mode applies to, n_chunk[r] = chunks
 * written (0 = mode does not apply); chunk c -> POS + pos_ref[c], REF[pos_ref, +len_ref), ALT[pos_alt, +len_alt). */
int32_t vto_blocksub_align(const vto_batch* in, uint32_t* chunk_beg, uint32_t* n_chunk, uint32_t* pos_ref, uint32_t* pos_alt, uint32_t* len_ref,
                           uint32_t* len_alt);

#ifdef __cplusplus
}
#endif
#endif
