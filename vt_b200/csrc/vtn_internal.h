// vtn_internal.h -- shared declarations of libvtn (host + device).  Not part of the ABI.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>
#include "../../include/vtn.h"

// ---------------------------------------------------------------------------------------------
// Resident reference layout (HBM).
// One 16-byte "group" per 32 bases: { u64 code (2 bits/base, A=0 C=1 T=2 G=3 = (ascii>>1)&3, base j at bits 2j..2j+1),
// u32 nonstd (bit j: raw FASTA byte is not one of upper-case ACGT), u32 lower (bit j: raw byte is one of
// acgt) }.  One LDG.128 therefore brings everything the hot path needs for 32 bases: the upper-cased
// base (faidx_fetch_uc_seq, hts_utils.cpp:41-51) and whether the raw byte (faidx_fetch_seq,
// faidx.c:875-881) differs from it.  Bytes that are neither ACGT nor acgt (N, IUPAC, ...) are kept
// verbatim in a run-length table (exc_*), keyed by global base position, searched only on the slow path.
// Contigs start on a group boundary; global position of base p0 of contig c = contig_goff[c]*32 + p0.
// ---------------------------------------------------------------------------------------------
struct VtnDevRef
{
    const uint4* groups;       // [n_groups]  (.x,.y = code lo/hi ; .z = nonstd ; .w = lower)
    const uint64_t* contig_goff; // [n_contig] first group of contig
    const int64_t* contig_len;   // [n_contig]
    const uint64_t* exc_start; // [n_exc] global base position of run start (sorted)
    const uint32_t* exc_len;   // [n_exc]
    const uint8_t* exc_byte;   // [n_exc] the raw byte repeated over the run
    uint32_t n_exc;
    int32_t n_contig;
};

struct VtnContig
{
    std::string name;
    int64_t len;
    uint64_t goff;
};

// device-side view of a batch and its result buffers
struct VtnDevBatchView
{
    uint32_t n_rec;
    uint32_t rec_base; // added to record indices written to chg_rec (chunked launches pass offset views)
    const int32_t* pos1;
    const int32_t* rid;
    const uint32_t* allele_beg;
    const uint32_t* allele_off;
    const uint8_t* bytes;
};

// device-side view of a compact biallelic batch (vtn_batch_bi); chunked launches pass offset views: pos1 / lens / tile_off
// are relative to the chunk, bytes / run_* / ext_* absolute (rec_base = first record of the chunk)
struct VtnDevBiView
{
    uint32_t n_rec;
    uint32_t rec_base;
    const int32_t* pos1;
    const uint16_t* lens;
    const uint8_t* bytes;
    const uint32_t* tile_off;
    const uint32_t* run_start;
    const int32_t* run_rid;
    uint32_t n_run;
    const uint32_t* ext_rec;
    const uint32_t* ext_len;
    uint32_t n_ext;
    uint32_t wide; // the batch averages more than 6 allele bytes per record (STR indels, C4): the fast launch stages a 24-byte-per-record window
};

struct VtnDevNotes
{
    uint32_t* rec;     // absolute record index
    uint8_t* status;
    uint32_t* count;
    uint32_t cap;
};

struct VtnDevNormOut
{
    uint8_t* status;
    uint8_t* vtype;
    uint32_t* chg_rec;
    int32_t* chg_pos1;
    uint32_t* chg_lt;
    uint32_t* chg_le;
    uint32_t* chg_rt;
    uint64_t* chg_off;   // [n_changed+1]
    uint8_t* chg_bytes;
    unsigned long long* totals; // [0]=n_changed [1]=n_changed_bytes
    uint32_t* counters;  // [12]
};

// per-tile scratch segments written by K1, gathered by K2 (vtn_normalize.cu)
struct VtnDevScratch
{
    uint4* q0;                // [n_rec] {record, new pos1, left_trimmed, left_extended}
    uint4* q1;                // [n_rec] {right_trimmed, byte offset of the new alleles inside the tile's byte window | bit 31: the
                              //          alleles are the 2-bit code stream in .z/.w (<= 32 bases) instead of bytes, code lo, code hi}
    uint8_t* bytes;           // [n_bytes] tile t owns the byte window of its records' old alleles
    unsigned long long* dir;  // [n_tiles] (changed count << 32) | new allele bytes
    uint32_t* blo;            // [n_tiles] start of the tile's byte window
};

struct vtn_uniq_buffers;  // vtn_uniq.cu
struct vtn_order_buffers; // vtn_order.cu
struct vtn_blocksub_buffers; // vtn_blocksub.cu
struct vtn_dbatch
{
    uint32_t n_rec = 0, n_allele = 0;
    uint64_t n_bytes = 0;
    uint32_t cap_rec = 0, cap_allele = 0; // allocated capacities (n_* <= cap_*)
    uint64_t cap_bytes = 0;
    // inputs
    int32_t* pos1 = nullptr;
    int32_t* rid = nullptr;
    uint32_t* allele_beg = nullptr;
    uint32_t* allele_off = nullptr;
    uint8_t* bytes = nullptr;
    // normalize outputs (capacity n_rec / n_bytes: new alleles never grow)
    uint8_t* status = nullptr;
    uint8_t* vtype = nullptr;
    uint32_t* chg_rec = nullptr;
    int32_t* chg_pos1 = nullptr;
    uint32_t* chg_lt = nullptr;
    uint32_t* chg_le = nullptr;
    uint32_t* chg_rt = nullptr;
    uint64_t* chg_off = nullptr;
    uint8_t* chg_bytes = nullptr;
    unsigned long long* totals = nullptr; // 2 (+ decompose: [2]=n_out)
    uint32_t* counters = nullptr;         // 12 normalize + 4 decompose
    VtnDevScratch scr{};
    // scan state
    unsigned long long* tile_state = nullptr; // [2*n_tiles]
    uint32_t* tile_counter = nullptr; // control block of vtn_launch_normalize: 2 * vtn_normalize_ctl_words(n_tiles) words
    uint32_t n_tiles = 0;
    // decompose outputs / fused intermediates
    uint32_t* src_rec = nullptr;
    uint16_t* alt_idx = nullptr;
    uint64_t out_bound = 0;    // sum(max(n_allele-1,1))
    struct vtn_dbatch* fused = nullptr; // biallelic batch materialised on device for the fused path
    bool owns_inputs = true;
    // compact biallelic layout (vtn_batch_bi): lens != nullptr; status doubles as sv, vtype / rid / allele_* stay null
    uint16_t* lens = nullptr;
    uint32_t* tile_off = nullptr;
    uint32_t* run_start = nullptr;
    int32_t* run_rid = nullptr;
    uint32_t* ext_rec = nullptr;
    uint32_t* ext_len = nullptr;
    uint32_t n_run = 0, n_ext = 0, cap_run = 0, cap_ext = 0;
    uint32_t* note_rec = nullptr;
    uint8_t* note_status = nullptr;
    uint32_t* note_count = nullptr;
    vtn_uniq_buffers* uniq = nullptr; // order / first / heap scratch of vtn_uniq_resident (allocated on first use)
    vtn_order_buffers* order = nullptr; // arena of vtn_writer_order_resident (allocated on first use)
    vtn_blocksub_buffers* blocksub = nullptr; // plan buffers of vtn_blocksub_resident (allocated on first use)
};

struct vtn_ctx
{
    int device = 0;
    cudaStream_t stream = nullptr;   // compute + primary copies
    cudaStream_t stream2 = nullptr;  // H2D stream of the pipelined host-buffer call
    cudaStream_t stream3 = nullptr;  // D2H stream of the pipelined host-buffer call
    std::vector<cudaEvent_t> chunk_events;
    unsigned long long* pinned_totals = nullptr;
    uint32_t chunk_rec = 1u << 20;   // records per chunk of the pipelined host-buffer call
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    int sm_count = 0;
    std::string err;
    uint64_t launches = 0;
    // reference
    std::vector<VtnContig> contigs;
    VtnDevRef dref{};
    void* d_groups = nullptr;
    void* d_contig_goff = nullptr;
    void* d_contig_len = nullptr;
    void* d_exc_start = nullptr;
    void* d_exc_len = nullptr;
    void* d_exc_byte = nullptr;
    int64_t ref_bytes = 0;
    uint64_t ref_n_groups = 0; // 16-byte groups behind d_groups (vtn_ref_clone)
    bool ref_loaded = false;
    struct VtnRefSource* ref_src = nullptr; // a FASTA opened by vtn_ref_open_fasta whose contigs load on first use (vtn_ref.cu)
    uint32_t ref_missing = 0;               // contigs of ref_src that are not resident yet; 0 for an eagerly loaded reference
    // reusable device arena for the host-buffer call
    vtn_dbatch* scratch = nullptr;
    vtn_dbatch* scratch_bi = nullptr; // the same for vtn_normalize_bi
    uint32_t scratch_rec = 0, scratch_allele = 0;
    uint64_t scratch_bytes = 0;
    // HMM family (vtn_hmm.cu): per-warp DP scratch (band boundary rows, back-pointers, reversed path) + work counter
    void* hmm_scratch = nullptr;
    size_t hmm_scratch_bytes = 0;
    unsigned* hmm_counter = nullptr;
    std::string annot_trace; // vtn_annot_set_trace
    void* order_cache = nullptr;               // arena of the host-buffer vtn_writer_order call, kept between calls (vtn_order.cu)
    struct vtn_vcf_dtext* tok_cache = nullptr; // device buffers of the host-buffer vtn_vcf_tokenize call, kept between calls
};
void vtn_hmm_ctx_release(vtn_ctx* ctx);
void vtn_order_ctx_release(vtn_ctx* ctx);
void vtn_ref_release(vtn_ctx* ctx); // frees the resident reference and a lazily opened FASTA (vtn_ctx_destroy)
int vtn_ref_clone(vtn_ctx* dst, const vtn_ctx* src);
int vtn_ref_ensure(vtn_ctx* ctx, uint64_t n, const int32_t* rid); // lazy reference: make every contig named in rid[] resident // replicate src's resident reference on dst's device (peer copies)
void vtn_uniq_dbatch_release(vtn_dbatch* d);
void vtn_order_dbatch_release(vtn_dbatch* d);
void vtn_blocksub_dbatch_release(vtn_dbatch* d);

int vtn_set_err(vtn_ctx* ctx, int code, const std::string& msg);
#define VTN_CUDA(ctx, call)                                                                              \
    do                                                                                                   \
    {                                                                                                    \
        cudaError_t e__ = (call);                                                                        \
        if (e__ != cudaSuccess)                                                                          \
            return vtn_set_err((ctx), VTN_E_CUDA, std::string(#call) + ": " + cudaGetErrorString(e__)); \
    } while (0)

// kernels' host launchers (vtn_normalize.cu / vtn_decompose.cu)
int vtn_launch_normalize(vtn_ctx* ctx, const VtnDevBatchView& in, const VtnDevNormOut& out, const VtnDevScratch& scr,
                         unsigned long long* tile_state, uint32_t* tile_counter, uint32_t n_tiles, bool accumulate = false);
int vtn_launch_normalize_bi(vtn_ctx* ctx, const VtnDevBiView& in, const VtnDevNormOut& out, const VtnDevNotes& notes, const VtnDevScratch& scr,
                            unsigned long long* tile_state, uint32_t* tile_counter, uint32_t n_tiles, bool accumulate = false);
int vtn_normalize_resident_bi(vtn_ctx* ctx, vtn_dbatch* d);
// one shard of a multi-device decompose -> normalize (vtn_api.cu; driven by vtn_multi.cu)
int vtn_shard_decompose_run(vtn_ctx* ctx, const vtn_batch* in, uint32_t r0, uint32_t r1, uint64_t* n_out, uint64_t* n_changed, uint64_t* n_changed_bytes);
int vtn_shard_decompose_fetch(vtn_ctx* ctx, uint32_t rec_base, uint64_t out_base, uint64_t chg_base, uint64_t byte_base, uint64_t n_out, uint64_t n_changed,
                              uint64_t n_changed_bytes, uint32_t* src_rec, uint16_t* alt_idx, vtn_decomp_counters* dc, vtn_norm_result* out);
uint32_t vtn_normalize_tile_records();
size_t vtn_normalize_ctl_words(uint32_t n_tiles); // 64-bit words behind vtn_dbatch::tile_counter
int vtn_launch_decompose(vtn_ctx* ctx, vtn_dbatch* db, bool materialise);
bool vtn_dbatch_resize(vtn_dbatch* d, uint32_t n_rec, uint32_t n_allele, uint64_t n_bytes);
int vtn_dbatch_alloc(vtn_ctx* ctx, uint32_t n_rec, uint32_t n_allele, uint64_t n_bytes, bool with_inputs, vtn_dbatch** out);
