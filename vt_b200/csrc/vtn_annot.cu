// vtn_annot.cu -- `vt annotate_indels` behind the C ABI (include/vtn.h: vtn_annotate_indels, vtn_ref_fetch_batch).
//
// What runs where:
//   device  vta_region_kernel   per record: classify_variant (variant_manip.cpp:227-508), is_normalized (:109-197), contains_N
//                               (:202-222), and for indels CandidateRegionExtractor::extract_regions_by_exact_alignment
//                               (candidate_region_extractor.cpp:162-267, trim :472-502): per ALT trim, right_align, left_align
//                               against the resident reference, merged into [min_beg1, max_end1]
//           vta_fetch_kernel    ReferenceSequence::fetch_seq (reference_sequence.cpp:150-168) for a whole list of windows
//           vtn_hmm_kernel      every RFHMM / LFHMM / AHMM alignment (vtn_hmm.cu), one batched launch per round
//   host    the per-indel control logic around them, which is string bookkeeping on tracts of tens of bases:
//           CandidateMotifPicker (candidate_motif_picker.cpp:54-150, :233-450) over MotifTree's k-mer census
//           (motif_tree.cpp:100-367, motif_map.cpp:86-185), FlankDetector's anchor clipping, phase choice, slen growth, purity
//           bookkeeping and composition / entropy (flank_detector.cpp:103-281, :382-507, :646-887), VNTR string helpers
//           (vntr.cpp:124-292).  libstdc++'s priority_queue decides ties between equally good motifs in the reference
//           (motif_tree.h:80-95); the same container with the same push order is used here.
// The reference walks one indel at a time (9 ms .. 150 ms each, nearly all of it in the three aligners); here every indel of
// the batch advances in lock step, so each FlankDetector round is one device call over all indels still open.
#include <math.h>
#include <stdio.h>
#include <string.h>
#include <algorithm>
#include <map>
#include <queue>
#include <string>
#include <vector>
#include "vtn_internal.h"
#include "vtn_ref_dev.cuh"
#include "vtn_classify.cuh"
#include "vtn_annot_host.h"

namespace
{

// ---------------------------------------------------------------------------------------------------------
// device
// ---------------------------------------------------------------------------------------------------------
// ReferenceSequence::fetch_base (reference_sequence.cpp:76-106): upper-cased base at 1-based p1, clipped like faidx
__device__ __forceinline__ uint8_t base1(const VtnDevRef& R, VtnRefCursor& cur, uint64_t gbase, int64_t clen, int64_t p1)
{
    int64_t p0 = p1 - 1;
    p0 = p0 < 0 ? 0 : (p0 >= clen ? clen - 1 : p0);
    return cur.uc(R, gbase + (uint64_t)p0);
}

struct AnnotDev
{
    uint8_t* flags;
    uint8_t* vtype;
    int32_t* reg_beg; // exact region before the anchor clipping, 1-based
    int32_t* reg_end;
    int32_t* max_dlen; // Variant::max_dlen (variant_manip.cpp:486)
    uint32_t* a_lt;    // per allele slot (ALTs): bases trim() took from the left / right, and right_align's steps
    uint32_t* a_rt;
    uint32_t* a_shift;
};

// what VariantManip says about a record from its alleles alone: classify_variant's type mask (variant_manip.cpp:337-490), is_normalized
// (:109-197) and contains_N (:202-222)
__device__ __forceinline__ void vta_site_props(const VtnDevBatchView& B, uint32_t ab, uint32_t n, const uint8_t* ref, uint32_t rlen, uint32_t& vtype,
                                               int32_t& max_dlen, bool& normalized, bool& has_n)
{
    // ---- classify_variant ----
    vtype = 0;
    bool homog = true;
    max_dlen = 0;
    for (uint32_t i = 1; i < n; ++i)
    {
        const uint32_t o = B.allele_off[ab + i], l = B.allele_off[ab + i + 1] - o;
        const uint32_t t = classify_alt(ref, rlen, B.bytes + o, l, homog);
        vtype |= t;
        // max_dlen only moves in the explicit-bases branch (variant_manip.cpp:337-490): not symbolic, not breakend, not REF
        bool sym = false;
        for (uint32_t j = 0; j < l; ++j) sym |= (B.bytes[o + j] == '<') | (B.bytes[o + j] == '[') | (B.bytes[o + j] == ']');
        bool same = (l == rlen);
        for (uint32_t j = 0; same && j < l; ++j) same = B.bytes[o + j] == ref[j];
        if (!sym && B.bytes[o] != '.' && !same)
        {
            const int32_t d = (int32_t)l - (int32_t)rlen;
            max_dlen = d > max_dlen ? d : max_dlen;
        }
    }
    if (!(vtype & (VTN_VT_VNTR | VTN_VT_SV)) && vtype != VTN_VT_REF && homog && rlen > 1 && n > 1) vtype |= VTN_VT_MNP;
    // ---- is_normalized (variant_manip.cpp:109-197) ----
    normalized = true;
    if (n >= 2)
    {
        bool exists_one = (rlen == 1), first_same = true, last_same = true, same = true;
        const uint8_t fb = vtn_toupper(ref[0]), lb = vtn_toupper(ref[rlen - 1]);
        uint32_t m = rlen, M = rlen;
        for (uint32_t i = 1; i < n; ++i)
        {
            const uint32_t o = B.allele_off[ab + i], l = B.allele_off[ab + i + 1] - o;
            const uint8_t* a = B.bytes + o;
            exists_one |= (l == 1);
            first_same &= (vtn_toupper(a[0]) == fb);
            last_same &= (vtn_toupper(a[l - 1]) == lb);
            if (l == rlen)
            {
                for (uint32_t j = 0; j < l && same; ++j) same = vtn_toupper(a[j]) == vtn_toupper(ref[j]);
            }
            else
                same = false;
            m = l < m ? l : m;
            M = l > M ? l : M;
        }
        if (n == 2 && m == 1 && M == 1) normalized = true;
        else if (same) normalized = true;
        else normalized = !(last_same || (!exists_one && first_same));
    }
    // ---- contains_N (variant_manip.cpp:202-222) ----
    has_n = false;
    for (uint32_t i = 0; i < n; ++i)
    {
        const uint32_t o = B.allele_off[ab + i], l = B.allele_off[ab + i + 1] - o;
        bool f = false;
        for (uint32_t j = 0; j < l; ++j) f |= (B.bytes[o + j] == 'N');
        if (f && !(i && B.bytes[o] == '<')) has_n = true;
    }
}

// VariantManip's allele-only predicates for a batch (vtn_variant_props): one thread per record
__global__ void __launch_bounds__(128) vta_props_kernel(VtnDevBatchView B, uint8_t* vtype_out, uint8_t* props_out)
{
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= B.n_rec) return;
    const uint32_t ab = B.allele_beg[r], n = B.allele_beg[r + 1] - ab;
    const uint32_t off0 = B.allele_off[ab], off1 = B.allele_off[ab + 1];
    uint32_t vtype;
    int32_t max_dlen;
    bool normalized, has_n;
    vta_site_props(B, ab, n, B.bytes + off0, off1 - off0, vtype, max_dlen, normalized, has_n);
    if (vtype_out) vtype_out[r] = (uint8_t)vtype;
    props_out[r] = (uint8_t)((normalized ? VTN_PROP_NORMALIZED : 0) | (has_n ? VTN_PROP_CONTAINS_N : 0));
}

__global__ void __launch_bounds__(128) vta_region_kernel(VtnDevRef R, VtnDevBatchView B, AnnotDev O)
{
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= B.n_rec) return;
    const uint32_t ab = B.allele_beg[r], n = B.allele_beg[r + 1] - ab;
    const int32_t pos1 = B.pos1[r], rid = B.rid[r];
    const uint32_t off0 = B.allele_off[ab], off1 = B.allele_off[ab + 1];
    const uint8_t* ref = B.bytes + off0;
    const uint32_t rlen = off1 - off0;
    uint32_t vtype;
    int32_t max_dlen;
    bool normalized, has_n;
    vta_site_props(B, ab, n, ref, rlen, vtype, max_dlen, normalized, has_n);
    O.vtype[r] = (uint8_t)vtype;
    O.max_dlen[r] = max_dlen;
    uint32_t flags = (normalized && !has_n) ? VTA_KEEP : 0u;
    int32_t min_beg1 = pos1, max_end1 = pos1;
    if (flags && (vtype & VTN_VT_INDEL))
    {
        flags |= VTA_INDEL;
        if (rid < 0 || rid >= R.n_contig || R.contig_len[rid] <= 0)
            flags |= VTA_NOCONTIG;
        else
        {
            const uint64_t gbase = R.contig_goff[rid] << 5;
            const int64_t clen = R.contig_len[rid];
            VtnRefCursor cur;
            for (uint32_t i = 1; i < n; ++i)
            {
                const uint32_t o = B.allele_off[ab + i], alen = B.allele_off[ab + i + 1] - o;
                const uint8_t* alt = B.bytes + o;
                // trim (candidate_region_extractor.cpp:472-502): one side at a time, right side first, raw bytes
                uint32_t rs = 0, rl = rlen, as = 0, al = alen;
                int64_t p1 = pos1;
                while (rl != 1 && al != 1)
                {
                    const bool last = ref[rs + rl - 1] == alt[as + al - 1], first = ref[rs] == alt[as];
                    if (!first && !last) break;
                    if (last) { --rl; --al; }
                    else { ++rs; ++as; --rl; --al; ++p1; }
                }
                // right_align (:256-267): while ref[0] == alt[0], both lose their first base and gain the next reference base
                const int64_t end1 = p1 + rl - 1;
                uint32_t s = 0;
                const uint32_t s_cap = (uint32_t)(clen > 0x3fffffff ? 0x3fffffff : clen) + rl + al;
                for (;;)
                {
                    const uint8_t cr = s < rl ? ref[rs + s] : base1(R, cur, gbase, clen, end1 + 1 + (int64_t)(s - rl));
                    const uint8_t ca = s < al ? alt[as + s] : base1(R, cur, gbase, clen, end1 + 1 + (int64_t)(s - al));
                    if (cr != ca) break;
                    if (++s > s_cap) { flags |= VTA_RUNAWAY; break; }
                }
                const int64_t end1r = end1 + s;
                // left_align (:239-251) on the right-aligned strings R'(i) = Rext(s+i): while the last bases agree and pos1 > 1, both lose
                // their last base and gain fetch_base(pos1-2) in front -- the base TWO to the left, as the reference has it
                const int64_t beg1 = end1r - rl + 1;
                uint32_t t = 0;
                for (;;)
                {
                    if (beg1 - (int64_t)t <= 1) break;
                    const int64_t ir = (int64_t)rl - 1 - t, ia = (int64_t)al - 1 - t; // index into R'L / A'L; negative = prepended bases
                    uint8_t cr, ca;
                    if (ir >= 0)
                    {
                        const uint32_t j = s + (uint32_t)ir;
                        cr = j < rl ? ref[rs + j] : base1(R, cur, gbase, clen, end1 + 1 + (int64_t)(j - rl));
                    }
                    else
                        cr = base1(R, cur, gbase, clen, beg1 + ir - 1);
                    if (ia >= 0)
                    {
                        const uint32_t j = s + (uint32_t)ia;
                        ca = j < al ? alt[as + j] : base1(R, cur, gbase, clen, end1 + 1 + (int64_t)(j - al));
                    }
                    else
                        ca = base1(R, cur, gbase, clen, beg1 + ia - 1);
                    if (cr != ca) break;
                    ++t;
                }
                const int64_t beg1l = beg1 - t;
                min_beg1 = beg1l < min_beg1 ? (int32_t)beg1l : min_beg1;
                max_end1 = end1r > max_end1 ? (int32_t)end1r : max_end1;
                O.a_lt[ab + i] = rs;
                O.a_rt[ab + i] = rlen - rs - rl;
                O.a_shift[ab + i] = s;
            }
        }
    }
    O.flags[r] = (uint8_t)flags;
    O.reg_beg[r] = min_beg1;
    O.reg_end[r] = max_end1;
}

// one warp per window: out[off[k] .. off[k+1]) = bases [beg0[k], beg0[k] + len) of contig rid[k]
__global__ void __launch_bounds__(256) vta_fetch_kernel(VtnDevRef R, uint32_t n, const int32_t* rid, const int64_t* beg0, const uint64_t* off, int upper,
                                                        uint8_t* out)
{
    const uint32_t k = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (k >= n) return;
    const uint64_t o0 = off[k], len = off[k + 1] - o0;
    const uint64_t g = (R.contig_goff[rid[k]] << 5) + (uint64_t)beg0[k];
    VtnRefCursor cur;
    for (uint64_t i = lane; i < len; i += 32) out[o0 + i] = upper ? cur.uc(R, g + i) : cur.raw(R, g + i);
}

int fetch_windows(vtn_ctx* ctx, const vta::FetchList& L, int upper, std::vector<uint8_t>& out)
{
    const uint32_t n = (uint32_t)L.size();
    out.resize((size_t)L.off.back() + 1);
    if (!n || !L.off.back()) return VTN_OK;
    int32_t* d_rid = nullptr;
    int64_t* d_beg = nullptr;
    uint64_t* d_off = nullptr;
    uint8_t* d_out = nullptr;
    cudaStream_t s = ctx->stream;
    cudaError_t e = cudaMalloc(&d_rid, 4 * (size_t)n);
    if (e == cudaSuccess) e = cudaMalloc(&d_beg, 8 * (size_t)n);
    if (e == cudaSuccess) e = cudaMalloc(&d_off, 8 * ((size_t)n + 1));
    if (e == cudaSuccess) e = cudaMalloc(&d_out, (size_t)L.off.back());
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_rid, L.rid.data(), 4 * (size_t)n, cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_beg, L.beg0.data(), 8 * (size_t)n, cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_off, L.off.data(), 8 * ((size_t)n + 1), cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess)
    {
        const uint32_t blocks = (uint32_t)(((uint64_t)n * 32 + 255) / 256);
        vta_fetch_kernel<<<blocks, 256, 0, s>>>(ctx->dref, n, d_rid, d_beg, d_off, upper, d_out);
        ctx->launches++;
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(out.data(), d_out, (size_t)L.off.back(), cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    cudaFree(d_rid); cudaFree(d_beg); cudaFree(d_off); cudaFree(d_out);
    if (e != cudaSuccess) return vtn_set_err(ctx, VTN_E_CUDA, std::string("reference window fetch: ") + cudaGetErrorString(e));
    return VTN_OK;
}


// the device behind the control logic
struct DeviceBackend : vta::AnnotBackend
{
    vtn_ctx* ctx;
    explicit DeviceBackend(vtn_ctx* c) : ctx(c) {}
    int64_t contig_len(int32_t rid) override { return (rid >= 0 && rid < (int32_t)ctx->contigs.size()) ? ctx->contigs[rid].len : -1; }
    int fail(int code, const std::string& msg) override { return vtn_set_err(ctx, code, msg); }
    int regions(const vtn_batch* in, vta::RegionOut& RO) override
    {
        const uint32_t n = in->n_rec;
        vtn_dbatch* db = nullptr;
        int rc = vtn_dbatch_create(ctx, in, &db);
        if (rc) return rc;
        RO.flags.resize(n); RO.vtype.resize(n); RO.reg_beg.resize(n); RO.reg_end.resize(n); RO.max_dlen.resize(n);
        RO.a_lt.resize(in->n_allele); RO.a_rt.resize(in->n_allele); RO.a_shift.resize(in->n_allele);
        std::vector<uint8_t>&flags = RO.flags, &vtype = RO.vtype;
        std::vector<int32_t>&reg_beg = RO.reg_beg, &reg_end = RO.reg_end, &max_dlen = RO.max_dlen;
        std::vector<uint32_t>&a_lt = RO.a_lt, &a_rt = RO.a_rt, &a_shift = RO.a_shift;
        cudaStream_t s = ctx->stream;
        {
            AnnotDev O{};
            cudaError_t e = cudaMalloc(&O.flags, n);
            if (e == cudaSuccess) e = cudaMalloc(&O.vtype, n);
            if (e == cudaSuccess) e = cudaMalloc(&O.reg_beg, 4 * (size_t)n);
            if (e == cudaSuccess) e = cudaMalloc(&O.reg_end, 4 * (size_t)n);
            if (e == cudaSuccess) e = cudaMalloc(&O.max_dlen, 4 * (size_t)n);
            if (e == cudaSuccess) e = cudaMalloc(&O.a_lt, 4 * (size_t)in->n_allele + 4);
            if (e == cudaSuccess) e = cudaMalloc(&O.a_rt, 4 * (size_t)in->n_allele + 4);
            if (e == cudaSuccess) e = cudaMalloc(&O.a_shift, 4 * (size_t)in->n_allele + 4);
            if (e == cudaSuccess) e = cudaMemsetAsync(O.a_lt, 0, 4 * (size_t)in->n_allele + 4, s);
            if (e == cudaSuccess) e = cudaMemsetAsync(O.a_rt, 0, 4 * (size_t)in->n_allele + 4, s);
            if (e == cudaSuccess) e = cudaMemsetAsync(O.a_shift, 0, 4 * (size_t)in->n_allele + 4, s);
            if (e == cudaSuccess)
            {
                VtnDevBatchView v{db->n_rec, 0u, db->pos1, db->rid, db->allele_beg, db->allele_off, db->bytes};
                vta_region_kernel<<<(n + 127) / 128, 128, 0, s>>>(ctx->dref, v, O);
                ctx->launches++;
                e = cudaGetLastError();
            }
            if (e == cudaSuccess) e = cudaMemcpyAsync(flags.data(), O.flags, n, cudaMemcpyDeviceToHost, s);
            if (e == cudaSuccess) e = cudaMemcpyAsync(vtype.data(), O.vtype, n, cudaMemcpyDeviceToHost, s);
            if (e == cudaSuccess) e = cudaMemcpyAsync(reg_beg.data(), O.reg_beg, 4 * (size_t)n, cudaMemcpyDeviceToHost, s);
            if (e == cudaSuccess) e = cudaMemcpyAsync(reg_end.data(), O.reg_end, 4 * (size_t)n, cudaMemcpyDeviceToHost, s);
            if (e == cudaSuccess) e = cudaMemcpyAsync(max_dlen.data(), O.max_dlen, 4 * (size_t)n, cudaMemcpyDeviceToHost, s);
            if (e == cudaSuccess && in->n_allele) e = cudaMemcpyAsync(a_lt.data(), O.a_lt, 4 * (size_t)in->n_allele, cudaMemcpyDeviceToHost, s);
            if (e == cudaSuccess && in->n_allele) e = cudaMemcpyAsync(a_rt.data(), O.a_rt, 4 * (size_t)in->n_allele, cudaMemcpyDeviceToHost, s);
            if (e == cudaSuccess && in->n_allele) e = cudaMemcpyAsync(a_shift.data(), O.a_shift, 4 * (size_t)in->n_allele, cudaMemcpyDeviceToHost, s);
            if (e == cudaSuccess) e = cudaStreamSynchronize(s);
            cudaFree(O.flags); cudaFree(O.vtype); cudaFree(O.reg_beg); cudaFree(O.reg_end); cudaFree(O.max_dlen); cudaFree(O.a_lt); cudaFree(O.a_rt); cudaFree(O.a_shift);
            vtn_dbatch_destroy(ctx, db);
            if (e != cudaSuccess) return vtn_set_err(ctx, VTN_E_CUDA, std::string("vtn_annotate_indels: ") + cudaGetErrorString(e));
        }
        return VTN_OK;
    }
    int fetch(const vta::FetchList& L, int upper, std::vector<uint8_t>& out) override { return fetch_windows(ctx, L, upper, out); }
    int hmm(const vta::HmmJob& J, std::vector<vtn_hmm_result>& res) override
    {
        res.assign(J.size(), vtn_hmm_result{});
        if (!J.size()) return VTN_OK;
        vtn_hmm_batch b{(uint32_t)J.size(), J.kind.data(), J.seq_off.data(), (const uint8_t*)J.bytes.data(), (uint64_t)J.bytes.size()};
        return vtn_hmm_align(ctx, &vta::FD_PARAMS, &b, res.data(), nullptr, 0);
    }
};

} // namespace

// ---------------------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------------------
int vtn_ref_fetch_batch(vtn_ctx* ctx, uint32_t n, const int32_t* rid, const int64_t* beg0, const int64_t* end0, int upper, uint64_t* off,
                        uint8_t* out, uint64_t cap)
{
    if (!ctx || !ctx->ref_loaded) return vtn_set_err(ctx, VTN_E_STATE, "no reference loaded");
    if (n && (!rid || !beg0 || !end0 || !off)) return vtn_set_err(ctx, VTN_E_ARG, "vtn_ref_fetch_batch: null arrays");
    VTN_CUDA(ctx, cudaSetDevice(ctx->device));
    if (ctx->ref_missing)
        if (int rc = vtn_ref_ensure(ctx, n, rid)) return rc;
    vta::FetchList L;
    for (uint32_t k = 0; k < n; ++k)
        L.add(rid[k], (rid[k] >= 0 && rid[k] < (int32_t)ctx->contigs.size()) ? ctx->contigs[rid[k]].len : -1, beg0[k] + 1, end0[k] + 1);
    for (uint32_t k = 0; k <= n; ++k) off[k] = L.off[k];
    if (L.off.back() > cap) return vtn_set_err(ctx, VTN_E_CAPACITY, "vtn_ref_fetch_batch: output too small (off[n] holds the need)");
    std::vector<uint8_t> buf;
    int rc = fetch_windows(ctx, L, upper, buf);
    if (rc) return rc;
    if (L.off.back()) memcpy(out, buf.data(), (size_t)L.off.back());
    return VTN_OK;
}

// VariantManip's allele-only predicates for a batch: classify_variant's mask, is_normalized, contains_N.  No reference needed;
// pos1 / rid of the batch are not read.
int vtn_variant_props(vtn_ctx* ctx, const vtn_batch* in, uint8_t* vtype, uint8_t* props)
{
    if (!ctx || !in || !props) return vtn_set_err(ctx, VTN_E_ARG, "vtn_variant_props: bad arguments");
    const uint32_t n = in->n_rec;
    if (!n) return VTN_OK;
    if (!in->allele_beg || !in->allele_off || (in->n_bytes && !in->bytes)) return vtn_set_err(ctx, VTN_E_ARG, "vtn_variant_props: null arrays");
    if (in->allele_beg[0] != 0 || in->allele_beg[n] != in->n_allele || in->allele_off[in->n_allele] != in->n_bytes)
        return vtn_set_err(ctx, VTN_E_ARG, "vtn_variant_props: allele_beg / allele_off do not describe the batch");
    for (uint32_t r = 0; r < n; ++r)
        if (in->allele_beg[r + 1] <= in->allele_beg[r]) return vtn_set_err(ctx, VTN_E_ARG, "vtn_variant_props: a record without alleles");
    for (uint32_t k = 0; k < in->n_allele; ++k)
        if (in->allele_off[k + 1] <= in->allele_off[k]) return vtn_set_err(ctx, VTN_E_ARG, "vtn_variant_props: an empty allele");
    VTN_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;
    uint32_t *d_beg = nullptr, *d_off = nullptr;
    uint8_t *d_bytes = nullptr, *d_vt = nullptr, *d_pr = nullptr;
    cudaError_t e = cudaMalloc(&d_beg, 4 * ((size_t)n + 1));
    if (e == cudaSuccess) e = cudaMalloc(&d_off, 4 * ((size_t)in->n_allele + 1));
    if (e == cudaSuccess) e = cudaMalloc(&d_bytes, (size_t)in->n_bytes + 1);
    if (e == cudaSuccess) e = cudaMalloc(&d_vt, n);
    if (e == cudaSuccess) e = cudaMalloc(&d_pr, n);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_beg, in->allele_beg, 4 * ((size_t)n + 1), cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_off, in->allele_off, 4 * ((size_t)in->n_allele + 1), cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_bytes, in->bytes, (size_t)in->n_bytes, cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess)
    {
        VtnDevBatchView v{n, 0u, nullptr, nullptr, d_beg, d_off, d_bytes};
        vta_props_kernel<<<(n + 127) / 128, 128, 0, s>>>(v, d_vt, d_pr);
        ctx->launches++;
        e = cudaGetLastError();
    }
    if (e == cudaSuccess && vtype) e = cudaMemcpyAsync(vtype, d_vt, n, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(props, d_pr, n, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    cudaFree(d_beg), cudaFree(d_off), cudaFree(d_bytes), cudaFree(d_vt), cudaFree(d_pr);
    if (e != cudaSuccess) return vtn_set_err(ctx, VTN_E_CUDA, std::string("vtn_variant_props: ") + cudaGetErrorString(e));
    return VTN_OK;
}

int vtn_annot_set_trace(vtn_ctx* ctx, const char* path)
{
    if (!ctx) return VTN_E_ARG;
    ctx->annot_trace = path ? path : "";
    return VTN_OK;
}

int vtn_annotate_indels(vtn_ctx* ctx, const vtn_batch* in, vtn_annot_result* out)
{
    if (!ctx || !in || !out) return VTN_E_ARG;
    if (!ctx->ref_loaded) return vtn_set_err(ctx, VTN_E_STATE, "vtn_annotate_indels: no reference loaded");
    if (in->n_rec && (!in->pos1 || !in->rid || !in->allele_beg || !in->allele_off || !out->flags || !out->vtype || !out->indel_idx))
        return vtn_set_err(ctx, VTN_E_ARG, "vtn_annotate_indels: null arrays");
    VTN_CUDA(ctx, cudaSetDevice(ctx->device));
    if (ctx->ref_missing)
        if (int rc = vtn_ref_ensure(ctx, in->n_rec, in->rid)) return rc;
    DeviceBackend be(ctx);
    return vta::annotate_core(be, in, out, ctx->annot_trace);
}
