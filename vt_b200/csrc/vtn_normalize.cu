// vtn_normalize.cu -- the normalize kernel (sm_100a) and its launcher.
//
// One thread owns one record; a CTA owns a tile of VTN_TILE consecutive records, fetched from a global
// tile counter (persistent CTAs).  Per record, in closed form instead of the reference's per-base loop:
//   classify_variant          variant_manip.cpp:227-508
//   is_not_ref_consistent     variant_manip.cpp:61-103      (skipped iff type == VT_SNP, normalize.cpp:205)
//   is_normalized             variant_manip.cpp:109-197
//   right_trim_or_left_extend variant_manip.cpp:513-583  ->  T = longest common suffix of E_i = prefix+allele_i
//   left_trim                 variant_manip.cpp:588-615
//   counters                  normalize.cpp:254-322
// Changed records are compacted in input order with a single-pass (decoupled look-back) scan over tiles of
// two quantities: number of changed records and bytes of their new alleles.
#include "vtn_internal.h"
#include "vtn_ref_dev.cuh"
#include "vtn_scan.cuh"

#define VTN_TILE 256

namespace
{

struct RecResult
{
    uint32_t T;      // right_trimmed
    uint32_t le;     // left_extended
    uint32_t lt;     // left_trimmed
    uint32_t nbytes; // total bytes of the new alleles (0 when unchanged)
    uint8_t status;
    uint8_t vtype;
};

__device__ __forceinline__ bool has_byte(const uint8_t* a, uint32_t l, uint8_t c)
{
    for (uint32_t i = 0; i < l; ++i)
        if (a[i] == c) return true;
    return false;
}

// classify one ALT against REF: variant_manip.cpp:254-491.  Returns VT_* bits, clears homog when lengths differ.
__device__ uint32_t classify_alt(const uint8_t* ref0, uint32_t rlen, const uint8_t* a, uint32_t l, bool& homog)
{
    bool lt = false, br = false;
    for (uint32_t i = 0; i < l; ++i)
    {
        uint8_t c = a[i];
        lt |= (c == '<');
        br |= (c == '[') | (c == ']');
    }
    if (lt)
    {
        uint32_t type = VTN_VT_REF;
        if (l >= 5)
        {
            if (a[0] == '<' && a[1] == 'V' && a[2] == 'N' && a[l - 1] == '>')
            {
                for (uint32_t j = 3; j < l - 1; ++j)
                    if (a[j] < '0' || a[j] > '9') type = VTN_VT_VNTR;
            }
            else if (l == 5 && a[0] == '<' && a[1] == 'S' && a[2] == 'T' && a[3] == 'R' && a[4] == '>')
                type = VTN_VT_VNTR;
            else if (a[0] == '<' && a[1] == 'S' && a[2] == 'T' && a[l - 1] == '>')
            {
                type = VTN_VT_VNTR;
                for (uint32_t j = 3; j < l - 1; ++j)
                    if ((a[j] < '0' || a[j] > '9') && a[j] != '.') type = VTN_VT_SV;
            }
        }
        return type == VTN_VT_VNTR ? VTN_VT_VNTR : VTN_VT_SV;
    }
    if (br) return VTN_VT_SV;
    if (a[0] == '.') return VTN_VT_REF;
    if (l == rlen)
    {
        bool same = true;
        for (uint32_t i = 0; i < l; ++i) same &= (a[i] == ref0[i]);
        if (same) return VTN_VT_REF;
    }
    else
        homog = false;
    // case-sensitive trim to a core (right first, then left), :360-390
    uint32_t rl = rlen, al = l, s = 0;
    while (rl != 1 && al != 1 && ref0[rl - 1] == a[al - 1]) { --rl; --al; }
    while (rl != 1 && al != 1 && ref0[s] == a[s]) { ++s; --rl; --al; }
    uint32_t mlen = rl < al ? rl : al;
    uint32_t diff = 0;
    if (mlen == 1 && rl != al)
    {
        uint8_t ls, le, ss;
        if (rl > al) { ls = ref0[s]; le = ref0[s + rl - 1]; ss = a[s]; }
        else { ls = a[s]; le = a[s + al - 1]; ss = ref0[s]; }
        if (ls != ss && le != ss) diff = 1;
    }
    else
    {
        for (uint32_t j = 0; j < mlen; ++j) diff += (ref0[s + j] != a[s + j]);
    }
    uint32_t type = 0;
    if (mlen == diff) type |= (mlen == 1) ? VTN_VT_SNP : VTN_VT_MNP;
    if (rl != al) type |= VTN_VT_INDEL;
    if (diff && diff < mlen) type |= VTN_VT_CLUMPED;
    return type;
}

// is_not_ref_consistent, variant_manip.cpp:61-103 with faidx clipping (faidx.c:832-860)
__device__ uint32_t ref_check(const VtnDevRef& R, uint64_t gbase, int64_t clen, int32_t pos1, const uint8_t* a0, uint32_t len0)
{
    int64_t beg = (int64_t)pos1 - 1, end = beg + (int64_t)len0 - 1;
    beg = beg < 0 ? 0 : (beg >= clen ? clen - 1 : beg);
    end = end < 0 ? 0 : (end >= clen ? clen - 1 : end);
    int64_t n = end - beg + 1;
    if (n > (int64_t)len0) n = len0;
    VtnRefCursor cur;
    uint32_t code = 0;
    for (int64_t i = 0; i < n; ++i)
    {
        uint8_t u = vtn_toupper(a0[i]);
        uint8_t f = cur.raw(R, gbase + (uint64_t)(beg + i));
        if (u != f)
        {
            if (f == 'N' || u == 'N') code = 1;
            else { code = 2; break; }
        }
    }
    return code;
}

// Byte k (counted from the right end, k = 0 is the last base) of E_i = contig prefix + upper(allele_i):
// inside the allele for k < len, else the upper-cased reference base left of POS (faidx_fetch_uc_seq at
// pos1-1-(k-len) in 1-based terms), clipped into the contig like faidx_adjust_position.
__device__ __forceinline__ uint8_t ext_byte(const VtnDevRef& R, VtnRefCursor& cur, uint64_t gbase, int64_t clen,
                                            int32_t pos1, const uint8_t* a, uint32_t len, uint32_t k)
{
    if (k < len) return vtn_toupper(a[len - 1 - k]);
    int64_t p0 = (int64_t)pos1 - 2 - (int64_t)(k - len);
    p0 = p0 < 0 ? 0 : (p0 >= clen ? clen - 1 : p0);
    return cur.uc(R, gbase + (uint64_t)p0);
}

// The whole per-record computation except writing the new alleles.
__device__ void normalize_record(const VtnDevRef& R, const VtnDevBatchView& B, uint32_t r, RecResult& out)
{
    const uint32_t ab = __ldg(B.allele_beg + r);
    const uint32_t n = __ldg(B.allele_beg + r + 1) - ab;
    const int32_t pos1 = __ldg(B.pos1 + r);
    const int32_t rid = __ldg(B.rid + r);
    const uint32_t off0 = __ldg(B.allele_off + ab);
    const uint32_t off1 = __ldg(B.allele_off + ab + 1);
    const uint8_t* a0 = B.bytes + off0;
    const uint32_t len0 = off1 - off0;
    out.T = out.le = out.lt = out.nbytes = 0;

    // ---- SNV fast path: biallelic, both alleles one byte ----
    if (n == 2)
    {
        const uint32_t off2 = __ldg(B.allele_off + ab + 2);
        if (len0 == 1 && off2 - off1 == 1)
        {
            uint8_t r0 = a0[0], c = a0[1];
            if (c != '<' && c != '[' && c != ']' && c != '.' && c != r0)
            {
                out.status = 0;
                out.vtype = VTN_VT_SNP;
                return;
            }
        }
    }

    // ---- classify_variant ----
    uint32_t vtype = 0;
    bool homog = true;
    uint32_t m = len0, M = len0; // min / max allele length
    {
        uint32_t o = off1;
        for (uint32_t i = 1; i < n; ++i)
        {
            uint32_t o2 = __ldg(B.allele_off + ab + i + 1);
            uint32_t l = o2 - o;
            vtype |= classify_alt(a0, len0, B.bytes + o, l, homog);
            m = l < m ? l : m;
            M = l > M ? l : M;
            o = o2;
        }
    }
    if (!(vtype & (VTN_VT_VNTR | VTN_VT_SV)) && vtype != VTN_VT_REF && homog && len0 > 1 && n > 1) vtype |= VTN_VT_MNP;
    out.vtype = (uint8_t)vtype;

    const bool rid_ok = rid >= 0 && rid < R.n_contig;
    uint64_t gbase = 0;
    int64_t clen = 0;
    if (rid_ok)
    {
        gbase = __ldg(R.contig_goff + rid) << 5;
        clen = __ldg(R.contig_len + rid);
    }

    // ---- REF consistency (normalize.cpp:205) ----
    uint32_t st = 0;
    if (vtype != VTN_VT_SNP)
    {
        if (!rid_ok || clen <= 0) st |= VTN_ST_FETCH_FAIL;
        else st |= ref_check(R, gbase, clen, pos1, a0, len0);
    }
    out.status = (uint8_t)st;
    if (st || (vtype & (VTN_VT_SV | VTN_VT_VNTR)) || n < 2) return;

    // ---- is_normalized (variant_manip.cpp:109-197) ----
    {
        bool exists_one = (len0 == 1), first_same = true, last_same = true, same = true;
        uint8_t fb = vtn_toupper(a0[0]), lb = vtn_toupper(a0[len0 - 1]);
        uint32_t o = off1;
        for (uint32_t i = 1; i < n; ++i)
        {
            uint32_t o2 = __ldg(B.allele_off + ab + i + 1);
            uint32_t l = o2 - o;
            const uint8_t* a = B.bytes + o;
            exists_one |= (l == 1);
            first_same &= (vtn_toupper(a[0]) == fb);
            last_same &= (vtn_toupper(a[l - 1]) == lb);
            if (l == len0)
            {
                for (uint32_t j = 0; j < l && same; ++j) same = vtn_toupper(a[j]) == vtn_toupper(a0[j]);
            }
            else
                same = false;
            o = o2;
        }
        if (n == 2 && len0 == 1 && m == 1 && M == 1) return;
        if (same) return;
        if (!(last_same || (!exists_one && first_same))) return;
    }
    if (!rid_ok || clen <= 0) // unreachable (a pure-SNP record is always normalized); kept for safety
    {
        out.status = VTN_ST_FETCH_FAIL;
        return;
    }
    st |= VTN_ST_ENTERED;

    // ---- right_trim_or_left_extend in closed form ----
    // cap = min_i len(E_i) - 1 : the pos1==1 guard of variant_manip.cpp:536-540
    const int64_t cap64 = (int64_t)pos1 - 2 + (int64_t)m;
    const uint32_t cap = cap64 < 0 ? 0u : (cap64 > 0xffffffffll ? 0xffffffffu : (uint32_t)cap64);
    uint32_t T = 0;
    VtnRefCursor cur;
    bool stop = false;
    // phase A+B per base, until every allele is exhausted (k >= M) or a mismatch
    while (!stop && T < cap && (T < M))
    {
        uint8_t c0 = ext_byte(R, cur, gbase, clen, pos1, a0, len0, T);
        uint32_t o = off1;
        for (uint32_t i = 1; i < n; ++i)
        {
            uint32_t o2 = __ldg(B.allele_off + ab + i + 1);
            uint8_t c = ext_byte(R, cur, gbase, clen, pos1, B.bytes + o, o2 - o, T);
            if (c != c0) { stop = true; break; }
            o = o2;
        }
        if (!stop) ++T;
    }
    // phase C: k >= M, every stream reads the reference; allele i lags allele 0 by len_i - len0 bases
    if (!stop && T < cap)
    {
        if (n == 2)
        {
            const uint32_t len1 = __ldg(B.allele_off + ab + 2) - off1;
            // 32-base chunks, descending: window of stream s ends (inclusive) at hi_s = pos1-2-T+len_s
            while (T < cap)
            {
                int64_t hi0 = (int64_t)pos1 - 2 - (int64_t)T + (int64_t)len0;
                int64_t hi1 = (int64_t)pos1 - 2 - (int64_t)T + (int64_t)len1;
                int64_t lo = (hi0 < hi1 ? hi0 : hi1) - 31, hi = hi0 > hi1 ? hi0 : hi1;
                bool fast = lo >= 0 && hi < clen;
                uint32_t nm = 0;
                if (fast)
                {
                    uint64_t c0, c1;
                    uint32_t e0, e1;
                    vtn_ref_window32(R, gbase + (uint64_t)(hi0 - 31), c0, e0);
                    vtn_ref_window32(R, gbase + (uint64_t)(hi1 - 31), c1, e1);
                    if ((e0 | e1) == 0)
                    {
                        uint64_t x = c0 ^ c1;
                        uint64_t mm = (x | (x >> 1)) & 0x5555555555555555ull;
                        nm = mm ? (uint32_t)(__clzll((long long)mm) >> 1) : 32u;
                    }
                    else
                        fast = false;
                }
                if (fast)
                {
                    uint32_t room = cap - T;
                    if (nm >= room) { T = cap; break; }
                    T += nm;
                    if (nm < 32) break;
                }
                else
                {
                    // per-base for up to 32 steps (contig edge, or N / IUPAC bases in the window)
                    uint32_t steps = 0;
                    bool mism = false;
                    while (T < cap && steps < 32)
                    {
                        uint8_t c0 = ext_byte(R, cur, gbase, clen, pos1, a0, len0, T);
                        uint8_t c1 = ext_byte(R, cur, gbase, clen, pos1, a0, len1, T); // k >= len: pointer unused
                        if (c0 != c1) { mism = true; break; }
                        ++T;
                        ++steps;
                    }
                    if (mism) break;
                }
            }
        }
        else
        {
            while (T < cap)
            {
                uint8_t c0 = ext_byte(R, cur, gbase, clen, pos1, a0, len0, T);
                uint32_t o = off1;
                bool mism = false;
                for (uint32_t i = 1; i < n; ++i)
                {
                    uint32_t o2 = __ldg(B.allele_off + ab + i + 1);
                    uint8_t c = ext_byte(R, cur, gbase, clen, pos1, B.bytes + o, o2 - o, T);
                    if (c != c0) { mism = true; break; }
                    o = o2;
                }
                if (mism) break;
                ++T;
            }
        }
    }
    const uint32_t le = (T >= m) ? T - m + 1 : 0;
    // ---- left_trim (only possible when nothing was left-extended) ----
    uint32_t lt = 0;
    if (le == 0)
    {
        const uint32_t room = m - T - 1; // every allele keeps >= 1 base
        while (lt < room)
        {
            uint8_t c0 = vtn_toupper(a0[lt]);
            uint32_t o = off1;
            bool mism = false;
            for (uint32_t i = 1; i < n; ++i)
            {
                if (vtn_toupper(B.bytes[o + lt]) != c0) { mism = true; break; }
                o = __ldg(B.allele_off + ab + i + 1);
            }
            if (mism) break;
            ++lt;
        }
    }
    out.T = T;
    out.le = le;
    out.lt = lt;
    if (T | le | lt)
    {
        st |= VTN_ST_CHANGED;
        uint32_t cat = 0; // normalize.cpp:254-311
        if (le) cat = (T > le) ? 4 : 5;
        else if (lt && T > le) cat = 3;
        else if (lt) cat = 1;
        else if (T > le) cat = 2;
        st |= cat << VTN_ST_CAT_SHIFT;
        // sum of new lengths = sum(len_i) - n*(T - le + lt)
        const uint32_t total = __ldg(B.allele_off + ab + n) - off0;
        out.nbytes = total + n * le - n * T - n * lt;
    }
    out.status = (uint8_t)st;
}

} // namespace

namespace
{

__global__ void __launch_bounds__(VTN_TILE) vtn_normalize_kernel(VtnDevRef R, VtnDevBatchView B, VtnDevNormOut O,
                                                                 unsigned long long* tile_state, uint32_t* tile_counter,
                                                                 uint32_t n_tiles)
{
    __shared__ uint32_t s_tile;
    __shared__ uint32_t s_wcnt[VTN_TILE / 32];
    __shared__ uint32_t s_wbytes[VTN_TILE / 32];
    __shared__ unsigned long long s_base[2];
    __shared__ uint32_t s_counters[12];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid < 12) s_counters[tid] = 0;

    unsigned long long* st_cnt = tile_state;
    unsigned long long* st_bytes = tile_state + n_tiles;

    for (;;)
    {
        __syncthreads(); // protects s_tile / s_w* / s_base reuse across iterations
        if (tid == 0) s_tile = atomicAdd(tile_counter, 1u);
        __syncthreads();
        const uint32_t tile = s_tile;
        if (tile >= n_tiles) break;
        const uint32_t r = tile * VTN_TILE + tid;
        const bool live = r < B.n_rec;

        RecResult res;
        res.T = res.le = res.lt = res.nbytes = 0;
        res.status = 0;
        res.vtype = 0;
        if (live)
        {
            normalize_record(R, B, r, res);
            O.status[r] = res.status;
            O.vtype[r] = res.vtype;
        }
        const bool changed = live && (res.status & VTN_ST_CHANGED);

        // counters (warp-aggregated): 10 categories + no_variants / no_refs
        {
            uint32_t cat = changed ? (res.status >> VTN_ST_CAT_SHIFT) : 0;
            uint32_t nall = live ? (__ldg(B.allele_beg + r + 1) - __ldg(B.allele_beg + r)) : 0;
            uint32_t slot = cat ? ((nall == 2 ? 0u : 5u) + cat - 1u) : 0xffu;
            unsigned any = __ballot_sync(0xffffffffu, cat != 0);
            while (any)
            {
                int leader = __ffs(any) - 1;
                uint32_t ls = __shfl_sync(0xffffffffu, slot, leader);
                unsigned peers = __ballot_sync(0xffffffffu, slot == ls);
                if (lane == leader) atomicAdd(&s_counters[ls], __popc(peers));
                any &= ~peers;
            }
            unsigned refs = __ballot_sync(0xffffffffu, live && res.vtype == VTN_VT_REF);
            unsigned vars = __ballot_sync(0xffffffffu, live && res.vtype != VTN_VT_REF);
            if (lane == 0)
            {
                if (vars) atomicAdd(&s_counters[10], __popc(vars));
                if (refs) atomicAdd(&s_counters[11], __popc(refs));
            }
        }

        // block exclusive scan of (changed, nbytes)
        uint32_t c = changed ? 1u : 0u, nb = res.nbytes;
        uint32_t ic = c, ib = nb;
        for (int o = 1; o < 32; o <<= 1)
        {
            uint32_t tc = __shfl_up_sync(0xffffffffu, ic, o);
            uint32_t tb = __shfl_up_sync(0xffffffffu, ib, o);
            if (lane >= o) { ic += tc; ib += tb; }
        }
        if (lane == 31) { s_wcnt[warp] = ic; s_wbytes[warp] = ib; }
        __syncthreads();
        uint32_t wc = 0, wb = 0, tot_c = 0, tot_b = 0;
#pragma unroll
        for (int w = 0; w < VTN_TILE / 32; ++w)
        {
            uint32_t xc = s_wcnt[w], xb = s_wbytes[w];
            if (w < warp) { wc += xc; wb += xb; }
            tot_c += xc;
            tot_b += xb;
        }
        // publish aggregate, look back, publish inclusive (warp 0)
        if (warp == 0)
        {
            if (lane == 0)
            {
                st_state(st_cnt + tile, (tile == 0 ? VTN_FLAG_INC : VTN_FLAG_AGG) | tot_c);
                st_state(st_bytes + tile, (tile == 0 ? VTN_FLAG_INC : VTN_FLAG_AGG) | tot_b);
            }
            unsigned long long pc = 0, pb = 0;
            if (tile)
            {
                pc = lookback(st_cnt, (int)tile, lane);
                pb = lookback(st_bytes, (int)tile, lane);
                if (lane == 0)
                {
                    st_state(st_cnt + tile, VTN_FLAG_INC | (pc + tot_c));
                    st_state(st_bytes + tile, VTN_FLAG_INC | (pb + tot_b));
                }
            }
            if (lane == 0)
            {
                s_base[0] = pc;
                s_base[1] = pb;
                if (tile == n_tiles - 1)
                {
                    O.totals[0] = pc + tot_c;
                    O.totals[1] = pb + tot_b;
                    O.chg_off[pc + tot_c] = pb + tot_b;
                }
            }
        }
        __syncthreads();
        if (changed)
        {
            const unsigned long long k = s_base[0] + wc + (ic - c);
            const unsigned long long ob = s_base[1] + wb + (ib - nb);
            O.chg_rec[k] = r;
            O.chg_pos1[k] = __ldg(B.pos1 + r) - (int32_t)res.le + (int32_t)res.lt;
            O.chg_lt[k] = res.lt;
            O.chg_le[k] = res.le;
            O.chg_rt[k] = res.T;
            O.chg_off[k] = ob;
            // new alleles: (le reference bases + upper(allele))[0 : le + len - T], or upper(allele)[lt : len - T]
            const uint32_t ab = __ldg(B.allele_beg + r);
            const uint32_t n = __ldg(B.allele_beg + r + 1) - ab;
            const int32_t pos1 = __ldg(B.pos1 + r);
            const int32_t rid = __ldg(B.rid + r);
            const uint64_t gbase = __ldg(R.contig_goff + rid) << 5;
            const int64_t clen = __ldg(R.contig_len + rid);
            uint8_t* dst = O.chg_bytes + ob;
            VtnRefCursor cur;
            uint32_t o = __ldg(B.allele_off + ab);
            for (uint32_t i = 0; i < n; ++i)
            {
                const uint32_t o2 = __ldg(B.allele_off + ab + i + 1);
                const uint32_t len = o2 - o;
                const uint32_t newlen = len + res.le - res.T - res.lt;
                const uint8_t* a = B.bytes + o;
                for (uint32_t j = 0; j < newlen; ++j)
                {
                    uint8_t ch;
                    if (j < res.le)
                    {
                        int64_t p0 = (int64_t)pos1 - 1 - (int64_t)res.le + (int64_t)j;
                        p0 = p0 < 0 ? 0 : (p0 >= clen ? clen - 1 : p0);
                        ch = cur.uc(R, gbase + (uint64_t)p0);
                    }
                    else
                        ch = vtn_toupper(a[j - res.le + res.lt]);
                    dst[j] = ch;
                }
                dst += newlen;
                o = o2;
            }
        }
    }
    __syncthreads();
    if (tid < 12 && s_counters[tid]) atomicAdd(O.counters + tid, s_counters[tid]);
}

} // namespace

uint32_t vtn_normalize_tile_records() { return VTN_TILE; }

int vtn_launch_normalize(vtn_ctx* ctx, const VtnDevBatchView& in, const VtnDevNormOut& out, unsigned long long* tile_state,
                         uint32_t* tile_counter, uint32_t n_tiles)
{
    VTN_CUDA(ctx, cudaMemsetAsync(out.totals, 0, 2 * sizeof(unsigned long long), ctx->stream));
    VTN_CUDA(ctx, cudaMemsetAsync(out.counters, 0, 12 * sizeof(uint32_t), ctx->stream));
    VTN_CUDA(ctx, cudaMemsetAsync(out.chg_off, 0, sizeof(uint64_t), ctx->stream));
    if (n_tiles == 0) return VTN_OK;
    VTN_CUDA(ctx, cudaMemsetAsync(tile_state, 0, 2ull * n_tiles * sizeof(unsigned long long), ctx->stream));
    VTN_CUDA(ctx, cudaMemsetAsync(tile_counter, 0, sizeof(uint32_t), ctx->stream));
    int per_sm = 0;
    VTN_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, vtn_normalize_kernel, VTN_TILE, 0));
    if (per_sm < 1) per_sm = 1;
    uint32_t grid = (uint32_t)(ctx->sm_count * per_sm);
    if (grid > n_tiles) grid = n_tiles;
    vtn_normalize_kernel<<<grid, VTN_TILE, 0, ctx->stream>>>(ctx->dref, in, out, tile_state, tile_counter, n_tiles);
    VTN_CUDA(ctx, cudaGetLastError());
    ctx->launches++;
    return VTN_OK;
}
