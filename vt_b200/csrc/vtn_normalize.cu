// vtn_normalize.cu -- the normalize kernel (sm_100a) and its launcher.
//
// One warp owns a tile of consecutive records, fetched from a global tile counter (persistent warps); see the
// kernel comment below.  Per record, in closed form instead of the reference's per-base loop:
//   classify_variant          variant_manip.cpp:227-508
//   is_not_ref_consistent     variant_manip.cpp:61-103      (skipped iff type == VT_SNP, normalize.cpp:205)
//   is_normalized             variant_manip.cpp:109-197
//   right_trim_or_left_extend variant_manip.cpp:513-583  ->  T = longest common suffix of E_i = prefix+allele_i
//   left_trim                 variant_manip.cpp:588-615
//   counters                  normalize.cpp:254-322
// Changed records are compacted in input order: per-tile scratch segments (K1) + scan/gather (K2).
#include "vtn_internal.h"
#include "vtn_ref_dev.cuh"
#include "vtn_scan.cuh"
#include "vtn_fast.cuh"
#include "vtn_classify.cuh"


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

// Phase C of the closed form for a biallelic record: every base of both alleles has been consumed
// (T >= max len), so both streams read the reference, at a lag of |len0-len1| bases.  Walks down in 32-base
// chunks (two packed windows, XOR, clz); per base only at contig edges or over N/IUPAC bases.
// Returns the final T (<= cap).
__device__ __noinline__ uint32_t lag_scan(const VtnDevRef& R, uint64_t gbase, int64_t clen, int32_t pos1, uint32_t len0, uint32_t len1,
                             uint32_t T, uint32_t cap)
{
    VtnRefCursor cur;
    while (T < cap)
    {
        // window of stream s ends (inclusive) at hi_s = pos1-2-T+len_s
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
            if (nm >= room) return cap;
            T += nm;
            if (nm < 32) return T;
        }
        else
        {
            // per base for up to 32 steps (contig edge, or N / IUPAC bases in the window)
            uint32_t steps = 0;
            while (T < cap && steps < 32)
            {
                uint8_t c0 = ext_byte(R, cur, gbase, clen, pos1, nullptr, len0, T);
                uint8_t c1 = ext_byte(R, cur, gbase, clen, pos1, nullptr, len1, T);
                if (c0 != c1) return T;
                ++T;
                ++steps;
            }
        }
    }
    return T;
}

// ---- fast path (see vtn_fast.cuh).  Returns false when the record is not eligible. ----
__device__ __forceinline__ bool fast_biallelic(const VtnDevRef& R, const uint8_t* bytes, uint32_t off0, uint32_t len0, uint32_t len1,
                                               int32_t pos1, int64_t clen, uint64_t gbase, VtnFastRes& o)
{
    if (len0 > 32 || len1 > 32 || pos1 < 34) return false;
    const int64_t p0 = (int64_t)pos1 - 1;
    if (p0 + 32 > clen) return false; // keeps every window inside the contig (records at contig ends -> generic)
    // ---- allele bytes -> 2-bit code stream (REF then ALT are contiguous) ----
    const uint32_t L = len0 + len1;
    uint64_t s0 = 0, s1 = 0, badacc = 0;
    for (uint32_t c = 0; 8 * c < L; ++c)
    {
        uint64_t bad;
        const uint32_t rem = L - 8 * c;
        const uint64_t code = vtn_codes8(vtn_fetch8(bytes, off0 + 8 * c), rem, bad);
        badacc |= bad;
        if (c < 4) s0 |= code << (16 * c);
        else s1 |= code << (16 * (c - 4));
    }
    if (badacc) return false;
    const uint64_t Rc = s0 & vtn_mask2(len0);
    const uint64_t Ac = (len0 == 32 ? s1 : ((s0 >> (2 * len0)) | (s1 << (64 - 2 * len0)))) & vtn_mask2(len1);
    const uint32_t m = len0 < len1 ? len0 : len1, M = len0 < len1 ? len1 : len0;

    // ---- classify_variant (variant_manip.cpp:337-505), all bases upper-case ACGT so compares are code compares ----
    const bool same = (len0 == len1) && (Rc == Ac);
    const uint64_t y = Rc ^ Ac;
    const uint64_t mmy = (y | (y >> 1)) & 0x5555555555555555ull;
    uint32_t cp_raw = mmy ? (uint32_t)(__ffsll((long long)mmy) - 1) >> 1 : 32u; // common prefix (may exceed m: capped below)
    if (cp_raw > m) cp_raw = m;
    uint32_t cs_raw = vtn_top_match(Rc << (64 - 2 * len0), Ac << (64 - 2 * len1)); // common suffix
    if (cs_raw > m) cs_raw = m;
    uint32_t vtype = VTN_VT_REF;
    if (!same)
    {
        const uint32_t cs = cs_raw < m - 1 ? cs_raw : m - 1;
        uint32_t rl = len0 - cs, al = len1 - cs;
        const uint32_t mla = rl < al ? rl : al;
        const uint32_t cp = cp_raw < mla - 1 ? cp_raw : mla - 1;
        rl -= cp;
        al -= cp;
        const uint32_t mlen = rl < al ? rl : al;
        uint32_t diff;
        if (mlen == 1 && rl != al)
        {
            const uint64_t lg = rl > al ? Rc : Ac, sh = rl > al ? Ac : Rc;
            const uint32_t ll = rl > al ? rl : al;
            const uint32_t ls = (uint32_t)(lg >> (2 * cp)) & 3u, le_ = (uint32_t)(lg >> (2 * (cp + ll - 1))) & 3u;
            const uint32_t ss = (uint32_t)(sh >> (2 * cp)) & 3u;
            diff = (ls != ss && le_ != ss) ? 1u : 0u;
        }
        else
            diff = (uint32_t)__popcll((mmy >> (2 * cp)) & vtn_mask2(mlen));
        if (mlen == diff) vtype |= (mlen == 1) ? VTN_VT_SNP : VTN_VT_MNP;
        if (rl != al) vtype |= VTN_VT_INDEL;
        if (diff && diff < mlen) vtype |= VTN_VT_CLUMPED;
        if (len0 == len1 && len0 > 1) vtype |= VTN_VT_MNP;
    }
    o.vtype = vtype;
    o.T = o.le = o.lt = 0;
    o.status = 0;
    o.nlr = o.nla = 0;
    o.newR = o.newA = 0;
    if (vtype == VTN_VT_SNP) return true; // no REF check (normalize.cpp:205); a pure SNP is always normalized

    // ---- reference windows and REF check (variant_manip.cpp:61-103) ----
    uint64_t V, W;
    uint32_t ev, nw;
    vtn_ref_window64(R, gbase + (uint64_t)(p0 - 32), V, W, ev, nw);
    if (nw & (uint32_t)(len0 >= 32 ? 0xffffffffu : ((1u << len0) - 1u))) return false; // N / lower-case under REF
    if ((W ^ Rc) & vtn_mask2(len0))
    {
        o.status = 2; // both sides are plain ACGT, so any mismatch is code 2
        return true;
    }
    // ---- is_normalized (variant_manip.cpp:109-197) ----
    if (same || (len0 == 1 && len1 == 1)) return true;
    const bool exists_one = (m == 1);
    const bool first_same = ((Rc ^ Ac) & 3ull) == 0;
    const bool last_same = cs_raw >= 1;
    if (!(last_same || (!exists_one && first_same))) return true;
    o.status = VTN_ST_ENTERED;

    // ---- right_trim_or_left_extend in closed form ----
    uint32_t T;
    if (cs_raw < m)
        T = cs_raw;
    else
    {
        // the shorter allele is a suffix of the longer one: the longer one's first D bases now face the
        // reference bases left of POS ([p0-D, p0), the top D bases of V)
        const uint32_t D = M - m;
        const uint64_t X = len0 > len1 ? Rc : Ac;
        const uint32_t dm = (D >= 32 ? 0xffffffffu : ((1u << D) - 1u));
        if (ev & (dm << (32 - D))) return false; // N / IUPAC bases there
        uint32_t e = vtn_top_match((X & vtn_mask2(D)) << (64 - 2 * D), V & ~vtn_mask2(32 - D));
        if (e > D) e = D;
        T = m + e;
        if (e == D)
        {
            // every base of both alleles is consumed: from here both streams read the reference, D bases apart
            // (lag_scan).  Its first 32-D comparisons are V against itself shifted by D bases: base index j of V
            // (= p0-32+j) faces index j-D, from j = 31 downwards -- no load.  pos1 >= 34 keeps the cap
            // (pos1-2+m >= M+32-D) out of reach of this step.
            const int64_t cap64 = (int64_t)pos1 - 2 + (int64_t)m;
            const uint32_t cap = cap64 > 0xffffffffll ? 0xffffffffu : (uint32_t)cap64;
            bool more = true;
            if (ev == 0)
            {
                const uint64_t xs = V ^ (V << (2 * D));
                const uint64_t ms = (xs | (xs >> 1)) & 0x5555555555555555ull & ~vtn_mask2(D);
                const uint32_t nm = ms ? (uint32_t)(__clzll((long long)ms) >> 1) : 32u - D;
                T += nm;
                more = (nm == 32u - D);
            }
            if (more) T = lag_scan(R, gbase, clen, pos1, len0, len1, T, cap);
        }
    }
    const uint32_t le = T >= m ? T - m + 1 : 0;
    uint32_t lt = 0;
    if (le == 0)
    {
        const uint32_t room = m - T - 1;
        lt = cp_raw < room ? cp_raw : room;
    }
    o.T = T;
    o.le = le;
    o.lt = lt;
    if (!(T | le | lt)) return true;
    uint32_t cat; // normalize.cpp:254-311
    if (le) cat = (T > le) ? 4 : 5;
    else if (lt && T > le) cat = 3;
    else if (lt) cat = 1;
    else cat = 2;
    o.status |= VTN_ST_CHANGED | (cat << VTN_ST_CAT_SHIFT);
    // ---- new alleles as codes ----
    if (le == 0)
    {
        o.nlr = len0 - T - lt;
        o.nla = len1 - T - lt;
        o.newR = (Rc >> (2 * lt)) & vtn_mask2(o.nlr);
        o.newA = (Ac >> (2 * lt)) & vtn_mask2(o.nla);
    }
    else
    {
        // new allele i = (le reference bases [p0-le, p0) ++ allele_i)[0 : le + len_i - T]
        const uint32_t D = M - m;
        uint64_t Vl; // codes of [p0-le, p0-le+32)
        uint32_t e2;
        if (le <= 32)
        {
            Vl = le == 32 ? V : (V >> (2 * (32 - le))); // only the low le bases are reference; the rest is overlaid below
            e2 = le == 32 ? ev : (ev >> (32 - le));
        }
        else
            vtn_ref_window32(R, gbase + (uint64_t)(p0 - (int64_t)le), Vl, e2);
        // the shift may have run through N / IUPAC bases (they compare like any base): those cannot be codes
        if (e2 & (D + 1 >= 32 ? 0xffffffffu : ((1u << (D + 1)) - 1u))) return false;
        const uint64_t X = len0 > len1 ? Rc : Ac;
        uint64_t lg; // the longer allele's new form, D+1 bases
        if (T < M && le < 32)
            lg = (Vl & vtn_mask2(le)) | ((X & vtn_mask2(M - T)) << (2 * le));
        else
            lg = Vl;
        lg &= vtn_mask2(D + 1);
        const uint64_t shn = Vl & 3ull; // the shorter allele's new form: the single anchor base
        if (D + 1 > 32) return false;   // cannot happen (D <= 31), keeps the compiler honest
        if (len0 > len1) { o.newR = lg; o.nlr = D + 1; o.newA = shn; o.nla = 1; }
        else { o.newA = lg; o.nla = D + 1; o.newR = shn; o.nlr = 1; }
    }
    return true;
}

// ---- the same fast path in 32-bit words, for records whose alleles are both <= 16 bases (nearly all of a real call set):
// every 2-bit operation is one instruction instead of a 64-bit pair, the window is 32 bases [p0-16, p0+16) = two groups,
// and the two loads are issued before the allele conversion so that their latency overlaps it.  Line references as in
// fast_biallelic; the two must return identical results wherever both accept a record.
__device__ __forceinline__ uint32_t vtn_m32(uint32_t nbases) { return __funnelshift_lc(0xffffffffu, 0u, 2u * nbases); } // low 2n bits, n in [0,16]

__device__ __forceinline__ uint32_t vtn_top_match32(uint32_t a, uint32_t b)
{
    const uint32_t x = a ^ b;
    const uint32_t mm = (x | (x >> 1)) & 0x55555555u;
    return mm ? (uint32_t)(__clz((int)mm) >> 1) : 16u;
}

__device__ __forceinline__ bool fast_biallelic16(const VtnDevRef& R, const uint8_t* bytes, uint32_t off0, uint32_t len0, uint32_t len1,
                                                 int32_t pos1, int64_t clen, uint64_t gbase, VtnFastRes& o)
{
    if (len0 - 1u > 15u || len1 - 1u > 15u || pos1 < 18) return false;
    const int64_t p0 = (int64_t)pos1 - 1;
    if (p0 + 16 > clen) return false; // keeps the window inside the contig
    const uint64_t gp = gbase + (uint64_t)(p0 - 16);
    const uint4 ga = vtn_ldg_group(R, gp >> 5), gb = vtn_ldg_group(R, (gp >> 5) + 1); // contigs carry two pad groups
    const uint32_t gs = (uint32_t)gp & 31u;
    // ---- allele bytes -> 2-bit code stream (REF then ALT are contiguous), <= 32 bases ----
    const uint32_t L = len0 + len1;
    uint64_t s0 = 0, badacc = 0;
    for (uint32_t c = 0; 8 * c < L; ++c)
    {
        uint64_t bad;
        const uint64_t code = vtn_codes8(vtn_fetch8(bytes, off0 + 8 * c), L - 8 * c, bad);
        badacc |= bad;
        s0 |= code << (16 * c);
    }
    if (badacc) return false;
    const uint32_t Rc = (uint32_t)s0 & vtn_m32(len0);
    const uint32_t Ac = (uint32_t)(s0 >> (2 * len0)) & vtn_m32(len1);
    const uint32_t m = len0 < len1 ? len0 : len1, M = len0 < len1 ? len1 : len0;

    // ---- classify_variant (variant_manip.cpp:337-505) ----
    const bool same = (len0 == len1) && (Rc == Ac);
    const uint32_t y = Rc ^ Ac;
    const uint32_t mmy = (y | (y >> 1)) & 0x55555555u;
    uint32_t cp_raw = mmy ? (uint32_t)(__ffs((int)mmy) - 1) >> 1 : 16u;
    if (cp_raw > m) cp_raw = m;
    uint32_t cs_raw = vtn_top_match32(Rc << (32 - 2 * len0), Ac << (32 - 2 * len1));
    if (cs_raw > m) cs_raw = m;
    uint32_t vtype = VTN_VT_REF;
    if (!same)
    {
        const uint32_t cs = cs_raw < m - 1 ? cs_raw : m - 1;
        uint32_t rl = len0 - cs, al = len1 - cs;
        const uint32_t mla = rl < al ? rl : al;
        const uint32_t cp = cp_raw < mla - 1 ? cp_raw : mla - 1;
        rl -= cp;
        al -= cp;
        const uint32_t mlen = rl < al ? rl : al;
        uint32_t diff;
        if (mlen == 1 && rl != al)
        {
            const uint32_t lg = rl > al ? Rc : Ac, sh = rl > al ? Ac : Rc;
            const uint32_t ll = rl > al ? rl : al;
            const uint32_t ls = (lg >> (2 * cp)) & 3u, le_ = (lg >> (2 * (cp + ll - 1))) & 3u;
            const uint32_t ss = (sh >> (2 * cp)) & 3u;
            diff = (ls != ss && le_ != ss) ? 1u : 0u;
        }
        else
            diff = (uint32_t)__popc((mmy >> (2 * cp)) & vtn_m32(mlen));
        if (mlen == diff) vtype |= (mlen == 1) ? VTN_VT_SNP : VTN_VT_MNP;
        if (rl != al) vtype |= VTN_VT_INDEL;
        if (diff && diff < mlen) vtype |= VTN_VT_CLUMPED;
        if (len0 == len1 && len0 > 1) vtype |= VTN_VT_MNP;
    }
    o.vtype = vtype;
    o.T = o.le = o.lt = 0;
    o.status = 0;
    o.nlr = o.nla = 0;
    o.newR = o.newA = 0;
    if (vtype == VTN_VT_SNP) return true; // no REF check (normalize.cpp:205)

    // ---- the window: V = codes of [p0-16, p0), W = codes of [p0, p0+16); ev / nw as in vtn_ref_window64 ----
    uint32_t V, W, ev, nw;
    {
        const uint32_t w0 = gs < 16 ? ga.x : ga.y, w1 = gs < 16 ? ga.y : gb.x, w2 = gs < 16 ? gb.x : gb.y;
        V = __funnelshift_r(w0, w1, 2 * gs); // the shift wraps at 32, which is what the word selection above expects
        W = __funnelshift_r(w1, w2, 2 * gs);
        ev = __funnelshift_r(ga.z & ~ga.w, gb.z & ~gb.w, gs) & 0xffffu;
        nw = __funnelshift_r(ga.z, gb.z, gs) >> 16;
    }
    if (nw & ((1u << len0) - 1u)) return false; // N / lower-case under REF
    if ((W ^ Rc) & vtn_m32(len0))
    {
        o.status = 2;
        return true;
    }
    // ---- is_normalized (variant_manip.cpp:109-197) ----
    if (same || (len0 == 1 && len1 == 1)) return true;
    const bool exists_one = (m == 1);
    const bool first_same = (y & 3u) == 0;
    const bool last_same = cs_raw >= 1;
    if (!(last_same || (!exists_one && first_same))) return true;
    o.status = VTN_ST_ENTERED;

    // ---- right_trim_or_left_extend in closed form ----
    const uint32_t D = M - m;
    const uint32_t X = len0 > len1 ? Rc : Ac;
    uint32_t T;
    if (cs_raw < m)
        T = cs_raw;
    else
    {
        if (ev & (((1u << D) - 1u) << (16 - D))) return false; // N / IUPAC bases among the D bases left of POS
        uint32_t e = vtn_top_match32((X & vtn_m32(D)) << (32 - 2 * D), V & ~vtn_m32(16 - D));
        if (e > D) e = D;
        T = m + e;
        if (e == D)
        {
            const int64_t cap64 = (int64_t)pos1 - 2 + (int64_t)m;
            const uint32_t cap = cap64 > 0xffffffffll ? 0xffffffffu : (uint32_t)cap64;
            bool more = true;
            if (ev == 0)
            { // first 16-D lag comparisons from V itself; pos1 >= 18 keeps the cap (>= m+16) out of reach
                const uint32_t xs = V ^ (V << (2 * D));
                const uint32_t ms = (xs | (xs >> 1)) & 0x55555555u & ~vtn_m32(D);
                const uint32_t nm = ms ? (uint32_t)(__clz((int)ms) >> 1) : 16u - D;
                T += nm;
                more = (nm == 16u - D);
            }
            if (more) T = lag_scan(R, gbase, clen, pos1, len0, len1, T, cap);
        }
    }
    const uint32_t le = T >= m ? T - m + 1 : 0;
    uint32_t lt = 0;
    if (le == 0)
    {
        const uint32_t room = m - T - 1;
        lt = cp_raw < room ? cp_raw : room;
    }
    o.T = T;
    o.le = le;
    o.lt = lt;
    if (!(T | le | lt)) return true;
    uint32_t cat; // normalize.cpp:254-311
    if (le) cat = (T > le) ? 4 : 5;
    else if (lt && T > le) cat = 3;
    else if (lt) cat = 1;
    else cat = 2;
    o.status |= VTN_ST_CHANGED | (cat << VTN_ST_CAT_SHIFT);
    // ---- new alleles as codes ----
    if (le == 0)
    {
        o.nlr = len0 - T - lt;
        o.nla = len1 - T - lt;
        o.newR = (Rc >> (2 * lt)) & vtn_m32(o.nlr);
        o.newA = (Ac >> (2 * lt)) & vtn_m32(o.nla);
    }
    else
    {
        // new allele i = (le reference bases [p0-le, p0) ++ allele_i)[0 : le + len_i - T]; the longer one has D+1 bases
        uint32_t Vl, e2; // codes / exception bits of [p0-le, p0-le+16)
        if (le <= 16)
        {
            Vl = le == 16 ? V : (V >> (2 * (16 - le)));
            e2 = le == 16 ? ev : (ev >> (16 - le));
        }
        else
        {
            uint64_t Vl64;
            vtn_ref_window32(R, gbase + (uint64_t)(p0 - (int64_t)le), Vl64, e2);
            Vl = (uint32_t)Vl64;
        }
        if (e2 & ((1u << (D + 1)) - 1u)) return false; // the shift ran through N / IUPAC bases: those cannot be codes
        uint32_t lg;
        if (T < M) // then le <= D <= 15
            lg = (Vl & vtn_m32(le)) | ((X & vtn_m32(M - T)) << (2 * le));
        else
            lg = Vl;
        lg &= vtn_m32(D + 1);
        const uint32_t shn = Vl & 3u;
        if (len0 > len1) { o.newR = lg; o.nlr = D + 1; o.newA = shn; o.nla = 1; }
        else { o.newA = lg; o.nla = D + 1; o.newR = shn; o.nlr = 1; }
    }
    return true;
}

// The whole per-record computation except writing the new alleles.
__device__ __noinline__ void normalize_record(const VtnDevRef& R, const VtnDevBatchView& B, uint32_t r, RecResult& out)
{
    const uint32_t ab = B.allele_beg[r];
    const uint32_t n = B.allele_beg[r + 1] - ab;
    const int32_t pos1 = B.pos1[r];
    const int32_t rid = B.rid[r];
    const uint32_t off0 = B.allele_off[ab];
    const uint32_t off1 = B.allele_off[ab + 1];
    const uint8_t* a0 = B.bytes + off0;
    const uint32_t len0 = off1 - off0;
    out.T = out.le = out.lt = out.nbytes = 0;

    // ---- SNV fast path: biallelic, both alleles one byte ----
    if (n == 2)
    {
        const uint32_t off2 = B.allele_off[ab + 2];
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
            uint32_t o2 = B.allele_off[ab + i + 1];
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
            uint32_t o2 = B.allele_off[ab + i + 1];
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
            uint32_t o2 = B.allele_off[ab + i + 1];
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
            const uint32_t len1 = B.allele_off[ab + 2] - off1;
            T = lag_scan(R, gbase, clen, pos1, len0, len1, T, cap);
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
                    uint32_t o2 = B.allele_off[ab + i + 1];
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
                o = B.allele_off[ab + i + 1];
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
        const uint32_t total = B.allele_off[ab + n] - off0;
        out.nbytes = total + n * le - n * T - n * lt;
    }
    out.status = (uint8_t)st;
}

} // namespace

namespace
{

// ---------------------------------------------------------------------------------------------------------
// K1  vtn_normalize_kernel: every WARP owns a tile of VTN_WTILE = 32 * VTN_RPL consecutive records and never
//     waits for another warp or tile (no __syncthreads, no inter-tile dependency):
//   phase 1  lane l owns records l*RPL .. l*RPL+RPL-1 of the tile: 128-bit loads of allele_beg; the tile's
//            allele_off window and allele byte window are staged in shared memory with coalesced 128-bit
//            loads; SNVs (80% of a real call set) are settled here; the rest is queued in record order.
//   phase 2  queued records are processed densely (lane i <- queue entry 32p+i): 2-bit register fast path,
//            generic byte-wise path for what it refuses.  Changed records are ranked with a warp scan (queue
//            order == record order) and written straight from registers to the tile's own scratch segment
//            (scratch index tile*WTILE + rank; new allele bytes inside the tile's old byte window, which they
//            can never outgrow).  The tile's (count, bytes) go to the tile directory.
//   phase 3  status / vtype of the tile as one 4-byte store per lane; counters stay in registers.
// K2  vtn_compact_kernel: single-pass scan (decoupled look-back over blocks of 256 tiles; all aggregates are
//     known up front so nobody waits on compute) of the tile directory + gather of the scratch segments
//     into the final chg_* arrays, in input order.
// ---------------------------------------------------------------------------------------------------------
#define VTN_RPL 4
#define VTN_WTILE (32 * VTN_RPL)
#define VTN_WARPS 4
#ifndef VTN_MIN_BLOCKS
#define VTN_MIN_BLOCKS 5
#endif
#ifndef VTN_MIN_BLOCKS_FAST
#define VTN_MIN_BLOCKS_FAST 7 // measured: 6 -> 0.297 ms, 7 -> 0.296 ms, 8 -> 0.304 ms per 10 M records (72 registers, 28 warps/SM)
#endif
#define VTN_AOFF_CAP (VTN_WTILE * 3 + 16)
#define VTN_BYTES_CAP (VTN_WTILE * 8 + 64)

struct __align__(16) WarpSmem
{
    uint32_t aoff[VTN_AOFF_CAP];
    uint8_t bytes[VTN_BYTES_CAP];
    int32_t pos1[VTN_WTILE], rid[VTN_WTILE];
    uint32_t queue[VTN_WTILE]; // record index in the tile | biallelic flag << 7 | (first allele slot - tile's first slot) << 8
    uint8_t st[VTN_WTILE], vt[VTN_WTILE];
};
#ifndef VTN_CONTIG_CAP
#define VTN_CONTIG_CAP 256
#endif
// K2's control block, in 64-bit words (see vtn_launch_normalize); vtn_normalize_ctl_words() sizes the allocation
#define VTC_CTL_BASE 2
#define VTC_CTL_AGG 4
#define VTC_AGG_SHIFT 44

__device__ __forceinline__ uint32_t warp_excl_scan(uint32_t v, int lane, uint32_t& total)
{
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1)
    {
        const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    total = __shfl_sync(0xffffffffu, inc, 31);
    return inc - v;
}

// new alleles of a record handled by the generic path: (le reference bases + upper(allele))[0 : le + len - T],
// or upper(allele)[lt : len - T]  (normalize.cpp:246-252 after variant_manip.cpp:513-615)
__device__ __noinline__ void write_generic_alleles(const VtnDevRef& R, const VtnDevBatchView& B, uint32_t rr, uint32_t T, uint32_t le,
                                      uint32_t lt, uint8_t* dst)
{
    const uint32_t ab2 = B.allele_beg[rr];
    const uint32_t n2 = B.allele_beg[rr + 1] - ab2;
    const int32_t pos1 = B.pos1[rr];
    const int32_t rid = B.rid[rr];
    const uint64_t gbase = __ldg(R.contig_goff + rid) << 5;
    const int64_t clen = __ldg(R.contig_len + rid);
    VtnRefCursor cur;
    uint32_t o = B.allele_off[ab2];
    for (uint32_t ai = 0; ai < n2; ++ai)
    {
        const uint32_t o2 = B.allele_off[ab2 + ai + 1];
        const uint32_t len = o2 - o;
        const uint32_t newlen = len + le - T - lt;
        const uint8_t* a = B.bytes + o;
        for (uint32_t j = 0; j < newlen; ++j)
        {
            uint8_t ch;
            if (j < le)
            {
                int64_t p0 = (int64_t)pos1 - 1 - (int64_t)le + (int64_t)j;
                p0 = p0 < 0 ? 0 : (p0 >= clen ? clen - 1 : p0);
                ch = cur.uc(R, gbase + (uint64_t)p0);
            }
            else
                ch = vtn_toupper(a[j - le + lt]);
            dst[j] = ch;
        }
        dst += newlen;
        o = o2;
    }
}

// FAST_ONLY = true : the common kernel.  Only the 2-bit register fast path is compiled in, which roughly halves the
//                     register footprint (more resident warps); a tile in which ANY queued record is refused by the
//                     fast path is abandoned and appended to defer_list.
// FAST_ONLY = false: the same code with the generic byte-wise path, run over the deferred tiles only.
// A tile is produced by exactly one of the two launches, so the tile directory / scratch contract of K2 is unchanged.
template <bool FAST_ONLY>
__global__ void __launch_bounds__(32 * VTN_WARPS, FAST_ONLY ? VTN_MIN_BLOCKS_FAST : VTN_MIN_BLOCKS) vtn_normalize_kernel(
    VtnDevRef R, VtnDevBatchView B, VtnDevNormOut O, VtnDevScratch X, uint32_t* defer_list, uint32_t* defer_count, uint32_t n_tiles_all,
    unsigned long long* k2_ctl, uint32_t tpb, int fresh)
{
    __shared__ WarpSmem SM[VTN_WARPS];
    __shared__ uint32_t s_lut[256]; // 4 packed 2-bit codes -> their 4 ASCII bytes (A=0 C=1 T=2 G=3)
    __shared__ uint64_t s_gbase[VTN_CONTIG_CAP];
    __shared__ int64_t s_clen[VTN_CONTIG_CAP];
    const int lane = threadIdx.x & 31;
    if (!FAST_ONLY && *defer_count == 0) return; // nothing was deferred (the usual case): leave before any staging
    WarpSmem& S = SM[threadIdx.x >> 5];
    uint32_t my_counter = 0; // lane k < 12 accumulates counter k of this warp
    const uint32_t cat_shift = 6u * (uint32_t)(lane < 5 ? lane : (lane < 10 ? lane - 5 : 0));
    // contig table (gbase, length) into shared memory: two dependent global loads fewer per queued record
    if (FAST_ONLY && blockIdx.x == 0)
    { // K2's base (it starts after this kernel): the totals of the previous chunk of a chunked host call, or 0
        if (threadIdx.x < 2) k2_ctl[VTC_CTL_BASE + threadIdx.x] = fresh ? 0ull : O.totals[threadIdx.x];
    }
    for (int v = threadIdx.x; v < 256; v += blockDim.x)
    {
        uint32_t w = 0;
        for (int q = 0; q < 4; ++q) w |= ((0x47544341u >> (8 * ((v >> (2 * q)) & 3))) & 0xffu) << (8 * q);
        s_lut[v] = w;
    }
    const bool contigs_staged = R.n_contig <= VTN_CONTIG_CAP;
    if (contigs_staged)
        for (int c = threadIdx.x; c < R.n_contig; c += blockDim.x)
        {
            s_gbase[c] = __ldg(R.contig_goff + c) << 5;
            s_clen[c] = __ldg(R.contig_len + c);
        }
    __syncthreads();

    // static round-robin tiles: tiles are independent, so no scheduling atomic is needed.  The deferred launch walks
    // defer_list instead of 0..n_tiles-1 (ti = position, tile = tile number).
    const uint32_t n_tiles = FAST_ONLY ? n_tiles_all : min(*defer_count, n_tiles_all);
    const uint32_t tile_stride = gridDim.x * VTN_WARPS;
    uint32_t ti = blockIdx.x * VTN_WARPS + (threadIdx.x >> 5);
#define TILE_OF(i) (FAST_ONLY ? (i) : __ldg(defer_list + (i)))
    uint32_t tile = ti < n_tiles ? TILE_OF(ti) : 0;
    // header vectors of the first tile; inside the loop the NEXT tile's are prefetched into registers
    uint4 h_ab = make_uint4(0, 0, 0, 0);
    int4 h_pos = make_int4(0, 0, 0, 0), h_rid = make_int4(0, 0, 0, 0);
    uint32_t h_abx = 0;
    if (ti < n_tiles)
    {
        const uint32_t r0 = tile * VTN_WTILE;
        h_ab = __ldg(reinterpret_cast<const uint4*>(B.allele_beg + r0) + lane); // arrays are padded
        h_pos = __ldg(reinterpret_cast<const int4*>(B.pos1 + r0) + lane);
        h_rid = __ldg(reinterpret_cast<const int4*>(B.rid + r0) + lane);
        if (lane == 31) h_abx = __ldg(B.allele_beg + r0 + VTN_WTILE);
    }
    while (ti < n_tiles)
    {
        const uint32_t next_ti = ti + tile_stride;
        const uint32_t next_tile = next_ti < n_tiles ? TILE_OF(next_ti) : 0;
        const uint32_t r0 = tile * VTN_WTILE;
        const uint32_t n_live = min((uint32_t)VTN_WTILE, B.n_rec - r0);
        // ---------------- phase 1 ----------------
        uint32_t ab[VTN_RPL + 1];
        {
            ab[0] = h_ab.x; ab[1] = h_ab.y; ab[2] = h_ab.z; ab[3] = h_ab.w;
            uint32_t nxt = __shfl_down_sync(0xffffffffu, h_ab.x, 1);
            if (lane == 31) nxt = h_abx;
            ab[4] = nxt;
            reinterpret_cast<int4*>(S.pos1)[lane] = h_pos;
            reinterpret_cast<int4*>(S.rid)[lane] = h_rid;
        }
        if (next_ti < n_tiles)
        {
            const uint32_t r1 = next_tile * VTN_WTILE;
            h_ab = __ldg(reinterpret_cast<const uint4*>(B.allele_beg + r1) + lane);
            h_pos = __ldg(reinterpret_cast<const int4*>(B.pos1 + r1) + lane);
            h_rid = __ldg(reinterpret_cast<const int4*>(B.rid + r1) + lane);
            if (lane == 31) h_abx = __ldg(B.allele_beg + r1 + VTN_WTILE);
        }
        const uint32_t s_lo = __shfl_sync(0xffffffffu, ab[0], 0);
        // one past the last live record's slots: lane 31 holds allele_beg[r0 + WTILE] for a full tile
        const uint32_t s_hi31 = __shfl_sync(0xffffffffu, ab[VTN_RPL], 31);
        const uint32_t s_hi = n_live == VTN_WTILE ? s_hi31 : __ldg(B.allele_beg + r0 + n_live);
        // stage allele_off[s_lo .. s_hi]
        const uint32_t a0 = s_lo & ~3u;
        const bool aoff_staged = (s_hi + 1 - a0) <= VTN_AOFF_CAP;
        if (aoff_staged)
        {
            for (uint32_t i = lane * 4; i < s_hi + 1 - a0; i += 128)
                *reinterpret_cast<uint4*>(&S.aoff[i]) = __ldg(reinterpret_cast<const uint4*>(B.allele_off + a0 + i));
        }
        __syncwarp();
#define AO(slot) (aoff_staged ? S.aoff[(slot) - a0] : __ldg(B.allele_off + (slot)))
        const uint32_t b_lo = AO(s_lo), b_hi = AO(s_hi);
        const uint32_t b0 = b_lo & ~15u;
        const bool bytes_staged = (b_hi + 16 - b0) <= VTN_BYTES_CAP;
        if (bytes_staged)
        {
            for (uint32_t i = lane * 16; i < b_hi + 16 - b0; i += 512)
                *reinterpret_cast<uint4*>(&S.bytes[i]) = __ldg(reinterpret_cast<const uint4*>(B.bytes + b0 + i));
        }
        __syncwarp();
        const uint8_t* BY = bytes_staged ? (S.bytes - b0) : B.bytes; // absolute offsets work on both

        uint32_t qmask = 0;
#pragma unroll
        for (int j = 0; j < VTN_RPL; ++j)
        {
            const uint32_t idx = lane * VTN_RPL + j;
            if (idx < n_live)
            {
                bool q = true;
                if (ab[j + 1] - ab[j] == 2)
                {
                    const uint32_t o0 = AO(ab[j]), o2 = AO(ab[j] + 2);
                    if (o2 - o0 == 2) // both alleles one byte
                    {
                        const uint8_t r0b = BY[o0], c = BY[o0 + 1];
                        q = (c == '<' || c == '[' || c == ']' || c == '.' || c == r0b);
                    }
                }
                if (q) qmask |= 1u << j;
            }
        }
        uint32_t nq;
        {
            uint32_t slot = warp_excl_scan(__popc(qmask), lane, nq);
#pragma unroll
            for (int j = 0; j < VTN_RPL; ++j)
                if (qmask & (1u << j)) S.queue[slot++] = (uint32_t)(lane * VTN_RPL + j) | (ab[j + 1] - ab[j] == 2 ? 0x80u : 0u) | ((ab[j] - s_lo) << 8);
        }
        __syncwarp();

        // ---------------- phase 2: dense passes over the queue ----------------
        uint32_t tot_c = 0, tot_b = 0;
        uint32_t cnt_vars = n_live - nq; // the SNVs settled in phase 1 (normalize.cpp:315-322)
        uint32_t cnt_refs = 0;
        uint32_t tile_cat = 0; // lane k < 10: category counter k of THIS tile (committed only if the tile is kept)
        bool deferred = false;
        for (uint32_t p = 0; p < nq; p += 32)
        {
            const uint32_t e = p + lane;
            const bool act = e < nq;
            uint32_t st = 0, vt = VTN_VT_SNP, T = 0, le = 0, lt = 0, nb = 0, n2 = 2, rr = 0;
            int32_t pos1 = 0;
            bool fast = false, elig = false;
            VtnFastRes f;
            uint32_t o0 = 0, l0 = 0, l1 = 0, idx = 0;
            int64_t clen = 0;
            uint64_t gbase = 0;
            if (act)
            {
                const uint32_t qe = S.queue[e];
                idx = qe & 0x7fu;
                rr = r0 + idx;
                const uint32_t ab2 = s_lo + (qe >> 8);
                n2 = (qe & 0x80u) ? 2u : 3u; // only "biallelic or not" matters below (eligibility, counter arity)
                pos1 = S.pos1[idx];
                const int32_t rid = S.rid[idx];
                if (n2 == 2 && rid >= 0 && rid < R.n_contig)
                {
                    elig = true;
                    o0 = AO(ab2);
                    const uint32_t o1 = AO(ab2 + 1), o2 = AO(ab2 + 2);
                    l0 = o1 - o0;
                    l1 = o2 - o1;
                    clen = contigs_staged ? s_clen[rid] : __ldg(R.contig_len + rid);
                    gbase = contigs_staged ? s_gbase[rid] : (__ldg(R.contig_goff + rid) << 5);
                }
            }
            // 32-bit words when every eligible record of this pass has alleles of <= 16 bases (warp-uniform choice)
            const bool use16 = __all_sync(0xffffffffu, !elig || (l0 - 1u < 16u && l1 - 1u < 16u));
            if (act)
            {
                if (elig)
                    fast = use16 ? fast_biallelic16(R, BY, o0, l0, l1, pos1, clen, gbase, f) : fast_biallelic(R, BY, o0, l0, l1, pos1, clen, gbase, f);
                if (fast)
                {
                    st = f.status; vt = f.vtype; T = f.T; le = f.le; lt = f.lt; nb = f.nlr + f.nla;
                }
                else if (!FAST_ONLY)
                {
                    RecResult res;
                    normalize_record(R, B, rr, res);
                    st = res.status; vt = res.vtype; T = res.T; le = res.le; lt = res.lt; nb = res.nbytes;
                }
                S.st[idx] = (uint8_t)st;
                S.vt[idx] = (uint8_t)vt;
            }
            if (FAST_ONLY && __any_sync(0xffffffffu, act && !fast))
            {
                deferred = true; // some record needs the generic path: the whole tile goes to the second launch
                break;
            }
            const uint32_t c = (st & VTN_ST_CHANGED) ? 1u : 0u;
            uint32_t pc, pb;
            const unsigned cmask = __ballot_sync(0xffffffffu, c);
            const uint32_t ec = __popc(cmask & ((1u << lane) - 1u));
            pc = __popc(cmask);
            const uint32_t eb = warp_excl_scan(c ? nb : 0u, lane, pb);
            if (c)
            {
                const uint32_t k = r0 + tot_c + ec;    // scratch slot: tile segment, rank in input order
                const uint32_t obl = tot_b + eb;       // byte offset inside the tile's byte window
                // new REF then new ALT as one 2-bit code stream: up to 32 bases travel inside the scratch entry and are
                // expanded to ASCII by K2; longer ones (and everything from the generic path) are written as bytes here
                const uint32_t ntot = fast ? f.nlr + f.nla : 0u;
                const bool coded = fast && ntot <= 32u;
                uint64_t lo = f.newR, hi = 0;
                if (fast)
                {
                    if (f.nlr >= 32)
                        hi = f.newA;
                    else
                    {
                        lo |= f.newA << (2 * f.nlr);
                        hi = f.nlr ? (f.newA >> (64 - 2 * f.nlr)) : 0;
                    }
                }
                X.q0[k] = make_uint4(B.rec_base + rr, (uint32_t)(pos1 - (int32_t)le + (int32_t)lt), lt, le);
                X.q1[k] = make_uint4(T, obl | (coded ? 0x80000000u : 0u), (uint32_t)lo, (uint32_t)(lo >> 32));
                uint8_t* dst = X.bytes + b_lo + obl;
                if (fast && !coded)
                {
                    for (uint32_t j = 0; j < ntot; j += 4)
                    {
                        const uint32_t w = s_lut[(uint32_t)lo & 0xffu];
                        dst[j] = (uint8_t)w;
                        if (j + 1 < ntot) dst[j + 1] = (uint8_t)(w >> 8);
                        if (j + 2 < ntot) dst[j + 2] = (uint8_t)(w >> 16);
                        if (j + 3 < ntot) dst[j + 3] = (uint8_t)(w >> 24);
                        lo = (lo >> 8) | (hi << 56);
                        hi >>= 8;
                    }
                }
                else if (!FAST_ONLY)
                    write_generic_alleles(R, B, rr, T, le, lt, dst);
            }
            tot_c += pc;
            tot_b += pb;
            // counters: category by arity (normalize.cpp:254-311), refs / variants (:315-322)
            // one-hot 6-bit fields (a pass adds at most 32 to any of them), summed over the warp by three REDUX
            const uint32_t cat = c ? (st >> VTN_ST_CAT_SHIFT) : 0u;
            const uint32_t fld = cat ? (1u << (6u * (cat - 1u))) : 0u;
            const uint32_t r_bi = __reduce_add_sync(0xffffffffu, n2 == 2 ? fld : 0u);
            const uint32_t r_mu = FAST_ONLY ? 0u : __reduce_add_sync(0xffffffffu, n2 == 2 ? 0u : fld); // the fast launch keeps biallelic tiles only
            const uint32_t r_rv = __reduce_add_sync(0xffffffffu, act ? (vt == VTN_VT_REF ? 1u : 0x100u) : 0u);
            if (lane < 10) tile_cat += ((lane < 5 ? r_bi : r_mu) >> cat_shift) & 63u;
            cnt_refs += r_rv & 0xffu;
            cnt_vars += r_rv >> 8;
        }
        if (deferred)
        {
            if (lane == 0) defer_list[atomicAdd(defer_count, 1u)] = tile;
            __syncwarp();
            ti = next_ti;
            tile = next_tile;
            continue;
        }
        my_counter += tile_cat;
        if (lane == 10) my_counter += cnt_vars;
        if (lane == 11) my_counter += cnt_refs;
        if (lane == 0)
        {
            X.dir[tile] = ((unsigned long long)tot_c << 32) | tot_b;
            X.blo[tile] = b_lo;
            // aggregate of the K2 block this tile belongs to (count << 44 | bytes): K2 sums its predecessors directly
            if (tot_c) atomicAdd(k2_ctl + VTC_CTL_AGG + tile / tpb, ((unsigned long long)tot_c << VTC_AGG_SHIFT) | tot_b);
        }
        __syncwarp();

        // ---------------- phase 3: status / vtype as one 4-byte store per lane ----------------
        {
            uint8_t sv[VTN_RPL], vv[VTN_RPL];
#pragma unroll
            for (int j = 0; j < VTN_RPL; ++j)
            {
                const uint32_t idx = lane * VTN_RPL + j;
                const bool q = qmask & (1u << j);
                sv[j] = q ? S.st[idx] : 0;
                vv[j] = q ? S.vt[idx] : VTN_VT_SNP;
            }
            if (lane * VTN_RPL + VTN_RPL <= n_live)
            {
                reinterpret_cast<uchar4*>(O.status + r0)[lane] = make_uchar4(sv[0], sv[1], sv[2], sv[3]);
                reinterpret_cast<uchar4*>(O.vtype + r0)[lane] = make_uchar4(vv[0], vv[1], vv[2], vv[3]);
            }
            else
            {
#pragma unroll
                for (int j = 0; j < VTN_RPL; ++j)
                    if (lane * VTN_RPL + j < n_live)
                    {
                        O.status[r0 + lane * VTN_RPL + j] = sv[j];
                        O.vtype[r0 + lane * VTN_RPL + j] = vv[j];
                    }
            }
        }
#undef AO
        __syncwarp(); // the next tile reuses this warp's shared arrays
        ti = next_ti;
        tile = next_tile;
    }
#undef TILE_OF
    if (lane < 12 && my_counter) atomicAdd(O.counters + lane, my_counter);
}

// ---- K2: scan of the tile directory + gather of the scratch segments into the final, input-ordered arrays ----
// Block b owns tiles [b*tpb, (b+1)*tpb).  Its base offsets are the plain sum of the aggregates K1 accumulated for
// blocks 0..b-1 (one 64-bit reduction per tile there), so no block waits on another: no look-back, no ticket.
#define VTC_TPB 512      // maximum tiles per block; the launch picks tpb <= VTC_TPB so that all blocks are resident at once, 2 per SM
#define VTC_THREADS 1024 // threads per block: the first tpb scan the directory, all of them gather
__global__ void __launch_bounds__(VTC_THREADS, 2) vtn_compact_kernel(VtnDevBatchView B, VtnDevScratch X, VtnDevNormOut O, uint32_t n_tiles,
                                                              const unsigned long long* __restrict__ ctl, uint32_t n_blocks, uint32_t tpb)
{
    __shared__ uint32_t s_wc[VTC_TPB / 32], s_wb[VTC_TPB / 32];
    __shared__ unsigned long long s_pc[VTC_THREADS / 32], s_pb[VTC_THREADS / 32];
    __shared__ uint32_t s_cnt[VTC_TPB], s_nb2[VTC_TPB], s_bc[VTC_TPB + 1];
    __shared__ unsigned long long s_bb[VTC_TPB];
    __shared__ uint32_t s_lut[256]; // 4 packed 2-bit codes -> their 4 ASCII bytes (A=0 C=1 T=2 G=3)
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid < 256)
    {
        uint32_t w = 0;
        for (int q = 0; q < 4; ++q) w |= ((0x47544341u >> (8 * ((tid >> (2 * q)) & 3))) & 0xffu) << (8 * q);
        s_lut[tid] = w;
    }
    const uint32_t blk = blockIdx.x;
    const uint32_t t = blk * tpb + tid;
    const unsigned long long d = (tid < (int)tpb && t < n_tiles) ? X.dir[t] : 0ull;
    // predecessors' aggregates (issued together with the directory load: one memory latency for both)
    unsigned long long pc = 0, pb = 0;
    for (uint32_t i = tid; i < blk; i += VTC_THREADS)
    {
        const unsigned long long a = ctl[VTC_CTL_AGG + i];
        pc += a >> VTC_AGG_SHIFT;
        pb += a & ((1ull << VTC_AGG_SHIFT) - 1ull);
    }
    const uint32_t c = (uint32_t)(d >> 32), nb = (uint32_t)d;
    uint32_t ic = c, ib = nb;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1)
    {
        const uint32_t tc = __shfl_up_sync(0xffffffffu, ic, o);
        const uint32_t tb = __shfl_up_sync(0xffffffffu, ib, o);
        if (lane >= o) { ic += tc; ib += tb; }
    }
    if (blk > (uint32_t)(warp * 32)) // warps that loaded no aggregate skip the reduction (uniform per warp)
    {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1)
        {
            pc += __shfl_xor_sync(0xffffffffu, pc, o);
            pb += __shfl_xor_sync(0xffffffffu, pb, o);
        }
    }
    if (lane == 31 && warp < (int)(tpb / 32)) { s_wc[warp] = ic; s_wb[warp] = ib; }
    if (lane == 0) { s_pc[warp] = pc; s_pb[warp] = pb; }
    __syncthreads();
    uint32_t wc = 0, wb = 0, tot_c = 0, tot_b = 0;
    for (int w = 0; w < (int)(tpb / 32); ++w)
    {
        const uint32_t xc = s_wc[w], xb = s_wb[w];
        if (w < warp) { wc += xc; wb += xb; }
        tot_c += xc;
        tot_b += xb;
    }
    unsigned long long base_c = ctl[VTC_CTL_BASE], base_b = ctl[VTC_CTL_BASE + 1];
    {
        const int nw = (int)min((blk + 31u) / 32u, (uint32_t)(VTC_THREADS / 32));
        for (int w = 0; w < nw; ++w) { base_c += s_pc[w]; base_b += s_pb[w]; }
    }
    if (tid == 0)
    {
        if (blk == 0) O.chg_off[base_c] = base_b; // a fresh run: chg_off[0] = 0; a later chunk rewrites the same value
        if (blk == n_blocks - 1)
        {
            O.totals[0] = base_c + tot_c;
            O.totals[1] = base_b + tot_b;
            O.chg_off[base_c + tot_c] = base_b + tot_b;
        }
    }
    if (tid < (int)tpb)
    {
        s_cnt[tid] = c;
        s_nb2[tid] = nb;
        s_bc[tid] = wc + (ic - c); // block-local exclusive prefix of changed records
        s_bb[tid] = base_b + wb + (ib - nb);
    }
    __syncthreads();
    // flat gather: entry j of this block (0 <= j < tot_c) belongs to the tile found by binary search on the
    // block-local prefix; every thread has many independent loads in flight
    for (uint32_t j = tid; j < tot_c; j += VTC_THREADS)
    {
        uint32_t lo = 0, hi = tpb - 1; // last tile tt with s_bc[tt] <= j
        while (lo < hi)
        {
            const uint32_t mid = (lo + hi + 1) >> 1;
            if (s_bc[mid] <= j) lo = mid;
            else hi = mid - 1;
        }
        const uint32_t tt = lo;
        const uint32_t i = j - s_bc[tt];
        const uint32_t tile = blk * tpb + tt;
        const uint32_t src = tile * VTN_WTILE + i;
        const unsigned long long k = base_c + j;
        const uint4 e0 = X.q0[src], e1 = X.q1[src];
        const uint32_t ob = e1.y & 0x7fffffffu;
        const uint32_t ob_next = (i + 1 < s_cnt[tt]) ? (X.q1[src + 1].y & 0x7fffffffu) : s_nb2[tt];
        const unsigned long long dst_off = s_bb[tt] + ob;
        uint8_t* __restrict__ db = O.chg_bytes + dst_off;
        const uint32_t n = ob_next - ob;
        O.chg_rec[k] = e0.x;
        O.chg_pos1[k] = (int32_t)e0.y;
        O.chg_lt[k] = e0.z;
        O.chg_le[k] = e0.w;
        O.chg_rt[k] = e1.x;
        O.chg_off[k] = dst_off;
        if (e1.y >> 31)
        { // the new alleles came as 2-bit codes (n <= 32 bases): four bases per table lookup
            uint32_t lo = e1.z, hi = e1.w;
            for (uint32_t q0 = 0; q0 < n; q0 += 4)
            {
                const uint32_t w = s_lut[lo & 0xffu];
                db[q0] = (uint8_t)w;
                if (q0 + 1 < n) db[q0 + 1] = (uint8_t)(w >> 8);
                if (q0 + 2 < n) db[q0 + 2] = (uint8_t)(w >> 16);
                if (q0 + 3 < n) db[q0 + 3] = (uint8_t)(w >> 24);
                lo = __funnelshift_r(lo, hi, 8);
                hi >>= 8;
            }
        }
        else
        {
            const uint8_t* __restrict__ sb = X.bytes + X.blo[tile] + ob;
            for (uint32_t q0 = 0; q0 < n; q0 += 8)
            { // eight byte loads in flight before the first store
                uint8_t tmp[8];
#pragma unroll
                for (int q = 0; q < 8; ++q) tmp[q] = q0 + q < n ? sb[q0 + q] : (uint8_t)0;
#pragma unroll
                for (int q = 0; q < 8; ++q)
                    if (q0 + q < n) db[q0 + q] = tmp[q];
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// K1 for the compact biallelic layout (vtn_batch_bi): the same tile scheme and the same per-record code (fast_biallelic16 /
// fast_biallelic / normalize_record) as vtn_normalize_kernel, with a front end that reads 6 header bytes per record instead
// of 20: lens as one 8-byte load per lane, the tile's byte offsets by a warp scan from tile_off[tile], contig ids from the
// run table (one binary search per tile).  Per record it leaves ONE byte (VT_* bits | REF-check code << 6); statuses that the
// changed list does not imply go to the notes list.  The scratch / tile-directory contract towards K2 is unchanged.
// FAST_ONLY as above: a tile with an escape length, more bytes than the staging window, or a record the register path
// refuses is appended to defer_list and redone by the <false> launch, which lays a small vtn_batch view of the tile
// (allele_beg / allele_off / rid in shared memory) under normalize_record.
// ---------------------------------------------------------------------------------------------------------
// staging window of the fast launch: mean allele bytes per record a tile may hold before it is deferred.  8 for ordinary call sets (C2
// averages 2.9), 24 for batches of long alleles (the STR indels of C4 average ~12; with the small window every tile was deferred and C4
// ran 2.6x slower than in the general layout).  The launcher picks the instantiation from the batch's own mean (VtnDevBiView::wide).
#define VTB_BPR_NARROW 8
#define VTB_BPR_WIDE 24
#ifndef VTB_MIN_BLOCKS_FAST
#define VTB_MIN_BLOCKS_FAST 8
#endif

template <int BPR>
struct __align__(16) BiWarpSmem
{
    enum { BYTES_CAP = VTN_WTILE * BPR + 64 };
    uint8_t bytes[BYTES_CAP];
    int32_t pos1[VTN_WTILE];
    uint32_t queue[VTN_WTILE]; // record index in the tile | REF length << 8 | ALT length << 16 (lengths < 255 on the fast launch)
    uint16_t qoff[VTN_WTILE];  // byte offset of the queued record inside the tile (fast launch: tile bytes <= BYTES_CAP)
    uint8_t sv[VTN_WTILE];
};
struct BiWarpGen // the generic launch's view of one tile as a vtn_batch
{
    uint32_t ao[2 * VTN_WTILE + 1];
    uint32_t ab[VTN_WTILE + 1];
    int32_t rid[VTN_WTILE];
};

__device__ __forceinline__ void vtn_note(const VtnDevNotes& N, uint32_t rec, uint32_t st)
{
    const uint32_t k = atomicAdd(N.count, 1u);
    if (k < N.cap)
    {
        N.rec[k] = rec;
        N.status[k] = (uint8_t)st;
    }
}

template <bool FAST_ONLY, int BPR>
__global__ void __launch_bounds__(32 * VTN_WARPS, FAST_ONLY ? (BPR > VTB_BPR_NARROW ? 6 : VTB_MIN_BLOCKS_FAST) : VTN_MIN_BLOCKS) vtn_normalize_bi_kernel(
    VtnDevRef R, VtnDevBiView B, VtnDevNormOut O, VtnDevNotes NT, VtnDevScratch X, uint32_t* defer_list, uint32_t* defer_count,
    uint32_t n_tiles_all, unsigned long long* k2_ctl, uint32_t tpb, int fresh)
{
    __shared__ BiWarpSmem<BPR> SM[VTN_WARPS];
    __shared__ BiWarpGen GM[FAST_ONLY ? 1 : VTN_WARPS];
    __shared__ uint32_t s_lut[256];
    __shared__ uint64_t s_gbase[VTN_CONTIG_CAP];
    __shared__ int64_t s_clen[VTN_CONTIG_CAP];
    const int lane = threadIdx.x & 31;
    if (!FAST_ONLY && blockIdx.x * VTN_WARPS >= *defer_count) return; // the list is usually empty or short: most blocks leave here
    BiWarpSmem<BPR>& S = SM[threadIdx.x >> 5];
    BiWarpGen& G = GM[FAST_ONLY ? 0 : (threadIdx.x >> 5)];
    uint32_t my_counter = 0;
    const uint32_t cat_shift = 6u * (uint32_t)(lane < 5 ? lane : 0);
    if (FAST_ONLY && blockIdx.x == 0)
    {
        if (threadIdx.x < 2) k2_ctl[VTC_CTL_BASE + threadIdx.x] = fresh ? 0ull : O.totals[threadIdx.x];
    }
    for (int v = threadIdx.x; v < 256; v += blockDim.x)
    {
        uint32_t w = 0;
        for (int q = 0; q < 4; ++q) w |= ((0x47544341u >> (8 * ((v >> (2 * q)) & 3))) & 0xffu) << (8 * q);
        s_lut[v] = w;
    }
    const bool contigs_staged = R.n_contig <= VTN_CONTIG_CAP;
    if (contigs_staged)
        for (int c = threadIdx.x; c < R.n_contig; c += blockDim.x)
        {
            s_gbase[c] = __ldg(R.contig_goff + c) << 5;
            s_clen[c] = __ldg(R.contig_len + c);
        }
    if (!FAST_ONLY)
        for (int i = lane; i <= VTN_WTILE; i += 32) G.ab[i] = 2u * (uint32_t)i;
    __syncthreads();

    const uint32_t n_tiles = FAST_ONLY ? n_tiles_all : min(*defer_count, n_tiles_all);
    const uint32_t tile_stride = gridDim.x * VTN_WARPS;
    uint32_t ti = blockIdx.x * VTN_WARPS + (threadIdx.x >> 5);
#define TILE_OF(i) (FAST_ONLY ? (i) : __ldg(defer_list + (i)))
    uint32_t tile = ti < n_tiles ? TILE_OF(ti) : 0;
    uint2 h_len = make_uint2(0, 0);
    int4 h_pos = make_int4(0, 0, 0, 0);
    uint32_t h_blo = 0;
    uint32_t run_k = 0, run_lo = 1, run_next = 0; // empty interval: the first tile searches
    if (ti < n_tiles)
    {
        const uint32_t r0 = tile * VTN_WTILE;
        h_len = __ldg(reinterpret_cast<const uint2*>(B.lens + r0) + lane); // arrays are padded
        h_pos = __ldg(reinterpret_cast<const int4*>(B.pos1 + r0) + lane);
        h_blo = __ldg(B.tile_off + tile);
    }
    while (ti < n_tiles)
    {
        const uint32_t next_ti = ti + tile_stride;
        const uint32_t next_tile = next_ti < n_tiles ? TILE_OF(next_ti) : 0;
        const uint32_t r0 = tile * VTN_WTILE;
        const uint32_t n_live = min((uint32_t)VTN_WTILE, B.n_rec - r0);
        const uint32_t abs0 = B.rec_base + r0;
        const uint32_t b_lo = h_blo;
        // ---------------- phase 1 ----------------
        uint32_t l0[VTN_RPL], l1[VTN_RPL];
        {
            const uint32_t w[2] = {h_len.x, h_len.y};
#pragma unroll
            for (int j = 0; j < VTN_RPL; ++j)
            {
                const uint32_t x = (w[j >> 1] >> (16 * (j & 1))) & 0xffffu;
                const bool live = (uint32_t)(lane * VTN_RPL + j) < n_live;
                l0[j] = live ? (x & 0xffu) : 0u;
                l1[j] = live ? (x >> 8) : 0u;
            }
            reinterpret_cast<int4*>(S.pos1)[lane] = h_pos;
        }
        if (next_ti < n_tiles)
        {
            const uint32_t r1 = next_tile * VTN_WTILE;
            h_len = __ldg(reinterpret_cast<const uint2*>(B.lens + r1) + lane);
            h_pos = __ldg(reinterpret_cast<const int4*>(B.pos1 + r1) + lane);
            h_blo = __ldg(B.tile_off + next_tile);
        }
        bool esc = false;
#pragma unroll
        for (int j = 0; j < VTN_RPL; ++j) esc |= (l0[j] == VTN_BI_LEN_ESC) | (l1[j] == VTN_BI_LEN_ESC);
        bool deferred = false;
        if (FAST_ONLY)
            deferred = __any_sync(0xffffffffu, esc);
        else if (esc)
        {
#pragma unroll
            for (int j = 0; j < VTN_RPL; ++j)
                if (l0[j] == VTN_BI_LEN_ESC || l1[j] == VTN_BI_LEN_ESC)
                {
                    const uint32_t ar = abs0 + lane * VTN_RPL + j;
                    uint32_t lo = 0, hi = B.n_ext; // first ext_rec >= ar
                    while (lo < hi)
                    {
                        const uint32_t mid = (lo + hi) >> 1;
                        if (__ldg(B.ext_rec + mid) < ar) lo = mid + 1;
                        else hi = mid;
                    }
                    if (lo < B.n_ext && __ldg(B.ext_rec + lo) == ar)
                    {
                        l0[j] = __ldg(B.ext_len + 2 * lo);
                        l1[j] = __ldg(B.ext_len + 2 * lo + 1);
                    }
                }
        }
        uint32_t tile_bytes = 0;
        uint32_t o[VTN_RPL]; // byte offset of my records inside the tile
        {
            uint32_t tot = 0;
#pragma unroll
            for (int j = 0; j < VTN_RPL; ++j) tot += l0[j] + l1[j];
            uint32_t e = warp_excl_scan(tot, lane, tile_bytes);
#pragma unroll
            for (int j = 0; j < VTN_RPL; ++j)
            {
                o[j] = e;
                e += l0[j] + l1[j];
            }
        }
        const uint32_t b0 = b_lo & ~15u;
        const bool bytes_staged = (b_lo - b0) + tile_bytes + 16 <= (uint32_t)BiWarpSmem<BPR>::BYTES_CAP;
        if (FAST_ONLY && !bytes_staged) deferred = true;
        if (FAST_ONLY && deferred)
        {
            if (lane == 0) defer_list[atomicAdd(defer_count, 1u)] = tile;
            __syncwarp();
            ti = next_ti;
            tile = next_tile;
            continue;
        }
        if (bytes_staged)
        {
            for (uint32_t i = lane * 16; i < (b_lo - b0) + tile_bytes + 16; i += 512)
                *reinterpret_cast<uint4*>(&S.bytes[i]) = __ldg(reinterpret_cast<const uint4*>(B.bytes + b0 + i));
        }
        // contig run of the tile's first record: last k with run_start[k] <= abs0 (uniform over the warp).  A warp's tiles
        // ascend, so the run of its previous tile is tried first
        if (abs0 < run_lo || abs0 >= run_next)
        {
            uint32_t lo = 0, hi = B.n_run - 1;
            while (lo < hi)
            {
                const uint32_t mid = (lo + hi + 1) >> 1;
                if (__ldg(B.run_start + mid) <= abs0) lo = mid;
                else hi = mid - 1;
            }
            run_k = lo;
            run_lo = __ldg(B.run_start + run_k);
            run_next = __ldg(B.run_start + run_k + 1);
        }
        __syncwarp();
        // absolute offsets work on both; the fast launch always stages, which keeps its loads in the shared window (LDS)
        const uint8_t* BY = FAST_ONLY ? (S.bytes - b0) : (bytes_staged ? (S.bytes - b0) : B.bytes);

        uint32_t qmask = 0;
#pragma unroll
        for (int j = 0; j < VTN_RPL; ++j)
        {
            const uint32_t idx = lane * VTN_RPL + j;
            if (idx < n_live)
            {
                bool q = true;
                if (l0[j] == 1 && l1[j] == 1)
                {
                    const uint8_t r0b = BY[b_lo + o[j]], c = BY[b_lo + o[j] + 1];
                    q = (c == '<' || c == '[' || c == ']' || c == '.' || c == r0b);
                }
                if (q) qmask |= 1u << j;
                if (!FAST_ONLY)
                {
                    G.ao[2 * idx] = b_lo + o[j];
                    G.ao[2 * idx + 1] = b_lo + o[j] + l0[j];
                    uint32_t k = run_k;
                    if (abs0 + idx >= run_next)
                        while (k + 1 < B.n_run && __ldg(B.run_start + k + 1) <= abs0 + idx) ++k;
                    G.rid[idx] = __ldg(B.run_rid + k);
                }
            }
        }
        if (!FAST_ONLY && lane == 0) G.ao[2 * n_live] = b_lo + tile_bytes;
        uint32_t nq;
        {
            uint32_t slot = warp_excl_scan(__popc(qmask), lane, nq);
#pragma unroll
            for (int j = 0; j < VTN_RPL; ++j)
                if (qmask & (1u << j))
                {
                    S.queue[slot] = (uint32_t)(lane * VTN_RPL + j) | (FAST_ONLY ? ((l0[j] << 8) | (l1[j] << 16)) : 0u);
                    S.qoff[slot] = (uint16_t)o[j];
                    ++slot;
                }
        }
        __syncwarp();

        // ---------------- phase 2: dense passes over the queue ----------------
        uint32_t tot_c = 0, tot_b = 0;
        uint32_t cnt_vars = n_live - nq; // the SNVs settled in phase 1 (normalize.cpp:315-322)
        uint32_t cnt_refs = 0;
        uint32_t tile_cat = 0; // lane k < 5: category counter k of THIS tile (committed only if the tile is kept)
        for (uint32_t p = 0; p < nq; p += 32)
        {
            const uint32_t e = p + lane;
            const bool act = e < nq;
            uint32_t st = 0, vt = VTN_VT_SNP, T = 0, le = 0, lt = 0, nb = 0, rr = 0;
            int32_t pos1 = 0;
            bool fast = false, elig = false;
            VtnFastRes f;
            uint32_t o0 = 0, ql0 = 0, ql1 = 0, idx = 0;
            int64_t clen = 0;
            uint64_t gbase = 0;
            if (act)
            {
                const uint32_t qe = S.queue[e];
                idx = qe & 0xffu;
                rr = r0 + idx;
                pos1 = S.pos1[idx];
                int32_t rid;
                if (FAST_ONLY)
                {
                    ql0 = (qe >> 8) & 0xffu;
                    ql1 = (qe >> 16) & 0xffu;
                    o0 = b_lo + S.qoff[e];
                    uint32_t k = run_k;
                    if (abs0 + idx >= run_next)
                        while (k + 1 < B.n_run && __ldg(B.run_start + k + 1) <= abs0 + idx) ++k;
                    rid = __ldg(B.run_rid + k);
                }
                else
                {
                    o0 = G.ao[2 * idx];
                    ql0 = G.ao[2 * idx + 1] - o0;
                    ql1 = G.ao[2 * idx + 2] - G.ao[2 * idx + 1];
                    rid = G.rid[idx];
                }
                if (rid >= 0 && rid < R.n_contig)
                {
                    elig = true;
                    clen = contigs_staged ? s_clen[rid] : __ldg(R.contig_len + rid);
                    gbase = contigs_staged ? s_gbase[rid] : (__ldg(R.contig_goff + rid) << 5);
                }
            }
            const bool use16 = __all_sync(0xffffffffu, !elig || (ql0 - 1u < 16u && ql1 - 1u < 16u));
            if (act)
            {
                if (elig)
                    fast = use16 ? fast_biallelic16(R, BY, o0, ql0, ql1, pos1, clen, gbase, f) : fast_biallelic(R, BY, o0, ql0, ql1, pos1, clen, gbase, f);
                if (fast)
                {
                    st = f.status; vt = f.vtype; T = f.T; le = f.le; lt = f.lt; nb = f.nlr + f.nla;
                }
                else if (!FAST_ONLY)
                {
                    VtnDevBatchView V{n_live, 0u, B.pos1 + r0, G.rid, G.ab, G.ao, B.bytes};
                    RecResult res;
                    normalize_record(R, V, idx, res);
                    st = res.status; vt = res.vtype; T = res.T; le = res.le; lt = res.lt; nb = res.nbytes;
                }
                S.sv[idx] = (uint8_t)(vt | ((st & VTN_ST_REFCODE_MASK) << 6));
            }
            if (FAST_ONLY && __any_sync(0xffffffffu, act && !fast))
            {
                deferred = true;
                break;
            }
            if (act && ((st & VTN_ST_FETCH_FAIL) || ((st & VTN_ST_ENTERED) && !(st & VTN_ST_CHANGED)))) vtn_note(NT, abs0 + idx, st);
            const uint32_t c = (st & VTN_ST_CHANGED) ? 1u : 0u;
            uint32_t pc, pb;
            const unsigned cmask = __ballot_sync(0xffffffffu, c);
            const uint32_t ec = __popc(cmask & ((1u << lane) - 1u));
            pc = __popc(cmask);
            const uint32_t eb = warp_excl_scan(c ? nb : 0u, lane, pb);
            if (c)
            {
                const uint32_t k = r0 + tot_c + ec;
                const uint32_t obl = tot_b + eb;
                const uint32_t ntot = fast ? f.nlr + f.nla : 0u;
                const bool coded = fast && ntot <= 32u;
                uint64_t lo = f.newR, hi = 0;
                if (fast)
                {
                    if (f.nlr >= 32)
                        hi = f.newA;
                    else
                    {
                        lo |= f.newA << (2 * f.nlr);
                        hi = f.nlr ? (f.newA >> (64 - 2 * f.nlr)) : 0;
                    }
                }
                X.q0[k] = make_uint4(B.rec_base + rr, (uint32_t)(pos1 - (int32_t)le + (int32_t)lt), lt, le);
                X.q1[k] = make_uint4(T, obl | (coded ? 0x80000000u : 0u), (uint32_t)lo, (uint32_t)(lo >> 32));
                uint8_t* dst = X.bytes + b_lo + obl;
                if (fast && !coded)
                {
                    for (uint32_t j = 0; j < ntot; j += 4)
                    {
                        const uint32_t w = s_lut[(uint32_t)lo & 0xffu];
                        dst[j] = (uint8_t)w;
                        if (j + 1 < ntot) dst[j + 1] = (uint8_t)(w >> 8);
                        if (j + 2 < ntot) dst[j + 2] = (uint8_t)(w >> 16);
                        if (j + 3 < ntot) dst[j + 3] = (uint8_t)(w >> 24);
                        lo = (lo >> 8) | (hi << 56);
                        hi >>= 8;
                    }
                }
                else if (!FAST_ONLY)
                {
                    VtnDevBatchView V{n_live, 0u, B.pos1 + r0, G.rid, G.ab, G.ao, B.bytes};
                    write_generic_alleles(R, V, idx, T, le, lt, dst);
                }
            }
            tot_c += pc;
            tot_b += pb;
            const uint32_t cat = c ? (st >> VTN_ST_CAT_SHIFT) : 0u;
            const uint32_t fld = cat ? (1u << (6u * (cat - 1u))) : 0u;
            const uint32_t r_bi = __reduce_add_sync(0xffffffffu, fld);
            const uint32_t r_rv = __reduce_add_sync(0xffffffffu, act ? (vt == VTN_VT_REF ? 1u : 0x100u) : 0u);
            if (lane < 5) tile_cat += (r_bi >> cat_shift) & 63u;
            cnt_refs += r_rv & 0xffu;
            cnt_vars += r_rv >> 8;
        }
        if (deferred)
        {
            if (lane == 0) defer_list[atomicAdd(defer_count, 1u)] = tile;
            __syncwarp();
            ti = next_ti;
            tile = next_tile;
            continue;
        }
        my_counter += tile_cat;
        if (lane == 10) my_counter += cnt_vars;
        if (lane == 11) my_counter += cnt_refs;
        if (lane == 0)
        {
            X.dir[tile] = ((unsigned long long)tot_c << 32) | tot_b;
            X.blo[tile] = b_lo;
            if (tot_c) atomicAdd(k2_ctl + VTC_CTL_AGG + tile / tpb, ((unsigned long long)tot_c << VTC_AGG_SHIFT) | tot_b);
        }
        __syncwarp();

        // ---------------- phase 3: one byte per record, four per lane ----------------
        {
            uint8_t sv[VTN_RPL];
#pragma unroll
            for (int j = 0; j < VTN_RPL; ++j) sv[j] = (qmask & (1u << j)) ? S.sv[lane * VTN_RPL + j] : (uint8_t)VTN_VT_SNP;
            if (lane * VTN_RPL + VTN_RPL <= n_live)
                reinterpret_cast<uchar4*>(O.status + r0)[lane] = make_uchar4(sv[0], sv[1], sv[2], sv[3]);
            else
            {
#pragma unroll
                for (int j = 0; j < VTN_RPL; ++j)
                    if (lane * VTN_RPL + j < n_live) O.status[r0 + lane * VTN_RPL + j] = sv[j];
            }
        }
        __syncwarp();
        ti = next_ti;
        tile = next_tile;
    }
#undef TILE_OF
    if (lane < 12 && my_counter) atomicAdd(O.counters + lane, my_counter);
}

} // namespace

uint32_t vtn_normalize_tile_records() { return VTN_WTILE; }
// 64-bit words of control block for a batch of n_tiles tiles: K2 never has more than n_tiles/32 + 1 blocks
size_t vtn_normalize_ctl_words(uint32_t n_tiles) { return VTC_CTL_AGG + (size_t)n_tiles / 32 + 2; }

int vtn_launch_normalize(vtn_ctx* ctx, const VtnDevBatchView& in, const VtnDevNormOut& out, const VtnDevScratch& scr,
                         unsigned long long* tile_state, uint32_t* tile_counter, uint32_t n_tiles, bool accumulate)
{
    cudaStream_t s = ctx->stream;
    if (!accumulate)
    {
        VTN_CUDA(ctx, cudaMemsetAsync(out.counters, 0, 12 * sizeof(uint32_t), s));
        if (n_tiles == 0)
        {
            VTN_CUDA(ctx, cudaMemsetAsync(out.totals, 0, 2 * sizeof(unsigned long long), s));
            VTN_CUDA(ctx, cudaMemsetAsync(out.chg_off, 0, sizeof(uint64_t), s));
        }
    }
    if (n_tiles == 0) return VTN_OK;
    // K2 geometry: tiles per block chosen so that every block is resident at once with two blocks per SM
    // (full occupancy, a single wave): tpb = ceil(n_tiles / (2 * SMs)) rounded up to a warp, within [32, VTC_TPB]
    uint32_t tpb = (n_tiles + 2 * (uint32_t)ctx->sm_count - 1) / (2 * (uint32_t)ctx->sm_count);
    tpb = (tpb + 31) & ~31u;
    if (tpb < 32) tpb = 32;
    if (tpb > VTC_TPB) tpb = VTC_TPB;
    const uint32_t n_blocks2 = (n_tiles + tpb - 1) / tpb;
    // control block (tile_counter viewed as 64-bit words): [0] low half = number of deferred tiles, [VTC_CTL_BASE..+1] =
    // totals before this launch (written by K1 block 0), [VTC_CTL_AGG + b] = aggregate of K2 block b (accumulated by K1).
    // One memset clears it; totals / chg_off[0] are written by the kernels themselves.
    unsigned long long* ctl = reinterpret_cast<unsigned long long*>(tile_counter);
    VTN_CUDA(ctx, cudaMemsetAsync(ctl, 0, (VTC_CTL_AGG + (size_t)n_blocks2) * sizeof(unsigned long long), s));
    static int per_sm = 0, per_sm_slow = 0;
    if (!per_sm)
    {
        VTN_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, vtn_normalize_kernel<true>, 32 * VTN_WARPS, 0));
        VTN_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm_slow, vtn_normalize_kernel<false>, 32 * VTN_WARPS, 0));
        if (per_sm < 1) per_sm = 1;
        if (per_sm_slow < 1) per_sm_slow = 1;
    }
    uint32_t grid = (uint32_t)(ctx->sm_count * per_sm);
    const uint32_t need = (n_tiles + VTN_WARPS - 1) / VTN_WARPS;
    if (grid > need) grid = need;
    uint32_t* defer_list = reinterpret_cast<uint32_t*>(tile_state); // at most n_tiles entries
    vtn_normalize_kernel<true><<<grid, 32 * VTN_WARPS, 0, s>>>(ctx->dref, in, out, scr, defer_list, tile_counter, n_tiles, ctl, tpb, accumulate ? 0 : 1);
    VTN_CUDA(ctx, cudaGetLastError());
    uint32_t grid2 = (uint32_t)(ctx->sm_count * per_sm_slow);
    if (grid2 > need) grid2 = need;
    vtn_normalize_kernel<false><<<grid2, 32 * VTN_WARPS, 0, s>>>(ctx->dref, in, out, scr, defer_list, tile_counter, n_tiles, ctl, tpb, 0);
    VTN_CUDA(ctx, cudaGetLastError());
    vtn_compact_kernel<<<n_blocks2, VTC_THREADS, 0, s>>>(in, scr, out, n_tiles, ctl, n_blocks2, tpb);
    VTN_CUDA(ctx, cudaGetLastError());
    ctx->launches += 3;
    return VTN_OK;
}

int vtn_launch_normalize_bi(vtn_ctx* ctx, const VtnDevBiView& in, const VtnDevNormOut& out, const VtnDevNotes& notes, const VtnDevScratch& scr,
                            unsigned long long* tile_state, uint32_t* tile_counter, uint32_t n_tiles, bool accumulate)
{
    cudaStream_t s = ctx->stream;
    if (!accumulate)
    {
        VTN_CUDA(ctx, cudaMemsetAsync(out.counters, 0, 12 * sizeof(uint32_t), s));
        VTN_CUDA(ctx, cudaMemsetAsync(notes.count, 0, sizeof(uint32_t), s));
        if (n_tiles == 0)
        {
            VTN_CUDA(ctx, cudaMemsetAsync(out.totals, 0, 2 * sizeof(unsigned long long), s));
            VTN_CUDA(ctx, cudaMemsetAsync(out.chg_off, 0, sizeof(uint64_t), s));
        }
    }
    if (n_tiles == 0) return VTN_OK;
    uint32_t tpb = (n_tiles + 2 * (uint32_t)ctx->sm_count - 1) / (2 * (uint32_t)ctx->sm_count);
    tpb = (tpb + 31) & ~31u;
    if (tpb < 32) tpb = 32;
    if (tpb > VTC_TPB) tpb = VTC_TPB;
    const uint32_t n_blocks2 = (n_tiles + tpb - 1) / tpb;
    unsigned long long* ctl = reinterpret_cast<unsigned long long*>(tile_counter);
    VTN_CUDA(ctx, cudaMemsetAsync(ctl, 0, (VTC_CTL_AGG + (size_t)n_blocks2) * sizeof(unsigned long long), s));
    static int per_sm = 0, per_sm_wide = 0, per_sm_slow = 0;
    if (!per_sm)
    {
        VTN_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, vtn_normalize_bi_kernel<true, VTB_BPR_NARROW>, 32 * VTN_WARPS, 0));
        VTN_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm_wide, vtn_normalize_bi_kernel<true, VTB_BPR_WIDE>, 32 * VTN_WARPS, 0));
        VTN_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm_slow, vtn_normalize_bi_kernel<false, VTB_BPR_NARROW>, 32 * VTN_WARPS, 0));
        if (per_sm < 1) per_sm = 1;
        if (per_sm_wide < 1) per_sm_wide = 1;
        if (per_sm_slow < 1) per_sm_slow = 1;
    }
    uint32_t grid = (uint32_t)(ctx->sm_count * (in.wide ? per_sm_wide : per_sm));
    const uint32_t need = (n_tiles + VTN_WARPS - 1) / VTN_WARPS;
    if (grid > need) grid = need;
    uint32_t* defer_list = reinterpret_cast<uint32_t*>(tile_state);
    if (in.wide)
        vtn_normalize_bi_kernel<true, VTB_BPR_WIDE><<<grid, 32 * VTN_WARPS, 0, s>>>(ctx->dref, in, out, notes, scr, defer_list, tile_counter, n_tiles, ctl, tpb, accumulate ? 0 : 1);
    else
        vtn_normalize_bi_kernel<true, VTB_BPR_NARROW><<<grid, 32 * VTN_WARPS, 0, s>>>(ctx->dref, in, out, notes, scr, defer_list, tile_counter, n_tiles, ctl, tpb, accumulate ? 0 : 1);
    VTN_CUDA(ctx, cudaGetLastError());
    // the deferred tiles: a full wave is launched, and a block whose first tile lies beyond the list's end returns at once
    uint32_t grid2 = (uint32_t)(ctx->sm_count * per_sm_slow);
    if (grid2 > need) grid2 = need;
    vtn_normalize_bi_kernel<false, VTB_BPR_NARROW><<<grid2, 32 * VTN_WARPS, 0, s>>>(ctx->dref, in, out, notes, scr, defer_list, tile_counter, n_tiles, ctl, tpb, 0);
    VTN_CUDA(ctx, cudaGetLastError());
    vtn_compact_kernel<<<n_blocks2, VTC_THREADS, 0, s>>>(VtnDevBatchView{}, scr, out, n_tiles, ctl, n_blocks2, tpb);
    VTN_CUDA(ctx, cudaGetLastError());
    ctx->launches += 3;
    return VTN_OK;
}
