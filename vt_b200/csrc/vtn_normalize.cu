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
#include "vtn_fast.cuh"

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

// Phase C of the closed form for a biallelic record: every base of both alleles has been consumed
// (T >= max len), so both streams read the reference, at a lag of |len0-len1| bases.  Walks down in 32-base
// chunks (two packed windows, XOR, clz); per base only at contig edges or over N/IUPAC bases.
// Returns the final T (<= cap).
__device__ uint32_t lag_scan(const VtnDevRef& R, uint64_t gbase, int64_t clen, int32_t pos1, uint32_t len0, uint32_t len1,
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
__device__ bool fast_biallelic(const VtnDevRef& R, const uint8_t* bytes, uint32_t off0, uint32_t len0, uint32_t len1,
                               int32_t pos1, int32_t rid, VtnFastRes& o)
{
    if (len0 > 32 || len1 > 32 || rid < 0 || rid >= R.n_contig || pos1 < 34) return false;
    const int64_t clen = __ldg(R.contig_len + rid);
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
    const uint64_t gbase = __ldg(R.contig_goff + rid) << 5;
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
            const int64_t cap64 = (int64_t)pos1 - 2 + (int64_t)m;
            const uint32_t cap = cap64 > 0xffffffffll ? 0xffffffffu : (uint32_t)cap64;
            T = lag_scan(R, gbase, clen, pos1, len0, len1, T, cap);
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

// Per-tile shared state.  Phase 1: every thread reads its record header and settles SNVs on the spot; other
// records are queued.  Phase 2: queued records are processed DENSELY (lane i takes queue entry i), fast path
// first, generic path for what it refuses.  Phase 3: block scan + look-back, then the changed records of
// the tile are written by consecutive threads so that every output array is written coalesced.
struct TileSmem
{
    uint64_t newR[VTN_TILE], newA[VTN_TILE];
    uint32_t T[VTN_TILE], le[VTN_TILE], lt[VTN_TILE], nbytes[VTN_TILE], ob[VTN_TILE];
    uint16_t queue[VTN_TILE], gqueue[VTN_TILE], cqueue[VTN_TILE];
    uint8_t status[VTN_TILE], vtype[VTN_TILE], nlr[VTN_TILE], nla[VTN_TILE], fast[VTN_TILE];
    uint32_t wcnt[VTN_TILE / 32], wbytes[VTN_TILE / 32];
    unsigned long long base[2];
    uint32_t counters[12];
    uint32_t tile, nq, ngq;
};

__device__ __forceinline__ uint32_t warp_push(uint32_t* counter, bool pred, int lane)
{
    // warp-aggregated append: returns this lane's slot (valid when pred)
    unsigned m = __ballot_sync(0xffffffffu, pred);
    uint32_t base = 0;
    if (lane == 0 && m) base = atomicAdd(counter, __popc(m));
    base = __shfl_sync(0xffffffffu, base, 0);
    return base + __popc(m & ((1u << lane) - 1u));
}

__global__ void __launch_bounds__(VTN_TILE) vtn_normalize_kernel(VtnDevRef R, VtnDevBatchView B, VtnDevNormOut O,
                                                                 unsigned long long* tile_state, uint32_t* tile_counter,
                                                                 uint32_t n_tiles)
{
    __shared__ TileSmem S;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid < 12) S.counters[tid] = 0;
    unsigned long long* st_cnt = tile_state;
    unsigned long long* st_bytes = tile_state + n_tiles;

    for (;;)
    {
        __syncthreads();
        if (tid == 0)
        {
            S.tile = atomicAdd(tile_counter, 1u);
            S.nq = 0;
            S.ngq = 0;
        }
        __syncthreads();
        const uint32_t tile = S.tile;
        if (tile >= n_tiles) break;
        const uint32_t r0 = tile * VTN_TILE;
        const uint32_t r = r0 + tid;
        const bool live = r < B.n_rec;

        // ---------------- phase 1: headers, SNVs ----------------
        uint32_t ab = 0, n = 0, off0 = 0, len0 = 0, len1 = 0;
        bool queued = false;
        uint8_t my_status = 0, my_vtype = 0;
        if (live)
        {
            ab = __ldg(B.allele_beg + r);
            n = __ldg(B.allele_beg + r + 1) - ab;
            off0 = __ldg(B.allele_off + ab);
            const uint32_t off1 = __ldg(B.allele_off + ab + 1);
            len0 = off1 - off0;
            queued = true;
            if (n == 2)
            {
                len1 = __ldg(B.allele_off + ab + 2) - off1;
                if (len0 == 1 && len1 == 1)
                {
                    const uint8_t r0b = B.bytes[off0], c = B.bytes[off0 + 1];
                    if (c != '<' && c != '[' && c != ']' && c != '.' && c != r0b)
                    {
                        my_vtype = VTN_VT_SNP; // variant_manip.cpp:462-465 with mlen == diff == 1; never REF-checked
                        queued = false;
                    }
                }
            }
        }
        S.T[tid] = 0;
        S.le[tid] = 0;
        S.lt[tid] = 0;
        S.nbytes[tid] = 0;
        S.fast[tid] = 0;
        {
            const uint32_t slot = warp_push(&S.nq, queued, lane);
            if (queued) S.queue[slot] = (uint16_t)tid;
        }
        if (!queued)
        {
            S.status[tid] = my_status;
            S.vtype[tid] = my_vtype;
        }
        __syncthreads();

        // ---------------- phase 2a: queued records, fast path, dense lanes ----------------
        const uint32_t nq = S.nq;
        for (uint32_t q = tid; q < ((nq + 31u) & ~31u); q += VTN_TILE)
        {
            bool generic = false;
            uint32_t t = 0;
            if (q < nq)
            {
                t = S.queue[q];
                const uint32_t rr = r0 + t;
                const uint32_t ab2 = __ldg(B.allele_beg + rr);
                const uint32_t n2 = __ldg(B.allele_beg + rr + 1) - ab2;
                generic = true;
                if (n2 == 2)
                {
                    const uint32_t o0 = __ldg(B.allele_off + ab2), o1 = __ldg(B.allele_off + ab2 + 1), o2 = __ldg(B.allele_off + ab2 + 2);
                    VtnFastRes f;
                    if (fast_biallelic(R, B.bytes, o0, o1 - o0, o2 - o1, __ldg(B.pos1 + rr), __ldg(B.rid + rr), f))
                    {
                        generic = false;
                        S.status[t] = (uint8_t)f.status;
                        S.vtype[t] = (uint8_t)f.vtype;
                        S.T[t] = f.T;
                        S.le[t] = f.le;
                        S.lt[t] = f.lt;
                        S.nbytes[t] = f.nlr + f.nla;
                        S.newR[t] = f.newR;
                        S.newA[t] = f.newA;
                        S.nlr[t] = (uint8_t)f.nlr;
                        S.nla[t] = (uint8_t)f.nla;
                        S.fast[t] = 1;
                    }
                }
            }
            const uint32_t slot = warp_push(&S.ngq, generic, lane);
            if (generic) S.gqueue[slot] = (uint16_t)t;
        }
        __syncthreads();
        // ---------------- phase 2b: what the fast path refused, generic byte-wise path ----------------
        const uint32_t ngq = S.ngq;
        for (uint32_t q = tid; q < ngq; q += VTN_TILE)
        {
            const uint32_t t = S.gqueue[q];
            RecResult res;
            normalize_record(R, B, r0 + t, res);
            S.status[t] = res.status;
            S.vtype[t] = res.vtype;
            S.T[t] = res.T;
            S.le[t] = res.le;
            S.lt[t] = res.lt;
            S.nbytes[t] = res.nbytes;
        }
        __syncthreads();

        // ---------------- phase 3: per-record outputs, counters, ordered compaction ----------------
        my_status = S.status[tid];
        my_vtype = S.vtype[tid];
        if (live)
        {
            O.status[r] = my_status;
            O.vtype[r] = my_vtype;
        }
        const bool changed = live && (my_status & VTN_ST_CHANGED);
        {
            const uint32_t cat = changed ? (my_status >> VTN_ST_CAT_SHIFT) : 0;
            const uint32_t slot = cat ? ((n == 2 ? 0u : 5u) + cat - 1u) : 0xffu;
            unsigned any = __ballot_sync(0xffffffffu, cat != 0);
            while (any)
            {
                const int leader = __ffs(any) - 1;
                const uint32_t ls = __shfl_sync(0xffffffffu, slot, leader);
                const unsigned peers = __ballot_sync(0xffffffffu, slot == ls);
                if (lane == leader) atomicAdd(&S.counters[ls], __popc(peers));
                any &= ~peers;
            }
            const unsigned refs = __ballot_sync(0xffffffffu, live && my_vtype == VTN_VT_REF);
            const unsigned vars = __ballot_sync(0xffffffffu, live && my_vtype != VTN_VT_REF);
            if (lane == 0)
            {
                if (vars) atomicAdd(&S.counters[10], __popc(vars));
                if (refs) atomicAdd(&S.counters[11], __popc(refs));
            }
        }
        const uint32_t c = changed ? 1u : 0u, nb = changed ? S.nbytes[tid] : 0u;
        uint32_t ic = c, ib = nb;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1)
        {
            const uint32_t tc = __shfl_up_sync(0xffffffffu, ic, o);
            const uint32_t tb = __shfl_up_sync(0xffffffffu, ib, o);
            if (lane >= o) { ic += tc; ib += tb; }
        }
        if (lane == 31) { S.wcnt[warp] = ic; S.wbytes[warp] = ib; }
        __syncthreads();
        uint32_t wc = 0, wb = 0, tot_c = 0, tot_b = 0;
#pragma unroll
        for (int w = 0; w < VTN_TILE / 32; ++w)
        {
            const uint32_t xc = S.wcnt[w], xb = S.wbytes[w];
            if (w < warp) { wc += xc; wb += xb; }
            tot_c += xc;
            tot_b += xb;
        }
        if (changed)
        {
            S.cqueue[wc + ic - 1] = (uint16_t)tid; // rank inside the tile, input order
            S.ob[tid] = wb + ib - nb;
        }
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
                S.base[0] = pc;
                S.base[1] = pb;
                if (tile == n_tiles - 1)
                {
                    O.totals[0] = pc + tot_c;
                    O.totals[1] = pb + tot_b;
                    O.chg_off[pc + tot_c] = pb + tot_b;
                }
            }
        }
        __syncthreads();
        // ---------------- phase 3b: changed records, one per consecutive thread ----------------
        for (uint32_t i = tid; i < tot_c; i += VTN_TILE)
        {
            const uint32_t t = S.cqueue[i];
            const uint32_t rr = r0 + t;
            const unsigned long long k = S.base[0] + i;
            const unsigned long long ob = S.base[1] + S.ob[t];
            const int32_t pos1 = __ldg(B.pos1 + rr);
            const uint32_t T = S.T[t], le = S.le[t], lt = S.lt[t];
            O.chg_rec[k] = rr;
            O.chg_pos1[k] = pos1 - (int32_t)le + (int32_t)lt;
            O.chg_lt[k] = lt;
            O.chg_le[k] = le;
            O.chg_rt[k] = T;
            O.chg_off[k] = ob;
            uint8_t* dst = O.chg_bytes + ob;
            if (S.fast[t])
            {
                uint64_t a = S.newR[t];
                const uint32_t l0 = S.nlr[t], l1 = S.nla[t];
                for (uint32_t j = 0; j < l0; ++j, a >>= 2) dst[j] = (uint8_t)(0x47544341u >> (8 * ((uint32_t)a & 3u)));
                a = S.newA[t];
                dst += l0;
                for (uint32_t j = 0; j < l1; ++j, a >>= 2) dst[j] = (uint8_t)(0x47544341u >> (8 * ((uint32_t)a & 3u)));
            }
            else
            {
                // generic: (le reference bases + upper(allele))[0 : le + len - T], or upper(allele)[lt : len - T]
                const uint32_t ab2 = __ldg(B.allele_beg + rr);
                const uint32_t n2 = __ldg(B.allele_beg + rr + 1) - ab2;
                const int32_t rid = __ldg(B.rid + rr);
                const uint64_t gbase = __ldg(R.contig_goff + rid) << 5;
                const int64_t clen = __ldg(R.contig_len + rid);
                VtnRefCursor cur;
                uint32_t o = __ldg(B.allele_off + ab2);
                for (uint32_t ai = 0; ai < n2; ++ai)
                {
                    const uint32_t o2 = __ldg(B.allele_off + ab2 + ai + 1);
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
        }
    }
    __syncthreads();
    if (tid < 12 && S.counters[tid]) atomicAdd(O.counters + tid, S.counters[tid]);
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
