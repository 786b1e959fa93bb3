// vtn_fast.cuh -- register-resident path for the common record shape: biallelic, both alleles <= 32 bases of
// upper-case ACGT, window inside the contig and free of N / IUPAC / soft-masked bases.  Alleles are converted
// to the reference's own 2-bit code ((ascii>>1)&3) so that classify_variant, the REF check, is_normalized,
// the common-suffix/prefix trims and the first 32 bases of left-shift are a handful of XOR / clz / ctz on
// 64-bit words -- no per-base loop.  Anything else returns false and goes to the generic path
// (normalize_record), which restates the same reference functions byte by byte.
#pragma once
#include "vtn_internal.h"
#include "vtn_ref_dev.cuh"

struct VtnFastRes
{
    uint32_t T, le, lt;   // right_trimmed, left_extended, left_trimmed
    uint32_t status;      // VTN_ST_* bits
    uint32_t vtype;       // VT_* bits
    uint64_t newR, newA;  // new alleles as 2-bit codes (valid when changed and both new lengths <= 32)
    uint32_t nlr, nla;    // new allele lengths
};

__device__ __forceinline__ uint64_t vtn_mask2(uint32_t nbases) { return nbases >= 32 ? ~0ull : ((1ull << (2 * nbases)) - 1ull); }

// 8 allele bytes starting at byte offset off (any alignment) of a 16-byte-padded array (global or staged in smem)
__device__ __forceinline__ uint64_t vtn_fetch8(const uint8_t* bytes, uint32_t off)
{
    const uint64_t* p = reinterpret_cast<const uint64_t*>(bytes + (off & ~7u));
    uint64_t w0 = *p;
    uint32_t s = (off & 7u) * 8u;
    if (s == 0) return w0;
    uint64_t w1 = p[1];
    return (w0 >> s) | (w1 << (64 - s));
}

// 8 ASCII bytes -> 16 bits of 2-bit codes; bad != 0 when one of the first nb bytes is not upper-case A/C/G/T
__device__ __forceinline__ uint32_t vtn_codes8(uint64_t w, uint32_t nb, uint64_t& bad)
{
    const uint64_t x = (w >> 1) & 0x0303030303030303ull;
    const uint64_t lo = x & 0x0101010101010101ull, hi = (x >> 1) & 0x0101010101010101ull;
    // expected ASCII of each code: A 0x41, C 0x43, T 0x54, G 0x47
    const uint64_t expect = 0x4141414141414141ull + (lo << 1) + ((hi & lo) << 2) + (hi & ~lo) * 0x13ull;
    const uint64_t bm = nb >= 8 ? ~0ull : ((1ull << (8 * nb)) - 1ull);
    bad = (expect ^ w) & bm;
    const uint32_t c0 = ((uint32_t)x * 0x01041040u) >> 24;
    const uint32_t c1 = ((uint32_t)(x >> 32) * 0x01041040u) >> 24;
    return c0 | (c1 << 8);
}

// number of equal leading (top) bases of two right-aligned code words
__device__ __forceinline__ uint32_t vtn_top_match(uint64_t a, uint64_t b)
{
    uint64_t x = a ^ b;
    uint64_t mm = (x | (x >> 1)) & 0x5555555555555555ull;
    return mm ? (uint32_t)(__clzll((long long)mm) >> 1) : 32u;
}

// 64 bases around 0-based contig position p0 (requires p0 >= 32): V = codes of [p0-32, p0), W = codes of
// [p0, p0+32); ev = exception bits (N/IUPAC) of V, nw = non-standard bits (anything but upper ACGT) of W.
__device__ __forceinline__ void vtn_ref_window64(const VtnDevRef& R, uint64_t gpos_lo, uint64_t& V, uint64_t& W, uint32_t& ev,
                                                 uint32_t& nw)
{
    const uint64_t g = gpos_lo >> 5;
    const uint32_t s = (uint32_t)gpos_lo & 31u;
    const uint4 a = vtn_ldg_group(R, g), b = vtn_ldg_group(R, g + 1);
    const uint64_t ca = vtn_group_code(a), cb = vtn_group_code(b);
    if (s == 0)
    {
        V = ca;
        W = cb;
        ev = a.z & ~a.w;
        nw = b.z;
        return;
    }
    const uint4 c = vtn_ldg_group(R, g + 2);
    const uint64_t cc = vtn_group_code(c);
    V = (ca >> (2 * s)) | (cb << (64 - 2 * s));
    W = (cb >> (2 * s)) | (cc << (64 - 2 * s));
    ev = __funnelshift_r(a.z & ~a.w, b.z & ~b.w, s);
    nw = __funnelshift_r(b.z, c.z, s);
}
