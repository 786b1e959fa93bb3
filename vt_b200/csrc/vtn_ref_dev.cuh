// vtn_ref_dev.cuh -- device-side accessors of the resident reference (layout in vtn_internal.h).
#pragma once
#include "vtn_internal.h"

__device__ __forceinline__ uint4 vtn_ldg_group(const VtnDevRef& R, uint64_t g) { return __ldg(R.groups + g); }

__device__ __forceinline__ uint64_t vtn_group_code(const uint4& q) { return ((uint64_t)q.y << 32) | q.x; }

// slow path: raw byte of an "exception" base (neither ACGT nor acgt) from the run-length table
static __device__ __noinline__ uint8_t vtn_exc_lookup(const VtnDevRef& R, uint64_t gpos)
{
    // last run with start <= gpos
    uint32_t lo = 0, hi = R.n_exc;
    while (lo < hi)
    {
        uint32_t mid = (lo + hi) >> 1;
        if (__ldg(R.exc_start + mid) <= gpos) lo = mid + 1;
        else hi = mid;
    }
    if (lo == 0) return 'N';
    uint32_t k = lo - 1;
    if (gpos - __ldg(R.exc_start + k) < __ldg(R.exc_len + k)) return __ldg(R.exc_byte + k);
    return 'N'; // unreachable for a consistent table
}

// A one-group cache: successive bases of one record almost always fall into the same 32-base group.
struct VtnRefCursor
{
    uint64_t g = ~0ull;
    uint4 q;
    __device__ __forceinline__ void seek(const VtnDevRef& R, uint64_t gpos)
    {
        uint64_t ng = gpos >> 5;
        if (ng != g)
        {
            g = ng;
            q = vtn_ldg_group(R, ng);
        }
    }
    // faidx_fetch_seq byte (raw case)
    __device__ __forceinline__ uint8_t raw(const VtnDevRef& R, uint64_t gpos)
    {
        seek(R, gpos);
        uint32_t j = (uint32_t)gpos & 31u;
        uint32_t c = (uint32_t)(vtn_group_code(q) >> (2 * j)) & 3u;
        uint8_t b = (uint8_t)((0x47544341u >> (8 * c)) & 0xff); // "ACTG": A=0 C=1 T=2 G=3 == (ascii >> 1) & 3
        if ((q.z >> j) & 1u)
        {
            if ((q.w >> j) & 1u) return b | 0x20;
            return vtn_exc_lookup(R, gpos);
        }
        return b;
    }
    // faidx_fetch_uc_seq byte: toupper() of every isgraph byte (hts_utils.cpp:41-51)
    __device__ __forceinline__ uint8_t uc(const VtnDevRef& R, uint64_t gpos)
    {
        seek(R, gpos);
        uint32_t j = (uint32_t)gpos & 31u;
        uint32_t c = (uint32_t)(vtn_group_code(q) >> (2 * j)) & 3u;
        uint8_t b = (uint8_t)((0x47544341u >> (8 * c)) & 0xff);
        if (((q.z & ~q.w) >> j) & 1u)
        {
            uint8_t e = vtn_exc_lookup(R, gpos);
            if (e >= 'a' && e <= 'z') e -= 0x20;
            return e;
        }
        return b;
    }
};

__device__ __forceinline__ uint8_t vtn_toupper(uint8_t c) { return (c >= 'a' && c <= 'z') ? (uint8_t)(c - 0x20) : c; }

// 32 upper-cased bases [gpos, gpos+32) as 2-bit codes (base gpos+j at bits 2j), plus the mask of bases in the
// window that are exceptions (need the run table).  Two LDG.128 at most.
__device__ __forceinline__ void vtn_ref_window32(const VtnDevRef& R, uint64_t gpos, uint64_t& code, uint32_t& exc)
{
    uint64_t g = gpos >> 5;
    uint32_t s = (uint32_t)gpos & 31u;
    uint4 a = vtn_ldg_group(R, g);
    uint64_t ca = vtn_group_code(a);
    uint32_t ea = a.z & ~a.w;
    if (s == 0)
    {
        code = ca;
        exc = ea;
        return;
    }
    uint4 b = vtn_ldg_group(R, g + 1);
    uint64_t cb = vtn_group_code(b);
    uint32_t eb = b.z & ~b.w;
    code = (ca >> (2 * s)) | (cb << (64 - 2 * s));
    exc = __funnelshift_r(ea, eb, s);
}
