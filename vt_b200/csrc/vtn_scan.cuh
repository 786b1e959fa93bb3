// vtn_scan.cuh -- single-pass ordered compaction across tiles (decoupled look-back), shared by the
// normalize and decompose kernels.
#pragma once
#include <stdint.h>

// ---- single-pass ordered compaction: tile descriptors for the decoupled look-back ----
// word = flag<<62 | value ; flag 0 = empty, 1 = tile aggregate, 2 = inclusive prefix.  Two words per tile
// (count of changed records, bytes of new alleles), each self-describing, so no cross-word ordering is needed.
#define VTN_FLAG_AGG (1ull << 62)
#define VTN_FLAG_INC (2ull << 62)
#define VTN_VAL_MASK ((1ull << 62) - 1)

__device__ __forceinline__ unsigned long long ld_state(const unsigned long long* p)
{
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_state(unsigned long long* p, unsigned long long v)
{
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// warp-parallel look-back over one descriptor array: returns exclusive prefix of tile `tile`
__device__ __forceinline__ unsigned long long lookback(const unsigned long long* state, int tile, int lane)
{
    unsigned long long prefix = 0;
    int base = tile - 1;
    while (base >= 0)
    {
        int j = base - lane;
        unsigned long long w = VTN_FLAG_INC; // virtual tile -1.. : inclusive prefix 0
        if (j >= 0)
        {
            do { w = ld_state(state + j); } while ((w >> 62) == 0);
        }
        unsigned inc_mask = __ballot_sync(0xffffffffu, (w >> 62) == 2);
        // lanes closer than (and including) the first inclusive descriptor contribute
        int first_inc = inc_mask ? (__ffs(inc_mask) - 1) : 32;
        unsigned long long v = (lane <= first_inc) ? (w & VTN_VAL_MASK) : 0ull;
        for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        prefix += v;
        if (inc_mask) break;
        base -= 32;
    }
    return prefix;
}

