// vtn_uniq.cu -- `vt uniq` on the device: which record of every set of duplicates survives, and the order
// in which the records leave the reader.
//
// Replaces the record loop of Igor::uniq() (uniq.cpp:110-156) over BCFSyncedReader(SYNC_BY_VAR) with one input file read
// sequentially: read_next_position (bcf_synced_reader.cpp:388-433), fill_buffer (:594-637), CompareBCFPtr
// (bcf_synced_reader.h:86-122) on (rid, pos1, strcmp of bcf_alleles2string_sorted, hts_utils.cpp:1136-1167).
//
// The reference walks the file with a std::priority_queue.  Records that compare equal have no defined order in a heap, so
// WHICH duplicate becomes crecs[0] (the one uniq writes) is decided by libstdc++'s sift order -- it is not "the first in
// the file".  Derivation used here (the oracle restates the sequential reader literally instead):
//   * fill_buffer only reads when fewer than two records are buffered, i.e. exactly when the LAST record p of a position
//     group is being popped (it is still the root of the heap during the pushes; bcf_synced_reader.cpp:411-415 removes it
//     from the buffer, fills, then pops), and it then pushes every record of the next position group q1..qm plus the first
//     record r of the position after that (:626-629 breaks on the first record that differs).
//   * p is smaller than everything pushed and r larger, and the input must be sorted (bcf_is_in_order, hts_utils.cpp:
//     1042-1050, else exit(1)), so the heap the group's pops see is a function of the group alone:
//         heap = [-inf, q1]; push q2 .. qm; push +inf (iff a later record exists); pop            (groups after the first)
//         heap = [];         push q1 .. qm; push +inf (iff a later record exists)                 (first group of the file)
//     followed by m times (output top; pop) -- the last pop is the next group's "-inf".
// Position groups are therefore independent: one thread simulates one group with libstdc++'s __push_heap / __adjust_heap
// (bits/stl_heap.h).  Singletons (nearly all records of a normalized call set) need no heap and no allele bytes at all.
#include "vtn_internal.h"

namespace
{

#define VTU_MINF 0xfffffffeu
#define VTU_PINF 0xffffffffu
#define VTU_LOCAL 40 // heap entries kept in local memory (groups of up to VTU_LOCAL-2 records); larger groups use global scratch

struct UqView
{
    uint32_t n_rec;
    const int32_t* pos1;
    const int32_t* rid;
    const uint32_t* allele_beg;
    const uint32_t* allele_off;
    const uint8_t* bytes;
};

// strcmp of two alleles given as byte ranges (unsigned bytes, shorter is smaller on a common prefix)
__device__ int allele_cmp(const UqView& B, uint32_t sa, uint32_t sb)
{
    const uint32_t oa = __ldg(B.allele_off + sa), la = __ldg(B.allele_off + sa + 1) - oa;
    const uint32_t ob = __ldg(B.allele_off + sb), lb = __ldg(B.allele_off + sb + 1) - ob;
    const uint32_t l = la < lb ? la : lb;
    for (uint32_t i = 0; i < l; ++i)
    {
        const int d = (int)B.bytes[oa + i] - (int)B.bytes[ob + i];
        if (d) return d;
    }
    return la < lb ? -1 : (la > lb ? 1 : 0);
}

// slot of the k-th (k >= 1) ALT of a record in qsort(cmpstr) order (hts_utils.cpp:1156); equal strings are interchangeable
__device__ uint32_t sorted_alt_slot(const UqView& B, uint32_t ab, uint32_t n, uint32_t k)
{
    if (n <= 2) return ab + k;
    for (uint32_t a = 1; a < n; ++a)
    {
        uint32_t rank = 1;
        for (uint32_t b = 1; b < n; ++b)
        {
            if (b == a) continue;
            const int d = allele_cmp(B, ab + b, ab + a);
            rank += (d < 0 || (d == 0 && b < a)) ? 1u : 0u;
        }
        if (rank == k) return ab + a;
    }
    return ab + k; // unreachable
}

// byte stream of bcf_alleles2string_sorted: REF , ALT(1) , ALT(2) ... then 0
struct KeyIter
{
    uint32_t ab, n, k, off, end;
    __device__ void open(const UqView& B, uint32_t r)
    {
        ab = __ldg(B.allele_beg + r);
        n = __ldg(B.allele_beg + r + 1) - ab;
        k = 0;
        off = __ldg(B.allele_off + ab);
        end = __ldg(B.allele_off + ab + 1);
    }
    __device__ uint32_t next(const UqView& B)
    {
        if (off < end) return B.bytes[off++];
        if (++k >= n)
        {
            k = n;
            return 0;
        }
        const uint32_t s = sorted_alt_slot(B, ab, n, k);
        off = __ldg(B.allele_off + s);
        end = __ldg(B.allele_off + s + 1);
        return ',';
    }
};

// strcmp of "REF,ALT" for two biallelic records without building the strings: the first difference inside the REFs
// decides; when one REF is a proper prefix of the other its ',' meets the other's next base; equal REFs leave the ALTs
__device__ __forceinline__ int key_cmp_biallelic(const UqView& B, uint32_t sa, uint32_t sb)
{
    const uint32_t a0 = __ldg(B.allele_off + sa), a1 = __ldg(B.allele_off + sa + 1), a2 = __ldg(B.allele_off + sa + 2);
    const uint32_t b0 = __ldg(B.allele_off + sb), b1 = __ldg(B.allele_off + sb + 1), b2 = __ldg(B.allele_off + sb + 2);
    const uint32_t la = a1 - a0, lb = b1 - b0, l = la < lb ? la : lb;
    for (uint32_t i = 0; i < l; ++i)
    {
        const int d = (int)B.bytes[a0 + i] - (int)B.bytes[b0 + i];
        if (d) return d;
    }
    if (la != lb) return la < lb ? (int)',' - (int)B.bytes[b0 + l] : (int)B.bytes[a0 + l] - (int)',';
    const uint32_t ma = a2 - a1, mb = b2 - b1, m = ma < mb ? ma : mb;
    for (uint32_t i = 0; i < m; ++i)
    {
        const int d = (int)B.bytes[a1 + i] - (int)B.bytes[b1 + i];
        if (d) return d;
    }
    return ma < mb ? -1 : (ma > mb ? 1 : 0);
}

// strcmp(a->alleles.s, b->alleles.s) for two records of the same position group
__device__ int key_cmp(const UqView& B, uint32_t ra, uint32_t rb)
{
    if (ra == rb) return 0;
    {
        const uint32_t sa = __ldg(B.allele_beg + ra), sb = __ldg(B.allele_beg + rb);
        if (__ldg(B.allele_beg + ra + 1) - sa == 2 && __ldg(B.allele_beg + rb + 1) - sb == 2) return key_cmp_biallelic(B, sa, sb);
    }
    KeyIter a, b;
    a.open(B, ra);
    b.open(B, rb);
    for (;;)
    {
        const uint32_t ca = a.next(B), cb = b.next(B);
        if (ca != cb) return (int)ca - (int)cb;
        if (ca == 0) return 0;
    }
}

// CompareBCFPtr::operator()(a, b) (bcf_synced_reader.h:89-121) within one position group plus the two sentinels: a > b
__device__ bool heap_comp(const UqView& B, uint32_t a, uint32_t b)
{
    if (a == VTU_MINF || b == VTU_PINF) return false;
    if (a == VTU_PINF) return b != VTU_PINF;
    if (b == VTU_MINF) return true;
    return key_cmp(B, a, b) > 0;
}

// libstdc++ bits/stl_heap.h __push_heap
__device__ void push_heap(const UqView& B, uint32_t* h, int hole, int top, uint32_t value)
{
    int parent = (hole - 1) / 2;
    while (hole > top && heap_comp(B, h[parent], value))
    {
        h[hole] = h[parent];
        hole = parent;
        parent = (hole - 1) / 2;
    }
    h[hole] = value;
}

// std::pop_heap + pop_back on a heap of len entries (-> len-1 entries)
__device__ void pop_heap(const UqView& B, uint32_t* h, int len)
{
    if (len > 1)
    {
        const uint32_t value = h[len - 1];
        h[len - 1] = h[0];
        const int n = len - 1; // __adjust_heap(first, 0, n, value)
        int hole = 0, child = 0;
        while (child < (n - 1) / 2)
        {
            child = 2 * (child + 1);
            if (heap_comp(B, h[child], h[child - 1])) child--;
            h[hole] = h[child];
            hole = child;
        }
        if ((n & 1) == 0 && child == (n - 2) / 2)
        {
            child = 2 * (child + 1);
            h[hole] = h[child - 1];
            hole = child - 1;
        }
        push_heap(B, h, hole, 0, value);
    }
}

// where a group's results go: records inside the block's tile are staged in shared memory (written out coalesced by
// phase 3), records of a group that runs past the tile's end go to global memory directly (the next block skips them)
struct UqOut
{
    uint32_t* s_order;
    uint8_t* s_first;
    uint32_t base, end; // tile = records [base, end)
    uint32_t* g_order;
    uint8_t* g_first;
    __device__ __forceinline__ void put(uint32_t idx, uint32_t r, bool f) const
    {
        if (idx < end)
        {
            s_order[idx - base] = r;
            s_first[idx - base] = f ? 1 : 0;
        }
        else
        {
            g_order[idx] = r;
            g_first[idx] = f ? 1 : 0;
        }
    }
};

// one position group [b, b+m), m >= 2
__device__ __noinline__ uint32_t simulate_group(const UqView& B, uint32_t b, uint32_t m, uint32_t* h, const UqOut& out)
{
    int len = 0;
    const bool trailing = b + m < B.n_rec;
    if (b > 0)
    {
        h[0] = VTU_MINF;
        h[1] = b;
        len = 2;
    }
    for (uint32_t i = (b > 0 ? 1 : 0); i < m; ++i)
    {
        h[len] = b + i;
        ++len;
        push_heap(B, h, len - 1, 0, h[len - 1]);
    }
    if (trailing)
    {
        h[len] = VTU_PINF;
        ++len;
        push_heap(B, h, len - 1, 0, VTU_PINF);
    }
    if (b > 0)
    {
        pop_heap(B, h, len);
        --len;
    }
    uint32_t uniq = 0, run_head = 0;
    for (uint32_t k = 0; k < m; ++k)
    {
        const uint32_t r = h[0];
        // bcfptr_cmp(cvariant, variant) == 0 continues the current read_next_position call (bcf_synced_reader.cpp:407)
        const bool f = (k == 0) || key_cmp(B, r, run_head) != 0;
        if (f)
        {
            run_head = r;
            ++uniq;
        }
        out.put(b + k, r, f);
        if (k + 1 < m)
        {
            pop_heap(B, h, len);
            --len;
        }
    }
    return uniq;
}

// A pair of records (the usual duplicate) between other records: heap = [-inf, q1, q2, +inf]; the pop moves the smaller
// child to the root, q2 when they tie (__adjust_heap takes the right child unless it compares greater), so
//   key(q2) > key(q1) -> q1, q2      else -> q2, q1.      (first / last group of the file go through simulate_group)
__device__ __forceinline__ uint32_t pair_group(const UqView& B, uint32_t b, const UqOut& out)
{
    const uint32_t ab = __ldg(B.allele_beg + b), a1 = __ldg(B.allele_beg + b + 1), a2 = __ldg(B.allele_beg + b + 2);
    int d; // strcmp(key(q2), key(q1))
    const uint32_t o0 = __ldg(B.allele_off + ab);
    if (a1 - ab == 2 && a2 - a1 == 2 && __ldg(B.allele_off + a2) - o0 == 4)
    { // four single-base alleles, contiguous: REF1 ALT1 REF2 ALT2
        const uint32_t r1 = B.bytes[o0], c1 = B.bytes[o0 + 1], r2 = B.bytes[o0 + 2], c2 = B.bytes[o0 + 3];
        d = r2 != r1 ? (int)r2 - (int)r1 : (int)c2 - (int)c1;
    }
    else
        d = key_cmp(B, b + 1, b);
    const bool q1_first = d > 0;
    out.put(b, q1_first ? b : b + 1, true);
    out.put(b + 1, q1_first ? b + 1 : b, d != 0);
    return d != 0 ? 2u : 1u;
}

// Every WARP owns tiles of VTU_WTILE = 256 consecutive records (8 per lane) and never waits for another warp:
//   phase 1  128-bit loads of pos / rid; neighbours by shuffle (one extra scalar load at each end of the tile); order check;
//            every record gets the singleton default (order = itself, first = 1) in the warp's shared tile; leaders of position
//            groups of >= 2 records are queued (warp scan, tile order).
//   phase 2  the queue is processed densely (lane l <- queue entry 32p+l): pair fast path or heap simulation.
//   phase 3  the tile is written out with 16-byte / 8-byte stores.  Records at the head of the tile that continue a group
//            led from an earlier tile are skipped (their leader's warp writes them).
// ctl[0] = no_unique_variants, ctl[1] = first record that violates bcf_is_in_order (0xffffffff: none)
#define VTU_WARPS 8
#define VTU_THREADS (32 * VTU_WARPS)
#define VTU_RPL 8
#define VTU_WTILE (32 * VTU_RPL)
struct __align__(16) UqWarpSmem
{
    uint32_t order[VTU_WTILE];
    uint8_t first[VTU_WTILE];
    uint32_t queue[VTU_WTILE / 2];
};

__global__ void __launch_bounds__(VTU_THREADS) vtn_uniq_kernel(UqView B, uint32_t* __restrict__ order, uint8_t* __restrict__ first,
                                                               uint32_t* __restrict__ big_heap, uint32_t* __restrict__ ctl, uint32_t n_tiles)
{
    __shared__ UqWarpSmem SM[VTU_WARPS];
    UqWarpSmem& S = SM[threadIdx.x >> 5];
    const uint32_t n = B.n_rec;
    const int lane = threadIdx.x & 31;
    uint32_t uniq = 0;
    for (uint32_t tile = blockIdx.x * VTU_WARPS + (threadIdx.x >> 5); tile < n_tiles; tile += gridDim.x * VTU_WARPS)
    {
        const uint32_t base = tile * VTU_WTILE;
        const uint32_t end = min(base + VTU_WTILE, n);
        // ---------------- phase 1 ----------------
        const uint32_t i0 = base + VTU_RPL * lane;
        int p[VTU_RPL + 2], c[VTU_RPL + 2]; // [0] = record i0-1, [1..8] = this lane's records, [9] = record i0+8
        {
            int4 pa = make_int4(0, 0, 0, 0), pb = pa, ca = pa, cb = pa;
            if (i0 < end)
            { // the arrays are padded (dalloc), so the last vectors of the batch may read a few entries past n_rec
                pa = __ldg(reinterpret_cast<const int4*>(B.pos1 + i0));
                pb = __ldg(reinterpret_cast<const int4*>(B.pos1 + i0) + 1);
                ca = __ldg(reinterpret_cast<const int4*>(B.rid + i0));
                cb = __ldg(reinterpret_cast<const int4*>(B.rid + i0) + 1);
            }
            p[1] = pa.x; p[2] = pa.y; p[3] = pa.z; p[4] = pa.w; p[5] = pb.x; p[6] = pb.y; p[7] = pb.z; p[8] = pb.w;
            c[1] = ca.x; c[2] = ca.y; c[3] = ca.z; c[4] = ca.w; c[5] = cb.x; c[6] = cb.y; c[7] = cb.z; c[8] = cb.w;
            p[0] = __shfl_up_sync(0xffffffffu, pb.w, 1);
            c[0] = __shfl_up_sync(0xffffffffu, cb.w, 1);
            p[9] = __shfl_down_sync(0xffffffffu, pa.x, 1);
            c[9] = __shfl_down_sync(0xffffffffu, ca.x, 1);
            if (lane == 0 && i0 > 0)
            {
                p[0] = __ldg(B.pos1 + i0 - 1);
                c[0] = __ldg(B.rid + i0 - 1);
            }
            if (lane == 31 && i0 + VTU_RPL < n)
            {
                p[9] = __ldg(B.pos1 + i0 + VTU_RPL);
                c[9] = __ldg(B.rid + i0 + VTU_RPL);
            }
        }
        uint32_t first_head = VTU_WTILE, lmask = 0;
#pragma unroll
        for (int j = 0; j < VTU_RPL; ++j)
        {
            const uint32_t i = i0 + j;
            if (i < end)
            {
                bool head = true;
                if (i > 0)
                {
                    head = p[j] != p[j + 1] || c[j] != c[j + 1];
                    const bool in_order = c[j] == c[j + 1] ? p[j] <= p[j + 1] : c[j] < c[j + 1]; // bcf_is_in_order(prev, this)
                    if (!in_order) atomicMin(ctl + 1, i);
                }
                const bool next_same = i + 1 < n && p[j + 2] == p[j + 1] && c[j + 2] == c[j + 1];
                if (head)
                {
                    if (first_head == VTU_WTILE) first_head = VTU_RPL * lane + j;
                    if (next_same) lmask |= 1u << j;
                    else ++uniq;
                }
            }
        }
        reinterpret_cast<uint4*>(S.order)[2 * lane] = make_uint4(i0, i0 + 1, i0 + 2, i0 + 3);
        reinterpret_cast<uint4*>(S.order)[2 * lane + 1] = make_uint4(i0 + 4, i0 + 5, i0 + 6, i0 + 7);
        reinterpret_cast<uint2*>(S.first)[lane] = make_uint2(0x01010101u, 0x01010101u);
        uint32_t qn;
        {
            uint32_t inc = __popc(lmask);
#pragma unroll
            for (int o = 1; o < 32; o <<= 1)
            {
                const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
                if (lane >= o) inc += t;
            }
            qn = __shfl_sync(0xffffffffu, inc, 31);
            uint32_t slot = inc - __popc(lmask);
#pragma unroll
            for (int j = 0; j < VTU_RPL; ++j)
                if (lmask & (1u << j)) S.queue[slot++] = i0 + j;
        }
        const uint32_t owned = __reduce_min_sync(0xffffffffu, first_head);
        __syncwarp();
        // ---------------- phase 2 ----------------
        const UqOut out{S.order, S.first, base, end, order, first};
        for (uint32_t q = lane; q < qn; q += 32)
        {
            const uint32_t i = S.queue[q];
            const int32_t pp = __ldg(B.pos1 + i), cc = __ldg(B.rid + i);
            uint32_t m = 2;
            while (i + m < n && __ldg(B.pos1 + i + m) == pp && __ldg(B.rid + i + m) == cc) ++m;
            if (m == 2 && i > 0 && i + 2 < n)
                uniq += pair_group(B, i, out);
            else if (m + 2 <= VTU_LOCAL)
            {
                uint32_t h[VTU_LOCAL];
                uniq += simulate_group(B, i, m, h, out);
            }
            else // groups of >= 39 records start >= 39 apart, so [i + 2*(i/39), +m+2) never overlaps the next big group's slice
                uniq += simulate_group(B, i, m, big_heap + i + 2 * (i / (VTU_LOCAL - 1)), out);
        }
        __syncwarp();
        // ---------------- phase 3 ----------------
        if ((uint32_t)(VTU_RPL * lane) >= owned && i0 + VTU_RPL <= end)
        {
            reinterpret_cast<uint4*>(order + i0)[0] = reinterpret_cast<const uint4*>(S.order)[2 * lane];
            reinterpret_cast<uint4*>(order + i0)[1] = reinterpret_cast<const uint4*>(S.order)[2 * lane + 1];
            *reinterpret_cast<uint2*>(first + i0) = reinterpret_cast<const uint2*>(S.first)[lane];
        }
        else
        {
#pragma unroll
            for (int j = 0; j < VTU_RPL; ++j)
                if ((uint32_t)(VTU_RPL * lane + j) >= owned && i0 + j < end)
                {
                    order[i0 + j] = S.order[VTU_RPL * lane + j];
                    first[i0 + j] = S.first[VTU_RPL * lane + j];
                }
        }
        __syncwarp(); // the next tile reuses this warp's shared arrays
    }
    uniq = __reduce_add_sync(0xffffffffu, uniq);
    if (lane == 0 && uniq) atomicAdd(ctl, uniq);
}

} // namespace

struct vtn_uniq_buffers
{
    uint32_t cap_rec = 0;
    uint32_t* order = nullptr;
    uint8_t* first = nullptr;
    uint32_t* big_heap = nullptr;
    uint32_t* ctl = nullptr;
};

static void uniq_release(vtn_uniq_buffers* u)
{
    if (!u) return;
    cudaFree(u->order);
    cudaFree(u->first);
    cudaFree(u->big_heap);
    cudaFree(u->ctl);
    delete u;
}

void vtn_uniq_dbatch_release(vtn_dbatch* d)
{
    uniq_release(d->uniq);
    d->uniq = nullptr;
}

int vtn_uniq_resident(vtn_ctx* ctx, vtn_dbatch* d)
{
    if (!ctx || !d) return VTN_E_ARG;
    VTN_CUDA(ctx, cudaSetDevice(ctx->device));
    if (!d->uniq || d->uniq->cap_rec < d->n_rec)
    {
        uniq_release(d->uniq);
        vtn_uniq_buffers* u = d->uniq = new vtn_uniq_buffers();
        const size_t n = (size_t)d->cap_rec + 64;
        u->cap_rec = d->cap_rec;
        VTN_CUDA(ctx, cudaMalloc((void**)&u->order, n * sizeof(uint32_t)));
        VTN_CUDA(ctx, cudaMalloc((void**)&u->first, n));
        VTN_CUDA(ctx, cudaMalloc((void**)&u->big_heap, (n + 2 * (n / (VTU_LOCAL - 1)) + 8) * sizeof(uint32_t)));
        VTN_CUDA(ctx, cudaMalloc((void**)&u->ctl, 2 * sizeof(uint32_t)));
    }
    vtn_uniq_buffers* u = d->uniq;
    cudaStream_t s = ctx->stream;
    VTN_CUDA(ctx, cudaMemsetAsync(u->ctl, 0, sizeof(uint32_t), s));
    VTN_CUDA(ctx, cudaMemsetAsync(u->ctl + 1, 0xff, sizeof(uint32_t), s));
    if (d->n_rec)
    {
        UqView v{d->n_rec, d->pos1, d->rid, d->allele_beg, d->allele_off, d->bytes};
        const uint32_t n_tiles = (d->n_rec + VTU_WTILE - 1) / VTU_WTILE;
        uint32_t grid = (n_tiles + VTU_WARPS - 1) / VTU_WARPS;
        const uint32_t cap = (uint32_t)ctx->sm_count * 8; // up to 8 blocks of 8 warps resident per SM; tile-stride beyond that
        if (grid > cap) grid = cap;
        vtn_uniq_kernel<<<grid, VTU_THREADS, 0, s>>>(v, u->order, u->first, u->big_heap, u->ctl, n_tiles);
        VTN_CUDA(ctx, cudaGetLastError());
        ctx->launches += 1;
    }
    return VTN_OK;
}

int vtn_dbatch_fetch_uniq(vtn_ctx* ctx, vtn_dbatch* d, uint32_t* order, uint8_t* first, vtn_uniq_counters* counters, int64_t* not_in_order)
{
    if (!ctx || !d || !d->uniq) return VTN_E_ARG;
    cudaStream_t s = ctx->stream;
    uint32_t ctl[2] = {0, 0xffffffffu};
    VTN_CUDA(ctx, cudaMemcpyAsync(ctl, d->uniq->ctl, sizeof ctl, cudaMemcpyDeviceToHost, s));
    if (d->n_rec && order) VTN_CUDA(ctx, cudaMemcpyAsync(order, d->uniq->order, sizeof(uint32_t) * d->n_rec, cudaMemcpyDeviceToHost, s));
    if (d->n_rec && first) VTN_CUDA(ctx, cudaMemcpyAsync(first, d->uniq->first, d->n_rec, cudaMemcpyDeviceToHost, s));
    VTN_CUDA(ctx, cudaStreamSynchronize(s));
    if (counters)
    {
        counters->no_total_variants = d->n_rec;
        counters->no_unique_variants = ctl[0];
    }
    if (not_in_order) *not_in_order = ctl[1] == 0xffffffffu ? -1 : (int64_t)ctl[1];
    return VTN_OK;
}
