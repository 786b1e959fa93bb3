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

// strcmp(a->alleles.s, b->alleles.s) for two records of the same position group
__device__ int key_cmp(const UqView& B, uint32_t ra, uint32_t rb)
{
    if (ra == rb) return 0;
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

// one position group [b, b+m), m >= 2
__device__ __noinline__ uint32_t simulate_group(const UqView& B, uint32_t b, uint32_t m, uint32_t* h, uint32_t* order, uint8_t* first)
{
    int len = 0;
    const bool trailing = b + m < B.n_rec;
    if (b > 0)
    {
        h[0] = VTU_MINF;
        h[1] = b;
        len = 2;
        for (uint32_t i = 1; i < m; ++i)
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
        pop_heap(B, h, len);
        --len;
    }
    else
    {
        for (uint32_t i = 0; i < m; ++i)
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
        order[b + k] = r;
        first[b + k] = f ? 1 : 0;
        if (k + 1 < m)
        {
            pop_heap(B, h, len);
            --len;
        }
    }
    return uniq;
}

// ctl[0] = no_unique_variants, ctl[1] = first record that violates bcf_is_in_order (0xffffffff: none)
__global__ void __launch_bounds__(256) vtn_uniq_kernel(UqView B, uint32_t* __restrict__ order, uint8_t* __restrict__ first,
                                                       uint32_t* __restrict__ big_heap, uint32_t* __restrict__ ctl)
{
    const uint32_t n = B.n_rec;
    uint32_t uniq = 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    {
        const int32_t p = __ldg(B.pos1 + i), c = __ldg(B.rid + i);
        bool head = true;
        if (i > 0)
        {
            const int32_t pp = __ldg(B.pos1 + i - 1), pc = __ldg(B.rid + i - 1);
            head = (pp != p) || (pc != c);
            const bool in_order = pc == c ? pp <= p : pc < c; // bcf_is_in_order(prev, this), hts_utils.cpp:1042-1050
            if (!in_order) atomicMin(ctl + 1, i);
        }
        if (!head) continue;
        uint32_t m = 1;
        while (i + m < n && __ldg(B.pos1 + i + m) == p && __ldg(B.rid + i + m) == c) ++m;
        if (m == 1)
        {
            order[i] = i;
            first[i] = 1;
            ++uniq;
        }
        else if (m + 2 <= VTU_LOCAL)
        {
            uint32_t h[VTU_LOCAL];
            uniq += simulate_group(B, i, m, h, order, first);
        }
        else // groups of >= 39 records start >= 39 apart, so [i + 2*(i/39), +m+2) never overlaps the next big group's slice
            uniq += simulate_group(B, i, m, big_heap + i + 2 * (i / (VTU_LOCAL - 1)), order, first);
    }
    uniq = __reduce_add_sync(0xffffffffu, uniq);
    if ((threadIdx.x & 31) == 0 && uniq) atomicAdd(ctl, uniq);
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
        uint32_t grid = (d->n_rec + 255) / 256;
        const uint32_t cap = (uint32_t)ctx->sm_count * 8; // 2048 threads per SM resident; grid-stride beyond that
        if (grid > cap) grid = cap;
        vtn_uniq_kernel<<<grid, 256, 0, s>>>(v, u->order, u->first, u->big_heap, u->ctl);
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
