// vtn_order.cu -- the output order of BCFOrderedWriter (bcf_ordered_writer.cpp:141-210 write, :252-300 flush) computed on the
// device for a whole batch at once: vt normalize re-sorts its output through that writer with a window of
// 10000 bp (normalize.cpp:101,139), and the order it produces is NOT a plain sort once a record lags more than the window.
//
// The writer keeps a list sorted by position (front = largest; a newcomer goes in front of the first element whose position
// is <= its own, so equal positions keep arrival order) and, after every insertion, writes from the back while
// back.pos <= max(front.pos - window, 1) and more than one record is buffered; a contig change writes everything.
// Closed form used here (the oracle emulates the list literally):
//   segment  = maximal run of records with the same rid (a rid change flushes, :185-189)
//   t        = arrival index inside the segment, M_s = max position of arrivals 0..s (the front after step s)
//   a record v leaves at step  e_v = min{ s >= max(t_v,1) : pos_v <= max(M_s - window, 1)  and  v is not the front },
//   or with the final flush of its segment when no such step exists; inside one step records leave by (pos, arrival).
//   For pos_v >= 2 the condition is M_s >= pos_v + window (M is monotone: galloping + binary search); the front can only
//   satisfy the cutoff when every position so far is <= 1, handled by a short forward walk (contig starts only).
//   The warning of :163-179 fires iff t_v >= 1 and pos_v < max(M_{t_v - 1} - window, 1) (and pos_v < M_{t_v - 1}).
// Output position = (number of records that leave at earlier steps) + rank by (pos, arrival) among those leaving at the same
// step: a histogram over steps, one exclusive scan, a scatter and a rank-by-counting inside each step's (short) list.
// Kernels: two reduce-then-scan passes (segmented running maximum + segment id/start; step histogram), all streaming.
#include <limits.h>
#include "vtn_internal.h"

namespace
{

#define VTO_THREADS 256
#define VTO_ITEMS 8
#define VTO_TILE (VTO_THREADS * VTO_ITEMS)

// ---- scan A element: (segments started, start index of the current segment, running maximum inside the segment) ----
struct SegMax
{
    uint32_t seg;   // number of segment starts in the range
    uint32_t start; // largest index at which a segment starts (0 if none)
    int32_t m;      // maximum position since the last segment start in the range
};
__device__ __forceinline__ SegMax seg_id() { return SegMax{0u, 0u, INT_MIN}; }
__device__ __forceinline__ SegMax seg_op(const SegMax& a, const SegMax& b)
{
    SegMax r;
    r.seg = a.seg + b.seg;
    r.start = a.start > b.start ? a.start : b.start;
    r.m = b.seg ? b.m : (a.m > b.m ? a.m : b.m);
    return r;
}
__device__ __forceinline__ SegMax seg_shfl_up(const SegMax& v, int o)
{
    SegMax r;
    r.seg = __shfl_up_sync(0xffffffffu, v.seg, o);
    r.start = __shfl_up_sync(0xffffffffu, v.start, o);
    r.m = __shfl_up_sync(0xffffffffu, v.m, o);
    return r;
}
__device__ __forceinline__ SegMax seg_load(const int32_t* rid, const int32_t* pos, uint32_t i)
{
    const bool flag = i == 0 || __ldg(rid + i) != __ldg(rid + i - 1);
    return SegMax{flag ? 1u : 0u, flag ? i : 0u, __ldg(pos + i)};
}
// the VTO_ITEMS consecutive elements of one thread, by 128-bit loads when the whole group is inside the batch
__device__ __forceinline__ void seg_load8(const int32_t* rid, const int32_t* pos, uint32_t base, uint32_t n, SegMax* e)
{
    if (base + VTO_ITEMS <= n)
    {
        const int4 r0 = __ldg(reinterpret_cast<const int4*>(rid + base)), r1 = __ldg(reinterpret_cast<const int4*>(rid + base) + 1);
        const int4 p0 = __ldg(reinterpret_cast<const int4*>(pos + base)), p1 = __ldg(reinterpret_cast<const int4*>(pos + base) + 1);
        const int32_t rr[9] = {base ? __ldg(rid + base - 1) : ~r0.x, r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
        const int32_t pp[8] = {p0.x, p0.y, p0.z, p0.w, p1.x, p1.y, p1.z, p1.w};
#pragma unroll
        for (int j = 0; j < VTO_ITEMS; ++j)
        {
            const bool flag = rr[j] != rr[j + 1];
            e[j] = SegMax{flag ? 1u : 0u, flag ? base + j : 0u, pp[j]};
        }
    }
    else
    {
#pragma unroll
        for (int j = 0; j < VTO_ITEMS; ++j) e[j] = base + j < n ? seg_load(rid, pos, base + j) : seg_id();
    }
}

// block-wide inclusive scan of one SegMax per thread (thread order); returns the exclusive prefix of the thread and the block total
__device__ __forceinline__ SegMax seg_block_excl(const SegMax& v, SegMax* s_warp, SegMax& total)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    SegMax inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1)
    {
        const SegMax t = seg_shfl_up(inc, o);
        if (lane >= o) inc = seg_op(t, inc);
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    SegMax wpre = seg_id(), tot = seg_id();
    for (int w = 0; w < VTO_THREADS / 32; ++w)
    {
        if (w < warp) wpre = seg_op(wpre, s_warp[w]);
        tot = seg_op(tot, s_warp[w]);
    }
    total = tot;
    SegMax up = seg_shfl_up(inc, 1);
    if (lane == 0) up = seg_id();
    __syncthreads();
    return seg_op(wpre, up);
}

// pass 1: aggregate of every tile
__global__ void __launch_bounds__(VTO_THREADS) vto_seg_reduce_kernel(const int32_t* rid, const int32_t* pos, uint32_t n, SegMax* agg)
{
    __shared__ SegMax s_warp[VTO_THREADS / 32];
    const uint32_t base = blockIdx.x * VTO_TILE + threadIdx.x * VTO_ITEMS;
    SegMax e[VTO_ITEMS];
    seg_load8(rid, pos, base, n, e);
    SegMax v = seg_id();
#pragma unroll
    for (int j = 0; j < VTO_ITEMS; ++j) v = seg_op(v, e[j]);
    SegMax total;
    seg_block_excl(v, s_warp, total);
    if (threadIdx.x == 0) agg[blockIdx.x] = total;
}

// pass 2: exclusive scan of the tile aggregates in place (one block)
__global__ void __launch_bounds__(1024) vto_seg_scan_kernel(SegMax* agg, uint32_t n_tiles)
{
    __shared__ SegMax s[1024];
    const uint32_t per = (n_tiles + 1023) / 1024;
    const uint32_t lo = threadIdx.x * per, hi = min(lo + per, n_tiles);
    SegMax mine = seg_id();
    for (uint32_t t = lo; t < hi; ++t) mine = seg_op(mine, agg[t]);
    s[threadIdx.x] = mine;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1)
    {
        SegMax x = seg_id();
        if ((int)threadIdx.x >= o) x = s[threadIdx.x - o];
        __syncthreads();
        if ((int)threadIdx.x >= o) s[threadIdx.x] = seg_op(x, s[threadIdx.x]);
        __syncthreads();
    }
    SegMax pre = threadIdx.x ? s[threadIdx.x - 1] : seg_id();
    for (uint32_t t = lo; t < hi; ++t)
    {
        const SegMax x = agg[t];
        agg[t] = pre;
        pre = seg_op(pre, x);
    }
}

// pass 3: rescan every tile with its carry; K[i] = segment id << 32 | biased running maximum (globally non-decreasing;
// a record is the first of its segment iff its segment id differs from its predecessor's)
__device__ __forceinline__ uint32_t bias(int32_t x) { return (uint32_t)x ^ 0x80000000u; }
__global__ void __launch_bounds__(VTO_THREADS) vto_seg_apply_kernel(const int32_t* rid, const int32_t* pos, uint32_t n, const SegMax* agg,
                                                                    unsigned long long* K)
{
    __shared__ SegMax s_warp[VTO_THREADS / 32];
    const uint32_t base = blockIdx.x * VTO_TILE + threadIdx.x * VTO_ITEMS;
    SegMax e[VTO_ITEMS];
    seg_load8(rid, pos, base, n, e);
    SegMax v = seg_id();
#pragma unroll
    for (int j = 0; j < VTO_ITEMS; ++j) v = seg_op(v, e[j]);
    SegMax total;
    SegMax pre = seg_op(agg[blockIdx.x], seg_block_excl(v, s_warp, total));
    unsigned long long k[VTO_ITEMS];
#pragma unroll
    for (int j = 0; j < VTO_ITEMS; ++j)
    {
        pre = seg_op(pre, e[j]);
        k[j] = ((unsigned long long)pre.seg << 32) | bias(pre.m);
    }
    if (base + VTO_ITEMS <= n)
    {
#pragma unroll
        for (int j = 0; j < VTO_ITEMS; j += 2) *reinterpret_cast<ulonglong2*>(K + base + j) = make_ulonglong2(k[j], k[j + 1]);
    }
    else
    {
#pragma unroll
        for (int j = 0; j < VTO_ITEMS; ++j)
            if (base + j < n) K[base + j] = k[j];
    }
}

// per record: the step at which it leaves -> slot (step s of a segment starting at b uses slot b+s-1, the final flush the
// segment's last slot), histogram of the slots, and the lateness warning
__global__ void __launch_bounds__(VTO_THREADS) vto_step_kernel(const int32_t* __restrict__ pos, uint32_t n, int32_t window,
                                                               const unsigned long long* __restrict__ K,
                                                               uint32_t* __restrict__ slot, uint32_t* __restrict__ cnt, int32_t* __restrict__ late_front)
{
    const uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= n) return;
    const int32_t p = __ldg(pos + v);
    const unsigned long long kv = K[v];
    const uint32_t seg = (uint32_t)(kv >> 32);
    const unsigned long long kprev = v ? K[v - 1] : 0ull; // segment ids start at 1
    const bool is_first = (uint32_t)(kprev >> 32) != seg; // first record of its segment (t_v = 0)
    const int32_t m_incl = (int32_t)((uint32_t)kv ^ 0x80000000u);
    // warning (:163-179): the buffer is not empty, v is smaller than everything in it and below the cutoff of the current front
    int32_t lf = INT_MIN;
    if (!is_first)
    {
        const int32_t m_prev = (int32_t)((uint32_t)kprev ^ 0x80000000u);
        long long cutoff = (long long)m_prev - window;
        if (cutoff < 1) cutoff = 1;
        if (p < cutoff && p < m_prev) lf = m_prev; // p < m_prev matters only while every position so far is <= 1 (the kept-back front)
    }
    late_front[v] = lf;
    uint32_t e; // absolute index of the step at which v leaves, or the index one past its segment (final flush)
    const unsigned long long seg_end_key = (unsigned long long)(seg + 1) << 32;
    if (p >= 2)
    {
        const long long want = (long long)p + window; // M_s >= pos + window
        const unsigned long long target = want > INT_MAX ? seg_end_key : (((unsigned long long)seg << 32) | bias((int32_t)want));
        uint32_t lo = is_first ? v + 1 : v; // max(t_v, 1) in absolute terms
        uint32_t hi = lo, step = 1;
        while (hi < n && K[hi] < target)
        { // gallop
            lo = hi + 1;
            hi = hi + step > hi ? hi + step : n;
            step <<= 1;
        }
        if (hi > n) hi = n;
        while (lo < hi)
        {
            const uint32_t mid = lo + (hi - lo) / 2;
            if (K[mid] < target) lo = mid + 1;
            else hi = mid;
        }
        e = lo;
    }
    else if (m_incl > p)
        e = v; // something larger arrived earlier: v is not the front and is below the cutoff (>= 1) at its own step
    else
    { // v is the front when it arrives: it stays until a record with pos >= its own goes in front of it
        e = v + 1;
        while (e < n && (uint32_t)(K[e] >> 32) == seg && __ldg(pos + e) < p) ++e;
    }
    // e inside the segment: step e - b uses slot e - 1; otherwise e is exactly the segment's end (the first index of the next
    // segment satisfies every search target of this one) and e - 1 is the segment's last slot = the final flush
    const uint32_t sl = e - 1;
    slot[v] = sl;
    atomicAdd(cnt + sl, 1u);
}

// ---- scan B: exclusive prefix sum of the step histogram (same three-pass shape) ----
__device__ __forceinline__ uint32_t sum_block_excl(uint32_t v, uint32_t* s_warp, uint32_t& total)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1)
    {
        const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    uint32_t wpre = 0, tot = 0;
    for (int w = 0; w < VTO_THREADS / 32; ++w)
    {
        if (w < warp) wpre += s_warp[w];
        tot += s_warp[w];
    }
    total = tot;
    __syncthreads();
    return wpre + inc - v;
}

__global__ void __launch_bounds__(VTO_THREADS) vto_sum_reduce_kernel(const uint32_t* cnt, uint32_t n, uint32_t* agg)
{
    __shared__ uint32_t s_warp[VTO_THREADS / 32];
    const uint32_t base = blockIdx.x * VTO_TILE + threadIdx.x * VTO_ITEMS;
    uint32_t v = 0;
    if (base + VTO_ITEMS <= n)
    {
        const uint4 a = __ldg(reinterpret_cast<const uint4*>(cnt + base)), b = __ldg(reinterpret_cast<const uint4*>(cnt + base) + 1);
        v = a.x + a.y + a.z + a.w + b.x + b.y + b.z + b.w;
    }
    else
        for (int j = 0; j < VTO_ITEMS; ++j)
            if (base + j < n) v += cnt[base + j];
    uint32_t total;
    sum_block_excl(v, s_warp, total);
    if (threadIdx.x == 0) agg[blockIdx.x] = total;
}

__global__ void __launch_bounds__(1024) vto_sum_scan_kernel(uint32_t* agg, uint32_t n_tiles)
{
    __shared__ uint32_t s[1024];
    const uint32_t per = (n_tiles + 1023) / 1024;
    const uint32_t lo = threadIdx.x * per, hi = min(lo + per, n_tiles);
    uint32_t mine = 0;
    for (uint32_t t = lo; t < hi; ++t) mine += agg[t];
    s[threadIdx.x] = mine;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1)
    {
        uint32_t x = 0;
        if ((int)threadIdx.x >= o) x = s[threadIdx.x - o];
        __syncthreads();
        s[threadIdx.x] += x;
        __syncthreads();
    }
    uint32_t pre = s[threadIdx.x] - mine;
    for (uint32_t t = lo; t < hi; ++t)
    {
        const uint32_t x = agg[t];
        agg[t] = pre;
        pre += x;
    }
}

__global__ void __launch_bounds__(VTO_THREADS) vto_sum_apply_kernel(const uint32_t* cnt, uint32_t n, const uint32_t* agg, uint32_t* base_out)
{
    __shared__ uint32_t s_warp[VTO_THREADS / 32];
    const uint32_t base = blockIdx.x * VTO_TILE + threadIdx.x * VTO_ITEMS;
    uint32_t e[VTO_ITEMS], v = 0;
    const bool full = base + VTO_ITEMS <= n;
    if (full)
    {
        const uint4 a = __ldg(reinterpret_cast<const uint4*>(cnt + base)), b = __ldg(reinterpret_cast<const uint4*>(cnt + base) + 1);
        e[0] = a.x; e[1] = a.y; e[2] = a.z; e[3] = a.w; e[4] = b.x; e[5] = b.y; e[6] = b.z; e[7] = b.w;
    }
    else
    {
#pragma unroll
        for (int j = 0; j < VTO_ITEMS; ++j) e[j] = base + j < n ? cnt[base + j] : 0u;
    }
#pragma unroll
    for (int j = 0; j < VTO_ITEMS; ++j) v += e[j];
    uint32_t total;
    uint32_t pre = agg[blockIdx.x] + sum_block_excl(v, s_warp, total);
    uint32_t o[VTO_ITEMS];
#pragma unroll
    for (int j = 0; j < VTO_ITEMS; ++j)
    {
        o[j] = pre;
        pre += e[j];
    }
    if (full)
    {
        reinterpret_cast<uint4*>(base_out + base)[0] = make_uint4(o[0], o[1], o[2], o[3]);
        reinterpret_cast<uint4*>(base_out + base)[1] = make_uint4(o[4], o[5], o[6], o[7]);
    }
    else
    {
#pragma unroll
        for (int j = 0; j < VTO_ITEMS; ++j)
            if (base + j < n) base_out[base + j] = o[j];
    }
}

// scatter: a record that leaves alone at its step goes straight to its output position; the others into their step's list
// (arbitrary order inside the list)
#define VTO_SC_ITEMS 4
// cnt[sl] is not read again: the records of step sl are base[sl + 1] - base[sl] (the same sector as base[sl]; every record has
// exactly one step, so the sum is n).  Four records per thread, their loads issued together: one record per thread left the
// kernel waiting out three dependent round trips (slot -> base -> list) with nothing else in flight.
__device__ __forceinline__ uint32_t step_end(const uint32_t* __restrict__ base, uint32_t sl, uint32_t n)
{
    return sl + 1 < n ? __ldg(base + sl + 1) : n;
}

__global__ void __launch_bounds__(VTO_THREADS) vto_scatter_kernel(uint32_t n, const uint32_t* __restrict__ slot, const uint32_t* __restrict__ base,
                                                                  uint32_t* __restrict__ fill, uint32_t* __restrict__ tmp,
                                                                  uint32_t* __restrict__ order)
{
    const uint32_t v0 = blockIdx.x * (VTO_THREADS * VTO_SC_ITEMS) + threadIdx.x; // rows of VTO_THREADS consecutive records
    uint32_t sl[VTO_SC_ITEMS], b[VTO_SC_ITEMS], e[VTO_SC_ITEMS];
#pragma unroll
    for (int j = 0; j < VTO_SC_ITEMS; ++j)
    {
        const uint32_t v = v0 + j * VTO_THREADS;
        sl[j] = v < n ? __ldg(slot + v) : 0u;
    }
#pragma unroll
    for (int j = 0; j < VTO_SC_ITEMS; ++j)
    {
        const uint32_t v = v0 + j * VTO_THREADS;
        b[j] = v < n ? __ldg(base + sl[j]) : 0u;
        e[j] = v < n ? step_end(base, sl[j], n) : 0u;
    }
#pragma unroll
    for (int j = 0; j < VTO_SC_ITEMS; ++j)
    {
        const uint32_t v = v0 + j * VTO_THREADS;
        if (v >= n) continue;
        if (e[j] - b[j] == 1) order[b[j]] = v;
        else tmp[b[j] + atomicAdd(fill + sl[j], 1u)] = v;
    }
}

// rank by (pos, arrival) inside the lists of more than one record: the writer empties its buffer from the back, smallest
// position first, equal positions in arrival order
__global__ void __launch_bounds__(VTO_THREADS) vto_rank_kernel(uint32_t n, const int32_t* __restrict__ pos, const uint32_t* __restrict__ slot,
                                                               const uint32_t* __restrict__ base, const uint32_t* __restrict__ tmp,
                                                               uint32_t* __restrict__ order)
{
    const uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= n) return;
    const uint32_t sl = slot[v];
    const uint32_t b = __ldg(base + sl), c = step_end(base, sl, n) - b;
    if (c <= 1) return;
    const int32_t p = __ldg(pos + v);
    uint32_t rank = 0;
    for (uint32_t k = b; k < b + c; ++k)
    {
        const uint32_t u = tmp[k];
        const int32_t q = __ldg(pos + u);
        rank += (q < p || (q == p && u < v)) ? 1u : 0u;
    }
    order[b + rank] = v;
}

} // namespace

// device arena of one ordering problem of up to cap records
struct vtn_order_buffers
{
    uint32_t cap = 0;
    size_t words = 0;
    uint32_t* arena = nullptr;
    unsigned long long* K = nullptr;
    int32_t *rid = nullptr, *pos = nullptr, *late = nullptr;
    uint32_t *start = nullptr, *slot = nullptr, *cnt = nullptr, *base = nullptr, *fill = nullptr, *tmp = nullptr, *order = nullptr;
    SegMax* agg = nullptr;
    uint32_t* agg2 = nullptr;
};

static void order_release(vtn_order_buffers* o)
{
    if (!o) return;
    cudaFree(o->arena);
    delete o;
}

static int order_alloc(vtn_ctx* ctx, uint32_t n, vtn_order_buffers** out)
{
    vtn_order_buffers* o = new vtn_order_buffers();
    const size_t pad = 64;
    const uint32_t n_tiles = (n + VTO_TILE - 1) / VTO_TILE;
    o->cap = n;
    o->words = ((size_t)n + pad + 3) & ~(size_t)3; // every array starts 16-byte aligned
    // one arena: K (8 B), rid, pos, start, slot, cnt, base, fill, tmp, order, late (4 B each), tile aggregates
    const size_t bytes = o->words * 4 * 10 + o->words * 8 + (size_t)(n_tiles + pad) * (sizeof(SegMax) + 4);
    cudaError_t e = cudaMalloc((void**)&o->arena, bytes);
    if (e != cudaSuccess)
    {
        delete o;
        return vtn_set_err(ctx, VTN_E_NOMEM, std::string("vtn_writer_order arena: ") + cudaGetErrorString(e));
    }
    o->K = reinterpret_cast<unsigned long long*>(o->arena);
    uint32_t* w = reinterpret_cast<uint32_t*>(o->K + o->words);
    o->rid = (int32_t*)w; w += o->words;
    o->pos = (int32_t*)w; w += o->words;
    o->start = w; w += o->words;
    o->slot = w; w += o->words;
    o->cnt = w; w += o->words;
    o->base = w; w += o->words;
    o->fill = w; w += o->words;
    o->tmp = w; w += o->words;
    o->order = w; w += o->words;
    o->late = (int32_t*)w; w += o->words;
    o->agg = reinterpret_cast<SegMax*>(w);
    o->agg2 = reinterpret_cast<uint32_t*>(o->agg + n_tiles + pad / 2);
    *out = o;
    return VTN_OK;
}

// the nine launches on the ctx stream; rid / pos already on the device
static int order_launch(vtn_ctx* ctx, vtn_order_buffers* o, const int32_t* d_rid, const int32_t* d_pos, uint32_t n, int32_t window)
{
    cudaStream_t s = ctx->stream;
    const uint32_t n_tiles = (n + VTO_TILE - 1) / VTO_TILE;
    const uint32_t g1 = (n + VTO_THREADS - 1) / VTO_THREADS;
    VTN_CUDA(ctx, cudaMemsetAsync(o->cnt, 0, o->words * 4 * 3, s)); // cnt, base, fill are adjacent
    vto_seg_reduce_kernel<<<n_tiles, VTO_THREADS, 0, s>>>(d_rid, d_pos, n, o->agg);
    vto_seg_scan_kernel<<<1, 1024, 0, s>>>(o->agg, n_tiles);
    vto_seg_apply_kernel<<<n_tiles, VTO_THREADS, 0, s>>>(d_rid, d_pos, n, o->agg, o->K);
    vto_step_kernel<<<g1, VTO_THREADS, 0, s>>>(d_pos, n, window, o->K, o->slot, o->cnt, o->late);
    vto_sum_reduce_kernel<<<n_tiles, VTO_THREADS, 0, s>>>(o->cnt, n, o->agg2);
    vto_sum_scan_kernel<<<1, 1024, 0, s>>>(o->agg2, n_tiles);
    vto_sum_apply_kernel<<<n_tiles, VTO_THREADS, 0, s>>>(o->cnt, n, o->agg2, o->base);
    vto_scatter_kernel<<<(n + VTO_THREADS * VTO_SC_ITEMS - 1) / (VTO_THREADS * VTO_SC_ITEMS), VTO_THREADS, 0, s>>>(n, o->slot, o->base, o->fill, o->tmp,
                                                                                                                   o->order);
    vto_rank_kernel<<<g1, VTO_THREADS, 0, s>>>(n, d_pos, o->slot, o->base, o->tmp, o->order);
    VTN_CUDA(ctx, cudaGetLastError());
    ctx->launches += 9;
    return VTN_OK;
}

void vtn_order_dbatch_release(vtn_dbatch* d)
{
    order_release(d->order);
    d->order = nullptr;
}

// Device-resident variant over the batch's own rid / pos1 arrays (bench / pipelines); window must be > 0.
int vtn_writer_order_resident(vtn_ctx* ctx, vtn_dbatch* d, int32_t window)
{
    if (!ctx || !d || window <= 0) return vtn_set_err(ctx, VTN_E_ARG, "vtn_writer_order_resident: bad arguments (window must be > 0)");
    if (!d->n_rec) return VTN_OK;
    VTN_CUDA(ctx, cudaSetDevice(ctx->device));
    if (!d->order || d->order->cap < d->n_rec)
    {
        order_release(d->order);
        d->order = nullptr;
        int rc = order_alloc(ctx, d->cap_rec, &d->order);
        if (rc) return rc;
    }
    return order_launch(ctx, d->order, d->rid, d->pos1, d->n_rec, window);
}

int vtn_dbatch_fetch_writer_order(vtn_ctx* ctx, vtn_dbatch* d, uint32_t* order, int32_t* late_front, uint32_t* n_pending)
{
    if (!ctx || !d || !d->order) return VTN_E_ARG;
    cudaStream_t s = ctx->stream;
    const uint32_t n = d->n_rec;
    if (n_pending) *n_pending = 0;
    if (n && order) VTN_CUDA(ctx, cudaMemcpyAsync(order, d->order->order, sizeof(uint32_t) * n, cudaMemcpyDeviceToHost, s));
    if (n && late_front) VTN_CUDA(ctx, cudaMemcpyAsync(late_front, d->order->late, sizeof(int32_t) * n, cudaMemcpyDeviceToHost, s));
    if (n && n_pending) VTN_CUDA(ctx, cudaMemcpyAsync(n_pending, d->order->cnt + (n - 1), sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    VTN_CUDA(ctx, cudaStreamSynchronize(s));
    return VTN_OK;
}

// Host-buffer call: H2D of rid / pos1, the kernels above, D2H of order (and late_front when asked for).
int vtn_writer_order(vtn_ctx* ctx, uint32_t n, const int32_t* rid, const int32_t* pos1, int32_t window, uint32_t* order, int32_t* late_front,
                     uint32_t* n_pending)
{
    if (!ctx || window < 0) return vtn_set_err(ctx, VTN_E_ARG, "vtn_writer_order: bad arguments (window must be >= 0)");
    if (n_pending) *n_pending = 0;
    if (n == 0) return VTN_OK;
    if (!rid || !pos1 || !order) return vtn_set_err(ctx, VTN_E_ARG, "vtn_writer_order: null arrays");
    if (window == 0)
    { // the writer bypasses its buffer (bcf_ordered_writer.cpp:193-203)
        for (uint32_t i = 0; i < n; ++i) order[i] = i;
        if (late_front)
            for (uint32_t i = 0; i < n; ++i) late_front[i] = INT_MIN;
        return VTN_OK;
    }
    VTN_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;
    // the arena lives in the context and is reused while it is large enough (a file is written block after block)
    vtn_order_buffers* o = static_cast<vtn_order_buffers*>(ctx->order_cache);
    int rc = VTN_OK;
    if (o && o->cap < n)
    {
        order_release(o);
        ctx->order_cache = o = nullptr;
    }
    if (!o)
    {
        rc = order_alloc(ctx, n + n / 4 + 1024, &o);
        if (rc) return rc;
        ctx->order_cache = o;
    }
    cudaError_t e = cudaMemcpyAsync(o->rid, rid, sizeof(int32_t) * n, cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(o->pos, pos1, sizeof(int32_t) * n, cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess) rc = order_launch(ctx, o, o->rid, o->pos, n, window);
    if (e == cudaSuccess && rc == VTN_OK) e = cudaMemcpyAsync(order, o->order, sizeof(uint32_t) * n, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess && rc == VTN_OK && late_front) e = cudaMemcpyAsync(late_front, o->late, sizeof(int32_t) * n, cudaMemcpyDeviceToHost, s);
    // records that only leave with the final flush of the LAST segment are still in the writer's buffer: slot n-1
    if (e == cudaSuccess && rc == VTN_OK && n_pending) e = cudaMemcpyAsync(n_pending, o->cnt + (n - 1), sizeof(uint32_t), cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) rc = vtn_set_err(ctx, VTN_E_CUDA, std::string("vtn_writer_order: ") + cudaGetErrorString(e));
    return rc;
}

void vtn_order_ctx_release(vtn_ctx* ctx)
{
    if (ctx->order_cache) order_release(static_cast<vtn_order_buffers*>(ctx->order_cache));
    ctx->order_cache = nullptr;
}
