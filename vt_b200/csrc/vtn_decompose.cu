// vtn_decompose.cu -- multiallelic fan-out (decompose.cpp:210-392 / 393-957, allele part) on device.
// A record with n alleles becomes max(n-1,1) output records {REF, ALT_i}; the kernel emits the plan
// (src_rec, alt_idx) in input order and, for the fused decompose->normalize path, materialises the
// biallelic batch directly in HBM so that vtn_normalize_kernel can run on it without a host round trip.
// Counters: no_variants, no_biallelic, no_multiallelic, no_additional_biallelic (decompose.cpp:139-143).
#include "vtn_internal.h"
#include "vtn_scan.cuh"

#define VTD_TILE 256
#define VTD_XCAP 2048 // outputs per tile covered by the output->source expansion table
#define VTD_INCAP 4096  // allele bytes of a tile's input records staged in shared memory (fallback: global loads)
#define VTD_OUTCAP 8192 // allele bytes of a tile's output records staged in shared memory (fallback: global stores)

namespace
{

struct FusedOut
{
    int32_t* pos1;
    int32_t* rid;
    uint32_t* allele_beg;
    uint32_t* allele_off;
    uint8_t* bytes;
};

// totals[0] = output records, totals[1] = output allele bytes, totals[2] = records with a single allele.
// Also publishes, per tile of VTD_TILE input records, the scan words the emit kernel needs:
//   agg_a[tile] = (output records << 31) | single-allele records,  agg_b[tile] = output allele bytes
// so that the emit kernel never waits on another tile (every aggregate is known before it starts).
__global__ void __launch_bounds__(VTD_TILE) vtd_count_kernel(VtnDevBatchView B, unsigned long long* totals, uint32_t* counters,
                                                              unsigned long long* agg_a, unsigned long long* agg_b, uint32_t n_tiles)
{
    __shared__ unsigned long long s_ra[VTD_TILE / 32], s_rb[VTD_TILE / 32];
    unsigned long long recs = 0, bytes = 0, singles = 0;
    uint32_t nbi = 0, nmulti = 0, nadd = 0;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (uint32_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x)
    {
        const uint32_t r = tile * VTD_TILE + threadIdx.x;
        unsigned long long va = 0, vb = 0;
        if (r < B.n_rec)
        {
            uint32_t ab = __ldg(B.allele_beg + r), ae = __ldg(B.allele_beg + r + 1);
            uint32_t n = ae - ab;
            uint32_t o0 = __ldg(B.allele_off + ab), o1 = __ldg(B.allele_off + ab + 1), oe = __ldg(B.allele_off + ae);
            uint32_t len0 = o1 - o0;
            if (n > 2) { ++nmulti; nadd += n - 2; }
            else ++nbi;
            if (n >= 2)
            {
                va = (unsigned long long)(n - 1) << 31;
                vb = (unsigned long long)(n - 1) * len0 + (oe - o1);
            }
            else
            {
                va = (1ull << 31) | 1ull;
                vb = len0;
            }
        }
        unsigned long long ra = va, rb = vb;
        for (int o = 16; o; o >>= 1)
        {
            ra += __shfl_xor_sync(0xffffffffu, ra, o);
            rb += __shfl_xor_sync(0xffffffffu, rb, o);
        }
        __syncthreads();
        if (lane == 0) { s_ra[warp] = ra; s_rb[warp] = rb; }
        __syncthreads();
        if (threadIdx.x == 0)
        {
            unsigned long long ta = 0, tb = 0;
            for (int w = 0; w < VTD_TILE / 32; ++w) { ta += s_ra[w]; tb += s_rb[w]; }
            agg_a[tile] = ta;
            agg_b[tile] = tb;
            recs += ta >> 31;
            singles += ta & 0x7fffffffull;
            bytes += tb;
        }
    }
    for (int o = 16; o; o >>= 1)
    {
        nbi += __shfl_xor_sync(0xffffffffu, nbi, o);
        nmulti += __shfl_xor_sync(0xffffffffu, nmulti, o);
        nadd += __shfl_xor_sync(0xffffffffu, nadd, o);
    }
    if (threadIdx.x == 0)
    {
        if (recs) atomicAdd(totals + 0, recs);
        if (bytes) atomicAdd(totals + 1, bytes);
        if (singles) atomicAdd(totals + 2, singles);
    }
    if (lane == 0)
    {
        if (nbi + nmulti) atomicAdd(counters + 0, nbi + nmulti);
        if (nbi) atomicAdd(counters + 1, nbi);
        if (nmulti) atomicAdd(counters + 2, nmulti);
        if (nadd) atomicAdd(counters + 3, nadd);
    }
}

// exclusive scan of the tile aggregates in place (one block; n_tiles is a few 10^4).  Every warp owns a contiguous chunk and
// walks it 32 tiles at a time with coalesced loads (four steps' loads in flight before the first warp scan); pass 1 reduces
// the chunk, the 32 chunk totals are scanned through shared memory, pass 2 rescans with the carry and writes.
__global__ void __launch_bounds__(1024) vtd_scan_kernel(unsigned long long* agg_a, unsigned long long* agg_b, uint32_t n_tiles)
{
    __shared__ unsigned long long s_a[32], s_b[32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t per = ((n_tiles + 31) / 32 + 31) & ~31u; // tiles per warp, a multiple of 32
    const uint32_t lo = min(warp * per, n_tiles), hi = min(lo + per, n_tiles);
    unsigned long long ta = 0, tb = 0;
    for (uint32_t t = lo; t < hi; t += 128)
    {
        unsigned long long xa[4], xb[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
        {
            const uint32_t i = t + 32 * u + lane;
            xa[u] = i < hi ? agg_a[i] : 0ull;
            xb[u] = i < hi ? agg_b[i] : 0ull;
        }
#pragma unroll
        for (int u = 0; u < 4; ++u)
        {
            ta += xa[u];
            tb += xb[u];
        }
    }
    for (int o = 16; o; o >>= 1)
    {
        ta += __shfl_xor_sync(0xffffffffu, ta, o);
        tb += __shfl_xor_sync(0xffffffffu, tb, o);
    }
    if (lane == 0)
    {
        s_a[warp] = ta;
        s_b[warp] = tb;
    }
    __syncthreads();
    unsigned long long ca = 0, cb = 0; // carry into this warp's chunk
    for (int w = 0; w < warp; ++w)
    {
        ca += s_a[w];
        cb += s_b[w];
    }
    for (uint32_t t = lo; t < hi; t += 128)
    {
        unsigned long long xa[4], xb[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
        {
            const uint32_t i = t + 32 * u + lane;
            xa[u] = i < hi ? agg_a[i] : 0ull;
            xb[u] = i < hi ? agg_b[i] : 0ull;
        }
#pragma unroll
        for (int u = 0; u < 4; ++u)
        {
            unsigned long long ia = xa[u], ib = xb[u];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1)
            {
                const unsigned long long pa = __shfl_up_sync(0xffffffffu, ia, o), pb = __shfl_up_sync(0xffffffffu, ib, o);
                if (lane >= o)
                {
                    ia += pa;
                    ib += pb;
                }
            }
            const uint32_t i = t + 32 * u + lane;
            if (i < hi)
            {
                agg_a[i] = ca + ia - xa[u];
                agg_b[i] = cb + ib - xb[u];
            }
            ca += __shfl_sync(0xffffffffu, ia, 31);
            cb += __shfl_sync(0xffffffffu, ib, 31);
        }
    }
}

// one allele: input bytes [src, src+len) -> output bytes [dst, dst+len), through the shared-memory windows when staged
__device__ __forceinline__ void copy_bytes(bool in_staged, bool out_staged, const uint8_t* s_in, uint8_t* s_out, const uint8_t* g_in, uint8_t* g_out,
                                           uint32_t in_lo, unsigned long long out_base, uint32_t out_shift, uint32_t src, unsigned long long dst,
                                           uint32_t len)
{
    if (in_staged && out_staged)
    {
        const uint8_t* a = s_in + (src - in_lo);
        uint8_t* b = s_out + out_shift + (uint32_t)(dst - out_base);
        for (uint32_t q = 0; q < len; ++q) b[q] = a[q];
    }
    else
    {
        for (uint32_t q = 0; q < len; ++q)
        {
            const uint8_t ch = in_staged ? s_in[src - in_lo + q] : g_in[src + q];
            if (out_staged) s_out[out_shift + (uint32_t)(dst - out_base) + q] = ch;
            else g_out[dst + q] = ch;
        }
    }
}

// scan word A: (out_recs << 31) | singles ; scan word B: out bytes
__global__ void __launch_bounds__(VTD_TILE) vtd_emit_kernel(VtnDevBatchView B, uint32_t* src_rec, uint16_t* alt_idx, FusedOut F,
                                                             int materialise, const unsigned long long* tile_state,
                                                             uint32_t n_tiles)
{
    __shared__ unsigned long long s_wa[VTD_TILE / 32];
    __shared__ unsigned long long s_wb[VTD_TILE / 32];
    __shared__ unsigned long long s_base[2];
    __shared__ unsigned long long s_ea[VTD_TILE], s_ob[VTD_TILE];
    __shared__ uint32_t s_ab[VTD_TILE], s_n[VTD_TILE], s_k[VTD_TILE];
    __shared__ uint16_t s_src[VTD_XCAP];
    __shared__ __align__(16) uint8_t s_in[VTD_INCAP + 16], s_out[VTD_OUTCAP + 16];
    __shared__ uint32_t s_io[2]; // first / one-past-last allele byte of the tile's input records
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned long long* st_a = tile_state;          // exclusive prefixes written by vtd_scan_kernel
    const unsigned long long* st_b = tile_state + n_tiles;
    for (uint32_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x)
    {
        __syncthreads();
        const uint32_t r = tile * VTD_TILE + tid;
        const bool live = r < B.n_rec;
        uint32_t ab = 0, n = 0, o0 = 0, o1 = 0, oe = 0;
        unsigned long long va = 0, vb = 0;
        if (live)
        {
            ab = __ldg(B.allele_beg + r);
            uint32_t ae = __ldg(B.allele_beg + r + 1);
            n = ae - ab;
            o0 = __ldg(B.allele_off + ab);
            o1 = __ldg(B.allele_off + ab + 1);
            oe = __ldg(B.allele_off + ae);
            if (n >= 2)
            {
                va = (unsigned long long)(n - 1) << 31;
                vb = (unsigned long long)(n - 1) * (o1 - o0) + (oe - o1);
            }
            else
            {
                va = (1ull << 31) | 1ull;
                vb = o1 - o0;
            }
        }
        unsigned long long ia = va, ib = vb;
        for (int o = 1; o < 32; o <<= 1)
        {
            unsigned long long ta = __shfl_up_sync(0xffffffffu, ia, o);
            unsigned long long tb = __shfl_up_sync(0xffffffffu, ib, o);
            if (lane >= o) { ia += ta; ib += tb; }
        }
        if (lane == 31) { s_wa[warp] = ia; s_wb[warp] = ib; }
        if (tid == 0) s_io[0] = o0;
        if (live && (r + 1 == B.n_rec || tid == VTD_TILE - 1)) s_io[1] = oe;
        __syncthreads();
        unsigned long long wa = 0, wb = 0, ta = 0, tb = 0;
#pragma unroll
        for (int w = 0; w < VTD_TILE / 32; ++w)
        {
            unsigned long long xa = s_wa[w], xb = s_wb[w];
            if (w < warp) { wa += xa; wb += xb; }
            ta += xa;
            tb += xb;
        }
        if (tid == 0)
        {
            const unsigned long long pa = __ldg(st_a + tile), pb = __ldg(st_b + tile);
            s_base[0] = pa;
            s_base[1] = pb;
            if (materialise && tile == n_tiles - 1)
            {
                unsigned long long nr = (pa + ta) >> 31, ns = (pa + ta) & 0x7fffffffull;
                F.allele_beg[nr] = (uint32_t)(2 * nr - ns);
                F.allele_off[2 * nr - ns] = (uint32_t)(pb + tb);
            }
        }
        __syncthreads();
        // ---- emission, OUTPUT-centric: thread j writes output record j of the tile, so that every store stream
        // (src_rec, alt_idx, pos1, rid, allele_beg, allele_off pairs, bytes) is coalesced; the source record is found
        // by binary search on the tile-local output prefix (every record emits at least one output).
        // the tile's input allele bytes -> shared memory (coalesced 128-bit loads; the byte-wise copies below then never wait on
        // global memory), and its output bytes are assembled in shared memory and leave with 128-bit stores
        const uint32_t in_lo = s_io[0] & ~15u, in_n = s_io[1] - in_lo;
        const bool in_staged = materialise && in_n <= VTD_INCAP;
        const unsigned long long out_base = s_base[1];
        const uint32_t out_shift = (uint32_t)out_base & 15u, out_n = (uint32_t)tb;
        const bool out_staged = materialise && tb + out_shift <= VTD_OUTCAP;
        if (in_staged)
            for (uint32_t i = tid * 16; i < in_n; i += VTD_TILE * 16)
                *reinterpret_cast<uint4*>(&s_in[i]) = __ldg(reinterpret_cast<const uint4*>(B.bytes + in_lo + i));
        const unsigned long long ea_t = s_base[0] + wa + (ia - va); // this thread's record: (outputs << 31 | singles) before it
        s_ea[tid] = live ? ea_t : ~0ull;
        s_ob[tid] = s_base[1] + wb + (ib - vb);
        s_ab[tid] = ab;
        s_n[tid] = n;
        const unsigned long long k0 = s_base[0] >> 31;                 // first output of the tile
        const uint32_t tot = (uint32_t)(ta >> 31);
        const uint32_t krel = (uint32_t)((ea_t >> 31) - k0);           // tile-local index of this record's first output
        s_k[tid] = live ? krel : 0xffffffffu;
        // expansion table output -> source record (tiles with at most VTD_XCAP outputs, i.e. practically all)
        const bool expanded = tot <= VTD_XCAP;
        if (expanded && live)
        {
            const uint32_t cnt = n >= 2 ? n - 1 : 1;
            for (uint32_t q = 0; q < cnt; ++q) s_src[krel + q] = (uint16_t)tid;
        }
        __syncthreads();
        const uint32_t n_in = min((uint32_t)VTD_TILE, B.n_rec - tile * VTD_TILE);
        for (uint32_t j = tid; j < tot; j += VTD_TILE)
        {
            uint32_t lo;
            if (expanded)
                lo = s_src[j];
            else
            {
                uint32_t hi = n_in - 1; // last record whose first output <= j
                lo = 0;
                while (lo < hi)
                {
                    const uint32_t mid = (lo + hi + 1) >> 1;
                    if (s_k[mid] <= j) lo = mid;
                    else hi = mid - 1;
                }
            }
            const uint32_t rec = lo;
            const unsigned long long ea = s_ea[rec];
            const uint32_t rr = tile * VTD_TILE + rec;
            const unsigned long long k = k0 + j;
            const uint32_t i = j - s_k[rec] + 1;                       // ALT index (1-based) for n >= 2
            const uint32_t rab = s_ab[rec], rn = s_n[rec];
            const uint32_t ro0 = __ldg(B.allele_off + rab), ro1 = __ldg(B.allele_off + rab + 1);
            const uint32_t len0 = ro1 - ro0;
            const unsigned long long slot = 2 * (ea >> 31) - (ea & 0x7fffffffull) + 2ull * (i - 1);
            src_rec[k] = rr;
            if (rn < 2)
            {
                alt_idx[k] = 0;
                if (materialise)
                {
                    const unsigned long long ob = s_ob[rec];
                    F.pos1[k] = __ldg(B.pos1 + rr);
                    F.rid[k] = __ldg(B.rid + rr);
                    F.allele_beg[k] = (uint32_t)slot;
                    F.allele_off[slot] = (uint32_t)ob;
                    copy_bytes(in_staged, out_staged, s_in, s_out, B.bytes, F.bytes, in_lo, out_base, out_shift, ro0, ob, len0);
                }
            }
            else
            {
                alt_idx[k] = (uint16_t)i;
                if (materialise)
                {
                    const uint32_t oa = __ldg(B.allele_off + rab + i), oa2 = __ldg(B.allele_off + rab + i + 1);
                    const unsigned long long ob = s_ob[rec] + (unsigned long long)(i - 1) * len0 + (oa - ro1);
                    F.pos1[k] = __ldg(B.pos1 + rr);
                    F.rid[k] = __ldg(B.rid + rr);
                    F.allele_beg[k] = (uint32_t)slot;
                    F.allele_off[slot] = (uint32_t)ob;
                    F.allele_off[slot + 1] = (uint32_t)(ob + len0);
                    copy_bytes(in_staged, out_staged, s_in, s_out, B.bytes, F.bytes, in_lo, out_base, out_shift, ro0, ob, len0);
                    copy_bytes(in_staged, out_staged, s_in, s_out, B.bytes, F.bytes, in_lo, out_base, out_shift, oa, ob + len0, oa2 - oa);
                }
            }
        }
        if (out_staged)
        {
            __syncthreads();
            const uint32_t total = out_shift + out_n;
            uint8_t* gdst = F.bytes + (out_base - out_shift); // 16-byte aligned
            for (uint32_t c = tid * 16; c < total; c += VTD_TILE * 16)
            {
                if (c >= out_shift && c + 16 <= total)
                    *reinterpret_cast<uint4*>(gdst + c) = *reinterpret_cast<const uint4*>(&s_out[c]);
                else
                    for (uint32_t q = (c > out_shift ? c : out_shift); q < c + 16 && q < total; ++q) gdst[q] = s_out[q];
            }
        }
    }
}

} // namespace

// Runs count (+ sync for sizes) and emit.  With materialise, db->fused holds the biallelic batch afterwards.
int vtn_launch_decompose(vtn_ctx* ctx, vtn_dbatch* d, bool materialise)
{
    cudaStream_t s = ctx->stream;
    VtnDevBatchView v{d->n_rec, 0u, d->pos1, d->rid, d->allele_beg, d->allele_off, d->bytes};
    VTN_CUDA(ctx, cudaMemsetAsync(d->totals, 0, 4 * sizeof(unsigned long long), s));
    VTN_CUDA(ctx, cudaMemsetAsync(d->counters + 12, 0, 4 * sizeof(uint32_t), s));
    unsigned long long tot[3] = {0, 0, 0};
    if (d->n_rec)
    {
        const uint32_t n_tiles = (d->n_rec + VTD_TILE - 1) / VTD_TILE;
        if (n_tiles > d->n_tiles) return vtn_set_err(ctx, VTN_E_STATE, "decompose: tile state too small");
        uint32_t grid = n_tiles;
        uint32_t maxg = (uint32_t)ctx->sm_count * 8;
        if (grid > maxg) grid = maxg;
        vtd_count_kernel<<<grid, VTD_TILE, 0, s>>>(v, d->totals, d->counters + 12, d->tile_state, d->tile_state + n_tiles, n_tiles);
        VTN_CUDA(ctx, cudaGetLastError());
        vtd_scan_kernel<<<1, 1024, 0, s>>>(d->tile_state, d->tile_state + n_tiles, n_tiles);
        VTN_CUDA(ctx, cudaGetLastError());
        ctx->launches += 2;
        VTN_CUDA(ctx, cudaMemcpyAsync(tot, d->totals, sizeof tot, cudaMemcpyDeviceToHost, s));
        VTN_CUDA(ctx, cudaStreamSynchronize(s));
    }
    if (tot[0] >= (1ull << 31) || tot[1] >= (1ull << 32))
        return vtn_set_err(ctx, VTN_E_ARG, "decompose output too large for one batch (>= 2^31 records or 4 GiB allele bytes)");
    const uint64_t n_out = tot[0];
    if (d->out_bound < n_out || !d->src_rec)
    {
        cudaFree(d->src_rec);
        cudaFree(d->alt_idx);
        d->src_rec = nullptr;
        d->alt_idx = nullptr;
        VTN_CUDA(ctx, cudaMalloc((void**)&d->src_rec, (n_out + 1) * sizeof(uint32_t)));
        VTN_CUDA(ctx, cudaMalloc((void**)&d->alt_idx, (n_out + 1) * sizeof(uint16_t)));
        d->out_bound = n_out;
    }
    FusedOut F{nullptr, nullptr, nullptr, nullptr, nullptr};
    if (materialise)
    {
        const uint32_t n_slots = (uint32_t)(2 * tot[0] - tot[2]);
        vtn_dbatch* f = d->fused;
        if (!f || !vtn_dbatch_resize(f, (uint32_t)n_out, n_slots, tot[1]))
        {
            if (f) vtn_dbatch_destroy(ctx, f);
            d->fused = nullptr;
            int rc = vtn_dbatch_alloc(ctx, (uint32_t)n_out, n_slots, tot[1], true, &d->fused);
            if (rc) return rc;
            f = d->fused;
        }
        F = FusedOut{f->pos1, f->rid, f->allele_beg, f->allele_off, f->bytes};
    }
    if (d->n_rec)
    {
        const uint32_t n_tiles = (d->n_rec + VTD_TILE - 1) / VTD_TILE;
        if (n_tiles > d->n_tiles) return vtn_set_err(ctx, VTN_E_STATE, "decompose: tile state too small");
        int per_sm = 0;
        VTN_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, vtd_emit_kernel, VTD_TILE, 0));
        if (per_sm < 1) per_sm = 1;
        uint32_t grid = (uint32_t)(ctx->sm_count * per_sm);
        if (grid > n_tiles) grid = n_tiles;
        vtd_emit_kernel<<<grid, VTD_TILE, 0, s>>>(v, d->src_rec, d->alt_idx, F, materialise ? 1 : 0, d->tile_state, n_tiles);
        VTN_CUDA(ctx, cudaGetLastError());
        ctx->launches++;
    }
    else if (materialise)
    {
        VTN_CUDA(ctx, cudaMemsetAsync(d->fused->allele_beg, 0, sizeof(uint32_t), s));
        VTN_CUDA(ctx, cudaMemsetAsync(d->fused->allele_off, 0, sizeof(uint32_t), s));
    }
    return VTN_OK;
}
