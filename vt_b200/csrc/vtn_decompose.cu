// vtn_decompose.cu -- multiallelic fan-out (decompose.cpp:210-392 / 393-957, allele part) on device.
// A record with n alleles becomes max(n-1,1) output records {REF, ALT_i}; the kernel emits the plan
// (src_rec, alt_idx) in input order and, for the fused decompose->normalize path, materialises the
// biallelic batch directly in HBM so that vtn_normalize_kernel can run on it without a host round trip.
// Counters: no_variants, no_biallelic, no_multiallelic, no_additional_biallelic (decompose.cpp:139-143).
#include "vtn_internal.h"
#include "vtn_scan.cuh"

#define VTD_TILE 256

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

// totals[0] = output records, totals[1] = output allele bytes, totals[2] = records with a single allele
__global__ void __launch_bounds__(256) vtd_count_kernel(VtnDevBatchView B, unsigned long long* totals, uint32_t* counters)
{
    unsigned long long recs = 0, bytes = 0, singles = 0;
    uint32_t nbi = 0, nmulti = 0, nadd = 0;
    for (uint32_t r = blockIdx.x * blockDim.x + threadIdx.x; r < B.n_rec; r += gridDim.x * blockDim.x)
    {
        uint32_t ab = __ldg(B.allele_beg + r), ae = __ldg(B.allele_beg + r + 1);
        uint32_t n = ae - ab;
        uint32_t o0 = __ldg(B.allele_off + ab), o1 = __ldg(B.allele_off + ab + 1), oe = __ldg(B.allele_off + ae);
        uint32_t len0 = o1 - o0;
        if (n > 2) { ++nmulti; nadd += n - 2; }
        else ++nbi;
        if (n >= 2)
        {
            recs += n - 1;
            bytes += (unsigned long long)(n - 1) * len0 + (oe - o1);
        }
        else
        {
            recs += 1;
            bytes += len0;
            singles += 1;
        }
    }
    for (int o = 16; o; o >>= 1)
    {
        recs += __shfl_xor_sync(0xffffffffu, recs, o);
        bytes += __shfl_xor_sync(0xffffffffu, bytes, o);
        singles += __shfl_xor_sync(0xffffffffu, singles, o);
        nbi += __shfl_xor_sync(0xffffffffu, nbi, o);
        nmulti += __shfl_xor_sync(0xffffffffu, nmulti, o);
        nadd += __shfl_xor_sync(0xffffffffu, nadd, o);
    }
    if ((threadIdx.x & 31) == 0)
    {
        if (recs) atomicAdd(totals + 0, recs);
        if (bytes) atomicAdd(totals + 1, bytes);
        if (singles) atomicAdd(totals + 2, singles);
        if (nbi + nmulti) atomicAdd(counters + 0, nbi + nmulti);
        if (nbi) atomicAdd(counters + 1, nbi);
        if (nmulti) atomicAdd(counters + 2, nmulti);
        if (nadd) atomicAdd(counters + 3, nadd);
    }
}

// scan word A: (out_recs << 31) | singles ; scan word B: out bytes
__global__ void __launch_bounds__(VTD_TILE) vtd_emit_kernel(VtnDevBatchView B, uint32_t* src_rec, uint16_t* alt_idx, FusedOut F,
                                                             int materialise, unsigned long long* tile_state,
                                                             uint32_t* tile_counter, uint32_t n_tiles)
{
    __shared__ uint32_t s_tile;
    __shared__ unsigned long long s_wa[VTD_TILE / 32];
    __shared__ unsigned long long s_wb[VTD_TILE / 32];
    __shared__ unsigned long long s_base[2];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    unsigned long long* st_a = tile_state;
    unsigned long long* st_b = tile_state + n_tiles;
    for (;;)
    {
        __syncthreads();
        if (tid == 0) s_tile = atomicAdd(tile_counter, 1u);
        __syncthreads();
        const uint32_t tile = s_tile;
        if (tile >= n_tiles) break;
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
        if (warp == 0)
        {
            if (lane == 0)
            {
                st_state(st_a + tile, (tile == 0 ? VTN_FLAG_INC : VTN_FLAG_AGG) | ta);
                st_state(st_b + tile, (tile == 0 ? VTN_FLAG_INC : VTN_FLAG_AGG) | tb);
            }
            unsigned long long pa = 0, pb = 0;
            if (tile)
            {
                pa = lookback(st_a, (int)tile, lane);
                pb = lookback(st_b, (int)tile, lane);
                if (lane == 0)
                {
                    st_state(st_a + tile, VTN_FLAG_INC | (pa + ta));
                    st_state(st_b + tile, VTN_FLAG_INC | (pb + tb));
                }
            }
            if (lane == 0)
            {
                s_base[0] = pa;
                s_base[1] = pb;
                if (materialise && tile == n_tiles - 1)
                {
                    unsigned long long nr = (pa + ta) >> 31, ns = (pa + ta) & 0x7fffffffull;
                    F.allele_beg[nr] = (uint32_t)(2 * nr - ns);
                    F.allele_off[2 * nr - ns] = (uint32_t)(pb + tb);
                }
            }
        }
        __syncthreads();
        if (live)
        {
            const unsigned long long ea = s_base[0] + wa + (ia - va);
            unsigned long long ob = s_base[1] + wb + (ib - vb);
            unsigned long long k = ea >> 31;
            unsigned long long slot = 2 * k - (ea & 0x7fffffffull);
            const int32_t pos1 = __ldg(B.pos1 + r), rid = __ldg(B.rid + r);
            const uint32_t len0 = o1 - o0;
            if (n < 2)
            {
                src_rec[k] = r;
                alt_idx[k] = 0;
                if (materialise)
                {
                    F.pos1[k] = pos1;
                    F.rid[k] = rid;
                    F.allele_beg[k] = (uint32_t)slot;
                    F.allele_off[slot] = (uint32_t)ob;
                    for (uint32_t j = 0; j < len0; ++j) F.bytes[ob + j] = B.bytes[o0 + j];
                }
            }
            else
            {
                uint32_t oa = o1;
                for (uint32_t i = 1; i < n; ++i, ++k, slot += 2)
                {
                    const uint32_t oa2 = __ldg(B.allele_off + ab + i + 1);
                    src_rec[k] = r;
                    alt_idx[k] = (uint16_t)i;
                    if (materialise)
                    {
                        F.pos1[k] = pos1;
                        F.rid[k] = rid;
                        F.allele_beg[k] = (uint32_t)slot;
                        F.allele_off[slot] = (uint32_t)ob;
                        F.allele_off[slot + 1] = (uint32_t)(ob + len0);
                        for (uint32_t j = 0; j < len0; ++j) F.bytes[ob + j] = B.bytes[o0 + j];
                        ob += len0;
                        const uint32_t la = oa2 - oa;
                        for (uint32_t j = 0; j < la; ++j) F.bytes[ob + j] = B.bytes[oa + j];
                        ob += la;
                    }
                    oa = oa2;
                }
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
        uint32_t grid = (d->n_rec + 255) / 256;
        uint32_t maxg = (uint32_t)ctx->sm_count * 8;
        if (grid > maxg) grid = maxg;
        vtd_count_kernel<<<grid, 256, 0, s>>>(v, d->totals, d->counters + 12);
        VTN_CUDA(ctx, cudaGetLastError());
        ctx->launches++;
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
        VTN_CUDA(ctx, cudaMemsetAsync(d->tile_state, 0, 2ull * n_tiles * sizeof(unsigned long long), s));
        VTN_CUDA(ctx, cudaMemsetAsync(d->tile_counter, 0, sizeof(uint32_t), s));
        int per_sm = 0;
        VTN_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, vtd_emit_kernel, VTD_TILE, 0));
        if (per_sm < 1) per_sm = 1;
        uint32_t grid = (uint32_t)(ctx->sm_count * per_sm);
        if (grid > n_tiles) grid = n_tiles;
        vtd_emit_kernel<<<grid, VTD_TILE, 0, s>>>(v, d->src_rec, d->alt_idx, F, materialise ? 1 : 0, d->tile_state, d->tile_counter, n_tiles);
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
