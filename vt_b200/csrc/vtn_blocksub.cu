// vtn_blocksub.cu -- vt decompose_blocksub on the device: the fan-out plan of block substitutions.
//
// Replaces the record loop of Igor::decompose_blocksub() (decompose_blocksub.cpp:185-421) except its alignment mode (-a,
// :198-303, Needleman-Wunsch over needle.cpp: not built).  A biallelic record whose REF and ALT have the same length > 1
// (:304) becomes one SNP per differing position (:341-375) or, with keep_mnv (-m), one record per run of differences at
// most max_mnv_dist apart (:322-337, :382-401, :405-410); every other record passes through (:413-418).
// Output k = (src_rec, off, len): new POS = pos1 + off, alleles REF[off, off+len) / ALT[off, off+len); len 0 = the record
// itself.  Counters as the reference counts them (:377-378 count differences also in -m mode).
//
// Kernels: vtb_count_kernel (one thread per record: outputs and counter contributions; the state machine of the reference
// loop is kept literally, including what it does for -d <= 0), reduce / scan / apply of the output counts, vtb_emit_kernel
// (writing at the record's offset: it replays the state machine on the difference mask the count pass left in cnt[], so it reads
// nothing of the record; only block substitutions longer than 25 bases walk their bytes again).  Records that are not block
// substitutions never touch allele bytes.  The counters are summed per tile and added up by the scan kernel (no atomics).
#include "vtn_internal.h"

namespace
{

#define VTB_THREADS 256
#define VTB_ITEMS 8
#define VTB_TILE (VTB_THREADS * VTB_ITEMS)

struct BsView
{
    uint32_t n_rec;
    const uint32_t* allele_beg;
    const uint32_t* allele_off;
    const uint8_t* bytes;
};

// Walks one record.  EMIT = false: returns the number of outputs, adds the differences to n_diff and tells whether the record is a
// block substitution.  EMIT = true: also writes the outputs at src/off/len + k0.
// what decides whether a record is walked: offset of REF and its length; ref_len 0 = not a block substitution
struct BsHead
{
    uint32_t o0, ref_len;
};

__device__ __forceinline__ BsHead head_of(const BsView& B, uint32_t ab, uint32_t n_allele)
{
    BsHead h{0u, 0u};
    if (n_allele == 2)
    {
        const uint32_t o0 = __ldg(B.allele_off + ab), o1 = __ldg(B.allele_off + ab + 1), o2 = __ldg(B.allele_off + ab + 2);
        const uint32_t ref_len = o1 - o0;
        if (ref_len != 1 && ref_len == o2 - o1) h = BsHead{o0, ref_len}; // ref_len 0 == alt_len 0 cannot come out of a VCF line
    }
    return h;
}

// The reference's loop over the positions of a block substitution (decompose_blocksub.cpp:306-411), the state machine kept
// literally; differs(i) tells whether REF[i] != ALT[i].  Returns the number of outputs; EMIT writes them at src/off/len + k0.
template <bool EMIT, class Differs>
__device__ __forceinline__ uint32_t walk_positions(uint32_t r, uint32_t n_pos, Differs differs, int keep_mnv, int max_dist, uint32_t& n_diff,
                                                   uint32_t* src, uint32_t* off, uint32_t* len, uint64_t k0)
{
    uint32_t k = 0;
    int prev = -1, start = -1;
    // the reference compares size_t(i - prev) with the int option: a negative -d converts to a huge unsigned value
    const unsigned long long md = (unsigned long long)(long long)max_dist;
    for (uint32_t i = 0; i < n_pos; ++i)
    {
        if (differs(i))
        {
            ++n_diff;
            if (keep_mnv)
            {
                if (prev == -1 || (unsigned long long)(i - (uint32_t)prev) <= md)
                {
                    prev = (int)i;
                    if (start == -1) start = (int)i;
                }
            }
            else
            {
                if (EMIT)
                {
                    src[k0 + k] = r;
                    off[k0 + k] = i;
                    len[k0 + k] = 1;
                }
                ++k;
            }
        }
        else if (keep_mnv && prev != -1 && !((unsigned long long)(i - (uint32_t)prev) < md))
        {
            if (EMIT)
            {
                src[k0 + k] = r;
                off[k0 + k] = (uint32_t)start;
                len[k0 + k] = (uint32_t)(prev - start + 1);
            }
            ++k;
            prev = -1;
            start = -1;
        }
    }
    if (keep_mnv && prev != -1)
    {
        if (EMIT)
        {
            src[k0 + k] = r;
            off[k0 + k] = (uint32_t)start;
            len[k0 + k] = (uint32_t)(prev - start + 1);
        }
        ++k;
    }
    return k;
}

// cnt[r] carries the count pass's findings to the emit pass:
//   mode (bits 31:30) 0: pass-through record, one output;
//                     2: block substitution of at most VTB_MASK_BITS bases: outputs in bits 29:25, difference mask in bits 24:0 --
//                        the emit pass replays the state machine on the mask and loads nothing of the record.  Positions past the
//                        last difference are equal ones: an open run is written either by the distance rule or after the loop, the
//                        same (start, prev) both ways, so the replay stops at the last set bit;
//                     3: longer block substitution: outputs in bits 29:0, the emit pass walks the bytes again.
#define VTB_MASK_BITS 25

// eight bytes from any address: three aligned words and two funnel shifts (the batch's byte buffer is padded past its end)
__device__ __forceinline__ unsigned long long load8(const uint8_t* p)
{
    const uintptr_t a = reinterpret_cast<uintptr_t>(p);
    const uint32_t* w = reinterpret_cast<const uint32_t*>(a & ~(uintptr_t)3);
    const uint32_t sh = (uint32_t)(a & 3u) * 8u;
    const uint32_t w0 = __ldg(w), w1 = __ldg(w + 1), w2 = __ldg(w + 2);
    return (unsigned long long)__funnelshift_r(w0, w1, sh) | ((unsigned long long)__funnelshift_r(w1, w2, sh) << 32);
}

// bit i = byte i of x is not zero
__device__ __forceinline__ uint32_t nonzero_bytes(unsigned long long x)
{
    const unsigned long long m7 = 0x7f7f7f7f7f7f7f7fULL;
    const unsigned long long t = ((((x & m7) + m7) | x) >> 7) & 0x0101010101010101ULL;
    return (uint32_t)((t * 0x0102040810204080ULL) >> 56);
}
__device__ __forceinline__ uint32_t outputs_of(uint32_t cf) { return (cf >> 30) == 2u ? (cf >> VTB_MASK_BITS) & 31u : cf & 0x3fffffffu; }

__device__ __forceinline__ uint32_t block_excl_sum_u32(uint32_t v, uint32_t* s_warp, uint32_t& total)
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
    for (int w = 0; w < VTB_THREADS / 32; ++w)
    {
        if (w < warp) wpre += s_warp[w];
        tot += s_warp[w];
    }
    total = tot;
    __syncthreads();
    return wpre + inc - v;
}

// per record: number of outputs -> cnt[r]; per tile: total outputs -> agg[tile]; counters
// counters: [0] no_variants [1] no_biallelic_blocksub [2] no_additional_snps [3] new_no_variants
__global__ void __launch_bounds__(VTB_THREADS) vtb_count_kernel(BsView B, int keep_mnv, int max_dist, uint32_t* __restrict__ cnt,
                                                                unsigned long long* __restrict__ agg, unsigned long long* __restrict__ tile_cnt)
{
    __shared__ uint32_t s_warp[VTB_THREADS / 32];
    __shared__ uint32_t s_bs[VTB_THREADS / 32], s_diff[VTB_THREADS / 32];
    uint32_t mine = 0, n_bs = 0, n_diff_all = 0, n_pass = 0;
    // record r of the tile is handled by thread r % VTB_THREADS (coalesced allele_beg loads); the tile total is all that is kept.
    // The loads that decide whether a record is walked are issued for all the thread's records before any walk (two rounds of
    // independent loads instead of VTB_ITEMS dependent chains).
    const int lane = threadIdx.x & 31;
    uint32_t ab[VTB_ITEMS], na[VTB_ITEMS];
    BsHead h[VTB_ITEMS];
#pragma unroll
    for (int j = 0; j < VTB_ITEMS; ++j)
    {
        const uint32_t r = blockIdx.x * VTB_TILE + j * VTB_THREADS + threadIdx.x;
        // allele_beg[r + 1] is the next lane's allele_beg[r]: one load per record, lane 31 (or the last record) loads its own end
        ab[j] = r < B.n_rec ? __ldg(B.allele_beg + r) : 0u;
        const uint32_t nxt = __shfl_down_sync(0xffffffffu, ab[j], 1);
        const bool own = lane == 31 || r + 1 >= B.n_rec;
        na[j] = r < B.n_rec ? (own ? __ldg(B.allele_beg + r + 1) : nxt) - ab[j] : 0u;
    }
#pragma unroll
    for (int j = 0; j < VTB_ITEMS; ++j)
    {
        // the same for allele_off: the end of a biallelic record's ALT is the next record's REF offset.  REF/ALT offsets are one
        // 8-byte load when the record's first allele index is even (always, in a batch of biallelic records).
        const uint32_t r = blockIdx.x * VTB_TILE + j * VTB_THREADS + threadIdx.x;
        const bool two = na[j] == 2u;
        uint32_t o0 = 0, o1 = 0;
        if (two)
        {
            if ((ab[j] & 1u) == 0u)
            {
                const uint2 o = __ldg(reinterpret_cast<const uint2*>(B.allele_off + ab[j]));
                o0 = o.x;
                o1 = o.y;
            }
            else
            {
                o0 = __ldg(B.allele_off + ab[j]);
                o1 = __ldg(B.allele_off + ab[j] + 1);
            }
        }
        const uint32_t nxt = __shfl_down_sync(0xffffffffu, o0, 1);
        const bool nxt_two = __shfl_down_sync(0xffffffffu, two ? 1 : 0, 1) != 0;
        h[j] = BsHead{0u, 0u};
        if (two)
        {
            const bool own = lane == 31 || r + 1 >= B.n_rec || !nxt_two;
            const uint32_t o2 = own ? __ldg(B.allele_off + ab[j] + 2) : nxt;
            const uint32_t ref_len = o1 - o0;
            if (ref_len != 1 && ref_len == o2 - o1) h[j] = BsHead{o0, ref_len};
        }
    }
#pragma unroll
    for (int j = 0; j < VTB_ITEMS; ++j)
    {
        const uint32_t r = blockIdx.x * VTB_TILE + j * VTB_THREADS + threadIdx.x;
        if (r < B.n_rec)
        {
            const bool is_bs = h[j].ref_len != 0;
            uint32_t nd = 0, k = 1, cf = 1;
            if (is_bs)
            {
                const uint32_t L = h[j].ref_len;
                const uint8_t* ref = B.bytes + h[j].o0;
                const uint8_t* alt = ref + L;
                if (L <= VTB_MASK_BITS)
                {
                    // difference mask eight bases at a time (unaligned 8-byte windows of REF and ALT, XOR, non-zero bytes -> bits);
                    // the outputs then follow from the mask alone, as in the emit pass
                    uint32_t mask = 0;
                    for (uint32_t c = 0; c < L; c += 8) mask |= nonzero_bytes(load8(ref + c) ^ load8(alt + c)) << c;
                    mask &= (1u << L) - 1u;
                    nd = (uint32_t)__popc(mask);
                    k = nd;
                    if (keep_mnv)
                    {
                        uint32_t unused = 0;
                        k = walk_positions<false>(r, 32u - (uint32_t)__clz(mask), [&](uint32_t i) { return ((mask >> i) & 1u) != 0u; }, keep_mnv, max_dist,
                                                  unused, nullptr, nullptr, nullptr, 0);
                    }
                    cf = 0x80000000u | (k << VTB_MASK_BITS) | mask;
                }
                else
                {
                    k = walk_positions<false>(r, L, [&](uint32_t i) { return ref[i] != alt[i]; }, keep_mnv, max_dist, nd, nullptr, nullptr, nullptr, 0);
                    cf = 0xC0000000u | k;
                }
            }
            cnt[r] = cf;
            mine += k;
            n_bs += is_bs ? 1u : 0u;
            n_pass += is_bs ? 0u : 1u;
            n_diff_all += nd;
        }
    }
    uint32_t total;
    block_excl_sum_u32(mine, s_warp, total);
    // the tile's counter contributions go to tile_cnt[] (block substitutions | differences << 32) and are summed by the scan
    // kernel: atomics on four shared words (one L2 sector) from every warp of 9766 blocks were what the whole pass waited for
    n_bs = __reduce_add_sync(0xffffffffu, n_bs);
    n_diff_all = __reduce_add_sync(0xffffffffu, n_diff_all);
    if ((threadIdx.x & 31) == 0)
    {
        s_bs[threadIdx.x >> 5] = n_bs;
        s_diff[threadIdx.x >> 5] = n_diff_all;
    }
    __syncthreads();
    if (threadIdx.x == 0)
    {
        uint32_t b = 0, d = 0;
        for (int w = 0; w < VTB_THREADS / 32; ++w)
        {
            b += s_bs[w];
            d += s_diff[w];
        }
        agg[blockIdx.x] = total;
        tile_cnt[blockIdx.x] = (unsigned long long)b | ((unsigned long long)d << 32);
    }
    (void)n_pass;
}

// exclusive scan of the tile totals in place (one block), grand total -> *total_out
// counters: [0] no_variants [1] no_biallelic_blocksub [2] no_additional_snps [3] new_no_variants
__global__ void __launch_bounds__(1024) vtb_scan_kernel(unsigned long long* agg, uint32_t n_tiles, unsigned long long* total_out,
                                                        const unsigned long long* __restrict__ tile_cnt, uint32_t n_rec, uint32_t* counters)
{
    __shared__ unsigned long long s[1024];
    __shared__ unsigned long long s_c[32];
    const uint32_t per = (n_tiles + 1023) / 1024;
    const uint32_t lo = threadIdx.x * per, hi = min(lo + per, n_tiles);
    unsigned long long mine = 0;
    uint32_t c_bs = 0, c_diff = 0;
    for (uint32_t t = lo; t < hi; ++t)
    {
        mine += agg[t];
        const unsigned long long x = tile_cnt[t];
        c_bs += (uint32_t)x;
        c_diff += (uint32_t)(x >> 32);
    }
    c_bs = __reduce_add_sync(0xffffffffu, c_bs);
    c_diff = __reduce_add_sync(0xffffffffu, c_diff);
    if ((threadIdx.x & 31) == 0) s_c[threadIdx.x >> 5] = (unsigned long long)c_bs | ((unsigned long long)c_diff << 32);
    s[threadIdx.x] = mine;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1)
    {
        unsigned long long x = 0;
        if ((int)threadIdx.x >= o) x = s[threadIdx.x - o];
        __syncthreads();
        s[threadIdx.x] += x;
        __syncthreads();
    }
    unsigned long long pre = s[threadIdx.x] - mine;
    for (uint32_t t = lo; t < hi; ++t)
    {
        const unsigned long long x = agg[t];
        agg[t] = pre;
        pre += x;
    }
    if (threadIdx.x == 1023) *total_out = s[1023];
    if (threadIdx.x == 0) // the barriers of the scan above order s_c
    {
        uint32_t n_bs = 0, n_diff = 0;
        for (int w = 0; w < 32; ++w)
        {
            n_bs += (uint32_t)s_c[w];
            n_diff += (uint32_t)(s_c[w] >> 32);
        }
        counters[0] = n_rec;
        counters[1] = n_bs;
        counters[2] = n_diff;
        counters[3] = n_diff + (n_rec - n_bs);
    }
}

// the same tiles again: exclusive prefix of every record inside its tile (record order!), then the walk that writes
__global__ void __launch_bounds__(VTB_THREADS) vtb_emit_kernel(BsView B, int keep_mnv, int max_dist, const uint32_t* __restrict__ cnt,
                                                               const unsigned long long* __restrict__ agg, uint32_t* __restrict__ src,
                                                               uint32_t* __restrict__ off, uint32_t* __restrict__ len, unsigned long long cap_out)
{
    // one scan per tile: every thread's eight counts are loaded together, scanned inside their warp (row j, warp w), the 64
    // (row, warp) totals are scanned by warp 0, and the write loop below only adds three numbers
    __shared__ uint32_t s_tot[VTB_ITEMS * (VTB_THREADS / 32)];
    __shared__ uint32_t s_pre[VTB_ITEMS][VTB_THREADS];
    static_assert(VTB_ITEMS * (VTB_THREADS / 32) == 64, "warp 0 scans two (row, warp) totals per lane");
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int j = 0; j < VTB_ITEMS; ++j)
    {
        const uint32_t r = blockIdx.x * VTB_TILE + j * VTB_THREADS + threadIdx.x;
        const uint32_t c = r < B.n_rec ? outputs_of(__ldg(cnt + r)) : 0u;
        uint32_t inc = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1)
        {
            const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        s_pre[j][threadIdx.x] = inc - c;
        if (lane == 31) s_tot[j * (VTB_THREADS / 32) + warp] = inc;
    }
    __syncthreads();
    if (warp == 0)
    {
        const uint32_t a = s_tot[2 * lane], b = s_tot[2 * lane + 1];
        uint32_t inc = a + b;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1)
        {
            const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        s_tot[2 * lane] = inc - a - b;
        s_tot[2 * lane + 1] = inc - b;
    }
    __syncthreads();
    const unsigned long long base = agg[blockIdx.x];
#pragma unroll 1
    for (int j = 0; j < VTB_ITEMS; ++j) // rows of VTB_THREADS consecutive records, in order
    {
        const uint32_t r = blockIdx.x * VTB_TILE + j * VTB_THREADS + threadIdx.x;
        const uint32_t cf = r < B.n_rec ? __ldg(cnt + r) : 0u; // second read of the tile's counts: L1
        const uint32_t c = outputs_of(cf);
        const uint32_t pre = s_tot[j * (VTB_THREADS / 32) + warp] + s_pre[j][threadIdx.x];
        if (r < B.n_rec && base + pre + c <= cap_out)
        {
            const uint32_t mode = cf >> 30;
            uint32_t nd = 0;
            if (mode == 2u)
            {
                const uint32_t mask = cf & ((1u << VTB_MASK_BITS) - 1u);
                if (!keep_mnv) // one SNP per set bit (:341-375): no need to step through the equal positions
                {
                    uint64_t k = base + pre;
                    for (uint32_t m = mask; m; m &= m - 1u, ++k)
                    {
                        src[k] = r;
                        off[k] = (uint32_t)__ffs((int)m) - 1u;
                        len[k] = 1;
                    }
                }
                else
                walk_positions<true>(r, 32u - (uint32_t)__clz(mask), [&](uint32_t i) { return ((mask >> i) & 1u) != 0u; }, keep_mnv, max_dist, nd,
                                     src, off, len, base + pre);
            }
            else if (mode == 3u)
            {
                const BsHead h = head_of(B, __ldg(B.allele_beg + r), 2u);
                const uint8_t* ref = B.bytes + h.o0;
                const uint8_t* alt = ref + h.ref_len;
                walk_positions<true>(r, h.ref_len, [&](uint32_t i) { return ref[i] != alt[i]; }, keep_mnv, max_dist, nd, src, off, len, base + pre);
            }
            else
            {
                src[base + pre] = r;
                off[base + pre] = 0;
                len[base + pre] = 0;
            }
        }
    }
}

} // namespace

struct vtn_blocksub_buffers
{
    uint32_t cap_rec = 0;
    unsigned long long cap_out = 0;
    uint32_t *cnt = nullptr, *src = nullptr, *off = nullptr, *len = nullptr, *counters = nullptr;
    unsigned long long *agg = nullptr, *total = nullptr, *tile_cnt = nullptr;
};

static void blocksub_release(vtn_blocksub_buffers* b)
{
    if (!b) return;
    cudaFree(b->cnt); cudaFree(b->src); cudaFree(b->off); cudaFree(b->len); cudaFree(b->counters); cudaFree(b->agg); cudaFree(b->total); cudaFree(b->tile_cnt);
    delete b;
}

void vtn_blocksub_dbatch_release(vtn_dbatch* d)
{
    blocksub_release(d->blocksub);
    d->blocksub = nullptr;
}

uint64_t vtn_blocksub_out_bound(const vtn_batch* in) { return in ? (uint64_t)in->n_rec + in->n_bytes / 2 + 1 : 0; }

int vtn_blocksub_resident(vtn_ctx* ctx, vtn_dbatch* d, int keep_mnv, int max_mnv_dist)
{
    if (!ctx || !d) return VTN_E_ARG;
    VTN_CUDA(ctx, cudaSetDevice(ctx->device));
    const unsigned long long need_out = (unsigned long long)d->cap_rec + d->cap_bytes / 2 + 1;
    if (!d->blocksub || d->blocksub->cap_rec < d->n_rec || d->blocksub->cap_out < need_out)
    {
        blocksub_release(d->blocksub);
        vtn_blocksub_buffers* b = d->blocksub = new vtn_blocksub_buffers();
        b->cap_rec = d->cap_rec;
        b->cap_out = need_out;
        const size_t n_tiles = ((size_t)d->cap_rec + VTB_TILE - 1) / VTB_TILE + 1;
        VTN_CUDA(ctx, cudaMalloc((void**)&b->cnt, ((size_t)d->cap_rec + 16) * 4));
        VTN_CUDA(ctx, cudaMalloc((void**)&b->src, (size_t)(need_out + 16) * 4));
        VTN_CUDA(ctx, cudaMalloc((void**)&b->off, (size_t)(need_out + 16) * 4));
        VTN_CUDA(ctx, cudaMalloc((void**)&b->len, (size_t)(need_out + 16) * 4));
        VTN_CUDA(ctx, cudaMalloc((void**)&b->counters, 4 * 4));
        VTN_CUDA(ctx, cudaMalloc((void**)&b->agg, n_tiles * 8));
        VTN_CUDA(ctx, cudaMalloc((void**)&b->tile_cnt, n_tiles * 8));
        VTN_CUDA(ctx, cudaMalloc((void**)&b->total, 8));
    }
    vtn_blocksub_buffers* b = d->blocksub;
    cudaStream_t s = ctx->stream;
    VTN_CUDA(ctx, cudaMemsetAsync(b->counters, 0, 16, s));
    VTN_CUDA(ctx, cudaMemsetAsync(b->total, 0, 8, s));
    if (d->n_rec)
    {
        const uint32_t n_tiles = (d->n_rec + VTB_TILE - 1) / VTB_TILE;
        BsView v{d->n_rec, d->allele_beg, d->allele_off, d->bytes};
        vtb_count_kernel<<<n_tiles, VTB_THREADS, 0, s>>>(v, keep_mnv, max_mnv_dist, b->cnt, b->agg, b->tile_cnt);
        vtb_scan_kernel<<<1, 1024, 0, s>>>(b->agg, n_tiles, b->total, b->tile_cnt, d->n_rec, b->counters);
        vtb_emit_kernel<<<n_tiles, VTB_THREADS, 0, s>>>(v, keep_mnv, max_mnv_dist, b->cnt, b->agg, b->src, b->off, b->len, b->cap_out);
        VTN_CUDA(ctx, cudaGetLastError());
        ctx->launches += 3;
    }
    return VTN_OK;
}

int vtn_dbatch_fetch_blocksub(vtn_ctx* ctx, vtn_dbatch* d, uint32_t* src_rec, uint32_t* off, uint32_t* len, uint64_t cap_out, uint64_t* n_out,
                              vtn_blocksub_counters* counters)
{
    if (!ctx || !d || !d->blocksub || !n_out) return VTN_E_ARG;
    cudaStream_t s = ctx->stream;
    unsigned long long total = 0;
    uint32_t c[4] = {0, 0, 0, 0};
    VTN_CUDA(ctx, cudaMemcpyAsync(&total, d->blocksub->total, 8, cudaMemcpyDeviceToHost, s));
    VTN_CUDA(ctx, cudaMemcpyAsync(c, d->blocksub->counters, 16, cudaMemcpyDeviceToHost, s));
    VTN_CUDA(ctx, cudaStreamSynchronize(s));
    *n_out = total;
    if (counters)
    {
        counters->no_variants = c[0];
        counters->no_biallelic_blocksub = c[1];
        counters->no_additional_snps = c[2];
        counters->new_no_variants = c[3];
    }
    if (total > cap_out || total > d->blocksub->cap_out) return vtn_set_err(ctx, VTN_E_CAPACITY, "vtn_blocksub: output capacity too small (n_out is set)");
    if (total)
    {
        VTN_CUDA(ctx, cudaMemcpyAsync(src_rec, d->blocksub->src, (size_t)total * 4, cudaMemcpyDeviceToHost, s));
        VTN_CUDA(ctx, cudaMemcpyAsync(off, d->blocksub->off, (size_t)total * 4, cudaMemcpyDeviceToHost, s));
        VTN_CUDA(ctx, cudaMemcpyAsync(len, d->blocksub->len, (size_t)total * 4, cudaMemcpyDeviceToHost, s));
        VTN_CUDA(ctx, cudaStreamSynchronize(s));
    }
    return VTN_OK;
}

// ---------------------------------------------------------------------------------------------------------
// vtn_blocksub_noalloc: the same three kernels for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
// serving loops, harnesses that forbid it).  No counterpart in the reference.
// The scratch (cnt[] per record, the per-tile aggregates) lives in a file-scope device array of fixed size, which every GPU
// that loads the library holds until exit: VTN_BLOCKSUB_NOALLOC_MAX_REC records -> 4 MiB + 8 KiB.  In host code the array's
// name is a host-side stand-in, not a device address; the address is looked up once per device, with that device current, by
// cudaGetSymbolAddress (below) and that pointer is what the kernels get.
// ---------------------------------------------------------------------------------------------------------
#define VTB_NA_MAX_TILES (VTN_BLOCKSUB_NOALLOC_MAX_REC / VTB_TILE)
#define VTB_NA_CNT_BYTES ((size_t)VTN_BLOCKSUB_NOALLOC_MAX_REC * 4)
#define VTB_NA_AGG_BYTES (((size_t)VTB_NA_MAX_TILES + 1) * 8)
static __device__ __align__(256) unsigned char vtb_noalloc_scratch[VTB_NA_CNT_BYTES + 2 * VTB_NA_AGG_BYTES];
static_assert(VTN_BLOCKSUB_NOALLOC_MAX_REC % VTB_TILE == 0, "the maximum is a whole number of tiles");

#include <mutex>
static std::mutex vtb_na_mutex;
static unsigned char* vtb_na_addr[64] = {nullptr}; // device address of vtb_noalloc_scratch, per device ordinal

int vtn_blocksub_noalloc(vtn_ctx* ctx, uint32_t n_rec, const uint32_t* d_allele_beg, const uint32_t* d_allele_off, const uint8_t* d_bytes, int keep_mnv,
                         int max_mnv_dist, uint32_t* d_src_rec, uint32_t* d_off, uint32_t* d_len, uint64_t cap_out, uint64_t* d_n_out, uint32_t* d_counters)
{
    if (!ctx) return VTN_E_ARG;
    if (n_rec > VTN_BLOCKSUB_NOALLOC_MAX_REC)
        return vtn_set_err(ctx, VTN_E_ARG, "vtn_blocksub_noalloc: " + std::to_string(n_rec) + " records exceed the fixed scratch (VTN_BLOCKSUB_NOALLOC_MAX_REC = " +
                                               std::to_string((unsigned)VTN_BLOCKSUB_NOALLOC_MAX_REC) + "); cut the batch or use vtn_blocksub");
    if (!d_n_out || !d_counters || (n_rec && (!d_allele_beg || !d_allele_off || !d_bytes || !d_src_rec || !d_off || !d_len)))
        return vtn_set_err(ctx, VTN_E_ARG, "vtn_blocksub_noalloc: null device pointer");
    if (ctx->device < 0 || ctx->device >= 64) return vtn_set_err(ctx, VTN_E_ARG, "vtn_blocksub_noalloc: device ordinal beyond 63");
    VTN_CUDA(ctx, cudaSetDevice(ctx->device));
    unsigned char* base = nullptr;
    {
        std::lock_guard<std::mutex> g(vtb_na_mutex);
        if (!vtb_na_addr[ctx->device])
        { // set-up, once per device, with that device current
            void* p = nullptr;
            VTN_CUDA(ctx, cudaGetSymbolAddress(&p, vtb_noalloc_scratch));
            vtb_na_addr[ctx->device] = (unsigned char*)p;
        }
        base = vtb_na_addr[ctx->device];
    }
    uint32_t* cnt = reinterpret_cast<uint32_t*>(base);
    unsigned long long* agg = reinterpret_cast<unsigned long long*>(base + VTB_NA_CNT_BYTES);
    unsigned long long* tile_cnt = reinterpret_cast<unsigned long long*>(base + VTB_NA_CNT_BYTES + VTB_NA_AGG_BYTES);
    cudaStream_t s = ctx->stream;
    // the shared array keeps the last call's contents: cnt[0, n_rec) and agg / tile_cnt[0, n_tiles) are all assigned by the count
    // kernel before anything reads them, the two results are cleared here (an empty batch runs no kernel)
    VTN_CUDA(ctx, cudaMemsetAsync(d_counters, 0, 4 * sizeof(uint32_t), s));
    VTN_CUDA(ctx, cudaMemsetAsync(d_n_out, 0, sizeof(uint64_t), s));
    if (n_rec)
    {
        const uint32_t n_tiles = (n_rec + VTB_TILE - 1) / VTB_TILE;
        BsView v{n_rec, d_allele_beg, d_allele_off, d_bytes};
        vtb_count_kernel<<<n_tiles, VTB_THREADS, 0, s>>>(v, keep_mnv, max_mnv_dist, cnt, agg, tile_cnt);
        vtb_scan_kernel<<<1, 1024, 0, s>>>(agg, n_tiles, reinterpret_cast<unsigned long long*>(d_n_out), tile_cnt, n_rec, d_counters);
        vtb_emit_kernel<<<n_tiles, VTB_THREADS, 0, s>>>(v, keep_mnv, max_mnv_dist, cnt, agg, d_src_rec, d_off, d_len, cap_out);
        VTN_CUDA(ctx, cudaGetLastError());
        ctx->launches += 3;
    }
    return VTN_OK;
}

// ---------------------------------------------------------------------------------------------------------
// Alignment mode (-a, decompose_blocksub.cpp:190-303): a biallelic record whose REF and ALT differ in length, REF longer than one base
// and ALT longer than one base, is cut along a global alignment of REF[1:] and ALT[1:] (NeedlemanWunsch, needle.cpp:36-101: match 0,
// mismatch -1, gap -1; a cell prefers the diagonal, then a deletion if strictly better, else an insertion if strictly better than the
// diagonal), the first bases force-aligned.  A chunk is one aligned column (match or mismatch) plus the gap columns that follow it; it
// is emitted when it holds a mismatch or a gap (:209-246).
// One thread per eligible record: these records are rare and short (tens of bases); the traceback matrix lives in global scratch.
// ---------------------------------------------------------------------------------------------------------
namespace
{
enum { NW_D = 1, NW_M = 2, NW_X = 3, NW_I = 4 }; // needle.h:44

struct AlignJob
{
    uint32_t rec;      // record index
    uint32_t out_base; // first chunk slot of the record
    uint64_t mat_off;  // byte offset of its traceback matrix in the scratch
    uint64_t row_off;  // int offset of its score row
};

__global__ void __launch_bounds__(64) vtb_align_kernel(const uint32_t* __restrict__ allele_beg, const uint32_t* __restrict__ allele_off,
                                                        const uint8_t* __restrict__ bytes, const AlignJob* __restrict__ jobs, uint32_t n_jobs,
                                                        uint8_t* __restrict__ mat_all, int* __restrict__ row_all, uint32_t* __restrict__ n_chunk,
                                                        uint32_t* __restrict__ pos_ref, uint32_t* __restrict__ pos_alt, uint32_t* __restrict__ len_ref,
                                                        uint32_t* __restrict__ len_alt)
{
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_jobs) return;
    const AlignJob J = jobs[t];
    const uint32_t ab = allele_beg[J.rec];
    const uint8_t* REF = bytes + allele_off[ab];
    const uint8_t* ALT = bytes + allele_off[ab + 1];
    const int lr = (int)(allele_off[ab + 1] - allele_off[ab]) - 1;     // len_ref of the alignment: REF[1:]
    const int la = (int)(allele_off[ab + 2] - allele_off[ab + 1]) - 1; // len_read: ALT[1:]
    const uint8_t* ref = REF + 1;
    const uint8_t* read = ALT + 1;
    uint8_t* mat = mat_all + J.mat_off;
    int* scores = row_all + J.row_off;
    const int W = la + 1;
    const int gap = -1, mism = -1; // needle.h:37
    for (int i = 0; i <= lr; ++i) mat[(size_t)i * W] = NW_D;
    for (int j = 0; j <= la; ++j)
    {
        mat[j] = NW_I;
        scores[j] = j * gap;
    }
    mat[0] = NW_M;
    int score_n = 0;
    for (int i = 1; i <= lr; ++i)
    {
        uint8_t* row = mat + (size_t)i * W;
        score_n += gap;
        int best_score = 0;
        for (int j = 1; j <= la; ++j)
        {
            const int score_nw = scores[j - 1], score_w = scores[j];
            int best_dir;
            int score_diag = score_nw;
            if (ref[i - 1] == read[j - 1])
                best_dir = NW_M; // + score_match (0)
            else
            {
                best_dir = NW_X;
                score_diag += mism;
            }
            best_score = score_diag;
            if (score_w + gap > best_score)
            {
                best_score = score_w + gap;
                best_dir = NW_D;
            }
            else if (score_n + gap > best_score)
            {
                best_score = score_n + gap;
                best_dir = NW_I;
            }
            scores[j - 1] = score_n;
            score_n = best_score;
            row[j] = (uint8_t)best_dir;
        }
        scores[la] = best_score;
    }
    // trace_path (needle.cpp:103-133) walked from the end; the chunk rule (decompose_blocksub.cpp:209-246) applied backwards: the gap
    // columns met since the last aligned column belong to the aligned column reached next
    int i = lr, j = la;
    uint32_t nd = 0, ni = 0, n = 0;
    const uint32_t base = J.out_base;
    while (i > 0 || j > 0)
    {
        const int d = mat[(size_t)i * W + j];
        if (d == NW_D)
        {
            ++nd;
            --i;
        }
        else if (d == NW_I)
        {
            ++ni;
            --j;
        }
        else
        {
            if (d == NW_X || nd || ni)
            { // column REF[i] / ALT[j] of the full alleles
                pos_ref[base + n] = (uint32_t)i, pos_alt[base + n] = (uint32_t)j, len_ref[base + n] = 1 + nd, len_alt[base + n] = 1 + ni;
                ++n;
            }
            nd = ni = 0;
            --i;
            --j;
        }
    }
    if (REF[0] != ALT[0] || nd || ni)
    { // the force-aligned first bases (:200-206)
        pos_ref[base + n] = 0, pos_alt[base + n] = 0, len_ref[base + n] = 1 + nd, len_alt[base + n] = 1 + ni;
        ++n;
    }
    for (uint32_t a = 0, b = n ? n - 1 : 0; a < b; ++a, --b)
    { // into ascending order
        uint32_t x;
        x = pos_ref[base + a], pos_ref[base + a] = pos_ref[base + b], pos_ref[base + b] = x;
        x = pos_alt[base + a], pos_alt[base + a] = pos_alt[base + b], pos_alt[base + b] = x;
        x = len_ref[base + a], len_ref[base + a] = len_ref[base + b], len_ref[base + b] = x;
        x = len_alt[base + a], len_alt[base + a] = len_alt[base + b], len_alt[base + b] = x;
    }
    n_chunk[J.rec] = n;
}

inline bool align_eligible(const vtn_batch* in, uint32_t r, uint32_t& rl, uint32_t& al)
{
    const uint32_t ab = in->allele_beg[r], n = in->allele_beg[r + 1] - ab;
    rl = in->allele_off[ab + 1] - in->allele_off[ab];
    al = n == 1 ? 0 : in->allele_off[ab + 2] - in->allele_off[ab + 1];
    return rl != al && rl != 1 && al > 1; // decompose_blocksub.cpp:190 (only ALT 1 is looked at)
}
} // namespace

uint64_t vtn_blocksub_align_bound(const vtn_batch* in)
{
    if (!in || !in->allele_beg || !in->allele_off) return 0;
    uint64_t k = 0;
    for (uint32_t r = 0; r < in->n_rec; ++r)
    {
        uint32_t rl, al;
        if (align_eligible(in, r, rl, al)) k += std::min(rl, al);
    }
    return k;
}

int vtn_blocksub_align(vtn_ctx* ctx, const vtn_batch* in, vtn_blocksub_align_plan* plan)
{
    if (!ctx || !in || !plan) return vtn_set_err(ctx, VTN_E_ARG, "vtn_blocksub_align: bad arguments");
    const uint32_t n = in->n_rec;
    if (n && (!in->allele_beg || !in->allele_off || !in->bytes || !plan->chunk_beg || !plan->n_chunk))
        return vtn_set_err(ctx, VTN_E_ARG, "vtn_blocksub_align: null arrays");
    std::vector<AlignJob> jobs;
    uint64_t out = 0;
    for (uint32_t r = 0; r < n; ++r)
    {
        uint32_t rl, al;
        plan->chunk_beg[r] = (uint32_t)out;
        plan->n_chunk[r] = 0;
        if (!align_eligible(in, r, rl, al)) continue;
        jobs.push_back(AlignJob{r, (uint32_t)out, (uint64_t)rl * al, (uint64_t)al}); // matrix bytes / row ints for now, offsets below
        out += std::min(rl, al);
    }
    if (n) plan->chunk_beg[n] = (uint32_t)out;
    plan->n_chunks = out;
    if (out > plan->cap_chunk) return vtn_set_err(ctx, VTN_E_CAPACITY, "vtn_blocksub_align: chunk arrays too small (vtn_blocksub_align_bound)");
    if (jobs.empty()) return VTN_OK;
    if (!plan->pos_ref || !plan->pos_alt || !plan->len_ref || !plan->len_alt) return vtn_set_err(ctx, VTN_E_ARG, "vtn_blocksub_align: null chunk arrays");
    VTN_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;
    uint32_t *d_beg = nullptr, *d_off = nullptr, *d_cnt = nullptr, *d_chunks = nullptr;
    uint8_t *d_bytes = nullptr, *d_mat = nullptr;
    int* d_row = nullptr;
    AlignJob* d_jobs = nullptr;
    const uint64_t budget = 512ull << 20; // traceback bytes per launch
    cudaError_t e = cudaMalloc(&d_beg, 4 * ((size_t)n + 1));
    if (e == cudaSuccess) e = cudaMalloc(&d_off, 4 * ((size_t)in->n_allele + 1));
    if (e == cudaSuccess) e = cudaMalloc(&d_bytes, (size_t)in->n_bytes + 1);
    if (e == cudaSuccess) e = cudaMalloc(&d_cnt, 4 * (size_t)n);
    if (e == cudaSuccess) e = cudaMalloc(&d_chunks, 16 * (size_t)out + 16);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_beg, in->allele_beg, 4 * ((size_t)n + 1), cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_off, in->allele_off, 4 * ((size_t)in->n_allele + 1), cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_bytes, in->bytes, (size_t)in->n_bytes, cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess) e = cudaMemsetAsync(d_cnt, 0, 4 * (size_t)n, s);
    uint32_t *d_pr = d_chunks, *d_pa = d_chunks + out, *d_lr = d_chunks + 2 * out, *d_la = d_chunks + 3 * out;
    size_t j0 = 0;
    while (e == cudaSuccess && j0 < jobs.size())
    { // slices of jobs whose traceback matrices fit the budget (a single larger one goes alone)
        size_t j1 = j0;
        uint64_t mat = 0, row = 0;
        while (j1 < jobs.size() && (j1 == j0 || mat + jobs[j1].mat_off <= budget))
        {
            const uint64_t mb = jobs[j1].mat_off, rb = jobs[j1].row_off;
            jobs[j1].mat_off = mat, jobs[j1].row_off = row;
            mat += mb, row += rb;
            ++j1;
        }
        const uint32_t nj = (uint32_t)(j1 - j0);
        e = cudaMalloc(&d_mat, (size_t)mat + 16);
        if (e == cudaSuccess) e = cudaMalloc(&d_row, 4 * (size_t)row + 16);
        if (e == cudaSuccess) e = cudaMalloc(&d_jobs, sizeof(AlignJob) * nj);
        if (e == cudaSuccess) e = cudaMemcpyAsync(d_jobs, jobs.data() + j0, sizeof(AlignJob) * nj, cudaMemcpyHostToDevice, s);
        if (e == cudaSuccess)
        {
            vtb_align_kernel<<<(nj + 63) / 64, 64, 0, s>>>(d_beg, d_off, d_bytes, d_jobs, nj, d_mat, d_row, d_cnt, d_pr, d_pa, d_lr, d_la);
            ctx->launches++;
            e = cudaGetLastError();
        }
        if (e == cudaSuccess) e = cudaStreamSynchronize(s);
        cudaFree(d_mat), cudaFree(d_row), cudaFree(d_jobs);
        d_mat = nullptr, d_row = nullptr, d_jobs = nullptr;
        j0 = j1;
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(plan->n_chunk, d_cnt, 4 * (size_t)n, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(plan->pos_ref, d_pr, 4 * (size_t)out, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(plan->pos_alt, d_pa, 4 * (size_t)out, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(plan->len_ref, d_lr, 4 * (size_t)out, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(plan->len_alt, d_la, 4 * (size_t)out, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    cudaFree(d_beg), cudaFree(d_off), cudaFree(d_bytes), cudaFree(d_cnt), cudaFree(d_chunks);
    if (e != cudaSuccess) return vtn_set_err(ctx, VTN_E_CUDA, std::string("vtn_blocksub_align: ") + cudaGetErrorString(e));
    return VTN_OK;
}
