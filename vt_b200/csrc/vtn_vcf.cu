// vtn_vcf.cu -- VCF record text -> the struct-of-arrays batch of include/vtn.h, on the device.
//
// Replaces, for the five site columns the hot path reads, htslib's per-line text parse: vcf_parse1 (lib/htslib/vcf.c: split on
// tabs, CHROM, POS as a decimal integer, "REF,ALT" into d.allele[] with ALT "." meaning no alternate allele) followed by
// bcf_unpack(BCF_UN_STR) / bcf_get_allele (normalize.cpp:196, decompose.cpp:205), i.e. what BCFOrderedReader::read
// (bcf_ordered_reader.cpp:238-298) hands to the record loops.  Everything right of ALT is not parsed: its offset is returned so
// that the caller passes it through untouched.
//
// Layout of the work (all streaming, byte-granular input read with 128-bit loads):
//   L1 vtk_count_kernel   '\n' per 8 KB tile (SWAR compare + popc)                      -> tile aggregates -> scan
//   L2 vtk_lines_kernel   the same tiles again: every '\n' writes the offset of the line that starts behind it
//   L3 vtk_fields_kernel  one thread per line: the first seven tabs, POS, CHROM length, allele count and bytes of REF + ALTs
//                         (lines are short and consecutive, so the byte walks of a warp share cache lines)
//   scan of (alleles << 32 | bytes) per record                                          -> allele_beg, byte offsets
//   L4 vtk_alleles_kernel one thread per line: allele_off entries and the allele bytes (commas dropped)
#include "vtn_internal.h"

namespace
{

#define VTK_THREADS 256
#define VTK_BPT 32 // bytes per thread per tile
#define VTK_TILE (VTK_THREADS * VTK_BPT)
#define VTK_ITEMS 8
#define VTK_RTILE (VTK_THREADS * VTK_ITEMS) // records per block in the record-level scans

__device__ __forceinline__ uint32_t nl_mask4(uint32_t w)
{ // bit 7 of every byte that equals '\n'
    const uint32_t x = w ^ 0x0a0a0a0au;
    return ~(((x & 0x7f7f7f7fu) + 0x7f7f7f7fu) | x) & 0x80808080u;
}

template <typename T>
__device__ __forceinline__ T block_excl_sum(T v, T* s_warp, T& total)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    T inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1)
    {
        const T t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    T wpre = 0, tot = 0;
    for (int w = 0; w < VTK_THREADS / 32; ++w)
    {
        if (w < warp) wpre += s_warp[w];
        tot += s_warp[w];
    }
    total = tot;
    __syncthreads();
    return wpre + inc - v;
}

// exclusive scan of tile aggregates in place (one block)
template <typename T>
__global__ void __launch_bounds__(1024) vtk_scan_kernel(T* agg, uint32_t n_tiles, T* total_out)
{
    __shared__ T s[1024];
    const uint32_t per = (n_tiles + 1023) / 1024;
    const uint32_t lo = threadIdx.x * per, hi = min(lo + per, n_tiles);
    T mine = 0;
    for (uint32_t t = lo; t < hi; ++t) mine += agg[t];
    s[threadIdx.x] = mine;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1)
    {
        T x = 0;
        if ((int)threadIdx.x >= o) x = s[threadIdx.x - o];
        __syncthreads();
        s[threadIdx.x] += x;
        __syncthreads();
    }
    T pre = s[threadIdx.x] - mine;
    for (uint32_t t = lo; t < hi; ++t)
    {
        const T x = agg[t];
        agg[t] = pre;
        pre += x;
    }
    if (threadIdx.x == 1023 && total_out) *total_out = s[1023];
}

// the 32 bytes of one thread as eight words (zero beyond the text)
__device__ __forceinline__ void load32(const uint8_t* text, uint64_t n, uint64_t off, uint32_t* w)
{
    if (off + VTK_BPT <= n)
    {
        const uint4 a = __ldg(reinterpret_cast<const uint4*>(text + off)), b = __ldg(reinterpret_cast<const uint4*>(text + off) + 1);
        w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w; w[4] = b.x; w[5] = b.y; w[6] = b.z; w[7] = b.w;
    }
    else
    {
#pragma unroll
        for (int k = 0; k < 8; ++k)
        {
            uint32_t x = 0;
            for (int q = 0; q < 4; ++q)
                if (off + 4 * k + q < n) x |= (uint32_t)text[off + 4 * k + q] << (8 * q);
            w[k] = x;
        }
    }
}

__global__ void __launch_bounds__(VTK_THREADS) vtk_count_kernel(const uint8_t* __restrict__ text, uint64_t n, uint32_t* __restrict__ agg)
{
    __shared__ uint32_t s_warp[VTK_THREADS / 32];
    const uint64_t off = (uint64_t)blockIdx.x * VTK_TILE + threadIdx.x * VTK_BPT;
    uint32_t w[8], c = 0;
    load32(text, n, off, w);
#pragma unroll
    for (int k = 0; k < 8; ++k) c += __popc(nl_mask4(w[k]));
    uint32_t total;
    block_excl_sum(c, s_warp, total);
    if (threadIdx.x == 0) agg[blockIdx.x] = total;
}

// line k+1 starts behind the k-th '\n'
__global__ void __launch_bounds__(VTK_THREADS) vtk_lines_kernel(const uint8_t* __restrict__ text, uint64_t n, const uint32_t* __restrict__ agg,
                                                                uint64_t* __restrict__ line_off, uint32_t cap_rec)
{
    __shared__ uint32_t s_warp[VTK_THREADS / 32];
    const uint64_t off = (uint64_t)blockIdx.x * VTK_TILE + threadIdx.x * VTK_BPT;
    uint32_t w[8], m[8], c = 0;
    load32(text, n, off, w);
#pragma unroll
    for (int k = 0; k < 8; ++k)
    {
        m[k] = nl_mask4(w[k]);
        c += __popc(m[k]);
    }
    uint32_t total;
    uint32_t idx = agg[blockIdx.x] + block_excl_sum(c, s_warp, total);
    if (!c) return;
#pragma unroll
    for (int k = 0; k < 8; ++k)
    {
        uint32_t mm = m[k];
        while (mm)
        {
            const int b = (__ffs(mm) - 1) >> 3;
            mm &= mm - 1;
            ++idx; // this '\n' ends line idx-1, line idx starts behind it
            if (idx <= cap_rec) line_off[idx] = off + 4 * k + b + 1;
        }
    }
}

struct VtkFields
{
    uint32_t* chrom_len;
    int32_t* pos1;
    uint32_t* tail_off;
    uint8_t* status;
    unsigned long long* cnt; // alleles << 32 | bytes, per record (scanned in place into the exclusive prefix)
    uint32_t* ref_off;       // offset of the REF column from the line start
};

// one thread per line: the first seven tabs
__global__ void __launch_bounds__(VTK_THREADS) vtk_fields_kernel(const uint8_t* __restrict__ text, uint64_t n_text, const uint64_t* __restrict__ line_off,
                                                                 uint32_t n_rec, VtkFields F)
{
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_rec) return;
    const uint64_t o = line_off[r];
    uint64_t end = line_off[r + 1];
    if (end > o && text[end - 1] == '\n') --end;
    if (end > o && text[end - 1] == '\r') --end;
    uint32_t tab[7];
    int nt = 0;
    uint32_t n_alt_commas = 0;
    const uint32_t len = (uint32_t)(end - o);
    for (uint32_t i = 0; i < len && nt < 7; ++i)
    {
        const uint8_t ch = text[o + i];
        if (ch == '\t') tab[nt++] = i;
        else if (nt == 4 && ch == ',') ++n_alt_commas;
    }
    uint32_t st = nt < 7 ? 1u : 0u, cl = 0, to = 0, ro = 0;
    int32_t p = 0;
    unsigned long long cnt = 0;
    if (!st)
    {
        cl = tab[0];
        // strtol(text + tab[0] + 1, 0, 10): optional blanks and sign, then digits
        uint32_t i = tab[0] + 1;
        while (i < tab[1] && (text[o + i] == ' ')) ++i;
        bool neg = false;
        if (i < tab[1] && (text[o + i] == '-' || text[o + i] == '+'))
        {
            neg = text[o + i] == '-';
            ++i;
        }
        unsigned long long v = 0; // strtol saturates at LONG_MAX / LONG_MIN; the caller's cast to int32 keeps the low bits
        bool sat = false;
        while (i < tab[1] && text[o + i] >= '0' && text[o + i] <= '9')
        {
            const unsigned d = text[o + i] - '0';
            if (v > (0x7fffffffffffffffull - d) / 10) sat = true;
            v = v * 10 + d;
            ++i;
        }
        const long long sv = sat ? (neg ? (long long)0x8000000000000000ull : 0x7fffffffffffffffll) : (neg ? -(long long)v : (long long)v);
        p = (int32_t)sv;
        to = tab[4] + 1;
        ro = tab[2] + 1;
        const uint32_t ref_len = tab[3] - (tab[2] + 1), alt_len = tab[4] - (tab[3] + 1);
        const bool no_alt = alt_len == 1 && text[o + tab[3] + 1] == '.';
        const uint32_t n_alt = no_alt ? 0u : n_alt_commas + 1u;
        const uint32_t alt_bytes = no_alt ? 0u : alt_len - n_alt_commas;
        cnt = ((unsigned long long)(1u + n_alt) << 32) | (unsigned long long)(ref_len + alt_bytes);
    }
    F.chrom_len[r] = cl;
    F.pos1[r] = p;
    F.tail_off[r] = to;
    F.status[r] = (uint8_t)st;
    F.cnt[r] = cnt;
    F.ref_off[r] = ro;
}

__global__ void __launch_bounds__(VTK_THREADS) vtk_rsum_reduce_kernel(const unsigned long long* __restrict__ cnt, uint32_t n, unsigned long long* __restrict__ agg)
{
    __shared__ unsigned long long s_warp[VTK_THREADS / 32];
    const uint32_t base = blockIdx.x * VTK_RTILE + threadIdx.x * VTK_ITEMS;
    unsigned long long v = 0;
#pragma unroll
    for (int j = 0; j < VTK_ITEMS; ++j)
        if (base + j < n) v += cnt[base + j];
    unsigned long long total;
    block_excl_sum(v, s_warp, total);
    if (threadIdx.x == 0) agg[blockIdx.x] = total;
}

// exclusive prefix per record: allele_beg[r] and the record's first byte offset (kept in cnt, in place)
__global__ void __launch_bounds__(VTK_THREADS) vtk_rsum_apply_kernel(unsigned long long* __restrict__ cnt, uint32_t n, const unsigned long long* __restrict__ agg,
                                                                     uint32_t* __restrict__ allele_beg)
{
    __shared__ unsigned long long s_warp[VTK_THREADS / 32];
    const uint32_t base = blockIdx.x * VTK_RTILE + threadIdx.x * VTK_ITEMS;
    unsigned long long e[VTK_ITEMS], v = 0;
#pragma unroll
    for (int j = 0; j < VTK_ITEMS; ++j)
    {
        e[j] = base + j < n ? cnt[base + j] : 0ull;
        v += e[j];
    }
    unsigned long long total;
    unsigned long long pre = agg[blockIdx.x] + block_excl_sum(v, s_warp, total);
#pragma unroll
    for (int j = 0; j < VTK_ITEMS; ++j)
    {
        if (base + j < n)
        {
            cnt[base + j] = pre;
            allele_beg[base + j] = (uint32_t)(pre >> 32);
        }
        pre += e[j];
        if (base + j + 1 == n) allele_beg[n] = (uint32_t)(pre >> 32);
    }
}

// one thread per line: allele_off entries and allele bytes (REF, then the ALTs without their commas)
__global__ void __launch_bounds__(VTK_THREADS) vtk_alleles_kernel(const uint8_t* __restrict__ text, const uint64_t* __restrict__ line_off, uint32_t n_rec,
                                                                  const unsigned long long* __restrict__ pre, const uint32_t* __restrict__ ref_off,
                                                                  const uint8_t* __restrict__ status, uint32_t* __restrict__ allele_off,
                                                                  uint8_t* __restrict__ bytes, uint32_t cap_allele, uint64_t cap_bytes,
                                                                  const unsigned long long* __restrict__ total_ptr)
{
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_rec) return;
    const unsigned long long total = *total_ptr;
    if (r == 0) allele_off[0] = 0;
    if (status[r]) return;
    const unsigned long long pr = pre[r];
    const unsigned long long nx = r + 1 < n_rec ? pre[r + 1] : total;
    uint32_t slot = (uint32_t)(pr >> 32);
    const uint32_t n_al = (uint32_t)(nx >> 32) - slot;
    uint64_t b = pr & ((1ull << 32) - 1);
    const uint64_t b_end = nx & ((1ull << 32) - 1);
    if ((uint64_t)slot + n_al > cap_allele || b_end > cap_bytes) return; // the host reports VTN_E_CAPACITY from the totals
    const uint8_t* p = text + line_off[r] + ref_off[r];
    // REF up to its tab
    while (*p != '\t') bytes[b++] = *p++;
    allele_off[++slot] = (uint32_t)b;
    ++p;
    if (n_al > 1)
    {
        for (;;)
        {
            const uint8_t ch = *p++;
            if (ch == ',' || ch == '\t')
            {
                allele_off[++slot] = (uint32_t)b;
                if (ch == '\t') break;
            }
            else
                bytes[b++] = ch;
        }
    }
}

} // namespace

// ---- device-resident text + output buffers sized by the caller's capacities ----
struct vtn_vcf_dtext
{
    uint64_t n_text = 0, cap_text = 0;
    bool open_tail = false;
    uint32_t cap_rec = 0, cap_allele = 0;
    uint64_t cap_bytes = 0;
    uint32_t n_rec = 0; // of the last run
    bool overflow_rec = false;
    uint64_t n_rec_needed = 0;
    uint8_t* text = nullptr;
    uint32_t *agg = nullptr, *nl = nullptr;
    uint64_t* line = nullptr;
    unsigned long long *cnt = nullptr, *agg2 = nullptr, *tot = nullptr;
    uint32_t *chrom = nullptr, *tail = nullptr, *ref = nullptr, *abeg = nullptr, *aoff = nullptr;
    int32_t* pos = nullptr;
    uint8_t *st = nullptr, *bytes = nullptr;
};

int vtn_vcf_dtext_destroy(vtn_ctx* ctx, vtn_vcf_dtext* h)
{
    if (!h) return VTN_OK;
    if (ctx) cudaSetDevice(ctx->device);
    cudaFree(h->text); cudaFree(h->agg); cudaFree(h->nl); cudaFree(h->line); cudaFree(h->cnt); cudaFree(h->agg2); cudaFree(h->tot);
    cudaFree(h->chrom); cudaFree(h->tail); cudaFree(h->ref); cudaFree(h->abeg); cudaFree(h->aoff); cudaFree(h->pos); cudaFree(h->st);
    cudaFree(h->bytes);
    delete h;
    return VTN_OK;
}

namespace
{
// device buffers for texts of up to cap_text bytes with the given record / allele capacities
int dtext_alloc(vtn_ctx* ctx, uint64_t cap_text, uint32_t cap_rec, uint32_t cap_allele, uint64_t cap_bytes, vtn_vcf_dtext** out)
{
    VTN_CUDA(ctx, cudaSetDevice(ctx->device));
    vtn_vcf_dtext* h = new vtn_vcf_dtext();
    h->cap_text = cap_text;
    h->cap_rec = cap_rec;
    h->cap_allele = cap_allele;
    h->cap_bytes = cap_bytes;
    const uint32_t n_tiles = (uint32_t)((cap_text + VTK_TILE - 1) / VTK_TILE);
    const uint32_t r_tiles = (cap_rec + VTK_RTILE - 1) / VTK_RTILE;
    const size_t nr = (size_t)cap_rec + 16;
    cudaError_t e = cudaSuccess;
#define A(p, bytes) if (e == cudaSuccess) e = cudaMalloc((void**)&(p), (bytes))
    A(h->text, cap_text + 64);
    A(h->agg, ((size_t)n_tiles + 1) * 4);
    A(h->nl, 4);
    A(h->line, nr * 8);
    A(h->cnt, nr * 8);
    A(h->agg2, ((size_t)r_tiles + 1) * 8);
    A(h->tot, 8);
    A(h->chrom, nr * 4);
    A(h->tail, nr * 4);
    A(h->ref, nr * 4);
    A(h->abeg, nr * 4);
    A(h->pos, nr * 4);
    A(h->st, nr);
    A(h->aoff, ((size_t)cap_allele + 16) * 4);
    A(h->bytes, (size_t)cap_bytes + 16);
#undef A
    if (e != cudaSuccess)
    {
        vtn_vcf_dtext_destroy(ctx, h);
        return vtn_set_err(ctx, VTN_E_NOMEM, std::string("vtn_vcf_dtext_create: ") + cudaGetErrorString(e));
    }
    *out = h;
    return VTN_OK;
}

int dtext_load(vtn_ctx* ctx, vtn_vcf_dtext* h, const char* text, uint64_t n_text)
{
    h->n_text = n_text;
    h->open_tail = n_text && text[n_text - 1] != '\n';
    if (n_text) VTN_CUDA(ctx, cudaMemcpyAsync(h->text, text, n_text, cudaMemcpyHostToDevice, ctx->stream));
    VTN_CUDA(ctx, cudaMemsetAsync(h->text + n_text, 0, 64, ctx->stream));
    VTN_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return VTN_OK;
}
} // namespace

int vtn_vcf_dtext_create(vtn_ctx* ctx, const char* text, uint64_t n_text, uint32_t cap_rec, uint32_t cap_allele, uint64_t cap_bytes,
                         vtn_vcf_dtext** out)
{
    if (!ctx || !out || (n_text && !text)) return vtn_set_err(ctx, VTN_E_ARG, "vtn_vcf_dtext_create: bad arguments");
    if (n_text >= (1ull << 32)) return vtn_set_err(ctx, VTN_E_ARG, "vtn_vcf_tokenize: text must be < 4 GiB per call (allele bytes of a batch are 32-bit offsets)");
    vtn_vcf_dtext* h = nullptr;
    int rc = dtext_alloc(ctx, n_text, cap_rec, cap_allele, cap_bytes, &h);
    if (rc) return rc;
    rc = dtext_load(ctx, h, text, n_text);
    if (rc)
    {
        vtn_vcf_dtext_destroy(ctx, h);
        return rc;
    }
    *out = h;
    return VTN_OK;
}

// The eight launches on the ctx stream.  One 4-byte readback (the line count) sizes the record-level grids; everything else,
// including the capacity checks, stays on the device until vtn_vcf_dtext_fetch.
int vtn_vcf_tokenize_resident(vtn_ctx* ctx, vtn_vcf_dtext* h)
{
    if (!ctx || !h) return VTN_E_ARG;
    cudaStream_t s = ctx->stream;
    h->n_rec = 0;
    h->overflow_rec = false;
    if (!h->n_text) return VTN_OK;
    const uint32_t n_tiles = (uint32_t)((h->n_text + VTK_TILE - 1) / VTK_TILE);
    vtk_count_kernel<<<n_tiles, VTK_THREADS, 0, s>>>(h->text, h->n_text, h->agg);
    vtk_scan_kernel<uint32_t><<<1, 1024, 0, s>>>(h->agg, n_tiles, h->nl);
    uint32_t n_nl = 0;
    VTN_CUDA(ctx, cudaMemcpyAsync(&n_nl, h->nl, sizeof n_nl, cudaMemcpyDeviceToHost, s));
    VTN_CUDA(ctx, cudaStreamSynchronize(s));
    ctx->launches += 2;
    h->n_rec_needed = (uint64_t)n_nl + (h->open_tail ? 1 : 0);
    if (h->n_rec_needed > h->cap_rec)
    {
        h->overflow_rec = true;
        return VTN_OK; // reported by fetch
    }
    const uint32_t n_rec = h->n_rec = (uint32_t)h->n_rec_needed;
    if (!n_rec) return VTN_OK;
    const uint32_t r_tiles = (n_rec + VTK_RTILE - 1) / VTK_RTILE;
    const uint64_t zero = 0;
    VTN_CUDA(ctx, cudaMemcpyAsync(h->line, &zero, 8, cudaMemcpyHostToDevice, s));
    VTN_CUDA(ctx, cudaMemcpyAsync(h->line + n_rec, &h->n_text, 8, cudaMemcpyHostToDevice, s)); // rewritten with the same value when the text ends in a line end
    vtk_lines_kernel<<<n_tiles, VTK_THREADS, 0, s>>>(h->text, h->n_text, h->agg, h->line, n_rec);
    const uint32_t g = (n_rec + VTK_THREADS - 1) / VTK_THREADS;
    VtkFields F{h->chrom, h->pos, h->tail, h->st, h->cnt, h->ref};
    vtk_fields_kernel<<<g, VTK_THREADS, 0, s>>>(h->text, h->n_text, h->line, n_rec, F);
    vtk_rsum_reduce_kernel<<<r_tiles, VTK_THREADS, 0, s>>>(h->cnt, n_rec, h->agg2);
    vtk_scan_kernel<unsigned long long><<<1, 1024, 0, s>>>(h->agg2, r_tiles, h->tot);
    vtk_rsum_apply_kernel<<<r_tiles, VTK_THREADS, 0, s>>>(h->cnt, n_rec, h->agg2, h->abeg);
    vtk_alleles_kernel<<<g, VTK_THREADS, 0, s>>>(h->text, h->line, n_rec, h->cnt, h->ref, h->st, h->aoff, h->bytes, h->cap_allele, h->cap_bytes, h->tot);
    VTN_CUDA(ctx, cudaGetLastError());
    ctx->launches += 6;
    return VTN_OK;
}

int vtn_vcf_dtext_fetch(vtn_ctx* ctx, vtn_vcf_dtext* h, vtn_vcf_fields* out)
{
    if (!ctx || !h || !out) return VTN_E_ARG;
    cudaStream_t s = ctx->stream;
    out->n_rec = 0;
    out->n_allele = 0;
    out->n_bytes = 0;
    if (h->overflow_rec)
    {
        out->n_rec = (uint32_t)(h->n_rec_needed > 0xffffffffull ? 0xffffffffu : h->n_rec_needed);
        return vtn_set_err(ctx, VTN_E_CAPACITY, "vtn_vcf_tokenize: more records than cap_rec (n_rec is set)");
    }
    const uint32_t n_rec = h->n_rec;
    if (!n_rec)
    {
        if (out->line_off) out->line_off[0] = 0;
        if (out->allele_beg) out->allele_beg[0] = 0;
        if (out->allele_off) out->allele_off[0] = 0;
        return VTN_OK;
    }
    if (n_rec > out->cap_rec) return vtn_set_err(ctx, VTN_E_CAPACITY, "vtn_vcf_dtext_fetch: cap_rec too small");
    unsigned long long total = 0;
    VTN_CUDA(ctx, cudaMemcpyAsync(&total, h->tot, 8, cudaMemcpyDeviceToHost, s));
    VTN_CUDA(ctx, cudaStreamSynchronize(s));
    const uint64_t n_allele = total >> 32, n_bytes = total & 0xffffffffull;
    out->n_rec = n_rec;
    out->n_allele = (uint32_t)n_allele;
    out->n_bytes = n_bytes;
    if (n_allele > h->cap_allele || n_bytes > h->cap_bytes || n_allele > out->cap_allele || n_bytes > out->cap_bytes)
        return vtn_set_err(ctx, VTN_E_CAPACITY, "vtn_vcf_tokenize: allele capacity too small (n_allele / n_bytes are set)");
    VTN_CUDA(ctx, cudaMemcpyAsync(out->line_off, h->line, ((size_t)n_rec + 1) * 8, cudaMemcpyDeviceToHost, s));
    VTN_CUDA(ctx, cudaMemcpyAsync(out->chrom_len, h->chrom, (size_t)n_rec * 4, cudaMemcpyDeviceToHost, s));
    VTN_CUDA(ctx, cudaMemcpyAsync(out->pos1, h->pos, (size_t)n_rec * 4, cudaMemcpyDeviceToHost, s));
    VTN_CUDA(ctx, cudaMemcpyAsync(out->tail_off, h->tail, (size_t)n_rec * 4, cudaMemcpyDeviceToHost, s));
    VTN_CUDA(ctx, cudaMemcpyAsync(out->status, h->st, (size_t)n_rec, cudaMemcpyDeviceToHost, s));
    VTN_CUDA(ctx, cudaMemcpyAsync(out->allele_beg, h->abeg, ((size_t)n_rec + 1) * 4, cudaMemcpyDeviceToHost, s));
    VTN_CUDA(ctx, cudaMemcpyAsync(out->allele_off, h->aoff, ((size_t)n_allele + 1) * 4, cudaMemcpyDeviceToHost, s));
    if (n_bytes) VTN_CUDA(ctx, cudaMemcpyAsync(out->bytes, h->bytes, (size_t)n_bytes, cudaMemcpyDeviceToHost, s));
    VTN_CUDA(ctx, cudaStreamSynchronize(s));
    return VTN_OK;
}

// Host-buffer call: H2D of the text, the kernels above, D2H of every output array.  The device buffers live in the context and are
// reused by the next call when they are large enough (a file is tokenized batch after batch with the same capacities).
int vtn_vcf_tokenize(vtn_ctx* ctx, const char* text, uint64_t n_text, vtn_vcf_fields* out)
{
    if (!ctx || !out || (n_text && !text)) return vtn_set_err(ctx, VTN_E_ARG, "vtn_vcf_tokenize: bad arguments");
    if (n_text >= (1ull << 32)) return vtn_set_err(ctx, VTN_E_ARG, "vtn_vcf_tokenize: text must be < 4 GiB per call (allele bytes of a batch are 32-bit offsets)");
    vtn_vcf_dtext*& h = ctx->tok_cache;
    if (h && (h->cap_text < n_text || h->cap_rec < out->cap_rec || h->cap_allele < out->cap_allele || h->cap_bytes < out->cap_bytes))
    {
        vtn_vcf_dtext_destroy(ctx, h);
        h = nullptr;
    }
    int rc = VTN_OK;
    if (!h)
    {
        auto grow = [](uint64_t v) { return v + v / 4 + 1024; };
        const uint64_t cr = std::min<uint64_t>(grow(out->cap_rec), 0xffffffffull), ca = std::min<uint64_t>(grow(out->cap_allele), 0xffffffffull);
        rc = dtext_alloc(ctx, grow(n_text), (uint32_t)cr, (uint32_t)ca, grow(out->cap_bytes), &h);
        if (rc) return rc;
    }
    rc = dtext_load(ctx, h, text, n_text);
    if (rc == VTN_OK) rc = vtn_vcf_tokenize_resident(ctx, h);
    if (rc == VTN_OK) rc = vtn_vcf_dtext_fetch(ctx, h, out);
    return rc;
}
