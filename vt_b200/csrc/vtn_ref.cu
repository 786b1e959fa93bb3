// vtn_ref.cu -- reference genome: FASTA/.fai reading, packing into the 16-byte-per-32-base group layout
// (vtn_internal.h), upload to HBM, and a faidx_fetch_seq-style read-back kernel.
// Replaces fai_load (variant_manip.cpp:35) and faidx_fetch_seq / faidx_fetch_uc_seq
// (lib/htslib/faidx.c:832-881, hts_utils.cpp:41-51) for the normalize path.
#include <ctype.h>
#include <stdio.h>
#include <string.h>
#include <zlib.h>
#include <algorithm>
#include <thread>
#include "vtn_internal.h"
#include "vtn_ref_dev.cuh"

namespace
{

struct ExcRun
{
    uint64_t start;
    uint32_t len;
    uint8_t byte;
};

inline int base_code(uint8_t c)
{
    switch (c)
    {
    case 'A': case 'a': return 0;
    case 'C': case 'c': return 1;
    case 'T': case 't': return 2; // (ascii >> 1) & 3, so allele bytes convert with one shift+mask
    case 'G': case 'g': return 3;
    default: return -1;
    }
}

// pack bases [b0, b1) of one contig (b0 multiple of 32) into groups; collects exception runs in order
void pack_range(const uint8_t* seq, int64_t b0, int64_t b1, uint64_t gpos0, uint4* groups, std::vector<ExcRun>& exc)
{
    for (int64_t g0 = b0; g0 < b1; g0 += 32)
    {
        uint64_t code = 0;
        uint32_t nonstd = 0, lower = 0;
        int64_t e = std::min<int64_t>(g0 + 32, b1);
        for (int64_t p = g0; p < e; ++p)
        {
            uint8_t c = seq[p];
            int j = (int)(p - g0);
            int k = base_code(c);
            if (k >= 0)
            {
                code |= (uint64_t)k << (2 * j);
                if (c >= 'a')
                {
                    nonstd |= 1u << j;
                    lower |= 1u << j;
                }
            }
            else
            {
                nonstd |= 1u << j;
                uint64_t gp = gpos0 + (uint64_t)p;
                if (!exc.empty() && exc.back().byte == c && exc.back().start + exc.back().len == gp && exc.back().len < 0xffffffffu)
                    exc.back().len++;
                else
                    exc.push_back(ExcRun{gp, 1u, c});
            }
        }
        uint4 q;
        q.x = (uint32_t)code;
        q.y = (uint32_t)(code >> 32);
        q.z = nonstd;
        q.w = lower;
        groups[g0 >> 5] = q;
    }
}

__global__ void vtn_ref_fetch_kernel(VtnDevRef R, int rid, int64_t beg, int64_t n, int upper, uint8_t* out)
{
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint64_t gbase = R.contig_goff[rid] << 5;
    VtnRefCursor cur;
    out[i] = upper ? cur.uc(R, gbase + (uint64_t)(beg + i)) : cur.raw(R, gbase + (uint64_t)(beg + i));
}

void free_ref(vtn_ctx* ctx)
{
    cudaFree(ctx->d_groups);
    cudaFree(ctx->d_contig_goff);
    cudaFree(ctx->d_contig_len);
    cudaFree(ctx->d_exc_start);
    cudaFree(ctx->d_exc_len);
    cudaFree(ctx->d_exc_byte);
    ctx->d_groups = ctx->d_contig_goff = ctx->d_contig_len = ctx->d_exc_start = ctx->d_exc_len = ctx->d_exc_byte = nullptr;
    ctx->contigs.clear();
    ctx->ref_loaded = false;
    ctx->ref_bytes = 0;
}

// read a whole (possibly gzip/bgzip) file into memory
bool slurp(const char* path, std::vector<uint8_t>& buf)
{
    gzFile f = gzopen(path, "rb");
    if (!f) return false;
    gzbuffer(f, 1 << 20);
    size_t cap = 1 << 26, n = 0;
    buf.resize(cap);
    for (;;)
    {
        if (n == cap)
        {
            cap *= 2;
            buf.resize(cap);
        }
        int want = (int)std::min<size_t>(cap - n, 1u << 30);
        int got = gzread(f, buf.data() + n, want);
        if (got < 0)
        {
            gzclose(f);
            return false;
        }
        if (got == 0) break;
        n += (size_t)got;
    }
    gzclose(f);
    buf.resize(n);
    return true;
}

struct FaiEntry
{
    std::string name;
    int64_t len, offset, line_blen, line_len;
};

bool read_fai(const std::string& path, std::vector<FaiEntry>& out)
{
    FILE* f = fopen(path.c_str(), "r");
    if (!f) return false;
    char line[65536];
    while (fgets(line, sizeof line, f))
    {
        FaiEntry e;
        char* tab = strchr(line, '\t');
        if (!tab) continue;
        e.name.assign(line, tab - line);
        long long a, b, c, d;
        if (sscanf(tab + 1, "%lld\t%lld\t%lld\t%lld", &a, &b, &c, &d) != 4) continue;
        e.len = a; e.offset = b; e.line_blen = c; e.line_len = d;
        out.push_back(e);
    }
    fclose(f);
    return true;
}

// build the index from the FASTA text itself (what fai_load does when no .fai exists; faidx.c fai_build_core)
void build_fai(const std::vector<uint8_t>& t, std::vector<FaiEntry>& out)
{
    size_t i = 0, n = t.size();
    while (i < n)
    {
        if (t[i] != '>') { ++i; continue; }
        size_t j = i + 1;
        while (j < n && !isspace(t[j])) ++j;
        FaiEntry e;
        e.name.assign((const char*)&t[i + 1], j - i - 1);
        while (j < n && t[j] != '\n') ++j;
        size_t p = j + 1;
        e.offset = (int64_t)p;
        e.len = 0; e.line_blen = 0; e.line_len = 0;
        bool first = true;
        while (p < n && t[p] != '>')
        {
            size_t q = p;
            int64_t bl = 0;
            while (q < n && t[q] != '\n') { if (isgraph(t[q])) ++bl; ++q; }
            int64_t ll = (int64_t)(q - p) + (q < n ? 1 : 0);
            if (first && bl) { e.line_blen = bl; e.line_len = ll; first = false; }
            e.len += bl;
            p = q + 1;
        }
        out.push_back(e);
        i = p;
    }
}

} // namespace

int vtn_ref_load_memory(vtn_ctx* ctx, int32_t n_contig, const char* const* names, const int64_t* len, const uint8_t* const* seq)
{
    if (!ctx || n_contig < 0 || (n_contig && (!len || !seq))) return vtn_set_err(ctx, VTN_E_ARG, "vtn_ref_load_memory: bad argument");
    VTN_CUDA(ctx, cudaSetDevice(ctx->device));
    free_ref(ctx);
    uint64_t n_groups = 0;
    for (int c = 0; c < n_contig; ++c)
    {
        VtnContig k;
        k.name = names && names[c] ? names[c] : std::to_string(c);
        k.len = len[c];
        k.goff = n_groups;
        n_groups += (uint64_t)((len[c] + 31) / 32) + 2; // +2 zero pad groups: window loads may touch g+1, g+2
        ctx->contigs.push_back(k);
    }
    n_groups += 2;
    uint4* h_groups = nullptr;
    VTN_CUDA(ctx, cudaMallocHost(&h_groups, n_groups * sizeof(uint4)));
    memset(h_groups, 0, n_groups * sizeof(uint4));
    // pack in parallel: split every contig into 32-base-aligned chunks
    unsigned nthr = std::max(1u, std::min(16u, std::thread::hardware_concurrency()));
    struct Job { int c; int64_t b0, b1; std::vector<ExcRun> exc; };
    std::vector<Job> jobs;
    const int64_t chunk = 1 << 24;
    for (int c = 0; c < n_contig; ++c)
        for (int64_t b = 0; b < len[c]; b += chunk) jobs.push_back(Job{c, b, std::min(len[c], b + chunk), {}});
    std::vector<std::thread> th;
    for (unsigned t = 0; t < nthr; ++t)
        th.emplace_back([&, t]() {
            for (size_t j = t; j < jobs.size(); j += nthr)
            {
                Job& J = jobs[j];
                uint64_t goff = ctx->contigs[J.c].goff;
                pack_range(seq[J.c], J.b0, J.b1, goff << 5, h_groups + goff, J.exc);
            }
        });
    for (auto& t : th) t.join();
    // merge exception runs (jobs are in global position order); fuse runs split at chunk borders
    std::vector<ExcRun> exc;
    for (auto& J : jobs)
        for (auto& e : J.exc)
        {
            if (!exc.empty() && exc.back().byte == e.byte && exc.back().start + exc.back().len == e.start &&
                (uint64_t)exc.back().len + e.len <= 0xffffffffull)
                exc.back().len += e.len;
            else
                exc.push_back(e);
        }
    std::vector<uint64_t> es(exc.size() + 1);
    std::vector<uint32_t> el(exc.size() + 1);
    std::vector<uint8_t> eb(exc.size() + 1);
    for (size_t i = 0; i < exc.size(); ++i) { es[i] = exc[i].start; el[i] = exc[i].len; eb[i] = exc[i].byte; }
    std::vector<uint64_t> goffs(n_contig + 1);
    std::vector<int64_t> lens(n_contig + 1);
    for (int c = 0; c < n_contig; ++c) { goffs[c] = ctx->contigs[c].goff; lens[c] = ctx->contigs[c].len; }

    cudaError_t e = cudaMalloc(&ctx->d_groups, n_groups * sizeof(uint4));
    if (e == cudaSuccess) e = cudaMemcpy(ctx->d_groups, h_groups, n_groups * sizeof(uint4), cudaMemcpyHostToDevice);
    cudaFreeHost(h_groups);
    if (e != cudaSuccess) return vtn_set_err(ctx, VTN_E_CUDA, std::string("reference upload: ") + cudaGetErrorString(e));
    VTN_CUDA(ctx, cudaMalloc(&ctx->d_contig_goff, (n_contig + 1) * sizeof(uint64_t)));
    VTN_CUDA(ctx, cudaMalloc(&ctx->d_contig_len, (n_contig + 1) * sizeof(int64_t)));
    VTN_CUDA(ctx, cudaMalloc(&ctx->d_exc_start, es.size() * sizeof(uint64_t)));
    VTN_CUDA(ctx, cudaMalloc(&ctx->d_exc_len, el.size() * sizeof(uint32_t)));
    VTN_CUDA(ctx, cudaMalloc(&ctx->d_exc_byte, eb.size()));
    VTN_CUDA(ctx, cudaMemcpy(ctx->d_contig_goff, goffs.data(), (n_contig + 1) * sizeof(uint64_t), cudaMemcpyHostToDevice));
    VTN_CUDA(ctx, cudaMemcpy(ctx->d_contig_len, lens.data(), (n_contig + 1) * sizeof(int64_t), cudaMemcpyHostToDevice));
    VTN_CUDA(ctx, cudaMemcpy(ctx->d_exc_start, es.data(), es.size() * sizeof(uint64_t), cudaMemcpyHostToDevice));
    VTN_CUDA(ctx, cudaMemcpy(ctx->d_exc_len, el.data(), el.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
    VTN_CUDA(ctx, cudaMemcpy(ctx->d_exc_byte, eb.data(), eb.size(), cudaMemcpyHostToDevice));
    ctx->dref.groups = (const uint4*)ctx->d_groups;
    ctx->dref.contig_goff = (const uint64_t*)ctx->d_contig_goff;
    ctx->dref.contig_len = (const int64_t*)ctx->d_contig_len;
    ctx->dref.exc_start = (const uint64_t*)ctx->d_exc_start;
    ctx->dref.exc_len = (const uint32_t*)ctx->d_exc_len;
    ctx->dref.exc_byte = (const uint8_t*)ctx->d_exc_byte;
    ctx->dref.n_exc = (uint32_t)exc.size();
    ctx->dref.n_contig = n_contig;
    ctx->ref_bytes = (int64_t)(n_groups * sizeof(uint4) + exc.size() * 13);
    ctx->ref_n_groups = n_groups;
    ctx->ref_loaded = true;
    return VTN_OK;
}

// The reference is packed once and uploaded once; the other devices of a multi-device run get it from the first one by peer
// copies (NVLink / NVSwitch when the devices can reach each other, staged through the host otherwise) -- SURVEY 8e.
int vtn_ref_clone(vtn_ctx* dst, const vtn_ctx* src)
{
    if (!dst || !src || !src->ref_loaded) return vtn_set_err(dst, VTN_E_STATE, "vtn_ref_clone: source has no reference");
    VTN_CUDA(dst, cudaSetDevice(dst->device));
    free_ref(dst);
    dst->contigs = src->contigs;
    const size_t nc = src->contigs.size() + 1, ne = (size_t)src->dref.n_exc + 1;
    const size_t sz[6] = {(size_t)src->ref_n_groups * sizeof(uint4), nc * sizeof(uint64_t), nc * sizeof(int64_t), ne * sizeof(uint64_t), ne * sizeof(uint32_t), ne};
    void** d[6] = {&dst->d_groups, &dst->d_contig_goff, &dst->d_contig_len, &dst->d_exc_start, &dst->d_exc_len, &dst->d_exc_byte};
    void* const sp[6] = {src->d_groups, src->d_contig_goff, src->d_contig_len, src->d_exc_start, src->d_exc_len, src->d_exc_byte};
    if (dst->device != src->device)
    {
        int can = 0;
        if (cudaDeviceCanAccessPeer(&can, dst->device, src->device) == cudaSuccess && can) cudaDeviceEnablePeerAccess(src->device, 0);
        cudaGetLastError(); // "already enabled" is fine; without peer access the copy below is staged by the driver
    }
    for (int i = 0; i < 6; ++i)
    {
        VTN_CUDA(dst, cudaMalloc(d[i], sz[i]));
        VTN_CUDA(dst, cudaMemcpyPeerAsync(*d[i], dst->device, sp[i], src->device, sz[i], dst->stream));
    }
    VTN_CUDA(dst, cudaStreamSynchronize(dst->stream));
    dst->dref = src->dref;
    dst->dref.groups = (const uint4*)dst->d_groups;
    dst->dref.contig_goff = (const uint64_t*)dst->d_contig_goff;
    dst->dref.contig_len = (const int64_t*)dst->d_contig_len;
    dst->dref.exc_start = (const uint64_t*)dst->d_exc_start;
    dst->dref.exc_len = (const uint32_t*)dst->d_exc_len;
    dst->dref.exc_byte = (const uint8_t*)dst->d_exc_byte;
    dst->ref_bytes = src->ref_bytes;
    dst->ref_n_groups = src->ref_n_groups;
    dst->ref_loaded = true;
    return VTN_OK;
}

int vtn_ref_load_fasta(vtn_ctx* ctx, const char* fasta_path)
{
    if (!ctx || !fasta_path) return vtn_set_err(ctx, VTN_E_ARG, "vtn_ref_load_fasta: bad argument");
    std::vector<uint8_t> text;
    if (!slurp(fasta_path, text)) return vtn_set_err(ctx, VTN_E_IO, std::string("Cannot load genome index: ") + fasta_path);
    std::vector<FaiEntry> fai;
    if (!read_fai(std::string(fasta_path) + ".fai", fai)) build_fai(text, fai);
    if (fai.empty()) return vtn_set_err(ctx, VTN_E_IO, std::string("Cannot load genome index: ") + fasta_path);
    // materialise contig bases exactly as fai_retrieve streams them: isgraph bytes from the seek offset
    std::vector<std::vector<uint8_t>> seqs(fai.size());
    std::vector<const char*> names(fai.size());
    std::vector<int64_t> lens(fai.size());
    std::vector<const uint8_t*> ptrs(fai.size());
    for (size_t c = 0; c < fai.size(); ++c)
    {
        const FaiEntry& e = fai[c];
        std::vector<uint8_t>& s = seqs[c];
        s.resize((size_t)std::max<int64_t>(e.len, 0));
        size_t p = (size_t)e.offset, k = 0;
        while (k < s.size() && p < text.size())
        {
            uint8_t ch = text[p++];
            if (isgraph(ch)) s[k++] = ch;
        }
        if (k < s.size()) return vtn_set_err(ctx, VTN_E_IO, std::string("FASTA shorter than its index: ") + e.name);
        names[c] = e.name.c_str();
        lens[c] = e.len;
        ptrs[c] = s.data();
    }
    text.clear();
    text.shrink_to_fit();
    return vtn_ref_load_memory(ctx, (int32_t)fai.size(), names.data(), lens.data(), ptrs.data());
}

int vtn_ref_n_contig(const vtn_ctx* ctx) { return ctx ? (int)ctx->contigs.size() : 0; }

int vtn_ref_contig_id(const vtn_ctx* ctx, const char* name)
{
    if (!ctx || !name) return -1;
    for (size_t i = 0; i < ctx->contigs.size(); ++i)
        if (ctx->contigs[i].name == name) return (int)i;
    return -1;
}

int64_t vtn_ref_contig_len(const vtn_ctx* ctx, int rid)
{
    if (!ctx || rid < 0 || rid >= (int)ctx->contigs.size()) return -1;
    return ctx->contigs[rid].len;
}

const char* vtn_ref_contig_name(const vtn_ctx* ctx, int rid)
{
    if (!ctx || rid < 0 || rid >= (int)ctx->contigs.size()) return nullptr;
    return ctx->contigs[rid].name.c_str();
}

int64_t vtn_ref_resident_bytes(const vtn_ctx* ctx) { return ctx ? ctx->ref_bytes : 0; }

int64_t vtn_ref_fetch(vtn_ctx* ctx, int rid, int64_t beg0, int64_t end0, int upper, uint8_t* out, int64_t cap)
{
    if (!ctx || !ctx->ref_loaded) return vtn_set_err(ctx, VTN_E_STATE, "no reference loaded");
    if (rid < 0 || rid >= (int)ctx->contigs.size()) return vtn_set_err(ctx, VTN_E_ARG, "unknown contig");
    int64_t clen = ctx->contigs[rid].len;
    if (clen <= 0) return vtn_set_err(ctx, VTN_E_ARG, "empty contig");
    if (end0 < beg0) beg0 = end0; // faidx.c:846-857
    beg0 = beg0 < 0 ? 0 : (beg0 >= clen ? clen - 1 : beg0);
    end0 = end0 < 0 ? 0 : (end0 >= clen ? clen - 1 : end0);
    int64_t n = end0 - beg0 + 1;
    if (n > cap) return vtn_set_err(ctx, VTN_E_CAPACITY, "vtn_ref_fetch: output too small");
    VTN_CUDA(ctx, cudaSetDevice(ctx->device));
    uint8_t* d = nullptr;
    VTN_CUDA(ctx, cudaMalloc(&d, (size_t)n));
    vtn_ref_fetch_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(ctx->dref, rid, beg0, n, upper, d);
    ctx->launches++;
    cudaError_t e = cudaMemcpyAsync(out, d, (size_t)n, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cudaFree(d);
    if (e != cudaSuccess) return vtn_set_err(ctx, VTN_E_CUDA, cudaGetErrorString(e));
    return n;
}
