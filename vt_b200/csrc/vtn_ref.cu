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
#include "vtn_fasta_pack.h"
#include "vtn_internal.h"
#include "vtn_ref_dev.cuh"

using vtf::ExcRun;

// what a lazily opened FASTA keeps on the host: the mapped text, its index, which contigs are resident, their exception runs
struct VtnRefSource
{
    vtf::FastaText text;
    std::vector<vtf::FaiEntry> fai;
    std::vector<char> loaded;
    std::vector<std::vector<ExcRun>> exc;
};

namespace
{

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
    ctx->ref_missing = 0;
    delete ctx->ref_src;
    ctx->ref_src = nullptr;
}


unsigned host_threads()
{
    unsigned n = std::max(1u, std::min(32u, std::thread::hardware_concurrency()));
    if (const char* e = getenv("VTN_HOST_THREADS"))
        if (atoi(e) > 0) n = (unsigned)atoi(e);
    return n;
}

// contig table + device arrays for a reference whose contig names and lengths are known; no bases resident yet
int ref_alloc_tables(vtn_ctx* ctx, int32_t n_contig, const char* const* names, const int64_t* len)
{
    VTN_CUDA(ctx, cudaSetDevice(ctx->device));
    free_ref(ctx);
    uint64_t n_groups = 0;
    for (int c = 0; c < n_contig; ++c)
    {
        VtnContig k;
        k.name = names && names[c] ? names[c] : std::to_string(c);
        k.len = len[c] > 0 ? len[c] : 0;
        k.goff = n_groups;
        n_groups += (uint64_t)((k.len + 31) / 32) + 2; // +2 zero pad groups: window loads may touch g+1, g+2
        ctx->contigs.push_back(k);
    }
    n_groups += 2;
    std::vector<uint64_t> goffs(n_contig + 1);
    std::vector<int64_t> lens(n_contig + 1);
    for (int c = 0; c < n_contig; ++c) goffs[c] = ctx->contigs[c].goff, lens[c] = ctx->contigs[c].len;
    cudaError_t e = cudaMalloc(&ctx->d_groups, n_groups * sizeof(uint4));
    if (e == cudaSuccess) e = cudaMemsetAsync(ctx->d_groups, 0, n_groups * sizeof(uint4), ctx->stream);
    if (e != cudaSuccess) return vtn_set_err(ctx, VTN_E_NOMEM, std::string("reference allocation: ") + cudaGetErrorString(e));
    VTN_CUDA(ctx, cudaMalloc(&ctx->d_contig_goff, (n_contig + 1) * sizeof(uint64_t)));
    VTN_CUDA(ctx, cudaMalloc(&ctx->d_contig_len, (n_contig + 1) * sizeof(int64_t)));
    VTN_CUDA(ctx, cudaMemcpyAsync(ctx->d_contig_goff, goffs.data(), (n_contig + 1) * sizeof(uint64_t), cudaMemcpyHostToDevice, ctx->stream));
    VTN_CUDA(ctx, cudaMemcpyAsync(ctx->d_contig_len, lens.data(), (n_contig + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, ctx->stream));
    VTN_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->dref.groups = (const uint4*)ctx->d_groups;
    ctx->dref.contig_goff = (const uint64_t*)ctx->d_contig_goff;
    ctx->dref.contig_len = (const int64_t*)ctx->d_contig_len;
    ctx->dref.exc_start = nullptr, ctx->dref.exc_len = nullptr, ctx->dref.exc_byte = nullptr;
    ctx->dref.n_exc = 0;
    ctx->dref.n_contig = n_contig;
    ctx->ref_n_groups = n_groups;
    ctx->ref_bytes = (int64_t)(n_groups * sizeof(uint4));
    return VTN_OK;
}

// replace the device exception-run table (all resident contigs' runs, in global position order)
int ref_upload_exc(vtn_ctx* ctx, const std::vector<ExcRun>& exc)
{
    std::vector<uint64_t> es(exc.size() + 1);
    std::vector<uint32_t> el(exc.size() + 1);
    std::vector<uint8_t> eb(exc.size() + 1);
    for (size_t i = 0; i < exc.size(); ++i) es[i] = exc[i].start, el[i] = exc[i].len, eb[i] = exc[i].byte;
    VTN_CUDA(ctx, cudaStreamSynchronize(ctx->stream)); // nothing in flight may still read the old table
    cudaFree(ctx->d_exc_start), cudaFree(ctx->d_exc_len), cudaFree(ctx->d_exc_byte);
    ctx->d_exc_start = ctx->d_exc_len = ctx->d_exc_byte = nullptr;
    VTN_CUDA(ctx, cudaMalloc(&ctx->d_exc_start, es.size() * sizeof(uint64_t)));
    VTN_CUDA(ctx, cudaMalloc(&ctx->d_exc_len, el.size() * sizeof(uint32_t)));
    VTN_CUDA(ctx, cudaMalloc(&ctx->d_exc_byte, eb.size()));
    VTN_CUDA(ctx, cudaMemcpy(ctx->d_exc_start, es.data(), es.size() * sizeof(uint64_t), cudaMemcpyHostToDevice));
    VTN_CUDA(ctx, cudaMemcpy(ctx->d_exc_len, el.data(), el.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
    VTN_CUDA(ctx, cudaMemcpy(ctx->d_exc_byte, eb.data(), eb.size(), cudaMemcpyHostToDevice));
    ctx->dref.exc_start = (const uint64_t*)ctx->d_exc_start;
    ctx->dref.exc_len = (const uint32_t*)ctx->d_exc_len;
    ctx->dref.exc_byte = (const uint8_t*)ctx->d_exc_byte;
    ctx->dref.n_exc = (uint32_t)exc.size();
    ctx->ref_bytes = (int64_t)(ctx->ref_n_groups * sizeof(uint4) + exc.size() * 13);
    return VTN_OK;
}

// pack the listed contigs of an opened FASTA (all worker threads on one contig after the other) and upload each one as soon as
// it is complete, while the workers pack the next
int ref_load_contigs(vtn_ctx* ctx, const std::vector<int>& which)
{
    VtnRefSource* S = ctx->ref_src;
    if (which.empty()) return VTN_OK;
    VTN_CUDA(ctx, cudaSetDevice(ctx->device));
    const size_t nc = ctx->contigs.size();
    std::vector<uint64_t> goff(nc), hoff(nc, 0);
    for (size_t c = 0; c < nc; ++c) goff[c] = ctx->contigs[c].goff;
    uint64_t n_host = 0;
    for (int c : which)
    {
        hoff[c] = n_host;
        n_host += (uint64_t)((ctx->contigs[c].len + 31) / 32);
    }
    vtf::Group* h = (vtf::Group*)malloc(std::max<uint64_t>(n_host, 1) * sizeof(vtf::Group));
    if (!h) return vtn_set_err(ctx, VTN_E_NOMEM, "reference packing: out of host memory");
    cudaError_t ce = cudaSuccess;
    const int bad = vtf::pack_contigs_ol(S->text.data, S->text.size, S->fai, which, goff, hoff, h, S->exc, host_threads(), [&](int c) {
        const uint64_t ng = (uint64_t)((ctx->contigs[c].len + 31) / 32);
        if (ng && ce == cudaSuccess)
            ce = cudaMemcpyAsync((uint4*)ctx->d_groups + goff[c], h + hoff[c], ng * sizeof(uint4), cudaMemcpyHostToDevice, ctx->stream);
    });
    if (ce == cudaSuccess) ce = cudaStreamSynchronize(ctx->stream);
    free(h);
    if (bad >= 0) return vtn_set_err(ctx, VTN_E_IO, std::string("FASTA shorter than its index: ") + S->fai[bad].name);
    if (ce != cudaSuccess) return vtn_set_err(ctx, VTN_E_CUDA, std::string("reference upload: ") + cudaGetErrorString(ce));
    for (int c : which)
        if (!S->loaded[c])
        {
            S->loaded[c] = 1;
            ctx->ref_missing--;
        }
    std::vector<ExcRun> all;
    for (size_t c = 0; c < nc; ++c)
        if (S->loaded[c]) all.insert(all.end(), S->exc[c].begin(), S->exc[c].end()); // contigs never touch: no fusing across them
    return ref_upload_exc(ctx, all);
}

int ref_open(vtn_ctx* ctx, const char* fasta_path)
{
    VtnRefSource* S = new VtnRefSource();
    if (!S->text.open(fasta_path))
    {
        delete S;
        return vtn_set_err(ctx, VTN_E_IO, std::string("Cannot load genome index: ") + fasta_path);
    }
    if (!vtf::read_fai(std::string(fasta_path) + ".fai", S->fai)) vtf::build_fai(S->text.data, S->text.size, S->fai);
    if (S->fai.empty())
    {
        delete S;
        return vtn_set_err(ctx, VTN_E_IO, std::string("Cannot load genome index: ") + fasta_path);
    }
    const size_t nc = S->fai.size();
    std::vector<const char*> names(nc);
    std::vector<int64_t> lens(nc);
    for (size_t c = 0; c < nc; ++c) names[c] = S->fai[c].name.c_str(), lens[c] = S->fai[c].len;
    int rc = ref_alloc_tables(ctx, (int32_t)nc, names.data(), lens.data());
    if (rc)
    {
        delete S;
        return rc;
    }
    S->loaded.assign(nc, 0);
    S->exc.resize(nc);
    ctx->ref_src = S;
    ctx->ref_missing = (uint32_t)nc;
    rc = ref_upload_exc(ctx, {});
    if (rc) return rc;
    ctx->ref_loaded = true;
    return VTN_OK;
}

} // namespace

void vtn_ref_release(vtn_ctx* ctx) { free_ref(ctx); }

int vtn_ref_load_memory(vtn_ctx* ctx, int32_t n_contig, const char* const* names, const int64_t* len, const uint8_t* const* seq)
{
    if (!ctx || n_contig < 0 || (n_contig && (!len || !seq))) return vtn_set_err(ctx, VTN_E_ARG, "vtn_ref_load_memory: bad argument");
    int rc = ref_alloc_tables(ctx, n_contig, names, len);
    if (rc) return rc;
    const uint64_t n_groups = ctx->ref_n_groups;
    vtf::Group* h_groups = (vtf::Group*)calloc(n_groups, sizeof(vtf::Group));
    if (!h_groups) return vtn_set_err(ctx, VTN_E_NOMEM, "reference packing: out of host memory");
    // pack in parallel: split every contig into 32-base-aligned chunks
    struct Job { int c; int64_t b0, b1; std::vector<ExcRun> exc; };
    std::vector<Job> jobs;
    const int64_t chunk = 1 << 22;
    for (int c = 0; c < n_contig; ++c)
        for (int64_t b = 0; b < len[c]; b += chunk) jobs.push_back(Job{c, b, std::min(len[c], b + chunk), {}});
    const unsigned nthr = std::max<unsigned>(1, std::min<size_t>(host_threads(), jobs.size()));
    std::vector<std::thread> th;
    for (unsigned t = 0; t < nthr; ++t)
        th.emplace_back([&, t]() {
            for (size_t j = t; j < jobs.size(); j += nthr)
            {
                Job& J = jobs[j];
                const uint64_t goff = ctx->contigs[J.c].goff;
                vtf::pack_bases_ol(seq[J.c] + J.b0, J.b0, J.b1, goff << 5, h_groups + goff, J.exc);
            }
        });
    for (auto& t : th) t.join();
    std::vector<ExcRun> exc; // jobs are in global position order
    for (auto& J : jobs) vtf::append_runs(exc, J.exc);
    cudaError_t e = cudaMemcpy(ctx->d_groups, h_groups, n_groups * sizeof(uint4), cudaMemcpyHostToDevice);
    free(h_groups);
    if (e != cudaSuccess) return vtn_set_err(ctx, VTN_E_CUDA, std::string("reference upload: ") + cudaGetErrorString(e));
    rc = ref_upload_exc(ctx, exc);
    if (rc) return rc;
    ctx->ref_loaded = true;
    return VTN_OK;
}

// The reference is packed once and uploaded once; the other devices of a multi-device run get it from the first one by peer
// copies (NVLink / NVSwitch when the devices can reach each other, staged through the host otherwise) -- SURVEY 8e.
int vtn_ref_clone(vtn_ctx* dst, const vtn_ctx* src)
{
    if (!dst || !src || !src->ref_loaded) return vtn_set_err(dst, VTN_E_STATE, "vtn_ref_clone: source has no reference");
    if (src->ref_missing) return vtn_set_err(dst, VTN_E_STATE, "vtn_ref_clone: the source reference is only partly resident (vtn_ref_require_all first)");
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

// Index only: the contig table comes from <fasta>.fai (built from the text when the file is absent), the group array is allocated
// for the whole genome, and a contig's bases are packed and uploaded the first time a batch names it (vtn_ref_require; the
// host-buffer entry points do that themselves).  This is the resident counterpart of what fai_load + faidx_fetch_seq give the
// reference: no cost for contigs a run never touches.
int vtn_ref_open_fasta(vtn_ctx* ctx, const char* fasta_path)
{
    if (!ctx || !fasta_path) return vtn_set_err(ctx, VTN_E_ARG, "vtn_ref_open_fasta: bad argument");
    return ref_open(ctx, fasta_path);
}

int vtn_ref_require(vtn_ctx* ctx, int rid)
{
    if (!ctx || !ctx->ref_loaded) return vtn_set_err(ctx, VTN_E_STATE, "no reference loaded");
    if (!ctx->ref_missing || rid < 0 || rid >= (int)ctx->contigs.size() || ctx->ref_src->loaded[rid]) return VTN_OK;
    return ref_load_contigs(ctx, std::vector<int>(1, rid));
}

int vtn_ref_require_all(vtn_ctx* ctx)
{
    if (!ctx || !ctx->ref_loaded) return vtn_set_err(ctx, VTN_E_STATE, "no reference loaded");
    if (!ctx->ref_missing) return VTN_OK;
    std::vector<int> which;
    for (size_t c = 0; c < ctx->contigs.size(); ++c)
        if (!ctx->ref_src->loaded[c]) which.push_back((int)c);
    int rc = ref_load_contigs(ctx, which);
    if (rc == VTN_OK)
    { // nothing left to read: give the mapping back
        delete ctx->ref_src;
        ctx->ref_src = nullptr;
    }
    return rc;
}

// every contig a batch names (runs of equal ids cost one comparison per record)
int vtn_ref_ensure(vtn_ctx* ctx, uint64_t n, const int32_t* rid)
{
    if (!ctx->ref_missing || !rid) return VTN_OK;
    std::vector<int> which;
    int32_t prev = -1;
    const int32_t nc = (int32_t)ctx->contigs.size();
    for (uint64_t r = 0; r < n; ++r)
    {
        const int32_t c = rid[r];
        if (c == prev) continue;
        prev = c;
        if (c >= 0 && c < nc && !ctx->ref_src->loaded[c] && std::find(which.begin(), which.end(), c) == which.end()) which.push_back(c);
    }
    if (which.empty()) return VTN_OK;
    std::sort(which.begin(), which.end());
    return ref_load_contigs(ctx, which);
}

int vtn_ref_load_fasta(vtn_ctx* ctx, const char* fasta_path)
{
    if (!ctx || !fasta_path) return vtn_set_err(ctx, VTN_E_ARG, "vtn_ref_load_fasta: bad argument");
    int rc = ref_open(ctx, fasta_path);
    if (rc) return rc;
    return vtn_ref_require_all(ctx);
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
    if (ctx->ref_missing)
        if (int rc = vtn_ref_require(ctx, rid)) return rc;
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
