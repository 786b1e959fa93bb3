// vtn_ref.cu -- reference genome: FASTA/.fai reading, packing into the 16-byte-per-32-base group layout
// (vtn_internal.h), upload to HBM, and a faidx_fetch_seq-style read-back kernel.
// Replaces fai_load (variant_manip.cpp:35) and faidx_fetch_seq / faidx_fetch_uc_seq
// (lib/htslib/faidx.c:832-881, hts_utils.cpp:41-51) for the normalize path.
#include <ctype.h>
#include <stdio.h>
#include <string.h>
#include <zlib.h>
#include <algorithm>
#include <map>
#include <thread>
#include "vtn_fasta_pack.h"
#include "vtn_internal.h"
#include "vtn_ref_dev.cuh"

using vtf::ExcRun;

// what a lazily opened FASTA keeps on the host: the mapped text, its index, which contigs are resident, their exception runs, and
// contigs that were packed on the host ahead of their upload (before a context existed, or by a background prefetch)
struct VtnRefSource
{
    vtf::FastaText text;
    std::vector<vtf::FaiEntry> fai;
    std::vector<uint64_t> goff; // group offset of every contig: the same layout ref_alloc_tables gives the device array
    std::vector<char> loaded;
    std::vector<std::vector<ExcRun>> exc;
    std::vector<std::vector<vtf::Group>> packed; // host-packed, not uploaded yet
    std::vector<char> packed_ok;                 // 1 packed, -1 the text was shorter than the index says
    std::vector<char> uploaded;                  // 1 a background task has copied the packed groups to the device (not yet accounted)
    std::vector<std::string> bg_err;             // what went wrong in a background task
    std::map<int, std::thread> inflight;         // background tasks: pack (unless packed already) and, once attached, upload
    ~VtnRefSource()
    {
        for (auto& kv : inflight)
            if (kv.second.joinable()) kv.second.join();
    }
};
struct vtn_fasta : VtnRefSource
{
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
    delete ctx->ref_src; // joins the background tasks first: they copy into d_groups
    ctx->ref_src = nullptr;
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

// host only: pack contig c of an opened FASTA into S->packed[c] (all host threads); the upload happens later
void src_pack(VtnRefSource* S, int c, unsigned nthr)
{
    const size_t nc = S->fai.size();
    std::vector<uint64_t> hoff(nc, 0);
    S->packed[c].assign((size_t)((std::max<int64_t>(S->fai[c].len, 0) + 31) / 32) + 1, vtf::Group{0, 0, 0, 0});
    const int bad = vtf::pack_contigs_ol(S->text.data, S->text.size, S->fai, std::vector<int>(1, c), S->goff, hoff, S->packed[c].data(), S->exc, nthr, [](int) {});
    S->packed_ok[c] = bad >= 0 ? -1 : 1;
}

// background task for contig c: pack it unless that was done ahead, then copy it to the device on a stream of its own (the copy
// comes from pageable memory: it keeps THIS thread busy, not the tool's).  d_groups = the context's group array.
void bg_pack_upload(VtnRefSource* S, int c, unsigned nthr, int device, void* d_groups)
{
    if (!S->packed_ok[c]) src_pack(S, c, nthr);
    if (S->packed_ok[c] <= 0 || !d_groups) return;
    const uint64_t ng = (uint64_t)((std::max<int64_t>(S->fai[c].len, 0) + 31) / 32);
    cudaError_t e = cudaSetDevice(device);
    cudaStream_t st = nullptr;
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking);
    if (e == cudaSuccess && ng) e = cudaMemcpyAsync((uint4*)d_groups + S->goff[c], S->packed[c].data(), ng * sizeof(uint4), cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (st) cudaStreamDestroy(st);
    if (e != cudaSuccess)
    {
        S->bg_err[c] = cudaGetErrorString(e);
        return;
    }
    std::vector<vtf::Group>().swap(S->packed[c]);
    S->packed_ok[c] = 0;
    S->uploaded[c] = 1;
}

VtnRefSource* src_open(const char* fasta_path, std::string& err)
{
    VtnRefSource* S = new vtn_fasta();
    if (!S->text.open(fasta_path))
    {
        delete S;
        err = std::string("Cannot load genome index: ") + fasta_path;
        return nullptr;
    }
    if (!vtf::read_fai(std::string(fasta_path) + ".fai", S->fai)) vtf::build_fai(S->text.data, S->text.size, S->fai);
    if (S->fai.empty())
    {
        delete S;
        err = std::string("Cannot load genome index: ") + fasta_path;
        return nullptr;
    }
    const size_t nc = S->fai.size();
    uint64_t g = 0;
    S->goff.resize(nc);
    for (size_t c = 0; c < nc; ++c)
    { // ref_alloc_tables' layout
        S->goff[c] = g;
        g += (uint64_t)((std::max<int64_t>(S->fai[c].len, 0) + 31) / 32) + 2;
    }
    S->loaded.assign(nc, 0);
    S->packed_ok.assign(nc, 0);
    S->uploaded.assign(nc, 0);
    S->bg_err.resize(nc);
    S->exc.resize(nc);
    S->packed.resize(nc);
    return S;
}

// Make the listed contigs resident.  A contig that was packed ahead (vtn_fasta_prepack, vtn_ref_prefetch) is only uploaded; the others
// are packed now (all worker threads on one contig after the other), each uploaded as soon as it is complete while the workers pack
// the next.
int ref_load_contigs(vtn_ctx* ctx, const std::vector<int>& which_all)
{
    VtnRefSource* S = ctx->ref_src;
    if (which_all.empty()) return VTN_OK;
    VTN_CUDA(ctx, cudaSetDevice(ctx->device));
    const size_t nc = ctx->contigs.size();
    std::vector<int> which, ahead;
    for (int c : which_all)
    {
        auto it = S->inflight.find(c);
        if (it != S->inflight.end())
        {
            if (it->second.joinable()) it->second.join();
            S->inflight.erase(it);
        }
        if (S->loaded[c]) continue;
        if (!S->bg_err[c].empty()) return vtn_set_err(ctx, VTN_E_CUDA, "reference upload (background): " + S->bg_err[c]);
        if (S->packed_ok[c] < 0) return vtn_set_err(ctx, VTN_E_IO, std::string("FASTA shorter than its index: ") + S->fai[c].name);
        if (S->uploaded[c])
        { // a background task brought it over: only the bookkeeping below is left
            S->uploaded[c] = 0;
            continue;
        }
        (S->packed_ok[c] > 0 ? ahead : which).push_back(c);
    }
    cudaError_t ce = cudaSuccess;
    for (int c : ahead)
    {
        const uint64_t ng = (uint64_t)((ctx->contigs[c].len + 31) / 32);
        if (ng && ce == cudaSuccess)
            ce = cudaMemcpyAsync((uint4*)ctx->d_groups + S->goff[c], S->packed[c].data(), ng * sizeof(uint4), cudaMemcpyHostToDevice, ctx->stream);
    }
    vtf::Group* h = nullptr;
    int bad = -1;
    if (!which.empty())
    {
        std::vector<uint64_t> hoff(nc, 0);
        uint64_t n_host = 0;
        for (int c : which)
        {
            hoff[c] = n_host;
            n_host += (uint64_t)((ctx->contigs[c].len + 31) / 32);
        }
        h = (vtf::Group*)malloc(std::max<uint64_t>(n_host, 1) * sizeof(vtf::Group));
        if (!h) return vtn_set_err(ctx, VTN_E_NOMEM, "reference packing: out of host memory");
        bad = vtf::pack_contigs_ol(S->text.data, S->text.size, S->fai, which, S->goff, hoff, h, S->exc, host_threads(), [&](int c) {
            const uint64_t ng = (uint64_t)((ctx->contigs[c].len + 31) / 32);
            if (ng && ce == cudaSuccess)
                ce = cudaMemcpyAsync((uint4*)ctx->d_groups + S->goff[c], h + hoff[c], ng * sizeof(uint4), cudaMemcpyHostToDevice, ctx->stream);
        });
    }
    if (ce == cudaSuccess) ce = cudaStreamSynchronize(ctx->stream);
    free(h);
    for (int c : ahead)
    {
        std::vector<vtf::Group>().swap(S->packed[c]);
        S->packed_ok[c] = 0;
    }
    if (bad >= 0) return vtn_set_err(ctx, VTN_E_IO, std::string("FASTA shorter than its index: ") + S->fai[bad].name);
    if (ce != cudaSuccess) return vtn_set_err(ctx, VTN_E_CUDA, std::string("reference upload: ") + cudaGetErrorString(ce));
    for (int c : which_all)
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

// device side of opening: contig table + group array for the whole genome; contigs packed ahead are uploaded at once
int ref_attach(vtn_ctx* ctx, VtnRefSource* S)
{
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
    ctx->ref_src = S;
    ctx->ref_missing = (uint32_t)nc;
    for (size_t c = 0; c < nc; ++c)
        if (S->loaded[c]) S->loaded[c] = 0; // a source comes in with nothing resident
    rc = ref_upload_exc(ctx, {});
    if (rc) return rc;
    ctx->ref_loaded = true;
    // contigs packed ahead go over in the background, each on its own thread and stream; whoever needs one joins its task
    for (size_t c = 0; c < nc; ++c)
        if (S->packed_ok[c] > 0 && !S->inflight.count((int)c))
            S->inflight[(int)c] = std::thread(bg_pack_upload, S, (int)c, 1u, ctx->device, ctx->d_groups);
    return VTN_OK;
}

int ref_open(vtn_ctx* ctx, const char* fasta_path)
{
    std::string err;
    VtnRefSource* S = src_open(fasta_path, err);
    if (!S) return vtn_set_err(ctx, VTN_E_IO, err);
    return ref_attach(ctx, S);
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

// ---- the host half on its own: open and pack before a context exists (CUDA context creation takes 0.25-2 s: a tool opens the FASTA,
// reads its input's first records and packs their contig meanwhile), then hand everything to the context ----
int vtn_fasta_open(const char* fasta_path, vtn_fasta** out)
{
    if (!fasta_path || !out) return VTN_E_ARG;
    std::string err;
    VtnRefSource* S = src_open(fasta_path, err);
    *out = static_cast<vtn_fasta*>(S);
    if (!S) return vtn_set_err(nullptr, VTN_E_IO, err);
    return VTN_OK;
}

int vtn_fasta_contig_id(const vtn_fasta* f, const char* name)
{
    if (!f || !name) return -1;
    for (size_t c = 0; c < f->fai.size(); ++c)
        if (f->fai[c].name == name) return (int)c;
    return -1;
}

int vtn_fasta_n_contig(const vtn_fasta* f) { return f ? (int)f->fai.size() : 0; }

int vtn_fasta_prepack(vtn_fasta* f, int rid)
{
    if (!f) return VTN_E_ARG;
    if (rid < 0 || rid >= (int)f->fai.size() || f->packed_ok[rid] || f->inflight.count(rid)) return VTN_OK;
    src_pack(f, rid, host_threads());
    return f->packed_ok[rid] > 0 ? VTN_OK : VTN_E_IO;
}

void vtn_fasta_close(vtn_fasta* f) { delete f; }

int vtn_ref_attach(vtn_ctx* ctx, vtn_fasta* f)
{
    if (!ctx || !f)
    {
        delete f;
        return vtn_set_err(ctx, VTN_E_ARG, "vtn_ref_attach: bad argument");
    }
    return ref_attach(ctx, f);
}

// background host pack of one contig of a lazily opened reference; vtn_ref_require (or any call that needs the contig) joins it and
// only has the upload left.  A sorted call set walks the contigs in order, so a tool prefetches contig c + 1 when c first appears.
int vtn_ref_prefetch(vtn_ctx* ctx, int rid)
{
    if (!ctx || !ctx->ref_loaded) return vtn_set_err(ctx, VTN_E_STATE, "no reference loaded");
    VtnRefSource* S = ctx->ref_src;
    if (!S || !ctx->ref_missing || rid < 0 || rid >= (int)S->fai.size() || S->inflight.count(rid) || S->loaded[rid] || S->packed_ok[rid]) return VTN_OK;
    const unsigned nthr = std::max(1u, host_threads() / 2); // the foreground keeps half of the cores
    S->inflight[rid] = std::thread(bg_pack_upload, S, rid, nthr, ctx->device, ctx->d_groups);
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
