// vtn_api.cu -- context, device batches and the host-buffer entry points of the C ABI (include/vtn.h).
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include "vtn_internal.h"

static thread_local std::string g_create_err;

int vtn_set_err(vtn_ctx* ctx, int code, const std::string& msg)
{
    if (ctx) ctx->err = msg;
    else g_create_err = msg;
    return code;
}

const char* vtn_last_error(const vtn_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_err.c_str(); }

int vtn_ctx_create(vtn_ctx** out, int device)
{
    if (!out) return vtn_set_err(nullptr, VTN_E_ARG, "vtn_ctx_create: null out");
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
        return vtn_set_err(nullptr, VTN_E_CUDA, std::string("no CUDA device (libvtn has no CPU fallback): ") + cudaGetErrorString(e));
    if (device < 0 || device >= n) return vtn_set_err(nullptr, VTN_E_ARG, "vtn_ctx_create: bad device index");
    vtn_ctx* ctx = new vtn_ctx();
    ctx->device = device;
    VTN_CUDA(nullptr, cudaSetDevice(device));
    cudaDeviceProp prop;
    VTN_CUDA(nullptr, cudaGetDeviceProperties(&prop, device));
    ctx->sm_count = prop.multiProcessorCount;
    if (const char* e = getenv("VTN_CHUNK_REC")) // records per pipelined chunk of the host-buffer call (tests use a small one)
    {
        long v = atol(e);
        if (v >= 128) ctx->chunk_rec = (uint32_t)(v / 128) * 128u;
    }
    if (const char* e = getenv("VTN_L2_FETCH")) // L2 fetch granularity in bytes (32 / 64 / 128): the reference gathers touch one or two 16-byte groups
    {
        long v = atol(e);
        if (v == 32 || v == 64 || v == 128) cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)v);
    }
    VTN_CUDA(nullptr, cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
    VTN_CUDA(nullptr, cudaStreamCreateWithFlags(&ctx->stream2, cudaStreamNonBlocking));
    VTN_CUDA(nullptr, cudaStreamCreateWithFlags(&ctx->stream3, cudaStreamNonBlocking));
    VTN_CUDA(nullptr, cudaEventCreate(&ctx->ev0));
    VTN_CUDA(nullptr, cudaEventCreate(&ctx->ev1));
    *out = ctx;
    return VTN_OK;
}

int vtn_ctx_destroy(vtn_ctx* ctx)
{
    if (!ctx) return VTN_OK;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    if (ctx->scratch) vtn_dbatch_destroy(ctx, ctx->scratch);
    if (ctx->scratch_bi) vtn_dbatch_destroy(ctx, ctx->scratch_bi);
    vtn_hmm_ctx_release(ctx);
    vtn_order_ctx_release(ctx);
    if (ctx->tok_cache) vtn_vcf_dtext_destroy(ctx, ctx->tok_cache);
    vtn_ref_release(ctx);
    cudaEventDestroy(ctx->ev0);
    cudaEventDestroy(ctx->ev1);
    cudaStreamDestroy(ctx->stream);
    cudaStreamDestroy(ctx->stream2);
    cudaStreamDestroy(ctx->stream3);
    for (auto e : ctx->chunk_events) cudaEventDestroy(e);
    if (ctx->pinned_totals) cudaFreeHost(ctx->pinned_totals);
    delete ctx;
    return VTN_OK;
}

int vtn_sync(vtn_ctx* ctx)
{
    if (!ctx) return VTN_E_ARG;
    VTN_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    VTN_CUDA(ctx, cudaStreamSynchronize(ctx->stream2));
    VTN_CUDA(ctx, cudaStreamSynchronize(ctx->stream3));
    return VTN_OK;
}

int vtn_host_alloc(vtn_ctx* ctx, size_t bytes, void** p)
{
    if (!ctx || !p) return VTN_E_ARG;
    VTN_CUDA(ctx, cudaSetDevice(ctx->device));
    VTN_CUDA(ctx, cudaMallocHost(p, bytes ? bytes : 1));
    return VTN_OK;
}

int vtn_host_free(vtn_ctx* ctx, void* p)
{
    if (!ctx) return VTN_E_ARG;
    VTN_CUDA(ctx, cudaFreeHost(p));
    return VTN_OK;
}

int vtn_timer_start(vtn_ctx* ctx)
{
    if (!ctx) return VTN_E_ARG;
    VTN_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    return VTN_OK;
}

int vtn_timer_stop(vtn_ctx* ctx, float* ms)
{
    if (!ctx || !ms) return VTN_E_ARG;
    VTN_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
    VTN_CUDA(ctx, cudaEventSynchronize(ctx->ev1));
    VTN_CUDA(ctx, cudaEventElapsedTime(ms, ctx->ev0, ctx->ev1));
    return VTN_OK;
}

uint64_t vtn_launch_count(const vtn_ctx* ctx) { return ctx ? ctx->launches : 0; }

// ---------------------------------------------------------------------------------------------------------

template <typename T>
static cudaError_t dalloc(T** p, size_t n)
{
    return cudaMalloc((void**)p, (n + 512) * sizeof(T)); // padded: tiles load whole 128-bit vectors
}

int vtn_dbatch_alloc(vtn_ctx* ctx, uint32_t n_rec, uint32_t n_allele, uint64_t n_bytes, bool with_inputs, vtn_dbatch** out)
{
    vtn_dbatch* d = new vtn_dbatch();
    d->n_rec = n_rec;
    d->n_allele = n_allele;
    d->n_bytes = n_bytes;
    d->owns_inputs = with_inputs;
    d->cap_rec = n_rec;
    d->cap_allele = n_allele;
    d->cap_bytes = n_bytes;
    const uint32_t tile = vtn_normalize_tile_records();
    d->n_tiles = (n_rec + tile - 1) / tile;
    cudaError_t e = cudaSuccess;
#define A(x) if (e == cudaSuccess) e = (x)
    if (with_inputs)
    {
        A(dalloc(&d->pos1, n_rec));
        A(dalloc(&d->rid, n_rec));
        A(dalloc(&d->allele_beg, (size_t)n_rec + 1));
        A(dalloc(&d->allele_off, (size_t)n_allele + 1));
        A(dalloc(&d->bytes, n_bytes + 16));
    }
    A(dalloc(&d->status, n_rec));
    A(dalloc(&d->vtype, n_rec));
    A(dalloc(&d->chg_rec, n_rec));
    A(dalloc(&d->chg_pos1, n_rec));
    A(dalloc(&d->chg_lt, n_rec));
    A(dalloc(&d->chg_le, n_rec));
    A(dalloc(&d->chg_rt, n_rec));
    A(dalloc(&d->chg_off, (size_t)n_rec + 1));
    A(dalloc(&d->chg_bytes, n_bytes + 16));
    A(dalloc(&d->totals, 4));
    A(dalloc(&d->counters, 16));
    A(dalloc(&d->scr.q0, n_rec));
    A(dalloc(&d->scr.q1, n_rec));
    A(dalloc(&d->scr.bytes, n_bytes + 16));
    A(dalloc(&d->scr.dir, (size_t)d->n_tiles));
    A(dalloc(&d->scr.blo, (size_t)d->n_tiles));
    A(dalloc(&d->tile_state, 2 * (size_t)d->n_tiles + 16)); // decompose's tile prefixes / the deferred-tile list (vtn_launch_normalize)
    A(dalloc(&d->tile_counter, 2 * vtn_normalize_ctl_words(d->n_tiles)));
#undef A
    if (e != cudaSuccess)
    {
        vtn_dbatch_destroy(ctx, d);
        return vtn_set_err(ctx, VTN_E_NOMEM, std::string("device batch allocation: ") + cudaGetErrorString(e));
    }
    *out = d;
    return VTN_OK;
}

bool vtn_dbatch_resize(vtn_dbatch* d, uint32_t n_rec, uint32_t n_allele, uint64_t n_bytes)
{
    if (n_rec > d->cap_rec || n_allele > d->cap_allele || n_bytes > d->cap_bytes) return false;
    d->n_rec = n_rec;
    d->n_allele = n_allele;
    d->n_bytes = n_bytes;
    const uint32_t tile = vtn_normalize_tile_records();
    d->n_tiles = (n_rec + tile - 1) / tile;
    return true;
}

int vtn_dbatch_destroy(vtn_ctx* ctx, vtn_dbatch* d)
{
    if (!d) return VTN_OK;
    if (ctx) cudaSetDevice(ctx->device);
    if (d->fused) vtn_dbatch_destroy(ctx, d->fused);
    vtn_uniq_dbatch_release(d);
    vtn_order_dbatch_release(d);
    vtn_blocksub_dbatch_release(d);
    if (d->owns_inputs)
    {
        cudaFree(d->pos1); cudaFree(d->rid); cudaFree(d->allele_beg); cudaFree(d->allele_off); cudaFree(d->bytes);
    }
    cudaFree(d->status); cudaFree(d->vtype); cudaFree(d->chg_rec); cudaFree(d->chg_pos1); cudaFree(d->chg_lt);
    cudaFree(d->chg_le); cudaFree(d->chg_rt); cudaFree(d->chg_off); cudaFree(d->chg_bytes); cudaFree(d->totals);
    cudaFree(d->scr.q0); cudaFree(d->scr.q1); cudaFree(d->scr.bytes); cudaFree(d->scr.dir); cudaFree(d->scr.blo);
    cudaFree(d->counters); cudaFree(d->tile_state); cudaFree(d->tile_counter); cudaFree(d->src_rec); cudaFree(d->alt_idx);
    cudaFree(d->lens); cudaFree(d->tile_off); cudaFree(d->run_start); cudaFree(d->run_rid); cudaFree(d->ext_rec); cudaFree(d->ext_len);
    cudaFree(d->note_rec); cudaFree(d->note_status); cudaFree(d->note_count);
    delete d;
    return VTN_OK;
}

static int check_batch(vtn_ctx* ctx, const vtn_batch* in)
{
    if (!in) return vtn_set_err(ctx, VTN_E_ARG, "null batch");
    if (in->n_rec && (!in->pos1 || !in->rid || !in->allele_beg || !in->allele_off || (in->n_bytes && !in->bytes)))
        return vtn_set_err(ctx, VTN_E_ARG, "batch has null arrays");
    if (in->n_bytes >= (1ull << 32)) return vtn_set_err(ctx, VTN_E_ARG, "batch allele bytes must be < 4 GiB");
    if (in->n_rec && (in->allele_beg[in->n_rec] != in->n_allele || in->allele_off[in->n_allele] != in->n_bytes))
        return vtn_set_err(ctx, VTN_E_ARG, "batch prefix arrays inconsistent with n_allele / n_bytes");
    return VTN_OK;
}

static int upload_inputs(vtn_ctx* ctx, const vtn_batch* in, vtn_dbatch* d, cudaStream_t s)
{
    if (!in->n_rec) return VTN_OK;
    VTN_CUDA(ctx, cudaMemcpyAsync(d->pos1, in->pos1, sizeof(int32_t) * in->n_rec, cudaMemcpyHostToDevice, s));
    VTN_CUDA(ctx, cudaMemcpyAsync(d->rid, in->rid, sizeof(int32_t) * in->n_rec, cudaMemcpyHostToDevice, s));
    VTN_CUDA(ctx, cudaMemcpyAsync(d->allele_beg, in->allele_beg, sizeof(uint32_t) * ((size_t)in->n_rec + 1), cudaMemcpyHostToDevice, s));
    VTN_CUDA(ctx, cudaMemcpyAsync(d->allele_off, in->allele_off, sizeof(uint32_t) * ((size_t)in->n_allele + 1), cudaMemcpyHostToDevice, s));
    if (in->n_bytes) VTN_CUDA(ctx, cudaMemcpyAsync(d->bytes, in->bytes, in->n_bytes, cudaMemcpyHostToDevice, s));
    return VTN_OK;
}

int vtn_dbatch_create(vtn_ctx* ctx, const vtn_batch* in, vtn_dbatch** db)
{
    if (!ctx || !db) return VTN_E_ARG;
    int rc = check_batch(ctx, in);
    if (rc) return rc;
    VTN_CUDA(ctx, cudaSetDevice(ctx->device));
    if (ctx->ref_missing && (rc = vtn_ref_ensure(ctx, in->n_rec, in->rid))) return rc;
    rc = vtn_dbatch_alloc(ctx, in->n_rec, in->n_allele, in->n_bytes, true, db);
    if (rc) return rc;
    rc = upload_inputs(ctx, in, *db, ctx->stream);
    if (rc) return rc;
    VTN_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return VTN_OK;
}

int vtn_normalize_resident(vtn_ctx* ctx, vtn_dbatch* d)
{
    if (!ctx || !d) return VTN_E_ARG;
    if (!ctx->ref_loaded) return vtn_set_err(ctx, VTN_E_STATE, "vtn_normalize: no reference loaded");
    if (d->lens) return vtn_normalize_resident_bi(ctx, d);
    VtnDevBatchView v{d->n_rec, 0u, d->pos1, d->rid, d->allele_beg, d->allele_off, d->bytes};
    VtnDevNormOut o{d->status, d->vtype, d->chg_rec, d->chg_pos1, d->chg_lt, d->chg_le, d->chg_rt, d->chg_off, d->chg_bytes, d->totals, d->counters};
    return vtn_launch_normalize(ctx, v, o, d->scr, d->tile_state, d->tile_counter, d->n_tiles);
}

static int fetch_norm(vtn_ctx* ctx, vtn_dbatch* d, vtn_norm_result* out)
{
    cudaStream_t s = ctx->stream;
    unsigned long long totals[2] = {0, 0};
    uint32_t counters[12];
    VTN_CUDA(ctx, cudaMemcpyAsync(totals, d->totals, sizeof totals, cudaMemcpyDeviceToHost, s));
    VTN_CUDA(ctx, cudaMemcpyAsync(counters, d->counters, sizeof counters, cudaMemcpyDeviceToHost, s));
    if (d->n_rec)
    {
        VTN_CUDA(ctx, cudaMemcpyAsync(out->status, d->status, d->n_rec, cudaMemcpyDeviceToHost, s));
        VTN_CUDA(ctx, cudaMemcpyAsync(out->vtype, d->vtype, d->n_rec, cudaMemcpyDeviceToHost, s));
    }
    VTN_CUDA(ctx, cudaStreamSynchronize(s));
    out->n_changed = (uint32_t)totals[0];
    out->n_changed_bytes = totals[1];
    memcpy(out->counters.bi, counters, 5 * sizeof(uint32_t));
    memcpy(out->counters.multi, counters + 5, 5 * sizeof(uint32_t));
    out->counters.no_variants = counters[10];
    out->counters.no_refs = counters[11];
    if (totals[0] > out->cap_changed || totals[1] > out->cap_changed_bytes)
        return vtn_set_err(ctx, VTN_E_CAPACITY, "vtn_normalize: result capacity too small (n_changed / n_changed_bytes are set)");
    const size_t k = (size_t)totals[0];
    if (k)
    {
        VTN_CUDA(ctx, cudaMemcpyAsync(out->chg_rec, d->chg_rec, k * 4, cudaMemcpyDeviceToHost, s));
        VTN_CUDA(ctx, cudaMemcpyAsync(out->chg_pos1, d->chg_pos1, k * 4, cudaMemcpyDeviceToHost, s));
        VTN_CUDA(ctx, cudaMemcpyAsync(out->chg_lt, d->chg_lt, k * 4, cudaMemcpyDeviceToHost, s));
        VTN_CUDA(ctx, cudaMemcpyAsync(out->chg_le, d->chg_le, k * 4, cudaMemcpyDeviceToHost, s));
        VTN_CUDA(ctx, cudaMemcpyAsync(out->chg_rt, d->chg_rt, k * 4, cudaMemcpyDeviceToHost, s));
        VTN_CUDA(ctx, cudaMemcpyAsync(out->chg_bytes, d->chg_bytes, (size_t)totals[1], cudaMemcpyDeviceToHost, s));
    }
    VTN_CUDA(ctx, cudaMemcpyAsync(out->chg_off, d->chg_off, (k + 1) * 8, cudaMemcpyDeviceToHost, s));
    VTN_CUDA(ctx, cudaStreamSynchronize(s));
    return VTN_OK;
}

int vtn_dbatch_fetch_result(vtn_ctx* ctx, vtn_dbatch* db, vtn_norm_result* out)
{
    if (!ctx || !db || !out) return VTN_E_ARG;
    return fetch_norm(ctx, db, out);
}

// make sure the ctx-owned scratch batch can hold (n_rec, n_allele, n_bytes)
static int ensure_scratch(vtn_ctx* ctx, uint32_t n_rec, uint32_t n_allele, uint64_t n_bytes)
{
    if (ctx->scratch && vtn_dbatch_resize(ctx->scratch, n_rec, n_allele, n_bytes)) return VTN_OK;
    if (ctx->scratch) vtn_dbatch_destroy(ctx, ctx->scratch);
    ctx->scratch = nullptr;
    uint32_t cr = n_rec + n_rec / 8 + 1024, ca = n_allele + n_allele / 8 + 2048;
    uint64_t cb = n_bytes + n_bytes / 8 + 4096;
    if (cr < n_rec) cr = n_rec;
    if (ca < n_allele) ca = n_allele;
    int rc = vtn_dbatch_alloc(ctx, cr, ca, cb, true, &ctx->scratch);
    if (rc) return rc;
    vtn_dbatch_resize(ctx->scratch, n_rec, n_allele, n_bytes);
    return VTN_OK;
}

// Host-buffer normalize, pipelined: the batch is cut into chunks of VTN_CHUNK_REC records; chunk c's H2D copies
// (stream2), kernels (stream) and status/vtype + changed-record D2H copies (stream3) overlap with the other
// chunks', so the call is bound by the PCIe direction that carries more bytes, not by the sum of the three.
// Device arrays hold the whole batch at absolute positions, so a chunk is just a view (offset pointers); K2
// continues the compaction after the previous chunk's totals, which keeps the output in input order.
int vtn_normalize(vtn_ctx* ctx, const vtn_batch* in, vtn_norm_result* out)
{
    if (!ctx || !out) return VTN_E_ARG;
    if (!ctx->ref_loaded) return vtn_set_err(ctx, VTN_E_STATE, "vtn_normalize: no reference loaded");
    int rc = check_batch(ctx, in);
    if (rc) return rc;
    VTN_CUDA(ctx, cudaSetDevice(ctx->device));
    if (ctx->ref_missing && (rc = vtn_ref_ensure(ctx, in->n_rec, in->rid))) return rc;
    rc = ensure_scratch(ctx, in->n_rec, in->n_allele, in->n_bytes);
    if (rc) return rc;
    vtn_dbatch* d = ctx->scratch;
    const uint32_t n = in->n_rec;
    const uint32_t tile = vtn_normalize_tile_records();
    const uint32_t VTN_CHUNK_REC = ctx->chunk_rec;
    const uint32_t n_chunks = (n + VTN_CHUNK_REC - 1) / VTN_CHUNK_REC;
    if (ctx->chunk_events.size() < 3 * (size_t)n_chunks + 3)
    {
        size_t old = ctx->chunk_events.size();
        ctx->chunk_events.resize(3 * (size_t)n_chunks + 3);
        for (size_t i = old; i < ctx->chunk_events.size(); ++i) VTN_CUDA(ctx, cudaEventCreateWithFlags(&ctx->chunk_events[i], cudaEventDisableTiming));
    }
    if (!ctx->pinned_totals) VTN_CUDA(ctx, cudaMallocHost((void**)&ctx->pinned_totals, 4096 * 2 * sizeof(unsigned long long)));
    if (n_chunks > 4096) return vtn_set_err(ctx, VTN_E_ARG, "vtn_normalize: batch too large (more than 2^32 records?)");
    cudaStream_t sc = ctx->stream, sh = ctx->stream2, sd = ctx->stream3;
    // everything on the side streams starts after what is already queued on the main stream
    cudaEvent_t ev_start = ctx->chunk_events[3 * (size_t)n_chunks];
    VTN_CUDA(ctx, cudaEventRecord(ev_start, sc));
    VTN_CUDA(ctx, cudaStreamWaitEvent(sh, ev_start, 0));
    VTN_CUDA(ctx, cudaStreamWaitEvent(sd, ev_start, 0));
    unsigned long long done_c = 0, done_b = 0; // changed records / bytes already copied back
    bool overflow = false;
    auto copy_changed = [&](unsigned long long tc, unsigned long long tb) -> int {
        if (tc > out->cap_changed || tb > out->cap_changed_bytes) { overflow = true; return VTN_OK; }
        if (tc > done_c)
        {
            const size_t k0 = (size_t)done_c, k = (size_t)(tc - done_c);
            VTN_CUDA(ctx, cudaMemcpyAsync(out->chg_rec + k0, d->chg_rec + k0, k * 4, cudaMemcpyDeviceToHost, sd));
            VTN_CUDA(ctx, cudaMemcpyAsync(out->chg_pos1 + k0, d->chg_pos1 + k0, k * 4, cudaMemcpyDeviceToHost, sd));
            VTN_CUDA(ctx, cudaMemcpyAsync(out->chg_lt + k0, d->chg_lt + k0, k * 4, cudaMemcpyDeviceToHost, sd));
            VTN_CUDA(ctx, cudaMemcpyAsync(out->chg_le + k0, d->chg_le + k0, k * 4, cudaMemcpyDeviceToHost, sd));
            VTN_CUDA(ctx, cudaMemcpyAsync(out->chg_rt + k0, d->chg_rt + k0, k * 4, cudaMemcpyDeviceToHost, sd));
            VTN_CUDA(ctx, cudaMemcpyAsync(out->chg_off + k0, d->chg_off + k0, k * 8, cudaMemcpyDeviceToHost, sd));
            if (tb > done_b)
                VTN_CUDA(ctx, cudaMemcpyAsync(out->chg_bytes + done_b, d->chg_bytes + done_b, (size_t)(tb - done_b), cudaMemcpyDeviceToHost, sd));
            done_c = tc;
            done_b = tb;
        }
        return VTN_OK;
    };
    if (n == 0)
    {
        rc = vtn_launch_normalize(ctx, VtnDevBatchView{0, 0, d->pos1, d->rid, d->allele_beg, d->allele_off, d->bytes},
                                  VtnDevNormOut{d->status, d->vtype, d->chg_rec, d->chg_pos1, d->chg_lt, d->chg_le, d->chg_rt, d->chg_off, d->chg_bytes, d->totals, d->counters},
                                  d->scr, d->tile_state, d->tile_counter, 0, false);
        if (rc) return rc;
    }
    for (uint32_t c = 0; c < n_chunks; ++c)
    {
        const uint32_t r0 = c * VTN_CHUNK_REC, r1 = (c + 1 == n_chunks) ? n : r0 + VTN_CHUNK_REC, nr = r1 - r0;
        const uint32_t s0 = in->allele_beg[r0], s1 = in->allele_beg[r1];
        const uint32_t b0 = in->allele_off[s0], b1 = in->allele_off[s1];
        cudaEvent_t ev_h = ctx->chunk_events[3 * (size_t)c], ev_k = ctx->chunk_events[3 * (size_t)c + 1], ev_t = ctx->chunk_events[3 * (size_t)c + 2];
        // ---- H2D of the chunk's slices ----
        VTN_CUDA(ctx, cudaMemcpyAsync(d->allele_beg + r0, in->allele_beg + r0, sizeof(uint32_t) * ((size_t)nr + 1), cudaMemcpyHostToDevice, sh));
        VTN_CUDA(ctx, cudaMemcpyAsync(d->allele_off + s0, in->allele_off + s0, sizeof(uint32_t) * ((size_t)(s1 - s0) + 1), cudaMemcpyHostToDevice, sh));
        if (b1 > b0) VTN_CUDA(ctx, cudaMemcpyAsync(d->bytes + b0, in->bytes + b0, (size_t)(b1 - b0), cudaMemcpyHostToDevice, sh));
        VTN_CUDA(ctx, cudaMemcpyAsync(d->pos1 + r0, in->pos1 + r0, sizeof(int32_t) * nr, cudaMemcpyHostToDevice, sh));
        VTN_CUDA(ctx, cudaMemcpyAsync(d->rid + r0, in->rid + r0, sizeof(int32_t) * nr, cudaMemcpyHostToDevice, sh));
        VTN_CUDA(ctx, cudaEventRecord(ev_h, sh));
        // ---- kernels on a view of the chunk ----
        VTN_CUDA(ctx, cudaStreamWaitEvent(sc, ev_h, 0));
        const uint32_t tile0 = r0 / tile, nt = (nr + tile - 1) / tile;
        VtnDevBatchView v{nr, r0, d->pos1 + r0, d->rid + r0, d->allele_beg + r0, d->allele_off, d->bytes};
        VtnDevNormOut o{d->status + r0, d->vtype + r0, d->chg_rec, d->chg_pos1, d->chg_lt, d->chg_le, d->chg_rt, d->chg_off, d->chg_bytes, d->totals, d->counters};
        VtnDevScratch x = d->scr;
        x.q0 += r0; x.q1 += r0; x.dir += tile0; x.blo += tile0;
        rc = vtn_launch_normalize(ctx, v, o, x, d->tile_state, d->tile_counter, nt, c != 0);
        if (rc) return rc;
        VTN_CUDA(ctx, cudaEventRecord(ev_k, sc));
        // ---- D2H: per-record outputs of this chunk, running totals ----
        VTN_CUDA(ctx, cudaStreamWaitEvent(sd, ev_k, 0));
        VTN_CUDA(ctx, cudaMemcpyAsync(out->status + r0, d->status + r0, nr, cudaMemcpyDeviceToHost, sd));
        VTN_CUDA(ctx, cudaMemcpyAsync(out->vtype + r0, d->vtype + r0, nr, cudaMemcpyDeviceToHost, sd));
        VTN_CUDA(ctx, cudaMemcpyAsync(ctx->pinned_totals + 2 * (size_t)c, d->totals, 2 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, sd));
        VTN_CUDA(ctx, cudaEventRecord(ev_t, sd));
        // the totals after chunk c-1 have surely landed by now: copy back that chunk's changed records
        if (c >= 1)
        {
            VTN_CUDA(ctx, cudaEventSynchronize(ctx->chunk_events[3 * (size_t)(c - 1) + 2]));
            rc = copy_changed(ctx->pinned_totals[2 * (size_t)(c - 1)], ctx->pinned_totals[2 * (size_t)(c - 1) + 1]);
            if (rc) return rc;
        }
    }
    uint32_t counters[12] = {0};
    unsigned long long totals[2] = {0, 0};
    VTN_CUDA(ctx, cudaStreamSynchronize(sc));
    VTN_CUDA(ctx, cudaStreamSynchronize(sd)); // pinned_totals of the last chunk is valid now
    VTN_CUDA(ctx, cudaMemcpyAsync(counters, d->counters, sizeof counters, cudaMemcpyDeviceToHost, sd));
    VTN_CUDA(ctx, cudaMemcpyAsync(totals, d->totals, sizeof totals, cudaMemcpyDeviceToHost, sd));
    VTN_CUDA(ctx, cudaStreamSynchronize(sd));
    rc = copy_changed(totals[0], totals[1]);
    if (rc) return rc;
    if (!overflow && totals[0] <= out->cap_changed)
        VTN_CUDA(ctx, cudaMemcpyAsync(out->chg_off + totals[0], d->chg_off + totals[0], 8, cudaMemcpyDeviceToHost, sd));
    VTN_CUDA(ctx, cudaStreamSynchronize(sd));
    out->n_changed = (uint32_t)totals[0];
    out->n_changed_bytes = totals[1];
    memcpy(out->counters.bi, counters, 5 * sizeof(uint32_t));
    memcpy(out->counters.multi, counters + 5, 5 * sizeof(uint32_t));
    out->counters.no_variants = counters[10];
    out->counters.no_refs = counters[11];
    if (overflow || totals[0] > out->cap_changed || totals[1] > out->cap_changed_bytes)
        return vtn_set_err(ctx, VTN_E_CAPACITY, "vtn_normalize: result capacity too small (n_changed / n_changed_bytes are set)");
    return VTN_OK;
}

// ---------------------------------------------------------------------------------------------------------
// decompose
uint64_t vtn_decompose_out_bound(const vtn_batch* in)
{
    if (!in || !in->n_rec) return 0;
    uint64_t k = 0;
    for (uint32_t r = 0; r < in->n_rec; ++r)
    {
        uint32_t n = in->allele_beg[r + 1] - in->allele_beg[r];
        k += n > 2 ? n - 1 : 1;
    }
    return k;
}

int vtn_decompose_resident(vtn_ctx* ctx, vtn_dbatch* d, int fuse_normalize)
{
    if (!ctx || !d) return VTN_E_ARG;
    if (fuse_normalize && !ctx->ref_loaded) return vtn_set_err(ctx, VTN_E_STATE, "vtn_decompose_normalize: no reference loaded");
    int rc = vtn_launch_decompose(ctx, d, fuse_normalize != 0);
    if (rc || !fuse_normalize) return rc;
    return vtn_normalize_resident(ctx, d->fused);
}

int vtn_dbatch_fetch_decompose(vtn_ctx* ctx, vtn_dbatch* d, uint32_t* src_rec, uint16_t* alt_idx, uint64_t* n_out,
                               vtn_decomp_counters* counters)
{
    if (!ctx || !d) return VTN_E_ARG;
    cudaStream_t s = ctx->stream;
    unsigned long long tot[3] = {0, 0, 0};
    uint32_t c[4] = {0, 0, 0, 0};
    VTN_CUDA(ctx, cudaMemcpyAsync(tot, d->totals, sizeof tot, cudaMemcpyDeviceToHost, s));
    VTN_CUDA(ctx, cudaMemcpyAsync(c, d->counters + 12, sizeof c, cudaMemcpyDeviceToHost, s));
    VTN_CUDA(ctx, cudaStreamSynchronize(s));
    if (n_out) *n_out = tot[0];
    if (counters)
    {
        counters->no_variants = c[0];
        counters->no_biallelic = c[1];
        counters->no_multiallelic = c[2];
        counters->no_additional_biallelic = c[3];
    }
    if (tot[0] && src_rec) VTN_CUDA(ctx, cudaMemcpyAsync(src_rec, d->src_rec, tot[0] * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    if (tot[0] && alt_idx) VTN_CUDA(ctx, cudaMemcpyAsync(alt_idx, d->alt_idx, tot[0] * sizeof(uint16_t), cudaMemcpyDeviceToHost, s));
    VTN_CUDA(ctx, cudaStreamSynchronize(s));
    return VTN_OK;
}

static int decompose_host(vtn_ctx* ctx, const vtn_batch* in, uint32_t* src_rec, uint16_t* alt_idx, uint64_t* n_out,
                          vtn_decomp_counters* counters, vtn_norm_result* nout)
{
    if (!ctx) return VTN_E_ARG;
    int rc = check_batch(ctx, in);
    if (rc) return rc;
    VTN_CUDA(ctx, cudaSetDevice(ctx->device));
    if (nout && ctx->ref_missing && (rc = vtn_ref_ensure(ctx, in->n_rec, in->rid))) return rc;
    rc = ensure_scratch(ctx, in->n_rec, in->n_allele, in->n_bytes);
    if (rc) return rc;
    vtn_dbatch* d = ctx->scratch;
    rc = upload_inputs(ctx, in, d, ctx->stream);
    if (rc) return rc;
    rc = vtn_decompose_resident(ctx, d, nout ? 1 : 0);
    if (rc) return rc;
    rc = vtn_dbatch_fetch_decompose(ctx, d, src_rec, alt_idx, n_out, counters);
    if (rc) return rc;
    if (nout) rc = fetch_norm(ctx, d->fused, nout);
    return rc;
}

int vtn_decompose(vtn_ctx* ctx, const vtn_batch* in, uint32_t* src_rec, uint16_t* alt_idx, uint64_t* n_out,
                  vtn_decomp_counters* counters)
{
    return decompose_host(ctx, in, src_rec, alt_idx, n_out, counters, nullptr);
}

int vtn_decompose_normalize(vtn_ctx* ctx, const vtn_batch* in, uint32_t* src_rec, uint16_t* alt_idx, uint64_t* n_out,
                            vtn_decomp_counters* dcounters, vtn_norm_result* out)
{
    if (!out) return VTN_E_ARG;
    if (!ctx || !ctx->ref_loaded) return vtn_set_err(ctx, VTN_E_STATE, "vtn_decompose_normalize: no reference loaded");
    return decompose_host(ctx, in, src_rec, alt_idx, n_out, dcounters, out);
}

// ---------------------------------------------------------------------------------------------------------
// One shard of a multi-device decompose -> normalize (vtn_multi.cu), in two phases so that every device copies its results
// straight into their final place in the caller's arrays: phase 1 uploads records [r0, r1) as slices of the caller's arrays,
// re-bases the two prefix arrays ON THE DEVICE (no host pass over the batch) and runs the kernels; once every shard's totals
// are known, phase 2 adds the shard's bases to the index arrays on the device and copies everything out.
// ---------------------------------------------------------------------------------------------------------
namespace
{
__global__ void vtn_add_u32_kernel(uint32_t* p, size_t n, uint32_t add)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] += add;
}
__global__ void vtn_add_u64_kernel(uint64_t* p, size_t n, uint64_t add)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] += add;
}
void add_u32(vtn_ctx* ctx, uint32_t* p, size_t n, uint32_t add)
{
    if (!n || !add) return;
    vtn_add_u32_kernel<<<(unsigned)std::min<size_t>((n + 255) / 256, 4096), 256, 0, ctx->stream>>>(p, n, add);
    ctx->launches++;
}
void add_u64(vtn_ctx* ctx, uint64_t* p, size_t n, uint64_t add)
{
    if (!n || !add) return;
    vtn_add_u64_kernel<<<(unsigned)std::min<size_t>((n + 255) / 256, 4096), 256, 0, ctx->stream>>>(p, n, add);
    ctx->launches++;
}
} // namespace

int vtn_shard_decompose_run(vtn_ctx* ctx, const vtn_batch* in, uint32_t r0, uint32_t r1, uint64_t* n_out, uint64_t* n_changed, uint64_t* n_changed_bytes)
{
    if (!ctx || !ctx->ref_loaded) return vtn_set_err(ctx, VTN_E_STATE, "vtn_multi_decompose_normalize: no reference loaded");
    VTN_CUDA(ctx, cudaSetDevice(ctx->device));
    const uint32_t ns = r1 - r0;
    *n_out = *n_changed = *n_changed_bytes = 0;
    if (!ns) return VTN_OK;
    const uint32_t s0 = in->allele_beg[r0], s1 = in->allele_beg[r1];
    const uint32_t b0 = in->allele_off[s0], b1 = in->allele_off[s1];
    int rc = ensure_scratch(ctx, ns, s1 - s0, b1 - b0);
    if (rc) return rc;
    vtn_dbatch* d = ctx->scratch;
    cudaStream_t s = ctx->stream;
    VTN_CUDA(ctx, cudaMemcpyAsync(d->pos1, in->pos1 + r0, sizeof(int32_t) * ns, cudaMemcpyHostToDevice, s));
    VTN_CUDA(ctx, cudaMemcpyAsync(d->rid, in->rid + r0, sizeof(int32_t) * ns, cudaMemcpyHostToDevice, s));
    VTN_CUDA(ctx, cudaMemcpyAsync(d->allele_beg, in->allele_beg + r0, sizeof(uint32_t) * ((size_t)ns + 1), cudaMemcpyHostToDevice, s));
    VTN_CUDA(ctx, cudaMemcpyAsync(d->allele_off, in->allele_off + s0, sizeof(uint32_t) * ((size_t)(s1 - s0) + 1), cudaMemcpyHostToDevice, s));
    if (b1 > b0) VTN_CUDA(ctx, cudaMemcpyAsync(d->bytes, in->bytes + b0, (size_t)(b1 - b0), cudaMemcpyHostToDevice, s));
    add_u32(ctx, d->allele_beg, (size_t)ns + 1, 0u - s0);
    add_u32(ctx, d->allele_off, (size_t)(s1 - s0) + 1, 0u - b0);
    VTN_CUDA(ctx, cudaGetLastError());
    rc = vtn_decompose_resident(ctx, d, 1);
    if (rc) return rc;
    unsigned long long tot[3] = {0, 0, 0}, nt[2] = {0, 0};
    VTN_CUDA(ctx, cudaMemcpyAsync(tot, d->totals, sizeof tot, cudaMemcpyDeviceToHost, s));
    VTN_CUDA(ctx, cudaMemcpyAsync(nt, d->fused->totals, sizeof nt, cudaMemcpyDeviceToHost, s));
    VTN_CUDA(ctx, cudaStreamSynchronize(s));
    *n_out = tot[0];
    *n_changed = nt[0];
    *n_changed_bytes = nt[1];
    return VTN_OK;
}

int vtn_shard_decompose_fetch(vtn_ctx* ctx, uint32_t rec_base, uint64_t out_base, uint64_t chg_base, uint64_t byte_base, uint64_t n_out, uint64_t n_changed,
                              uint64_t n_changed_bytes, uint32_t* src_rec, uint16_t* alt_idx, vtn_decomp_counters* dc, vtn_norm_result* out)
{
    VTN_CUDA(ctx, cudaSetDevice(ctx->device));
    memset(dc, 0, sizeof *dc);
    memset(&out->counters, 0, sizeof out->counters);
    if (!n_out) return VTN_OK; // an empty shard ran nothing
    vtn_dbatch* d = ctx->scratch;
    vtn_dbatch* f = d->fused;
    cudaStream_t s = ctx->stream;
    uint32_t c[4] = {0, 0, 0, 0}, nc[12];
    VTN_CUDA(ctx, cudaMemcpyAsync(c, d->counters + 12, sizeof c, cudaMemcpyDeviceToHost, s));
    VTN_CUDA(ctx, cudaMemcpyAsync(nc, f->counters, sizeof nc, cudaMemcpyDeviceToHost, s));
    const size_t no = (size_t)n_out, k = (size_t)n_changed;
    add_u32(ctx, d->src_rec, no, rec_base);
    VTN_CUDA(ctx, cudaMemcpyAsync(src_rec + out_base, d->src_rec, no * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    VTN_CUDA(ctx, cudaMemcpyAsync(alt_idx + out_base, d->alt_idx, no * sizeof(uint16_t), cudaMemcpyDeviceToHost, s));
    VTN_CUDA(ctx, cudaMemcpyAsync(out->status + out_base, f->status, no, cudaMemcpyDeviceToHost, s));
    VTN_CUDA(ctx, cudaMemcpyAsync(out->vtype + out_base, f->vtype, no, cudaMemcpyDeviceToHost, s));
    if (k)
    {
        add_u32(ctx, f->chg_rec, k, (uint32_t)out_base);
        add_u64(ctx, f->chg_off, k, byte_base);
        VTN_CUDA(ctx, cudaMemcpyAsync(out->chg_rec + chg_base, f->chg_rec, k * 4, cudaMemcpyDeviceToHost, s));
        VTN_CUDA(ctx, cudaMemcpyAsync(out->chg_pos1 + chg_base, f->chg_pos1, k * 4, cudaMemcpyDeviceToHost, s));
        VTN_CUDA(ctx, cudaMemcpyAsync(out->chg_lt + chg_base, f->chg_lt, k * 4, cudaMemcpyDeviceToHost, s));
        VTN_CUDA(ctx, cudaMemcpyAsync(out->chg_le + chg_base, f->chg_le, k * 4, cudaMemcpyDeviceToHost, s));
        VTN_CUDA(ctx, cudaMemcpyAsync(out->chg_rt + chg_base, f->chg_rt, k * 4, cudaMemcpyDeviceToHost, s));
        VTN_CUDA(ctx, cudaMemcpyAsync(out->chg_off + chg_base, f->chg_off, k * 8, cudaMemcpyDeviceToHost, s));
        VTN_CUDA(ctx, cudaMemcpyAsync(out->chg_bytes + byte_base, f->chg_bytes, (size_t)n_changed_bytes, cudaMemcpyDeviceToHost, s));
    }
    VTN_CUDA(ctx, cudaStreamSynchronize(s));
    dc->no_variants = c[0];
    dc->no_biallelic = c[1];
    dc->no_multiallelic = c[2];
    dc->no_additional_biallelic = c[3];
    memcpy(out->counters.bi, nc, 5 * sizeof(uint32_t));
    memcpy(out->counters.multi, nc + 5, 5 * sizeof(uint32_t));
    out->counters.no_variants = nc[10];
    out->counters.no_refs = nc[11];
    return VTN_OK;
}

// Host-buffer uniq: H2D of the batch, one kernel, D2H of order / first and the two counters.
int vtn_uniq(vtn_ctx* ctx, const vtn_batch* in, uint32_t* order, uint8_t* first, vtn_uniq_counters* counters, int64_t* not_in_order)
{
    if (!ctx) return VTN_E_ARG;
    int rc = check_batch(ctx, in);
    if (rc) return rc;
    if (in->n_rec && (!order || !first)) return vtn_set_err(ctx, VTN_E_ARG, "vtn_uniq: null output arrays");
    VTN_CUDA(ctx, cudaSetDevice(ctx->device));
    rc = ensure_scratch(ctx, in->n_rec, in->n_allele, in->n_bytes);
    if (rc) return rc;
    rc = upload_inputs(ctx, in, ctx->scratch, ctx->stream);
    if (rc) return rc;
    rc = vtn_uniq_resident(ctx, ctx->scratch);
    if (rc) return rc;
    return vtn_dbatch_fetch_uniq(ctx, ctx->scratch, order, first, counters, not_in_order);
}

// Host-buffer decompose_blocksub plan: H2D of the batch, three kernels, D2H of the plan and the counters.
int vtn_blocksub(vtn_ctx* ctx, const vtn_batch* in, int keep_mnv, int max_mnv_dist, uint32_t* src_rec, uint32_t* off, uint32_t* len,
                 uint64_t cap_out, uint64_t* n_out, vtn_blocksub_counters* counters)
{
    if (!ctx || !n_out) return VTN_E_ARG;
    int rc = check_batch(ctx, in);
    if (rc) return rc;
    VTN_CUDA(ctx, cudaSetDevice(ctx->device));
    rc = ensure_scratch(ctx, in->n_rec, in->n_allele, in->n_bytes);
    if (rc) return rc;
    rc = upload_inputs(ctx, in, ctx->scratch, ctx->stream);
    if (rc) return rc;
    rc = vtn_blocksub_resident(ctx, ctx->scratch, keep_mnv, max_mnv_dist);
    if (rc) return rc;
    return vtn_dbatch_fetch_blocksub(ctx, ctx->scratch, src_rec, off, len, cap_out, n_out, counters);
}
