// vtn_bi.cu -- host side of the compact biallelic layout (vtn_batch_bi, include/vtn.h): the re-packing helper, device
// batches, the resident launch and the pipelined host-buffer call vtn_normalize_bi.  The kernels are in vtn_normalize.cu.
#include <algorithm>
#include <string.h>
#include "vtn_internal.h"

struct vtn_bi_host
{
    vtn_batch_bi v{};
    void* mem[8] = {nullptr};
};

static uint32_t bi_tiles(uint32_t n_rec) { return (n_rec + VTN_BI_TILE - 1) / VTN_BI_TILE; }

int vtn_bi_host_create(vtn_ctx* ctx, const vtn_batch* in, vtn_bi_host** out)
{
    if (!ctx || !in || !out) return VTN_E_ARG;
    *out = nullptr;
    const uint32_t n = in->n_rec;
    if (n && (in->n_allele != 2ull * n)) return vtn_set_err(ctx, VTN_E_ARG, "vtn_bi_host_create: batch is not biallelic");
    for (uint32_t r = 0; r < n; ++r)
        if (in->allele_beg[r] != 2 * r) return vtn_set_err(ctx, VTN_E_ARG, "vtn_bi_host_create: batch is not biallelic");
    if (n && in->allele_off[0] != 0) return vtn_set_err(ctx, VTN_E_ARG, "vtn_bi_host_create: allele_off[0] != 0");
    // first pass: runs and long records
    uint32_t n_run = 0, n_ext = 0;
    for (uint32_t r = 0; r < n; ++r)
    {
        if (r == 0 || in->rid[r] != in->rid[r - 1]) ++n_run;
        const uint32_t a = in->allele_off[2 * r + 1] - in->allele_off[2 * r], b = in->allele_off[2 * r + 2] - in->allele_off[2 * r + 1];
        if (a >= VTN_BI_LEN_ESC || b >= VTN_BI_LEN_ESC) ++n_ext;
    }
    vtn_bi_host* h = new vtn_bi_host();
    const uint32_t nt = bi_tiles(n);
    const size_t sz[8] = {sizeof(int32_t) * ((size_t)n + 1), sizeof(uint16_t) * ((size_t)n + 1), (size_t)in->n_bytes + 1, sizeof(uint32_t) * ((size_t)nt + 1),
                          sizeof(uint32_t) * ((size_t)n_run + 1), sizeof(int32_t) * ((size_t)n_run + 1), sizeof(uint32_t) * ((size_t)n_ext + 1),
                          sizeof(uint32_t) * (2 * (size_t)n_ext + 1)};
    cudaSetDevice(ctx->device);
    for (int i = 0; i < 8; ++i)
        if (cudaMallocHost(&h->mem[i], sz[i]) != cudaSuccess)
        {
            vtn_bi_host_destroy(ctx, h);
            return vtn_set_err(ctx, VTN_E_NOMEM, "vtn_bi_host_create: pinned allocation failed");
        }
    int32_t* pos1 = (int32_t*)h->mem[0];
    uint16_t* lens = (uint16_t*)h->mem[1];
    uint8_t* bytes = (uint8_t*)h->mem[2];
    uint32_t* tile_off = (uint32_t*)h->mem[3];
    uint32_t* run_start = (uint32_t*)h->mem[4];
    int32_t* run_rid = (int32_t*)h->mem[5];
    uint32_t* ext_rec = (uint32_t*)h->mem[6];
    uint32_t* ext_len = (uint32_t*)h->mem[7];
    if (n) memcpy(pos1, in->pos1, sizeof(int32_t) * n);
    if (in->n_bytes) memcpy(bytes, in->bytes, (size_t)in->n_bytes);
    uint32_t kr = 0, ke = 0;
    for (uint32_t r = 0; r < n; ++r)
    {
        if (r == 0 || in->rid[r] != in->rid[r - 1])
        {
            run_start[kr] = r;
            run_rid[kr++] = in->rid[r];
        }
        if (r % VTN_BI_TILE == 0) tile_off[r / VTN_BI_TILE] = in->allele_off[2 * r];
        const uint32_t a = in->allele_off[2 * r + 1] - in->allele_off[2 * r], b = in->allele_off[2 * r + 2] - in->allele_off[2 * r + 1];
        if (a >= VTN_BI_LEN_ESC || b >= VTN_BI_LEN_ESC)
        {
            ext_rec[ke] = r;
            ext_len[2 * ke] = a;
            ext_len[2 * ke + 1] = b;
            ++ke;
            lens[r] = (uint16_t)(VTN_BI_LEN_ESC | (VTN_BI_LEN_ESC << 8));
        }
        else
            lens[r] = (uint16_t)(a | (b << 8));
    }
    run_start[kr] = n;
    tile_off[nt] = (uint32_t)in->n_bytes;
    h->v.n_rec = n;
    h->v.n_run = n_run;
    h->v.n_ext = n_ext;
    h->v.n_bytes = in->n_bytes;
    h->v.pos1 = pos1;
    h->v.lens = lens;
    h->v.bytes = bytes;
    h->v.tile_off = tile_off;
    h->v.run_start = run_start;
    h->v.run_rid = run_rid;
    h->v.ext_rec = ext_rec;
    h->v.ext_len = ext_len;
    *out = h;
    return VTN_OK;
}

const vtn_batch_bi* vtn_bi_host_view(const vtn_bi_host* h) { return h ? &h->v : nullptr; }

int vtn_bi_host_destroy(vtn_ctx* ctx, vtn_bi_host* h)
{
    if (!h) return VTN_OK;
    if (ctx) cudaSetDevice(ctx->device);
    for (int i = 0; i < 8; ++i)
        if (h->mem[i]) cudaFreeHost(h->mem[i]);
    delete h;
    return VTN_OK;
}

static int check_bi(vtn_ctx* ctx, const vtn_batch_bi* in)
{
    if (!in) return vtn_set_err(ctx, VTN_E_ARG, "null batch");
    if (!in->n_rec) return VTN_OK;
    if (!in->pos1 || !in->lens || !in->tile_off || !in->run_start || !in->run_rid || (in->n_bytes && !in->bytes) || (in->n_ext && (!in->ext_rec || !in->ext_len)))
        return vtn_set_err(ctx, VTN_E_ARG, "biallelic batch has null arrays");
    if (in->n_bytes >= (1ull << 32)) return vtn_set_err(ctx, VTN_E_ARG, "batch allele bytes must be < 4 GiB");
    if (in->n_run == 0 || in->run_start[0] != 0 || in->run_start[in->n_run] != in->n_rec)
        return vtn_set_err(ctx, VTN_E_ARG, "biallelic batch: contig runs must cover [0, n_rec)");
    for (uint32_t k = 0; k < in->n_run; ++k)
        if (in->run_start[k] >= in->run_start[k + 1]) return vtn_set_err(ctx, VTN_E_ARG, "biallelic batch: run_start must be strictly increasing");
    for (uint32_t k = 1; k < in->n_ext; ++k)
        if (in->ext_rec[k - 1] >= in->ext_rec[k]) return vtn_set_err(ctx, VTN_E_ARG, "biallelic batch: ext_rec must be strictly increasing");
    const uint32_t nt = bi_tiles(in->n_rec);
    if (in->tile_off[0] != 0 || in->tile_off[nt] != in->n_bytes) return vtn_set_err(ctx, VTN_E_ARG, "biallelic batch: tile_off inconsistent with n_bytes");
    return VTN_OK;
}

template <typename T>
static cudaError_t dalloc(T** p, size_t n)
{
    return cudaMalloc((void**)p, (n + 512) * sizeof(T)); // padded: tiles load whole 128-bit vectors
}

static int alloc_bi(vtn_ctx* ctx, uint32_t n_rec, uint64_t n_bytes, uint32_t n_run, uint32_t n_ext, vtn_dbatch** out)
{
    vtn_dbatch* d = new vtn_dbatch();
    d->n_rec = d->cap_rec = n_rec;
    d->n_allele = d->cap_allele = 2 * n_rec;
    d->n_bytes = d->cap_bytes = n_bytes;
    d->n_run = d->cap_run = n_run;
    d->n_ext = d->cap_ext = n_ext;
    const uint32_t tile = vtn_normalize_tile_records();
    d->n_tiles = (n_rec + tile - 1) / tile;
    cudaError_t e = cudaSuccess;
#define A(x) if (e == cudaSuccess) e = (x)
    A(dalloc(&d->pos1, n_rec));
    A(dalloc(&d->lens, n_rec));
    A(dalloc(&d->bytes, n_bytes + 16));
    A(dalloc(&d->tile_off, (size_t)d->n_tiles + 1));
    A(dalloc(&d->run_start, (size_t)n_run + 1));
    A(dalloc(&d->run_rid, (size_t)n_run + 1));
    A(dalloc(&d->ext_rec, (size_t)n_ext + 1));
    A(dalloc(&d->ext_len, 2 * (size_t)n_ext + 1));
    A(dalloc(&d->status, n_rec)); // sv
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
    A(dalloc(&d->tile_state, 2 * (size_t)d->n_tiles + 16));
    A(dalloc(&d->tile_counter, 2 * vtn_normalize_ctl_words(d->n_tiles)));
    A(dalloc(&d->note_rec, n_rec));
    A(dalloc(&d->note_status, n_rec));
    A(dalloc(&d->note_count, 4));
#undef A
    if (e != cudaSuccess)
    {
        vtn_dbatch_destroy(ctx, d);
        return vtn_set_err(ctx, VTN_E_NOMEM, std::string("device batch allocation: ") + cudaGetErrorString(e));
    }
    *out = d;
    return VTN_OK;
}

static int upload_bi_small(vtn_ctx* ctx, const vtn_batch_bi* in, vtn_dbatch* d, cudaStream_t s)
{ // the arrays every chunk needs whole: contig runs, long-record table
    VTN_CUDA(ctx, cudaMemcpyAsync(d->run_start, in->run_start, sizeof(uint32_t) * ((size_t)in->n_run + 1), cudaMemcpyHostToDevice, s));
    VTN_CUDA(ctx, cudaMemcpyAsync(d->run_rid, in->run_rid, sizeof(int32_t) * in->n_run, cudaMemcpyHostToDevice, s));
    if (in->n_ext)
    {
        VTN_CUDA(ctx, cudaMemcpyAsync(d->ext_rec, in->ext_rec, sizeof(uint32_t) * in->n_ext, cudaMemcpyHostToDevice, s));
        VTN_CUDA(ctx, cudaMemcpyAsync(d->ext_len, in->ext_len, sizeof(uint32_t) * 2 * in->n_ext, cudaMemcpyHostToDevice, s));
    }
    return VTN_OK;
}

int vtn_dbatch_create_bi(vtn_ctx* ctx, const vtn_batch_bi* in, vtn_dbatch** db)
{
    if (!ctx || !db) return VTN_E_ARG;
    int rc = check_bi(ctx, in);
    if (rc) return rc;
    VTN_CUDA(ctx, cudaSetDevice(ctx->device));
    if (ctx->ref_missing && (rc = vtn_ref_ensure(ctx, in->n_run, in->run_rid))) return rc;
    rc = alloc_bi(ctx, in->n_rec, in->n_bytes, in->n_run, in->n_ext, db);
    if (rc) return rc;
    vtn_dbatch* d = *db;
    cudaStream_t s = ctx->stream;
    if (in->n_rec)
    {
        VTN_CUDA(ctx, cudaMemcpyAsync(d->pos1, in->pos1, sizeof(int32_t) * in->n_rec, cudaMemcpyHostToDevice, s));
        VTN_CUDA(ctx, cudaMemcpyAsync(d->lens, in->lens, sizeof(uint16_t) * in->n_rec, cudaMemcpyHostToDevice, s));
        VTN_CUDA(ctx, cudaMemcpyAsync(d->tile_off, in->tile_off, sizeof(uint32_t) * ((size_t)d->n_tiles + 1), cudaMemcpyHostToDevice, s));
        if (in->n_bytes) VTN_CUDA(ctx, cudaMemcpyAsync(d->bytes, in->bytes, (size_t)in->n_bytes, cudaMemcpyHostToDevice, s));
        rc = upload_bi_small(ctx, in, d, s);
        if (rc) return rc;
    }
    VTN_CUDA(ctx, cudaStreamSynchronize(s));
    return VTN_OK;
}

static VtnDevBiView bi_view(const vtn_dbatch* d, uint32_t r0, uint32_t nr)
{
    const uint32_t tile = vtn_normalize_tile_records();
    return VtnDevBiView{nr, r0, d->pos1 + r0, d->lens + r0, d->bytes, d->tile_off + r0 / tile, d->run_start, d->run_rid, d->n_run, d->ext_rec, d->ext_len, d->n_ext,
                        d->n_bytes > 6ull * d->n_rec ? 1u : 0u};
}

int vtn_normalize_resident_bi(vtn_ctx* ctx, vtn_dbatch* d)
{
    VtnDevNormOut o{d->status, nullptr, d->chg_rec, d->chg_pos1, d->chg_lt, d->chg_le, d->chg_rt, d->chg_off, d->chg_bytes, d->totals, d->counters};
    VtnDevNotes nt{d->note_rec, d->note_status, d->note_count, d->cap_rec};
    return vtn_launch_normalize_bi(ctx, bi_view(d, 0, d->n_rec), o, nt, d->scr, d->tile_state, d->tile_counter, d->n_tiles);
}

static void sort_notes(uint32_t n, uint32_t* rec, uint8_t* st)
{
    if (n < 2) return;
    std::vector<std::pair<uint32_t, uint8_t>> v(n);
    for (uint32_t i = 0; i < n; ++i) v[i] = {rec[i], st[i]};
    std::sort(v.begin(), v.end());
    for (uint32_t i = 0; i < n; ++i)
    {
        rec[i] = v[i].first;
        st[i] = v[i].second;
    }
}

static int fetch_tail(vtn_ctx* ctx, vtn_dbatch* d, vtn_norm_result_bi* out, cudaStream_t s, unsigned long long totals[2])
{ // counters + notes (after the kernels of the last chunk)
    uint32_t counters[12];
    uint32_t n_notes = 0;
    VTN_CUDA(ctx, cudaMemcpyAsync(totals, d->totals, 2 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));
    VTN_CUDA(ctx, cudaMemcpyAsync(counters, d->counters, sizeof counters, cudaMemcpyDeviceToHost, s));
    VTN_CUDA(ctx, cudaMemcpyAsync(&n_notes, d->note_count, sizeof n_notes, cudaMemcpyDeviceToHost, s));
    VTN_CUDA(ctx, cudaStreamSynchronize(s));
    out->n_changed = (uint32_t)totals[0];
    out->n_changed_bytes = totals[1];
    memcpy(out->counters.bi, counters, 5 * sizeof(uint32_t));
    memcpy(out->counters.multi, counters + 5, 5 * sizeof(uint32_t));
    out->counters.no_variants = counters[10];
    out->counters.no_refs = counters[11];
    out->n_notes = n_notes;
    if (n_notes > out->cap_notes) return vtn_set_err(ctx, VTN_E_CAPACITY, "vtn_normalize_bi: note capacity too small (n_notes is set)");
    if (n_notes)
    {
        VTN_CUDA(ctx, cudaMemcpyAsync(out->note_rec, d->note_rec, sizeof(uint32_t) * n_notes, cudaMemcpyDeviceToHost, s));
        VTN_CUDA(ctx, cudaMemcpyAsync(out->note_status, d->note_status, n_notes, cudaMemcpyDeviceToHost, s));
        VTN_CUDA(ctx, cudaStreamSynchronize(s));
        sort_notes(n_notes, out->note_rec, out->note_status);
    }
    return VTN_OK;
}

int vtn_dbatch_fetch_result_bi(vtn_ctx* ctx, vtn_dbatch* d, vtn_norm_result_bi* out)
{
    if (!ctx || !d || !out || !d->lens) return VTN_E_ARG;
    cudaStream_t s = ctx->stream;
    unsigned long long totals[2] = {0, 0};
    int rc = fetch_tail(ctx, d, out, s, totals);
    if (rc) return rc;
    if (d->n_rec && out->sv) VTN_CUDA(ctx, cudaMemcpyAsync(out->sv, d->status, d->n_rec, cudaMemcpyDeviceToHost, s));
    if (totals[0] > out->cap_changed || totals[1] > out->cap_changed_bytes)
        return vtn_set_err(ctx, VTN_E_CAPACITY, "vtn_normalize_bi: result capacity too small (n_changed / n_changed_bytes are set)");
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

static bool bi_fits(const vtn_dbatch* d, uint32_t n_rec, uint64_t n_bytes, uint32_t n_run, uint32_t n_ext)
{
    return d && n_rec <= d->cap_rec && n_bytes <= d->cap_bytes && n_run <= d->cap_run && n_ext <= d->cap_ext;
}

// Host-buffer normalize over the compact layout, pipelined exactly like vtn_normalize (vtn_api.cu): chunk c's H2D copies
// (stream2), kernels (stream) and sv + changed-record D2H copies (stream3) overlap with the other chunks'.
int vtn_normalize_bi(vtn_ctx* ctx, const vtn_batch_bi* in, vtn_norm_result_bi* out)
{
    if (!ctx || !out) return VTN_E_ARG;
    if (!ctx->ref_loaded) return vtn_set_err(ctx, VTN_E_STATE, "vtn_normalize_bi: no reference loaded");
    int rc = check_bi(ctx, in);
    if (rc) return rc;
    VTN_CUDA(ctx, cudaSetDevice(ctx->device));
    if (ctx->ref_missing && (rc = vtn_ref_ensure(ctx, in->n_run, in->run_rid))) return rc;
    if (!bi_fits(ctx->scratch_bi, in->n_rec, in->n_bytes, in->n_run, in->n_ext))
    {
        if (ctx->scratch_bi) vtn_dbatch_destroy(ctx, ctx->scratch_bi);
        ctx->scratch_bi = nullptr;
        rc = alloc_bi(ctx, in->n_rec + in->n_rec / 8 + 1024, in->n_bytes + in->n_bytes / 8 + 4096, in->n_run + in->n_run / 8 + 64, in->n_ext + in->n_ext / 8 + 64, &ctx->scratch_bi);
        if (rc) return rc;
    }
    vtn_dbatch* d = ctx->scratch_bi;
    const uint32_t n = in->n_rec;
    const uint32_t tile = vtn_normalize_tile_records();
    d->n_rec = n;
    d->n_bytes = in->n_bytes;
    d->n_run = in->n_run;
    d->n_ext = in->n_ext;
    d->n_tiles = (n + tile - 1) / tile;
    const uint32_t CH = ctx->chunk_rec;
    const uint32_t n_chunks = (n + CH - 1) / CH;
    if (ctx->chunk_events.size() < 3 * (size_t)n_chunks + 3)
    {
        size_t old = ctx->chunk_events.size();
        ctx->chunk_events.resize(3 * (size_t)n_chunks + 3);
        for (size_t i = old; i < ctx->chunk_events.size(); ++i) VTN_CUDA(ctx, cudaEventCreateWithFlags(&ctx->chunk_events[i], cudaEventDisableTiming));
    }
    if (!ctx->pinned_totals) VTN_CUDA(ctx, cudaMallocHost((void**)&ctx->pinned_totals, 4096 * 2 * sizeof(unsigned long long)));
    if (n_chunks > 4096) return vtn_set_err(ctx, VTN_E_ARG, "vtn_normalize_bi: batch too large");
    cudaStream_t sc = ctx->stream, sh = ctx->stream2, sd = ctx->stream3;
    cudaEvent_t ev_start = ctx->chunk_events[3 * (size_t)n_chunks];
    VTN_CUDA(ctx, cudaEventRecord(ev_start, sc));
    VTN_CUDA(ctx, cudaStreamWaitEvent(sh, ev_start, 0));
    VTN_CUDA(ctx, cudaStreamWaitEvent(sd, ev_start, 0));
    VtnDevNormOut o_all{d->status, nullptr, d->chg_rec, d->chg_pos1, d->chg_lt, d->chg_le, d->chg_rt, d->chg_off, d->chg_bytes, d->totals, d->counters};
    VtnDevNotes nt{d->note_rec, d->note_status, d->note_count, d->cap_rec};
    unsigned long long done_c = 0, done_b = 0;
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
        rc = vtn_launch_normalize_bi(ctx, bi_view(d, 0, 0), o_all, nt, d->scr, d->tile_state, d->tile_counter, 0, false);
        if (rc) return rc;
    }
    else
    {
        rc = upload_bi_small(ctx, in, d, sh);
        if (rc) return rc;
        VTN_CUDA(ctx, cudaMemcpyAsync(d->tile_off, in->tile_off, sizeof(uint32_t) * ((size_t)d->n_tiles + 1), cudaMemcpyHostToDevice, sh));
    }
    for (uint32_t c = 0; c < n_chunks; ++c)
    {
        const uint32_t r0 = c * CH, r1 = (c + 1 == n_chunks) ? n : r0 + CH, nr = r1 - r0;
        const uint32_t tile0 = r0 / tile, ntl = (nr + tile - 1) / tile;
        const uint32_t b0 = in->tile_off[tile0], b1 = in->tile_off[tile0 + ntl];
        cudaEvent_t ev_h = ctx->chunk_events[3 * (size_t)c], ev_k = ctx->chunk_events[3 * (size_t)c + 1], ev_t = ctx->chunk_events[3 * (size_t)c + 2];
        VTN_CUDA(ctx, cudaMemcpyAsync(d->lens + r0, in->lens + r0, sizeof(uint16_t) * nr, cudaMemcpyHostToDevice, sh));
        if (b1 > b0) VTN_CUDA(ctx, cudaMemcpyAsync(d->bytes + b0, in->bytes + b0, (size_t)(b1 - b0), cudaMemcpyHostToDevice, sh));
        VTN_CUDA(ctx, cudaMemcpyAsync(d->pos1 + r0, in->pos1 + r0, sizeof(int32_t) * nr, cudaMemcpyHostToDevice, sh));
        VTN_CUDA(ctx, cudaEventRecord(ev_h, sh));
        VTN_CUDA(ctx, cudaStreamWaitEvent(sc, ev_h, 0));
        VtnDevNormOut o = o_all;
        o.status = d->status + r0;
        VtnDevScratch x = d->scr;
        x.q0 += r0; x.q1 += r0; x.dir += tile0; x.blo += tile0;
        rc = vtn_launch_normalize_bi(ctx, bi_view(d, r0, nr), o, nt, x, d->tile_state, d->tile_counter, ntl, c != 0);
        if (rc) return rc;
        VTN_CUDA(ctx, cudaEventRecord(ev_k, sc));
        VTN_CUDA(ctx, cudaStreamWaitEvent(sd, ev_k, 0));
        if (out->sv) VTN_CUDA(ctx, cudaMemcpyAsync(out->sv + r0, d->status + r0, nr, cudaMemcpyDeviceToHost, sd));
        VTN_CUDA(ctx, cudaMemcpyAsync(ctx->pinned_totals + 2 * (size_t)c, d->totals, 2 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, sd));
        VTN_CUDA(ctx, cudaEventRecord(ev_t, sd));
        if (c >= 1)
        {
            VTN_CUDA(ctx, cudaEventSynchronize(ctx->chunk_events[3 * (size_t)(c - 1) + 2]));
            rc = copy_changed(ctx->pinned_totals[2 * (size_t)(c - 1)], ctx->pinned_totals[2 * (size_t)(c - 1) + 1]);
            if (rc) return rc;
        }
    }
    VTN_CUDA(ctx, cudaStreamSynchronize(sc));
    VTN_CUDA(ctx, cudaStreamSynchronize(sd));
    unsigned long long totals[2] = {0, 0};
    rc = fetch_tail(ctx, d, out, sd, totals);
    const int rc_tail = rc;
    if (rc && rc != VTN_E_CAPACITY) return rc;
    rc = copy_changed(totals[0], totals[1]);
    if (rc) return rc;
    if (!overflow && totals[0] <= out->cap_changed)
        VTN_CUDA(ctx, cudaMemcpyAsync(out->chg_off + totals[0], d->chg_off + totals[0], 8, cudaMemcpyDeviceToHost, sd));
    VTN_CUDA(ctx, cudaStreamSynchronize(sd));
    if (overflow || totals[0] > out->cap_changed || totals[1] > out->cap_changed_bytes)
        return vtn_set_err(ctx, VTN_E_CAPACITY, "vtn_normalize_bi: result capacity too small (n_changed / n_changed_bytes are set)");
    return rc_tail;
}
