// vtn_multi.cu -- several GPUs of one box behind one call (include/vtn.h: vtn_multi_*).
//
// The path shards trivially (records are independent, the reference is read-only), so there is no collective and no NCCL:
// one vtn_ctx per device, the reference packed once and replicated by peer copies (vtn_ref_clone), one host thread per device
// for the duration of a call, the batch cut into contiguous record ranges (vtn_shard_plan: by contig, then by record range --
// the reference's own way to go parallel is one process per interval, program.cpp:152-190), every range through the
// single-device path on its device, and the results put back together in input order.
#include <string.h>
#include <algorithm>
#include <condition_variable>
#include <mutex>
#include <thread>
#include "vtn_internal.h"

struct ShardScratch // pinned staging of one shard's compacted results (grown on demand, kept across calls)
{
    void* mem[12] = {nullptr};
    size_t cap[12] = {0};
    void* get(int i, size_t bytes)
    {
        if (bytes > cap[i])
        {
            if (mem[i]) cudaFreeHost(mem[i]);
            mem[i] = nullptr;
            cap[i] = 0;
            const size_t want = bytes + bytes / 8 + 4096;
            if (cudaMallocHost(&mem[i], want) != cudaSuccess) return nullptr;
            cap[i] = want;
        }
        return mem[i];
    }
    void release()
    {
        for (int i = 0; i < 12; ++i)
            if (mem[i]) cudaFreeHost(mem[i]);
    }
};

struct vtn_multi
{
    std::vector<vtn_ctx*> ctx;
    std::vector<ShardScratch> scr;
    std::string err;
};

static thread_local std::string g_multi_err;
static int multi_err(vtn_multi* m, int code, const std::string& msg)
{
    if (m) m->err = msg;
    else g_multi_err = msg;
    return code;
}

const char* vtn_multi_last_error(const vtn_multi* m) { return m ? m->err.c_str() : g_multi_err.c_str(); }

int vtn_multi_create(vtn_multi** out, const int* device_ids, int n_dev)
{
    if (!out || !device_ids || n_dev < 1) return multi_err(nullptr, VTN_E_ARG, "vtn_multi_create: bad argument");
    *out = nullptr;
    vtn_multi* m = new vtn_multi();
    for (int i = 0; i < n_dev; ++i)
    {
        vtn_ctx* c = nullptr;
        int rc = vtn_ctx_create(&c, device_ids[i]);
        if (rc)
        {
            const std::string e = vtn_last_error(nullptr);
            vtn_multi_destroy(m);
            return multi_err(nullptr, rc, "vtn_multi_create: device " + std::to_string(device_ids[i]) + ": " + e);
        }
        m->ctx.push_back(c);
    }
    m->scr.resize(n_dev);
    *out = m;
    return VTN_OK;
}

int vtn_multi_destroy(vtn_multi* m)
{
    if (!m) return VTN_OK;
    for (size_t i = 0; i < m->ctx.size(); ++i)
    {
        cudaSetDevice(m->ctx[i]->device);
        if (i < m->scr.size()) m->scr[i].release();
        vtn_ctx_destroy(m->ctx[i]);
    }
    delete m;
    return VTN_OK;
}

int vtn_multi_n_dev(const vtn_multi* m) { return m ? (int)m->ctx.size() : 0; }
vtn_ctx* vtn_multi_ctx(vtn_multi* m, int slot) { return (m && slot >= 0 && slot < (int)m->ctx.size()) ? m->ctx[slot] : nullptr; }

static int replicate(vtn_multi* m)
{
    for (size_t i = 1; i < m->ctx.size(); ++i)
    {
        int rc = vtn_ref_clone(m->ctx[i], m->ctx[0]);
        if (rc) return multi_err(m, rc, std::string("replicating the reference: ") + vtn_last_error(m->ctx[i]));
    }
    return VTN_OK;
}

int vtn_multi_ref_load_fasta(vtn_multi* m, const char* path)
{
    if (!m) return VTN_E_ARG;
    int rc = vtn_ref_load_fasta(m->ctx[0], path);
    if (rc) return multi_err(m, rc, vtn_last_error(m->ctx[0]));
    return replicate(m);
}

int vtn_multi_ref_load_memory(vtn_multi* m, int32_t n_contig, const char* const* names, const int64_t* len, const uint8_t* const* seq)
{
    if (!m) return VTN_E_ARG;
    int rc = vtn_ref_load_memory(m->ctx[0], n_contig, names, len, seq);
    if (rc) return multi_err(m, rc, vtn_last_error(m->ctx[0]));
    return replicate(m);
}

// Contiguous record ranges, one per shard: equal record counts (cuts on multiples of `align`), and a cut that falls within a
// sixteenth of a shard of a contig change is moved onto it, so that a sorted call set is cut "by contig, then by record range".
int vtn_shard_plan(const int32_t* rid, uint32_t n_rec, uint32_t n_shard, uint32_t align, uint32_t* cut)
{
    if (!cut || n_shard < 1 || (n_rec && !rid)) return VTN_E_ARG;
    if (align < 1) align = 1;
    cut[0] = 0;
    cut[n_shard] = n_rec;
    const uint64_t per = ((uint64_t)n_rec + n_shard - 1) / n_shard;
    const uint32_t slack = (uint32_t)(per / 16);
    for (uint32_t s = 1; s < n_shard; ++s)
    {
        uint64_t c = (uint64_t)n_rec * s / n_shard;
        c = (c + align / 2) / align * align;
        if (c > n_rec) c = n_rec;
        uint32_t best = (uint32_t)c;
        for (uint32_t d = 0; d <= slack; ++d) // nearest contig change within the slack (rounded up to the alignment)
        {
            bool found = false;
            for (int sign = -1; sign <= 1 && !found; sign += 2)
            {
                const int64_t p = (int64_t)c + sign * (int64_t)d;
                if (p <= 0 || p >= (int64_t)n_rec) continue;
                if (rid[p] != rid[p - 1])
                {
                    best = (uint32_t)(((uint64_t)p + align - 1) / align * align);
                    found = true;
                }
            }
            if (found) break;
        }
        if (best > n_rec) best = n_rec;
        cut[s] = best < cut[s - 1] ? cut[s - 1] : best;
    }
    return VTN_OK;
}

static void add_counters(vtn_norm_counters& a, const vtn_norm_counters& b)
{
    for (int i = 0; i < 5; ++i) a.bi[i] += b.bi[i], a.multi[i] += b.multi[i];
    a.no_variants += b.no_variants;
    a.no_refs += b.no_refs;
}

// ---------------------------------------------------------------------------------------------------------
// compact biallelic layout over several devices
// ---------------------------------------------------------------------------------------------------------
int vtn_multi_normalize_bi(vtn_multi* m, const vtn_batch_bi* in, vtn_norm_result_bi* out)
{
    if (!m || !in || !out) return VTN_E_ARG;
    const uint32_t S = (uint32_t)m->ctx.size(), n = in->n_rec;
    if (S == 1)
    { // nothing to shard or to put together
        int rc = vtn_normalize_bi(m->ctx[0], in, out);
        if (rc) return multi_err(m, rc, vtn_last_error(m->ctx[0]));
        return VTN_OK;
    }
    if (n && (!in->pos1 || !in->lens || !in->tile_off || !in->run_start || !in->run_rid || !in->n_run)) return multi_err(m, VTN_E_ARG, "biallelic batch has null arrays");
    std::vector<uint32_t> cut(S + 1);
    cut[0] = 0;
    cut[S] = n;
    { // contig ids are runs here: cut on tiles, and let a run boundary near a cut attract it
        const uint64_t per = ((uint64_t)n + S - 1) / S;
        for (uint32_t s = 1; s < S; ++s)
        {
            uint64_t c = (uint64_t)n * s / S;
            c = (c + VTN_BI_TILE / 2) / VTN_BI_TILE * VTN_BI_TILE;
            if (c > n) c = n;
            for (uint32_t k = 1; k < in->n_run; ++k)
            {
                const uint64_t b = ((uint64_t)in->run_start[k] + VTN_BI_TILE - 1) / VTN_BI_TILE * VTN_BI_TILE;
                const uint64_t dist = b > c ? b - c : c - b;
                if (dist <= per / 16 && b <= n)
                {
                    c = b;
                    break;
                }
            }
            cut[s] = (uint32_t)std::max<uint64_t>(c, cut[s - 1]);
        }
    }
    struct Shard
    {
        vtn_batch_bi v{};
        std::vector<uint32_t> tile_off, run_start, ext_rec;
        std::vector<int32_t> run_rid;
        vtn_norm_result_bi r{};
        int rc = 0;
        std::string err;
    };
    std::vector<Shard> sh(S);
    std::vector<std::thread> th;
    for (uint32_t s = 0; s < S; ++s)
        th.emplace_back([&, s]() {
            Shard& H = sh[s];
            vtn_ctx* ctx = m->ctx[s];
            cudaSetDevice(ctx->device);
            const uint32_t r0 = cut[s], r1 = cut[s + 1], ns = r1 - r0;
            const uint32_t t0 = r0 / VTN_BI_TILE, nt = (ns + VTN_BI_TILE - 1) / VTN_BI_TILE;
            const uint32_t b0 = ns ? in->tile_off[t0] : 0, b1 = ns ? in->tile_off[t0 + nt] : 0;
            // the shard as a batch of its own: pos1 / lens / bytes are slices, the small tables are re-based
            H.tile_off.resize(nt + 1);
            for (uint32_t t = 0; t <= nt; ++t) H.tile_off[t] = (ns ? in->tile_off[t0 + t] : 0) - b0;
            for (uint32_t k = 0; k < in->n_run; ++k)
            {
                const uint32_t a = std::max(in->run_start[k], r0), b = std::min(in->run_start[k + 1], r1);
                if (a < b)
                {
                    H.run_start.push_back(a - r0);
                    H.run_rid.push_back(in->run_rid[k]);
                }
            }
            H.run_start.push_back(ns);
            const uint32_t* e0 = std::lower_bound(in->ext_rec, in->ext_rec + in->n_ext, r0);
            const uint32_t* e1 = std::lower_bound(in->ext_rec, in->ext_rec + in->n_ext, r1);
            for (const uint32_t* e = e0; e < e1; ++e) H.ext_rec.push_back(*e - r0);
            H.ext_rec.push_back(0);
            H.v.n_rec = ns;
            H.v.n_run = (uint32_t)H.run_rid.size();
            H.v.n_ext = (uint32_t)(e1 - e0);
            H.v.n_bytes = b1 - b0;
            H.v.pos1 = in->pos1 + r0;
            H.v.lens = in->lens + r0;
            H.v.bytes = in->bytes + b0;
            H.v.tile_off = H.tile_off.data();
            H.v.run_start = H.run_start.data();
            H.v.run_rid = H.run_rid.data();
            H.v.ext_rec = H.ext_rec.data();
            H.v.ext_len = in->ext_len + 2 * (e0 - in->ext_rec);
            ShardScratch& X = m->scr[s];
            vtn_norm_result_bi& R = H.r;
            R.sv = out->sv ? out->sv + r0 : nullptr;
            R.cap_changed = ns;
            R.cap_changed_bytes = H.v.n_bytes;
            R.cap_notes = ns;
            R.chg_rec = (uint32_t*)X.get(0, 4 * (size_t)ns + 4);
            R.chg_pos1 = (int32_t*)X.get(1, 4 * (size_t)ns + 4);
            R.chg_lt = (uint32_t*)X.get(2, 4 * (size_t)ns + 4);
            R.chg_le = (uint32_t*)X.get(3, 4 * (size_t)ns + 4);
            R.chg_rt = (uint32_t*)X.get(4, 4 * (size_t)ns + 4);
            R.chg_off = (uint64_t*)X.get(5, 8 * ((size_t)ns + 1));
            R.chg_bytes = (uint8_t*)X.get(6, (size_t)H.v.n_bytes + 16);
            R.note_rec = (uint32_t*)X.get(7, 4 * (size_t)ns + 4);
            R.note_status = (uint8_t*)X.get(8, (size_t)ns + 4);
            if (!R.chg_rec || !R.chg_pos1 || !R.chg_lt || !R.chg_le || !R.chg_rt || !R.chg_off || !R.chg_bytes || !R.note_rec || !R.note_status)
            {
                H.rc = VTN_E_NOMEM;
                H.err = "pinned staging allocation failed";
                return;
            }
            H.rc = vtn_normalize_bi(ctx, &H.v, &R);
            if (H.rc) H.err = vtn_last_error(ctx);
        });
    for (auto& t : th) t.join();
    th.clear();
    for (uint32_t s = 0; s < S; ++s)
        if (sh[s].rc) return multi_err(m, sh[s].rc, "device " + std::to_string(m->ctx[s]->device) + ": " + sh[s].err);
    // ---- put the shards' results together in input order ----
    std::vector<uint64_t> kc(S + 1, 0), kb(S + 1, 0), kn(S + 1, 0);
    memset(&out->counters, 0, sizeof out->counters);
    for (uint32_t s = 0; s < S; ++s)
    {
        kc[s + 1] = kc[s] + sh[s].r.n_changed;
        kb[s + 1] = kb[s] + sh[s].r.n_changed_bytes;
        kn[s + 1] = kn[s] + sh[s].r.n_notes;
        add_counters(out->counters, sh[s].r.counters);
    }
    out->n_changed = (uint32_t)kc[S];
    out->n_changed_bytes = kb[S];
    out->n_notes = (uint32_t)kn[S];
    if (kc[S] > out->cap_changed || kb[S] > out->cap_changed_bytes) return multi_err(m, VTN_E_CAPACITY, "vtn_multi_normalize_bi: result capacity too small (n_changed / n_changed_bytes are set)");
    if (kn[S] > out->cap_notes) return multi_err(m, VTN_E_CAPACITY, "vtn_multi_normalize_bi: note capacity too small (n_notes is set)");
    for (uint32_t s = 0; s < S; ++s)
        th.emplace_back([&, s]() {
            const vtn_norm_result_bi& R = sh[s].r;
            const size_t k0 = (size_t)kc[s], k = R.n_changed;
            const uint32_t r0 = cut[s];
            for (size_t j = 0; j < k; ++j) out->chg_rec[k0 + j] = R.chg_rec[j] + r0;
            if (k)
            {
                memcpy(out->chg_pos1 + k0, R.chg_pos1, 4 * k);
                memcpy(out->chg_lt + k0, R.chg_lt, 4 * k);
                memcpy(out->chg_le + k0, R.chg_le, 4 * k);
                memcpy(out->chg_rt + k0, R.chg_rt, 4 * k);
                memcpy(out->chg_bytes + kb[s], R.chg_bytes, (size_t)R.n_changed_bytes);
            }
            for (size_t j = 0; j < k; ++j) out->chg_off[k0 + j] = R.chg_off[j] + kb[s];
            for (size_t j = 0; j < R.n_notes; ++j)
            {
                out->note_rec[kn[s] + j] = R.note_rec[j] + r0;
                out->note_status[kn[s] + j] = R.note_status[j];
            }
        });
    for (auto& t : th) t.join();
    out->chg_off[kc[S]] = kb[S];
    return VTN_OK;
}

// ---------------------------------------------------------------------------------------------------------
// decompose -> normalize (fused on each device) over several devices: configuration C3.  Two phases (vtn_api.cu:
// vtn_shard_decompose_run / _fetch): every device runs its record range and reports its totals; with all totals known each
// device adds its bases to its index arrays on the device and copies straight into the caller's arrays -- no host merge pass.
// ---------------------------------------------------------------------------------------------------------
int vtn_multi_decompose_normalize(vtn_multi* m, const vtn_batch* in, uint32_t* src_rec, uint16_t* alt_idx, uint64_t* n_out, vtn_decomp_counters* dcounters,
                                  vtn_norm_result* out)
{
    if (!m || !in || !out || !src_rec || !alt_idx || !n_out) return VTN_E_ARG;
    const uint32_t S = (uint32_t)m->ctx.size(), n = in->n_rec;
    if (S == 1)
    {
        int rc = vtn_decompose_normalize(m->ctx[0], in, src_rec, alt_idx, n_out, dcounters, out);
        if (rc) return multi_err(m, rc, vtn_last_error(m->ctx[0]));
        return VTN_OK;
    }
    if (n && (!in->pos1 || !in->rid || !in->allele_beg || !in->allele_off || (in->n_bytes && !in->bytes))) return multi_err(m, VTN_E_ARG, "batch has null arrays");
    // More chunks than slots: slot s takes chunks s, s + S, ... in order.  A chunk's kernels need nothing from the others; only its
    // copy-out needs the totals of every chunk before it (its place in the caller's arrays), so a slot waits for the earlier chunks'
    // RUNS, not for their copies: one slot's H2D overlaps another slot's D2H, on one device (repeat its id in vtn_multi_create) as
    // well as across devices.
    // One chunk per slot unless VTN_MULTI_CHUNK_REC asks for smaller ones (tests): measured on B200, more rounds of smaller chunks cost
    // more in per-chunk synchronisation than their finer overlap gains (12.5 M records over 2 GPUs: 10.5 ms with 2 chunks, 14.1 ms
    // with 6); the overlap comes from giving a device two or three slots.
    uint32_t rounds = 1;
    if (const char* e = getenv("VTN_MULTI_CHUNK_REC"))
        if (atol(e) > 0)
        {
            const uint64_t per_round = (uint64_t)atol(e) * S;
            rounds = (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>(64, (n + per_round / 2) / per_round));
        }
    const uint32_t C = S * rounds;
    std::vector<uint32_t> cut(C + 1);
    int rc = vtn_shard_plan(in->rid, n, C, 1, cut.data());
    if (rc) return multi_err(m, rc, "vtn_shard_plan failed");
    struct Shard
    {
        uint64_t n_out = 0, n_changed = 0, n_changed_bytes = 0;
        vtn_decomp_counters dc{};
        vtn_norm_counters nc{};
        bool run_done = false;
    };
    std::vector<Shard> sh(C);
    std::mutex mu;
    std::condition_variable cv;
    int fail_rc = 0;
    std::string fail_msg;
    auto fail = [&](int code, const std::string& msg) {
        std::lock_guard<std::mutex> g(mu);
        if (!fail_rc) fail_rc = code, fail_msg = msg;
        cv.notify_all();
    };
    std::vector<std::thread> th;
    for (uint32_t s = 0; s < S; ++s)
        th.emplace_back([&, s]() {
            for (uint32_t c = s; c < C; c += S)
            {
                Shard& H = sh[c];
                const int r1 = vtn_shard_decompose_run(m->ctx[s], in, cut[c], cut[c + 1], &H.n_out, &H.n_changed, &H.n_changed_bytes);
                if (r1)
                {
                    fail(r1, "device " + std::to_string(m->ctx[s]->device) + ": " + vtn_last_error(m->ctx[s]));
                    return;
                }
                uint64_t ko = 0, kc = 0, kb = 0;
                {
                    std::unique_lock<std::mutex> g(mu);
                    H.run_done = true;
                    cv.notify_all();
                    cv.wait(g, [&]() {
                        if (fail_rc) return true;
                        for (uint32_t j = 0; j < c; ++j)
                            if (!sh[j].run_done) return false;
                        return true;
                    });
                    if (fail_rc) return;
                    for (uint32_t j = 0; j < c; ++j) ko += sh[j].n_out, kc += sh[j].n_changed, kb += sh[j].n_changed_bytes;
                }
                if (kc + H.n_changed > out->cap_changed || kb + H.n_changed_bytes > out->cap_changed_bytes)
                {
                    fail(VTN_E_CAPACITY, "vtn_multi_decompose_normalize: result capacity too small");
                    return;
                }
                vtn_norm_result view = *out;
                const int r2 = vtn_shard_decompose_fetch(m->ctx[s], cut[c], ko, kc, kb, H.n_out, H.n_changed, H.n_changed_bytes, src_rec, alt_idx, &H.dc, &view);
                H.nc = view.counters;
                if (r2)
                {
                    fail(r2, "device " + std::to_string(m->ctx[s]->device) + ": " + vtn_last_error(m->ctx[s]));
                    return;
                }
            }
        });
    for (auto& t : th) t.join();
    if (fail_rc) return multi_err(m, fail_rc, fail_msg);
    uint64_t ko = 0, kc = 0, kb = 0;
    vtn_decomp_counters dtot{};
    memset(&out->counters, 0, sizeof out->counters);
    for (uint32_t c = 0; c < C; ++c)
    {
        ko += sh[c].n_out, kc += sh[c].n_changed, kb += sh[c].n_changed_bytes;
        add_counters(out->counters, sh[c].nc);
        dtot.no_variants += sh[c].dc.no_variants;
        dtot.no_biallelic += sh[c].dc.no_biallelic;
        dtot.no_multiallelic += sh[c].dc.no_multiallelic;
        dtot.no_additional_biallelic += sh[c].dc.no_additional_biallelic;
    }
    *n_out = ko;
    out->n_changed = (uint32_t)kc;
    out->n_changed_bytes = kb;
    out->chg_off[kc] = kb;
    if (dcounters) *dcounters = dtot;
    return VTN_OK;
}
