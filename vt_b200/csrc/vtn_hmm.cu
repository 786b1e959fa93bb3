// vtn_hmm.cu -- AHMM / LFHMM / RFHMM Viterbi realignment on sm_100a: one alignment per warp.
//
// Replaces <model>::align (ahmm.cpp:468-587, lfhmm.cpp:489-654, rfhmm.cpp:516-648) + trace_path + collect_statistics.
// Layout of the DP: the model rows are cut into bands of 32; lane k owns row 32b+k+1 of band b and sweeps the read
// columns on a skewed wavefront (at step t it fills column t-k+1).  The three neighbours of a cell are then
//   left  = the lane's own previous cell (registers),
//   up    = lane k-1's previous cell  (one __shfl_up per state),
//   diag  = what the lane received as "up" one step earlier (registers).
// Lane 0 takes its "up" from the last row of the previous band, kept per warp in a boundary row (L1/L2-resident
// global scratch, prefetched one step ahead); lane 31 writes that row.  Back-pointers are one byte per cell
// (vtn_hmm_core.cuh), four steps packed per 32-bit word, stored [band][step/4][lane] so that every store is a
// coalesced 128-byte line.  The end cell is reduced over lanes with the reference's tie rules; lane 0 walks the
// path back and gathers the statistics; the path is written out forward by the whole warp.
// Work is distributed dynamically (atomic counter) over a list sorted by DP size, largest first.
// No tensor cores: the recurrence is FADD/compare/select on a dependent chain (issue-bound).
#include <algorithm>
#include <cstring>
#include "vtn_internal.h"
#include "vtn_hmm_core.cuh"

namespace
{
constexpr int HMM_MAXLEN = 1024;       // ahmm.cpp:26
constexpr int HMM_MAX_MOTIF = 255;
constexpr int HMM_MAX_FLANK = 31;
constexpr int HMM_WARPS_PER_BLOCK = 4;
constexpr int HMM_BLOCKS_PER_SM = 4;
constexpr int HMM_MAX_BANDS = 33;                                   // ceil((1023 + 31) / 32)
constexpr int HMM_MAX_S4 = (HMM_MAXLEN + 31 + 3) / 4;               // step words per band
// per-warp global scratch
constexpr size_t HMM_BND_BYTES = (size_t)HMM_MAXLEN * (16 + 4);     // float4 {m,d,i,x} + int pmr per column
constexpr size_t HMM_BP_BYTES = (size_t)HMM_MAX_BANDS * HMM_MAX_S4 * 32 * 4;
constexpr size_t HMM_REV_BYTES = 4096 * 4;
constexpr size_t HMM_WARP_SCRATCH = HMM_BND_BYTES + HMM_BP_BYTES + HMM_REV_BYTES;

struct HmmDevBatch
{
    uint32_t n;
    const uint8_t* kind;
    const uint32_t* seq_off;
    const uint8_t* bytes;
    const uint32_t* order;
};

struct HmmWarpSmem
{
    uint8_t read[HMM_MAXLEN];
    uint8_t motif[HMM_MAX_MOTIF + 1];
    uint8_t flank[HMM_MAX_FLANK + 1];
    uint16_t disc[HMM_MAXLEN + 8];
    float4 bchunk[32]; // 32 columns of the previous band's last row {m,d,i,x}
    int bpchunk[32];   // ... and its MR positions (RFHMM)
    vtn_hmm_result res;
};

__device__ __forceinline__ HmmCell shfl_up_cell(const HmmCell& c, bool want_x, bool want_p)
{
    HmmCell o;
    o.m = __shfl_up_sync(0xffffffffu, c.m, 1);
    o.d = __shfl_up_sync(0xffffffffu, c.d, 1);
    o.i = __shfl_up_sync(0xffffffffu, c.i, 1);
    o.x = want_x ? __shfl_up_sync(0xffffffffu, c.x, 1) : HMM_NEG_INF;
    o.pmr = want_p ? __shfl_up_sync(0xffffffffu, c.pmr, 1) : 0;
    o.z = HMM_NEG_INF;
    return o;
}

// combine two end-cell candidates with the scan's tie rule (hmm_end_consider)
__device__ __forceinline__ void end_merge(HmmEnd& a, float s, int rank, bool last_wins)
{
    if (rank < 0) return;
    hmm_end_consider(a, last_wins, s, rank);
}

struct BpReader
{
    const uint32_t* bpw;
    int S4;
    __device__ __forceinline__ uint8_t operator()(int i, int j) const
    {
        const int b = (i - 1) >> 5, lane = (i - 1) & 31, t = j - 1 + lane;
        const uint32_t w = bpw[((size_t)b * S4 + (t >> 2)) * 32 + lane];
        return (uint8_t)(w >> (8 * (t & 3)));
    }
};

template <int KIND>
__device__ __forceinline__ void hmm_fill(const HmmConsts& T, const HmmDims& D, HmmWarpSmem& sm, float4* bnd,
                                         int* bndp, uint32_t* bpw, int S4, int lane, HmmEnd& end)
{
    const int rlen = D.rlen, plen = D.plen, mlen = D.mlen;
    const int lflen = KIND == VTN_HMM_LFHMM ? D.flen : 0;
    const int rflen = KIND == VTN_HMM_RFHMM ? D.flen : 0;
    const int body = rlen - rflen;
    const int nbands = (plen + 31) >> 5;
    const int nsteps4 = S4 << 2; // rlen + 31 rounded up to a multiple of 4: the extra steps fall outside every row
    constexpr bool WANT_X = KIND != VTN_HMM_AHMM;
    constexpr bool WANT_P = KIND == VTN_HMM_RFHMM;
    const bool lane0 = lane == 0;
    for (int b = 0; b < nbands; ++b)
    {
        const int i = (b << 5) + lane + 1;
        const bool row_on = i <= plen;
        const int k = i - lflen;
        int p_row = 1;
        uint8_t base = 0;
        if (row_on)
        {
            if (KIND == VTN_HMM_LFHMM && i <= lflen)
                base = sm.flank[i - 1];
            else
            {
                p_row = hmm_row_p(k, mlen);
                base = sm.motif[p_row - 1];
            }
        }
        const bool p_is_mlen = row_on && k >= 1 && p_row == mlen;
        // motif position of row i-1 == mlen  <=>  this row starts a new copy (p_row == 1), for i >= 2
        const bool dg_p_is_mlen = row_on && i >= 2 && p_row == 1;
        const unsigned ncol = row_on ? (unsigned)rlen : 0u; // columns this lane fills: 1 <= j <= ncol
        HmmCell cur = hmm_dead_cell(), dg = hmm_dead_cell();
        const bool last_band = b == nbands - 1;
        uint32_t* bpo = bpw + (size_t)b * S4 * 32 + lane;
        for (int t4 = 0; t4 < nsteps4; t4 += 4)
        {
            // every 32 steps: the next 32 columns of the previous band's last row -> shared memory (coalesced);
            // lane 0 then takes its "up" neighbour from there with a broadcast LDS, no divergent branch
            if ((t4 & 31) == 0)
            {
                __syncwarp();
                const int col = t4 + 1 + lane;
                float4 v = make_float4(HMM_NEG_INF, HMM_NEG_INF, HMM_NEG_INF, HMM_NEG_INF);
                int vp = 0;
                if (b > 0 && col <= rlen)
                {
                    v = bnd[col];
                    if (WANT_P) vp = bndp[col];
                }
                sm.bchunk[lane] = v;
                if (WANT_P) sm.bpchunk[lane] = vp;
                __syncwarp();
            }
            uint32_t w = 0;
#pragma unroll
            for (int u = 0; u < 4; ++u)
            {
                const int t = t4 + u;
                const int j = t - lane + 1;
                HmmCell up = shfl_up_cell(cur, WANT_X, WANT_P);
                {
                    const float4 nb = sm.bchunk[t & 31];
                    up.m = lane0 ? nb.x : up.m;
                    up.d = lane0 ? nb.y : up.d;
                    up.i = lane0 ? nb.z : up.i;
                    if (WANT_X) up.x = lane0 ? nb.w : up.x;
                    if (WANT_P) up.pmr = lane0 ? sm.bpchunk[t & 31] : up.pmr;
                }
                uint8_t bpb = 0;
                if ((unsigned)(j - 1) < ncol)
                {
                    const uint8_t rb = sm.read[j - 1];
                    const float emis = base != rb ? T.e_mismatch : T.e_match;
                    HmmCell nc;
                    if (KIND == VTN_HMM_AHMM)
                        nc = hmm_cell_ahmm(T, i, j, emis, dg, up, cur, bpb);
                    else if (KIND == VTN_HMM_LFHMM)
                        nc = hmm_cell_lfhmm(T, i, j, lflen, p_is_mlen, emis, dg, up, cur, bpb);
                    else
                        nc = hmm_cell_rfhmm(T, i, j, body, rflen, dg_p_is_mlen, emis, sm.flank, rb, dg, up, cur, bpb);
                    cur = nc;
                    if (j == rlen) hmm_end_row(end, KIND, i, lflen, cur);
                    if (lane == 31 && !last_band)
                    {
                        bnd[j] = make_float4(cur.m, cur.d, cur.i, cur.x);
                        if (WANT_P) bndp[j] = cur.pmr;
                    }
                }
                dg = up;
                w |= (uint32_t)bpb << (8 * u);
            }
            bpo[(size_t)(t4 >> 2) * 32] = w;
        }
        __syncwarp();
    }
}

__device__ __forceinline__ void warp_copy_bytes(uint8_t* dst, const uint8_t* src, int n, int lane)
{
    for (int k = lane; k < n; k += 32) dst[k] = src[k];
}

__global__ void __launch_bounds__(HMM_WARPS_PER_BLOCK * 32)
vtn_hmm_kernel(const HmmConsts T, const HmmDevBatch B, vtn_hmm_result* __restrict__ out, int32_t* __restrict__ path,
               uint32_t path_stride, uint8_t* __restrict__ scratch, unsigned* __restrict__ counter)
{
    __shared__ HmmWarpSmem s_all[HMM_WARPS_PER_BLOCK];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    HmmWarpSmem& sm = s_all[wib];
    const size_t gw = (size_t)blockIdx.x * HMM_WARPS_PER_BLOCK + wib;
    uint8_t* ws = scratch + gw * HMM_WARP_SCRATCH;
    float4* bnd = reinterpret_cast<float4*>(ws);
    int* bndp = reinterpret_cast<int*>(ws + (size_t)HMM_MAXLEN * 16);
    uint32_t* bpw = reinterpret_cast<uint32_t*>(ws + HMM_BND_BYTES);
    int32_t* rev = reinterpret_cast<int32_t*>(ws + HMM_BND_BYTES + HMM_BP_BYTES);

    for (;;)
    {
        unsigned idx = 0;
        if (lane == 0) idx = atomicAdd(counter, 1u);
        idx = __shfl_sync(0xffffffffu, idx, 0);
        if (idx >= B.n) break;
        const uint32_t a = B.order[idx];
        const int kind = B.kind[a];
        const uint32_t o0 = B.seq_off[3 * a], o1 = B.seq_off[3 * a + 1], o2 = B.seq_off[3 * a + 2], o3 = B.seq_off[3 * a + 3];
        int mlen = (int)(o1 - o0), flen = (int)(o2 - o1), rlen = (int)(o3 - o2);
        int status = VTN_HMM_ST_OK;
        if (kind == VTN_HMM_AHMM)
        {
            flen = 0;
            if (rlen >= HMM_MAXLEN) rlen = HMM_MAXLEN - 1; // ahmm.cpp:476-480
        }
        else if (rlen >= HMM_MAXLEN || rlen + flen >= HMM_MAXLEN)
            status = VTN_HMM_ST_TOO_LONG; // lfhmm.cpp:498-502, rfhmm.cpp:524-528 (rows >= 1024 leave the arrays)
        if (kind > VTN_HMM_RFHMM || rlen < 1 || mlen < 1 || mlen > HMM_MAX_MOTIF || flen > HMM_MAX_FLANK ||
            (kind != VTN_HMM_AHMM && flen < 1))
            status = status ? status : VTN_HMM_ST_BAD_MODEL;
        __syncwarp();
        if (lane == 0)
        {
            vtn_hmm_result r;
            r.optimal_score = HMM_NEG_INF, r.optimal_state = 0, r.optimal_probe_len = 0, r.path_len = 0;
            for (int k = 0; k < 12; ++k) r.pos[k] = -1; // clear_statistics, ahmm.cpp:735-755
            r.motif_count = 0, r.exact_motif_count = 0, r.motif_concordance = 0, r.trf_score = 0;
            r.frac_no_repeats = 0, r.last_motif_pos = -1, r.status = status, r.reserved = 0;
            sm.res = r;
        }
        if (status == VTN_HMM_ST_OK)
        {
            warp_copy_bytes(sm.motif, B.bytes + o0, mlen, lane);
            warp_copy_bytes(sm.flank, B.bytes + o1, flen, lane);
            warp_copy_bytes(sm.read, B.bytes + o2, rlen, lane);
            __syncwarp();
            HmmDims D;
            D.kind = kind, D.rlen = rlen, D.mlen = mlen, D.flen = flen, D.plen = rlen + flen;
            const int S4 = (rlen + 31 + 3) >> 2;
            HmmEnd end;
            end.score = 0, end.rank = -1;
            if (kind == VTN_HMM_AHMM)
                hmm_fill<VTN_HMM_AHMM>(T, D, sm, bnd, bndp, bpw, S4, lane, end);
            else if (kind == VTN_HMM_LFHMM)
                hmm_fill<VTN_HMM_LFHMM>(T, D, sm, bnd, bndp, bpw, S4, lane, end);
            else
                hmm_fill<VTN_HMM_RFHMM>(T, D, sm, bnd, bndp, bpw, S4, lane, end);
            // reduce the end cell over lanes with the scan's tie rule
            const bool last_wins = kind != VTN_HMM_RFHMM;
#pragma unroll
            for (int o = 16; o >= 1; o >>= 1)
            {
                const float s2 = __shfl_xor_sync(0xffffffffu, end.score, o);
                const int r2 = __shfl_xor_sync(0xffffffffu, end.rank, o);
                end_merge(end, s2, r2, last_wins);
            }
            __syncwarp();
            if (lane == 0)
            {
                if (end.rank < 0 || !(end.score > HMM_NEG_INF))
                    sm.res.status = VTN_HMM_ST_NOPATH;
                else
                {
                    sm.res.optimal_score = end.score;
                    sm.res.optimal_state = hmm_end_state(kind, end.rank);
                    sm.res.optimal_probe_len = end.rank >> 3;
                    BpReader rd;
                    rd.bpw = bpw, rd.S4 = S4;
                    sm.res.status = hmm_traceback(D, sm.read, sm.motif, sm.res.optimal_state, sm.res.optimal_probe_len, rd,
                                                  sm.disc, &sm.res, path ? rev : nullptr, 4096);
                }
            }
            __syncwarp();
            if (path && sm.res.status == VTN_HMM_ST_OK)
            {
                const int n = sm.res.path_len;
                int32_t* dst = path + (size_t)a * path_stride;
                for (int k = lane; k < n && k < (int)path_stride; k += 32) dst[k] = rev[n - 1 - k];
            }
        }
        __syncwarp();
        // 96-byte result, one 4-byte word per lane
        if (lane < (int)(sizeof(vtn_hmm_result) / 4))
            reinterpret_cast<int32_t*>(out + a)[lane] = reinterpret_cast<const int32_t*>(&sm.res)[lane];
        __syncwarp();
    }
}
} // namespace

// ------------------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------------------
struct vtn_hmm_dbatch
{
    uint32_t n = 0;
    uint32_t path_stride = 0;
    HmmConsts consts{};
    uint8_t* d_kind = nullptr;
    uint32_t* d_seq_off = nullptr;
    uint8_t* d_bytes = nullptr;
    uint32_t* d_order = nullptr;
    vtn_hmm_result* d_out = nullptr;
    int32_t* d_path = nullptr;
    uint64_t cells = 0, proc_comps = 0;
};

void vtn_hmm_ctx_release(vtn_ctx* ctx)
{
    if (!ctx) return;
    cudaFree(ctx->hmm_scratch);
    cudaFree(ctx->hmm_counter);
    ctx->hmm_scratch = nullptr;
    ctx->hmm_counter = nullptr;
    ctx->hmm_scratch_bytes = 0;
}

static int hmm_grid(const vtn_ctx* ctx) { return ctx->sm_count * HMM_BLOCKS_PER_SM; }

static int hmm_ensure_scratch(vtn_ctx* ctx)
{
    const size_t need = (size_t)hmm_grid(ctx) * HMM_WARPS_PER_BLOCK * HMM_WARP_SCRATCH;
    if (ctx->hmm_scratch_bytes >= need) return VTN_OK;
    vtn_hmm_ctx_release(ctx);
    cudaError_t e = cudaMalloc(&ctx->hmm_scratch, need);
    if (e == cudaSuccess) e = cudaMalloc((void**)&ctx->hmm_counter, 64);
    if (e != cudaSuccess)
    {
        vtn_hmm_ctx_release(ctx);
        return vtn_set_err(ctx, VTN_E_NOMEM, std::string("HMM scratch allocation: ") + cudaGetErrorString(e));
    }
    ctx->hmm_scratch_bytes = need;
    return VTN_OK;
}

extern "C" int vtn_hmm_transitions(int kind, const vtn_hmm_params* par, float* T49)
{
    if (!par || !T49 || kind < 0 || kind > VTN_HMM_RFHMM) return VTN_E_ARG;
    HmmConsts c;
    vtn_hmm_make_consts(par, &c);
    vtn_hmm_table(kind, c, T49);
    return VTN_OK;
}

extern "C" int vtn_hmm_dbatch_destroy(vtn_ctx* ctx, vtn_hmm_dbatch* db)
{
    if (!db) return VTN_OK;
    if (ctx) cudaSetDevice(ctx->device);
    cudaFree(db->d_kind);
    cudaFree(db->d_seq_off);
    cudaFree(db->d_bytes);
    cudaFree(db->d_order);
    cudaFree(db->d_out);
    cudaFree(db->d_path);
    delete db;
    return VTN_OK;
}

extern "C" int vtn_hmm_dbatch_create(vtn_ctx* ctx, const vtn_hmm_params* par, const vtn_hmm_batch* in,
                                     uint32_t path_stride, vtn_hmm_dbatch** out)
{
    if (!ctx) return VTN_E_ARG;
    if (!par || !in || !out) return vtn_set_err(ctx, VTN_E_ARG, "vtn_hmm: null argument");
    if (in->n && (!in->kind || !in->seq_off || !in->bytes)) return vtn_set_err(ctx, VTN_E_ARG, "vtn_hmm: batch has null arrays");
    if (in->n >= (1u << 30)) return vtn_set_err(ctx, VTN_E_ARG, "vtn_hmm: too many alignments in one batch");
    if (in->n && in->seq_off[3 * (size_t)in->n] != in->n_bytes)
        return vtn_set_err(ctx, VTN_E_ARG, "vtn_hmm: seq_off inconsistent with n_bytes");
    VTN_CUDA(ctx, cudaSetDevice(ctx->device));
    int rc = hmm_ensure_scratch(ctx);
    if (rc) return rc;
    vtn_hmm_dbatch* db = new vtn_hmm_dbatch();
    db->n = in->n;
    db->path_stride = path_stride;
    vtn_hmm_make_consts(par, &db->consts);
    // work list: largest DP first (counting sort on cells/1024), so the dynamic scheduler ends on small items
    const uint32_t n = in->n;
    std::vector<uint32_t> order(n), key(n);
    constexpr uint32_t NB = 1100;
    std::vector<uint32_t> cnt(NB + 1, 0);
    for (uint32_t a = 0; a < n; ++a)
    {
        uint64_t mlen = in->seq_off[3 * a + 1] - in->seq_off[3 * a], flen = in->seq_off[3 * a + 2] - in->seq_off[3 * a + 1];
        uint64_t rlen = in->seq_off[3 * a + 3] - in->seq_off[3 * a + 2];
        const int kind = in->kind[a];
        (void)mlen;
        if (kind == VTN_HMM_AHMM)
        {
            flen = 0;
            if (rlen > 1023) rlen = 1023;
        }
        uint64_t cells = 0, pc = 0;
        if (rlen < 1024 && rlen + flen < 1024)
        {
            cells = (rlen + flen) * rlen;
            // SURVEY 8d "W": predecessor evaluations of the reference per cell
            pc = kind == VTN_HMM_AHMM ? 10 * cells : kind == VTN_HMM_LFHMM ? 2 * flen * rlen + 13 * rlen * rlen : 19 * cells;
        }
        db->cells += cells;
        db->proc_comps += pc;
        uint32_t k = (uint32_t)(cells >> 10);
        if (k >= NB) k = NB - 1;
        key[a] = NB - 1 - k; // descending
        ++cnt[key[a] + 1];
    }
    for (uint32_t k = 0; k < NB; ++k) cnt[k + 1] += cnt[k];
    for (uint32_t a = 0; a < n; ++a) order[cnt[key[a]]++] = a;

    cudaError_t e = cudaSuccess;
    auto alloc = [&](void** p, size_t bytes) {
        if (e == cudaSuccess) e = cudaMalloc(p, bytes ? bytes : 16);
    };
    alloc((void**)&db->d_kind, n);
    alloc((void**)&db->d_seq_off, (3 * (size_t)n + 1) * 4);
    alloc((void**)&db->d_bytes, in->n_bytes + 16);
    alloc((void**)&db->d_order, (size_t)n * 4);
    alloc((void**)&db->d_out, (size_t)n * sizeof(vtn_hmm_result));
    if (path_stride) alloc((void**)&db->d_path, (size_t)n * path_stride * 4);
    if (e != cudaSuccess)
    {
        vtn_hmm_dbatch_destroy(ctx, db);
        return vtn_set_err(ctx, VTN_E_NOMEM, std::string("vtn_hmm device batch allocation: ") + cudaGetErrorString(e));
    }
    if (n)
    {
        cudaMemcpyAsync(db->d_kind, in->kind, n, cudaMemcpyHostToDevice, ctx->stream);
        cudaMemcpyAsync(db->d_seq_off, in->seq_off, (3 * (size_t)n + 1) * 4, cudaMemcpyHostToDevice, ctx->stream);
        cudaMemcpyAsync(db->d_bytes, in->bytes, in->n_bytes, cudaMemcpyHostToDevice, ctx->stream);
        cudaMemcpyAsync(db->d_order, order.data(), (size_t)n * 4, cudaMemcpyHostToDevice, ctx->stream);
    }
    e = cudaStreamSynchronize(ctx->stream); // order[] is a local
    if (e != cudaSuccess)
    {
        vtn_hmm_dbatch_destroy(ctx, db);
        return vtn_set_err(ctx, VTN_E_CUDA, std::string("vtn_hmm upload: ") + cudaGetErrorString(e));
    }
    *out = db;
    return VTN_OK;
}

extern "C" int vtn_hmm_align_resident(vtn_ctx* ctx, vtn_hmm_dbatch* db)
{
    if (!ctx || !db) return VTN_E_ARG;
    VTN_CUDA(ctx, cudaSetDevice(ctx->device));
    if (db->n == 0) return VTN_OK;
    VTN_CUDA(ctx, cudaMemsetAsync(ctx->hmm_counter, 0, 4, ctx->stream));
    HmmDevBatch B;
    B.n = db->n, B.kind = db->d_kind, B.seq_off = db->d_seq_off, B.bytes = db->d_bytes, B.order = db->d_order;
    int grid = hmm_grid(ctx);
    const int need = (int)((db->n + HMM_WARPS_PER_BLOCK - 1) / HMM_WARPS_PER_BLOCK);
    if (grid > need) grid = need;
    vtn_hmm_kernel<<<grid, HMM_WARPS_PER_BLOCK * 32, 0, ctx->stream>>>(db->consts, B, db->d_out, db->d_path, db->path_stride,
                                                                      (uint8_t*)ctx->hmm_scratch, ctx->hmm_counter);
    VTN_CUDA(ctx, cudaGetLastError());
    ctx->launches += 1;
    return VTN_OK;
}

extern "C" int vtn_hmm_dbatch_fetch(vtn_ctx* ctx, vtn_hmm_dbatch* db, vtn_hmm_result* out, int32_t* path)
{
    if (!ctx || !db || !out) return VTN_E_ARG;
    VTN_CUDA(ctx, cudaSetDevice(ctx->device));
    if (db->n)
    {
        VTN_CUDA(ctx, cudaMemcpyAsync(out, db->d_out, (size_t)db->n * sizeof(vtn_hmm_result), cudaMemcpyDeviceToHost, ctx->stream));
        if (path && db->d_path)
            VTN_CUDA(ctx, cudaMemcpyAsync(path, db->d_path, (size_t)db->n * db->path_stride * 4, cudaMemcpyDeviceToHost, ctx->stream));
    }
    VTN_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return VTN_OK;
}

extern "C" int vtn_hmm_dbatch_work(vtn_ctx* ctx, vtn_hmm_dbatch* db, uint64_t* cells, uint64_t* proc_comps)
{
    if (!ctx || !db) return VTN_E_ARG;
    if (cells) *cells = db->cells;
    if (proc_comps) *proc_comps = db->proc_comps;
    return VTN_OK;
}

extern "C" int vtn_hmm_align(vtn_ctx* ctx, const vtn_hmm_params* par, const vtn_hmm_batch* in, vtn_hmm_result* out,
                             int32_t* path, uint32_t path_stride)
{
    if (!ctx) return VTN_E_ARG;
    if (!out) return vtn_set_err(ctx, VTN_E_ARG, "vtn_hmm_align: null result array");
    vtn_hmm_dbatch* db = nullptr;
    int rc = vtn_hmm_dbatch_create(ctx, par, in, path ? path_stride : 0, &db);
    if (rc) return rc;
    rc = vtn_hmm_align_resident(ctx, db);
    if (rc == VTN_OK) rc = vtn_hmm_dbatch_fetch(ctx, db, out, path);
    vtn_hmm_dbatch_destroy(ctx, db);
    return rc;
}
