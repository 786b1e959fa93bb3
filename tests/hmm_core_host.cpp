// tests/hmm_core_host.cpp -- TEST INFRASTRUCTURE ONLY (built by tests/test_hmm_core_cpu.py with g++).
// Drives the host+device header vt_b200/csrc/vtn_hmm_core.cuh (the recurrences and the traceback the CUDA
// kernel uses) with plain row-major loops, so that the REDUCED form -- 1-byte back-pointers, tracks rebuilt
// from positions -- is checked against the oracle / the reference's known answers without a GPU.
// Not part of the product: libvtn.so neither contains nor calls this.
#include <vector>
#include <cstring>
#include "../vt_b200/csrc/vtn_hmm_core.cuh"

extern "C" int hmmcore_transitions(int kind, const vtn_hmm_params* par, float* T)
{
    HmmConsts c;
    vtn_hmm_make_consts(par, &c);
    vtn_hmm_table(kind, c, T);
    return 0;
}

extern "C" int hmmcore_align(int kind, const vtn_hmm_params* par, const uint8_t* motif, int mlen, const uint8_t* flank,
                             int flen, const uint8_t* read, int rlen, vtn_hmm_result* r, int32_t* path, int path_cap)
{
    HmmConsts T;
    vtn_hmm_make_consts(par, &T);
    memset(r, 0, sizeof(*r));
    for (int k = 0; k < 12; ++k) r->pos[k] = -1;
    r->last_motif_pos = -1;
    if (kind == VTN_HMM_AHMM && rlen >= 1024) rlen = 1023;
    if (kind != VTN_HMM_AHMM && (rlen >= 1024 || rlen + flen >= 1024)) { r->status = VTN_HMM_ST_TOO_LONG; return 0; }
    if (rlen < 1 || mlen < 1 || (kind != VTN_HMM_AHMM && flen < 1)) { r->status = VTN_HMM_ST_BAD_MODEL; return 0; }
    HmmDims D;
    D.kind = kind, D.rlen = rlen, D.mlen = mlen, D.flen = kind == VTN_HMM_AHMM ? 0 : flen;
    D.plen = kind == VTN_HMM_AHMM ? rlen : rlen + flen;
    const int lflen = kind == VTN_HMM_LFHMM ? flen : 0, W = rlen + 1;
    std::vector<HmmCell> prev(W, hmm_dead_cell()), cur(W, hmm_dead_cell());
    std::vector<uint8_t> bp((size_t)(D.plen + 1) * W, 0);
    HmmEnd end;
    end.score = 0, end.rank = -1;
    if (kind != VTN_HMM_LFHMM || 0 >= lflen) hmm_end_row(end, kind, 0, lflen, prev[rlen]);
    for (int i = 1; i <= D.plen; ++i)
    {
        cur[0] = hmm_dead_cell();
        const int k = i - lflen;
        const uint8_t base = (kind == VTN_HMM_LFHMM && i <= lflen) ? flank[i - 1] : motif[hmm_row_p(k, mlen) - 1];
        for (int j = 1; j <= rlen; ++j)
        {
            const float emis = base != read[j - 1] ? T.e_mismatch : T.e_match;
            uint8_t b;
            if (kind == VTN_HMM_AHMM)
                cur[j] = hmm_cell_ahmm(T, i, j, emis, prev[j - 1], prev[j], cur[j - 1], b);
            else if (kind == VTN_HMM_LFHMM)
                cur[j] = hmm_cell_lfhmm(T, i, j, lflen, i > lflen && hmm_row_p(k, mlen) == mlen, emis, prev[j - 1],
                                        prev[j], cur[j - 1], b);
            else
                cur[j] = hmm_cell_rfhmm(T, i, j, rlen - flen, flen, i >= 2 && hmm_row_p(i - 1, mlen) == mlen, emis,
                                        flank, read[j - 1], prev[j - 1], prev[j], cur[j - 1], b);
            bp[(size_t)i * W + j] = b;
        }
        hmm_end_row(end, kind, i, lflen, cur[rlen]);
        prev.swap(cur);
    }
    if (end.rank < 0 || !(end.score > HMM_NEG_INF)) { r->status = VTN_HMM_ST_NOPATH; return 0; }
    r->optimal_score = end.score;
    r->optimal_state = hmm_end_state(kind, end.rank);
    r->optimal_probe_len = end.rank >> 3;
    std::vector<uint16_t> disc(1100, 0);
    std::vector<int32_t> rev(4096);
    const uint8_t* bpp = bp.data();
    int rc = hmm_traceback(D, read, motif, r->optimal_state, r->optimal_probe_len,
                           [bpp, W](int i, int j) { return bpp[(size_t)i * W + j]; }, disc.data(), r, rev.data(), 4096);
    r->status = rc;
    if (rc == 0 && path)
        for (int k = 0; k < r->path_len && k < path_cap; ++k) path[k] = rev[r->path_len - 1 - k];
    return 0;
}
