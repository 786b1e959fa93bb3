// vtn_hmm_core.cuh -- cell recurrences and traceback of the AHMM / LFHMM / RFHMM Viterbi aligners in the
// REDUCED form the kernels use.  Host+device inline code: the CUDA kernel (vtn_hmm.cu) and the CPU unit test of
// this header (tests/hmm_core_host.cpp, row-major loops) share it, so the recurrences are checked against the
// oracle without a GPU; the warp wavefront around them is checked on the GPU.
//
// Reduction (each point derived from the reference, file:line into atks/vt):
//  * A cell is "live" iff its V is finite.  proc_comp (ahmm.cpp:427-449) takes a predecessor only when
//    score > max_score with max_score starting at -inf, so dead predecessors never win and their tracks are
//    irrelevant; every track condition that can fail for a LIVE predecessor is kept below.
//  * The track word [u|d|c|p] (ahmm.h:57-67) of a live cell is a function of its row:  every M/D move advances
//    the motif position by one (ahmm.h:428-526), I/Z moves copy it (ahmm.h:538-556, lfhmm.h:572-636), so in
//    row i the motif position is p=((k-1)%mlen)+1, copy c=(k-1)/mlen+1 with k = i (AHMM, RFHMM) or i-lflen
//    (LFHMM); LFHMM's ML is live only on the diagonal i==j<=lflen (lfhmm.h:418-436); RFHMM's Y only in row 0
//    (rfhmm.cpp:232-249, "invariant" :553-560).  Only RFHMM's MR position depends on the path and is carried.
//    So one byte of back-pointers per cell replaces 3-5 track words.
//  * Float expression order is the reference's: ((V + T) + emission) (+ valid = 0), left to right.
#pragma once
#include <stdint.h>
#include <math.h>
#include "../../include/vtn.h"

#ifdef __CUDACC__
#define VTN_HD __host__ __device__ __forceinline__
#else
#define VTN_HD inline
#endif

#define HMM_NEG_INF (-INFINITY)

// transition + emission constants of one parameter set (all three models), computed on the host by
// vtn_hmm_make_consts with the reference's expressions
struct HmmConsts
{
    // AHMM (ahmm.cpp:172-182)
    float a_sm, a_mm, a_dm, a_im, a_md, a_dd, a_mi, a_ii;
    // LFHMM (lfhmm.cpp:179-199)
    float l_mlm, l_mm, l_dm, l_im, l_mld, l_md, l_dd, l_mi, l_ii, l_mz, l_dz, l_zz, l_mlz;
    // RFHMM (rfhmm.cpp:190-214); I->M stays -inf (:197)
    float r_sm, r_ym, r_mm, r_dm, r_yd, r_md, r_dd, r_mi, r_ii, r_smr, r_ymr, r_mmr, r_dmr, r_imr, r_mrmr;
    float e_match, e_mismatch; // ahmm.cpp:797-812 with LogTool::pl2log10_varp (log_tool.cpp:90-108)
};

// state codes of the reference (ahmm.cpp:29-36, lfhmm.cpp:28-37, rfhmm.cpp:29-37)
enum { HA_S = 0, HA_M = 1, HA_D = 2, HA_I = 3, HA_E = 5 };
enum { HL_S = 0, HL_ML = 1, HL_M = 2, HL_D = 3, HL_I = 4, HL_Z = 5, HL_E = 6 };
enum { HR_S = 0, HR_Y = 1, HR_M = 2, HR_D = 3, HR_I = 4, HR_MR = 5, HR_E = 6 };
enum { HC_LFLANK = 0, HC_MOTIF = 1, HC_RFLANK = 2, HC_UNMODELED = 3 };

VTN_HD int32_t hmm_make_track(int u, int d, int c, int p)
{
    return (int32_t)(((uint32_t)u << 27) | ((uint32_t)d << 24) | ((uint32_t)c << 12) | (uint32_t)p);
}

struct HmmDims
{
    int kind, rlen, plen, mlen, flen; // flen = lflen (LFHMM) / rflen (RFHMM) / 0
};

// motif position / copy of row i (1-based); k = i - lflen for LFHMM
VTN_HD int hmm_row_p(int k, int mlen) { return (k - 1) % mlen + 1; }
VTN_HD int hmm_row_c(int k, int mlen) { return (k - 1) / mlen + 1; }

#define HMM_TRY(score_expr, code)       \
    {                                   \
        float s__ = (score_expr);       \
        if (s__ > best)                 \
        {                               \
            best = s__;                 \
            bp = (code);                \
        }                               \
    }

// values of one cell's states; x = ML (LFHMM) or MR (RFHMM); z = Z (LFHMM); pmr = MR model position
struct HmmCell
{
    float m, d, i, x, z;
    int pmr;
};

VTN_HD HmmCell hmm_dead_cell()
{
    HmmCell c;
    c.m = c.d = c.i = c.x = c.z = HMM_NEG_INF;
    c.pmr = 0;
    return c;
}

// ---- AHMM: ahmm.cpp:490-539.  dg = (i-1,j-1), up = (i-1,j), lf = (i,j-1); emis = emission of (row base, read[j-1])
VTN_HD HmmCell hmm_cell_ahmm(const HmmConsts& T, int i, int j, float emis, const HmmCell& dg, const HmmCell& up,
                             const HmmCell& lf, uint8_t& bpbyte)
{
    HmmCell o = hmm_dead_cell();
    float best;
    int bp;
    best = HMM_NEG_INF, bp = 0;
    if (i == 1 && j == 1) HMM_TRY((0.0f + T.a_sm) + emis, HA_S) // S is live only at (0,0)
    HMM_TRY((dg.m + T.a_mm) + emis, HA_M)
    HMM_TRY((dg.d + T.a_dm) + emis, HA_D)
    HMM_TRY((dg.i + T.a_im) + emis, HA_I)
    o.m = best;
    uint8_t b = (uint8_t)bp;
    best = HMM_NEG_INF, bp = 0;
    HMM_TRY(up.m + T.a_md, 0)
    HMM_TRY(up.d + T.a_dd, 1)
    o.d = best;
    b |= (uint8_t)(bp << 2);
    best = HMM_NEG_INF, bp = 0;
    HMM_TRY(lf.m + T.a_mi, 0)
    HMM_TRY(lf.i + T.a_ii, 1)
    o.i = best;
    b |= (uint8_t)(bp << 3);
    bpbyte = b;
    return o;
}

// ---- LFHMM: lfhmm.cpp:511-597.  x = ML.  p_is_mlen: motif position of THIS row == mlen (D->Z, lfhmm.h:600-609)
VTN_HD HmmCell hmm_cell_lfhmm(const HmmConsts& T, int i, int j, int lflen, bool p_is_mlen, float emis,
                              const HmmCell& dg, const HmmCell& up, const HmmCell& lf, uint8_t& bpbyte)
{
    HmmCell o = hmm_dead_cell();
    float best;
    int bp;
    uint8_t b = 0;
    if (i <= lflen)
    {
        best = HMM_NEG_INF, bp = 0;
        if (i == 1 && j == 1) HMM_TRY((0.0f + 0.0f) + emis, 0)
        HMM_TRY((dg.x + 0.0f) + emis, 1) // ML->ML: source position i-1 < lflen holds here
        o.x = best;
    }
    else
    {
        best = HMM_NEG_INF, bp = 0;
        if (i - 1 == lflen) HMM_TRY((dg.x + T.l_mlm) + emis, 0) // ML->M needs p==lflen (lfhmm.h:453-461)
        HMM_TRY((dg.m + T.l_mm) + emis, 1)
        HMM_TRY((dg.d + T.l_dm) + emis, 2)
        HMM_TRY((dg.i + T.l_im) + emis, 3)
        o.m = best;
        b = (uint8_t)bp;
        best = HMM_NEG_INF, bp = 0;
        if (i - 1 == lflen) HMM_TRY(up.x + T.l_mld, 0) // ML->D needs p==lflen (lfhmm.h:515-523)
        HMM_TRY(up.m + T.l_md, 1)
        HMM_TRY(up.d + T.l_dd, 2)
        o.d = best;
        b |= (uint8_t)(bp << 2);
        best = HMM_NEG_INF, bp = 0; // ML->I: ML is dead in rows > lflen
        HMM_TRY(lf.m + T.l_mi, 0)
        HMM_TRY(lf.i + T.l_ii, 1)
        o.i = best;
        b |= (uint8_t)(bp << 4);
        best = HMM_NEG_INF, bp = 0;
        HMM_TRY(lf.m + T.l_mz, 0)
        if (p_is_mlen) HMM_TRY(lf.d + T.l_dz, 1)
        HMM_TRY(lf.z + T.l_zz, 2)
        o.z = best;
        b |= (uint8_t)(bp << 5);
    }
    bpbyte = b;
    return o;
}

// ---- RFHMM: rfhmm.cpp:535-611.  x = MR, pmr = its model position.  body = rlen - rflen.
// dg_p_is_mlen: motif position of row i-1 == mlen (M/D/I -> MR, rfhmm.h:617-651).  emis_m: emission against this
// row's motif base; rf = right flank bases; rb = read[j-1]
VTN_HD HmmCell hmm_cell_rfhmm(const HmmConsts& T, int i, int j, int body, int rflen, bool dg_p_is_mlen, float emis_m,
                              const uint8_t* rf, uint8_t rb, const HmmCell& dg, const HmmCell& up, const HmmCell& lf,
                              uint8_t& bpbyte)
{
    HmmCell o = hmm_dead_cell();
    float best;
    int bp;
    const int js = j - 1; // source read position of the diagonal / left moves
    best = HMM_NEG_INF, bp = 0;
    if (i == 1)
    {
        if (js == 0) HMM_TRY((0.0f + T.r_sm) + emis_m, 0)   // S->M: j==0 (rfhmm.h:424-431)
        if (js < body) HMM_TRY((0.0f + T.r_ym) + emis_m, 1) // Y->M from row 0 (rfhmm.h:432-440)
    }
    if (js < body) HMM_TRY((dg.m + T.r_mm) + emis_m, 2) // M->M: j < rlen-rflen (rfhmm.h:445-460)
    HMM_TRY((dg.d + T.r_dm) + emis_m, 3)
    // I->M: T[I][M] = -inf
    o.m = best;
    uint8_t b = (uint8_t)bp;
    best = HMM_NEG_INF, bp = 0;
    if (i == 1 && j < body) HMM_TRY(0.0f + T.r_yd, 0) // Y->D (rfhmm.h:503-510)
    HMM_TRY(up.m + T.r_md, 1)
    HMM_TRY(up.d + T.r_dd, 2)
    o.d = best;
    b |= (uint8_t)(bp << 2);
    best = HMM_NEG_INF, bp = 0; // S->I, Y->I: sources dead in rows >= 1
    HMM_TRY(lf.m + T.r_mi, 0)
    HMM_TRY(lf.i + T.r_ii, 1)
    o.i = best;
    b |= (uint8_t)(bp << 4);
    // MR: every entry lands on right-flank position 1 (rfhmm.h:593-651); MR->MR continues it (:653-661)
    best = HMM_NEG_INF, bp = 0;
    int pm = 0;
    float e1 = (rflen > 0 && rf[0] != rb) ? T.e_mismatch : T.e_match;
    if (rflen > 0)
    {
        if (i == 1 && js == 0) HMM_TRY((0.0f + T.r_smr) + e1, 0)
        if (i == 1 && js == body) HMM_TRY((0.0f + T.r_ymr) + e1, 1)
        if (dg_p_is_mlen && js == body)
        {
            HMM_TRY((dg.m + T.r_mmr) + e1, 2)
            HMM_TRY((dg.d + T.r_dmr) + e1, 3)
            HMM_TRY((dg.i + T.r_imr) + e1, 4)
        }
    }
    pm = 1;
    if (dg.pmr < rflen && dg.x > HMM_NEG_INF)
    {
        float e = (rf[dg.pmr] != rb) ? T.e_mismatch : T.e_match;
        float s = (dg.x + T.r_mrmr) + e;
        if (s > best)
        {
            best = s;
            bp = 5;
            pm = dg.pmr + 1;
        }
    }
    o.x = best;
    o.pmr = best > HMM_NEG_INF ? pm : 0;
    b |= (uint8_t)(bp << 5);
    bpbyte = b;
    return o;
}

// ---- end cell: trace_path's scan over the last column (ahmm.cpp:601-628, lfhmm.cpp:667-702, rfhmm.cpp:659-669).
// AHMM / LFHMM accept with >= (the LAST maximal candidate in scan order wins), RFHMM with > (the FIRST).
// A candidate is ranked by i*8 + position of its state in the scan, so the scan reduces to (score, rank).
struct HmmEnd
{
    float score;
    int rank; // i*8 + order; -1 = none
};

VTN_HD void hmm_end_consider(HmmEnd& e, bool last_wins, float s, int rank)
{
    if (e.rank < 0 || s > e.score || (s == e.score && (last_wins ? rank > e.rank : rank < e.rank)))
    {
        e.score = s;
        e.rank = rank;
    }
}

// candidates of row i (cell at column rlen)
VTN_HD void hmm_end_row(HmmEnd& e, int kind, int i, int lflen, const HmmCell& c)
{
    if (kind == VTN_HMM_AHMM)
    {
        hmm_end_consider(e, true, c.m, i * 8 + 0);
        hmm_end_consider(e, true, c.d, i * 8 + 1);
    }
    else if (kind == VTN_HMM_LFHMM)
    {
        if (i < lflen) return;
        hmm_end_consider(e, true, c.z, i * 8 + 0);
        hmm_end_consider(e, true, c.m, i * 8 + 1);
        hmm_end_consider(e, true, c.d, i * 8 + 2);
        hmm_end_consider(e, true, c.i, i * 8 + 3);
    }
    else
        hmm_end_consider(e, false, c.x, i * 8);
}

VTN_HD int hmm_end_state(int kind, int rank)
{
    const int o = rank & 7;
    if (kind == VTN_HMM_AHMM) return o == 0 ? HA_M : HA_D;
    if (kind == VTN_HMM_LFHMM) return o == 0 ? HL_Z : o == 1 ? HL_M : o == 2 ? HL_D : HL_I;
    return HR_MR;
}

// ---- collect_statistics (ahmm.cpp:676-730, lfhmm.cpp:750-861, rfhmm.cpp:727-840) on track FIELDS: src = (u,c,p)
// of the previous (later) path element, des = (u,c,p) of this one, j = read position before the step,
// mism = des is a motif position whose base differs from read[j-1].
enum { HP_LFS = 0, HP_LFE = 2, HP_MOS = 4, HP_MOE = 6, HP_RFS = 8, HP_RFE = 10 }; // +0 model, +1 read
VTN_HD void hmm_stats(int kind, vtn_hmm_result* r, uint16_t* disc, int src_u, int src_c, int src_p, int des_u,
                      int des_c, int des_p, int j, bool mism)
{
    int32_t* pos = r->pos;
    if (kind == VTN_HMM_AHMM)
    {
        if (src_u == HA_E && (des_u == HA_M || des_u == HA_D || des_u == HA_I))
        {
            pos[HP_RFS] = pos[HP_RFS + 1] = pos[HP_RFE] = pos[HP_RFE + 1] = -1;
            pos[HP_MOE] = des_c;
            r->motif_count = des_c;
            r->last_motif_pos = des_p;
            pos[HP_MOE + 1] = j;
            for (int k = 1; k <= des_c; ++k) disc[k] = 0;
            if (des_u == HA_D || mism) ++disc[des_c];
        }
        if (des_u == HA_M)
        {
            if (mism)
            {
                r->trf_score -= 7;
                ++disc[des_c];
            }
            else
                r->trf_score += 2;
        }
        if (des_u == HA_D || des_u == HA_I)
        {
            r->trf_score -= 7;
            ++disc[des_c];
        }
    }
    else if (kind == VTN_HMM_LFHMM)
    {
        if (src_u == HL_E)
        {
            if (des_u == HL_Z)
            {
                pos[HP_RFE] = -1;
                pos[HP_RFE + 1] = j + 1;
            }
            else if (des_u == HL_M || des_u == HL_D || des_u == HL_I)
            {
                pos[HP_RFS] = 0, pos[HP_RFS + 1] = j, pos[HP_RFE] = 0, pos[HP_RFE + 1] = j;
                pos[HP_MOE] = des_c;
                r->motif_count = des_c;
                pos[HP_MOE + 1] = j;
                r->trf_score = 0;
                for (int k = 1; k <= des_c; ++k) disc[k] = 0;
                if (des_u == HL_D || des_u == HL_I || mism) ++disc[des_c];
            }
        }
        else if (src_u == HL_Z)
        {
            if (des_u == HL_M || des_u == HL_D)
            {
                pos[HP_RFS] = src_p;
                pos[HP_RFS + 1] = j + 1;
                pos[HP_MOE] = des_c;
                r->motif_count = des_c;
                r->last_motif_pos = des_p;
                pos[HP_MOE + 1] = j;
                r->trf_score = 0;
                for (int k = 1; k <= des_c; ++k) disc[k] = 0;
                if (des_u == HL_D || mism) ++disc[des_c];
            }
        }
        else if (src_u == HL_M)
        {
            if (des_u == HL_ML)
            {
                pos[HP_MOS] = src_c;
                pos[HP_MOS + 1] = j + 1;
                pos[HP_LFE] = des_p;
                pos[HP_LFE + 1] = j;
            }
        }
        else if (src_u == HL_ML)
        {
            if (des_u == HL_S)
            {
                pos[HP_LFS] = src_p;
                pos[HP_LFS + 1] = j + 1;
            }
        }
        if (des_u == HL_M)
        {
            if (mism)
            {
                r->trf_score -= 7;
                ++disc[des_c];
            }
            else
                r->trf_score += 2;
        }
        if (des_u == HL_D || des_u == HL_I)
        {
            r->trf_score -= 7;
            ++disc[des_c];
        }
    }
    else
    {
        if (src_u == HR_E)
        {
            if (des_u == HR_MR)
            {
                pos[HP_RFE] = des_p;
                pos[HP_RFE + 1] = j;
            }
        }
        else if (src_u == HR_MR)
        {
            if (des_u == HR_M || des_u == HR_D)
            {
                pos[HP_RFS] = src_p;
                pos[HP_RFS + 1] = j + 1;
                pos[HP_MOE] = des_c;
                r->motif_count = des_c;
                pos[HP_MOE + 1] = j;
            }
            else if (des_u == HR_Y)
            {
                pos[HP_RFS] = src_p;
                pos[HP_RFS + 1] = j + 1;
                pos[HP_LFE] = des_p;
                pos[HP_LFE + 1] = j;
            }
        }
        else if (src_u == HR_M)
        {
            if (des_u == HR_Y)
            {
                pos[HP_MOS] = src_c;
                pos[HP_MOS + 1] = j + 1;
                pos[HP_LFE] = des_p;
                pos[HP_LFE + 1] = j;
            }
        }
        else if (src_u == HR_Y)
        {
            if (des_u == HR_S)
            {
                pos[HP_LFS] = des_p;
                pos[HP_LFS + 1] = j + 1;
            }
        }
        // RFHMM's per-copy discordance counts (rfhmm.cpp:823-835) feed nothing that align() reports
    }
}

// ---- traceback + statistics: ahmm.cpp:630-730, lfhmm.cpp:704-861, rfhmm.cpp:671-840, then the concordance
// pass ahmm.cpp:569-583 / lfhmm.cpp:637-653.  BP: uint8_t operator()(int i, int j) = back-pointer byte of a cell.
// disc: >= plen+2 counters (motif_discordance); rev: reversed path (capacity rev_cap) or null.
// Returns 0, or VTN_HMM_ST_NOPATH when the walk leaves the matrix (cannot happen from a live end cell).
template <class BP>
VTN_HD int hmm_traceback(const HmmDims& D, const uint8_t* read, const uint8_t* motif, int end_state, int end_i, BP bpf,
                         uint16_t* disc, vtn_hmm_result* r, int32_t* rev, int rev_cap)
{
    const int kind = D.kind, mlen = D.mlen, lflen = kind == VTN_HMM_LFHMM ? D.flen : 0;
    const int S = 0;
    const int E = kind == VTN_HMM_AHMM ? (int)HA_E : (int)HL_E;
    int i = end_i, j = D.rlen, n = 0;
    int state = end_state;
    int src_u = E, src_c = 0, src_p = kind == VTN_HMM_RFHMM ? D.flen + 1 : 0;
    int des_d = 0, des_c = 0, des_p = 0, des_u = 0;
    if (kind == VTN_HMM_AHMM) r->trf_score = 0; // ahmm.cpp:600
    // RFHMM: length of the MR run that ends the path = right-flank position of the end cell
    int mr_left = 0;
    if (kind == VTN_HMM_RFHMM)
    {
        int ii = i, jj = j;
        mr_left = 1;
        while (ii >= 1 && jj >= 1 && ((bpf(ii, jj) >> 5) & 7) == 5)
        {
            ++mr_left;
            --ii;
            --jj;
        }
    }
    int prev;
    do
    {
        if (i < 0 || j < 0 || (i == 0 && !(kind == VTN_HMM_RFHMM && state == HR_Y)) || n > 4095) return VTN_HMM_ST_NOPATH;
        const int u = state;
        const uint8_t b = (i >= 1 && j >= 1) ? bpf(i, j) : 0;
        // track of this cell and its predecessor state
        if (kind == VTN_HMM_AHMM)
        {
            des_d = HC_MOTIF, des_c = hmm_row_c(i, mlen), des_p = hmm_row_p(i, mlen);
            prev = u == HA_M ? (b & 3) : u == HA_D ? (((b >> 2) & 1) ? HA_D : HA_M) : (((b >> 3) & 1) ? HA_I : HA_M);
        }
        else if (kind == VTN_HMM_LFHMM)
        {
            if (u == HL_ML)
            {
                des_d = HC_LFLANK, des_c = 0, des_p = i;
                prev = i == 1 ? HL_S : HL_ML;
            }
            else
            {
                des_d = HC_MOTIF, des_c = hmm_row_c(i - lflen, mlen), des_p = hmm_row_p(i - lflen, mlen);
                if (u == HL_M)
                    prev = HL_ML + (b & 3);
                else if (u == HL_D)
                    prev = HL_ML + ((b >> 2) & 3);
                else if (u == HL_I)
                    prev = ((b >> 4) & 1) ? HL_I : HL_M;
                else
                {
                    const int z = (b >> 5) & 3;
                    prev = z == 0 ? HL_M : z == 1 ? HL_D : HL_Z;
                }
            }
        }
        else
        {
            if (u == HR_Y)
            {
                des_d = HC_LFLANK, des_c = 0, des_p = j;
                prev = j == 1 ? HR_S : HR_Y;
            }
            else if (u == HR_MR)
            {
                prev = (b >> 5) & 7; // codes 0..5 = S,Y,M,D,I,MR
                des_d = HC_RFLANK, des_c = (prev <= HR_M) ? 1 : 0, des_p = mr_left--;
            }
            else
            {
                des_d = HC_MOTIF, des_c = hmm_row_c(i, mlen), des_p = hmm_row_p(i, mlen);
                if (u == HR_M)
                    prev = b & 3; // 0..3 = S,Y,M,D
                else if (u == HR_D)
                    prev = HR_Y + ((b >> 2) & 3);
                else
                    prev = ((b >> 4) & 1) ? HR_I : HR_M;
            }
        }
        des_u = u;
        if (rev && n < rev_cap) rev[n] = hmm_make_track(des_u, des_d, des_c, des_p);
        const bool mism = des_d == HC_MOTIF && des_p >= 1 && motif[des_p - 1] != read[j - 1];
        hmm_stats(kind, r, disc, src_u, src_c, src_p, des_u, des_c, des_p, j, mism);
        src_u = des_u, src_c = des_c, src_p = des_p;
        // step
        if (kind == VTN_HMM_AHMM)
        {
            if (u == HA_M) { --i; --j; }
            else if (u == HA_D) --i;
            else --j;
        }
        else if (kind == VTN_HMM_LFHMM)
        {
            if (u == HL_ML || u == HL_M) { --i; --j; }
            else if (u == HL_D) --i;
            else --j;
        }
        else
        {
            if (u == HR_M || u == HR_MR) { --i; --j; }
            else if (u == HR_D) --i;
            else --j;
        }
        ++n;
        state = prev;
    } while (prev != S);
    // closing call collect_statistics(src_t, last_t, j): last_t is the final cell's track with u = S
    des_u = S;
    hmm_stats(kind, r, disc, src_u, src_c, src_p, des_u, des_c, des_p, j, false);
    r->path_len = n;
    if (kind != VTN_HMM_RFHMM)
    {
        r->frac_no_repeats = r->motif_count - (mlen - r->last_motif_pos) / ((float)mlen);
        int exact = r->motif_count;
        float conc = 0;
        for (int k = 1; k <= r->motif_count; ++k)
        {
            const int dk = disc[k];
            if (dk) --exact;
            if (mlen >= dk) conc += (float)(mlen - dk) / mlen;
        }
        r->exact_motif_count = exact;
        if (kind == VTN_HMM_AHMM)
            conc /= r->motif_count;
        else
            conc = (float)(conc * (1.0 / r->motif_count));
        r->motif_concordance = conc;
    }
    return 0;
}

// ---- host: constants from the parameters, with the reference's expressions.  The sub-expressions are float
// (float members, int literals), log10 is the double one and the result is stored to float -- verified
// bit-identical to the tables the reference classes build (tests/test_hmm_oracle.py, tests/golden/hmm_answers.txt).
inline float hmm_l10(float x) { return (float)log10((double)x); }
inline void vtn_hmm_make_consts(const vtn_hmm_params* par, HmmConsts* c)
{
    const float delta = par->delta, epsilon = par->epsilon, tau = par->tau, eta = par->eta;
    c->a_sm = hmm_l10(((1 - delta)) / ((1 - eta) * (1 - eta)));
    c->a_mm = hmm_l10(((1 - 2 * delta - tau)) / ((1 - eta) * (1 - eta)));
    c->a_dm = hmm_l10(((1 - epsilon - tau)) / ((1 - eta) * (1 - eta)));
    c->a_im = c->a_dm;
    c->a_md = hmm_l10(delta / (1 - eta));
    c->a_dd = hmm_l10(epsilon / (1 - eta));
    c->a_mi = c->a_md;
    c->a_ii = c->a_dd;
    c->l_mlm = hmm_l10(((1 - 2 * delta - tau)) / (eta * (1 - eta) * (1 - eta)));
    c->l_mm = c->a_mm;
    c->l_dm = c->a_dm;
    c->l_im = c->l_dm;
    c->l_mld = hmm_l10((tau * delta) / (eta * (1 - eta)));
    c->l_md = c->a_md;
    c->l_dd = c->a_dd;
    c->l_mi = c->l_md;
    c->l_ii = c->l_dd;
    c->l_mz = hmm_l10((tau * (1 - eta)) / (eta * (1 - eta)));
    c->l_dz = c->l_mz;
    c->l_zz = hmm_l10((1 - eta) / (1 - eta));
    c->l_mlz = hmm_l10((tau * tau * (1 - eta)) / (eta * eta * eta * (1 - eta)));
    c->r_sm = hmm_l10((tau * (1 - 2 * delta - tau)) / (tau * eta * (1 - eta) * (1 - eta)));
    c->r_ym = c->r_sm;
    c->r_mm = c->a_mm;
    c->r_dm = c->a_dm;
    c->r_yd = hmm_l10((tau * delta) / (eta * (1 - eta)));
    c->r_md = c->a_md;
    c->r_dd = c->a_dd;
    c->r_mi = hmm_l10(delta / (1 - eta));
    c->r_ii = hmm_l10(epsilon / (1 - eta));
    c->r_smr = hmm_l10((tau * tau * (1 - tau)) / (tau * eta * eta * eta * (1 - eta) * (1 - eta)));
    c->r_ymr = hmm_l10((eta * tau) / (eta * eta * eta));
    c->r_mmr = hmm_l10((tau * (1 - tau)) / (eta * eta * (1 - eta) * (1 - eta)));
    c->r_dmr = c->r_mmr;
    c->r_imr = c->r_dmr;
    c->r_mrmr = c->a_mm;
    // LogTool::pl2log10_varp (log_tool.cpp:90-108) and log10_emission_odds (ahmm.cpp:797-812)
    uint32_t pl = (uint32_t)par->pl;
    if (pl > 3236) pl = 3236;
    const double e = pow(10, -((double)pl) / 10.0);
    const double v = log10(e * (1 - e)) / 2;
    c->e_mismatch = (float)(v - par->mismatch_penalty);
    c->e_match = (float)(-(v));
}

// the 7x7 table in the reference's state numbering (for vtn_hmm_transitions)
inline void vtn_hmm_table(int kind, const HmmConsts& c, float* T)
{
    for (int k = 0; k < 49; ++k) T[k] = -INFINITY;
#define TT(a, b) T[(a)*7 + (b)]
    if (kind == VTN_HMM_AHMM)
    {
        TT(HA_S, HA_M) = c.a_sm, TT(HA_M, HA_M) = c.a_mm, TT(HA_D, HA_M) = c.a_dm, TT(HA_I, HA_M) = c.a_im;
        TT(HA_S, HA_D) = c.a_md, TT(HA_M, HA_D) = c.a_md, TT(HA_D, HA_D) = c.a_dd;
        TT(HA_M, HA_I) = c.a_mi, TT(HA_I, HA_I) = c.a_ii;
    }
    else if (kind == VTN_HMM_LFHMM)
    {
        TT(HL_S, HL_ML) = 0, TT(HL_ML, HL_ML) = 0;
        TT(HL_ML, HL_M) = c.l_mlm, TT(HL_M, HL_M) = c.l_mm, TT(HL_D, HL_M) = c.l_dm, TT(HL_I, HL_M) = c.l_im;
        TT(HL_ML, HL_D) = c.l_mld, TT(HL_M, HL_D) = c.l_md, TT(HL_D, HL_D) = c.l_dd;
        TT(HL_ML, HL_I) = c.l_mld, TT(HL_M, HL_I) = c.l_mi, TT(HL_I, HL_I) = c.l_ii;
        TT(HL_ML, HL_Z) = c.l_mlz;
        TT(HL_M, HL_Z) = c.l_mz, TT(HL_D, HL_Z) = c.l_dz, TT(HL_I, HL_Z) = c.l_mz, TT(HL_Z, HL_Z) = c.l_zz;
    }
    else
    {
        TT(HR_S, HR_Y) = 0, TT(HR_Y, HR_Y) = 0;
        TT(HR_S, HR_M) = c.r_sm, TT(HR_Y, HR_M) = c.r_ym, TT(HR_M, HR_M) = c.r_mm, TT(HR_D, HR_M) = c.r_dm;
        TT(HR_S, HR_D) = c.r_yd, TT(HR_Y, HR_D) = c.r_yd, TT(HR_M, HR_D) = c.r_md, TT(HR_D, HR_D) = c.r_dd;
        TT(HR_S, HR_I) = c.r_yd, TT(HR_Y, HR_I) = c.r_yd, TT(HR_M, HR_I) = c.r_mi, TT(HR_I, HR_I) = c.r_ii;
        TT(HR_S, HR_MR) = c.r_smr, TT(HR_Y, HR_MR) = c.r_ymr, TT(HR_M, HR_MR) = c.r_mmr, TT(HR_D, HR_MR) = c.r_dmr;
        TT(HR_I, HR_MR) = c.r_imr, TT(HR_MR, HR_MR) = c.r_mrmr;
    }
#undef TT
}
