/* oracle/hmm_oracle.c -- TEST INFRASTRUCTURE ONLY.  See hmm_oracle.h.
 * Loop-for-loop CPU restatement of the reference's three Viterbi aligners: full V/U matrices, 32-bit track
 * words, one proc_comp per predecessor, the same float expression order.  Deliberately NOT the
 * reduced form the CUDA kernels use (2-bit back-pointers, position-derived tracks).
 *   AHMM  : ahmm.cpp:157-183 (T), :188-235 (U/V), :427-463 (proc_comp), :468-587 (align), :592-671 (trace),
 *           :676-730 (statistics), moves ahmm.h:418-556
 *   LFHMM : lfhmm.cpp:164-200, :205-276, :448-484, :489-654, :659-745, :750-861, moves lfhmm.h:413-638
 *   RFHMM : rfhmm.cpp:175-215, :220-302, :475-503, :516-648, :650-725, :727-840, moves rfhmm.h:407-668
 * U/V are allocated per call at (plen+1)x(rlen+1) instead of the reference's persistent 1024x1024: cells
 * outside the current problem are never read and row/column 0 keep their constructor values. */
#include "hmm_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define LFLANK 0
#define MOTIF 1
#define RFLANK 2
#define UNMODELED 3
#define UNCERTAIN 4
#define MT_MODEL 0
#define MT_READ 1
#define MT_MATCH 2

#define track_get_u(t) ((((uint32_t)(t)) & 0xF8000000u) >> 27)
#define track_get_d(t) ((((uint32_t)(t)) & 0x07000000u) >> 24)
#define track_get_c(t) ((((uint32_t)(t)) & 0x00FFF000u) >> 12)
#define track_get_p(t) ((((uint32_t)(t)) & 0x00000FFFu))
#define track_set_u(t, u) ((int32_t)((((uint32_t)(t)) & 0x07FFFFFFu) | (((uint32_t)(u)) << 27)))
#define make_track(u, d, c, p) ((int32_t)((((uint32_t)(u)) << 27) | (((uint32_t)(d)) << 24) | (((uint32_t)(c)) << 12) | ((uint32_t)(p))))
#define NULL_TRACK 0x7C000000
#define START_TRACK 0x78000000

/* state codes per model: ahmm.cpp:29-36, lfhmm.cpp:28-37, rfhmm.cpp:29-37 */
enum { A_S = 0, A_M = 1, A_D = 2, A_I = 3, A_E = 5, A_N = 6, A_TBD = 7 };
enum { L_S = 0, L_ML = 1, L_M = 2, L_D = 3, L_I = 4, L_Z = 5, L_E = 6, L_N = 7, L_TBD = 8 };
enum { R_S = 0, R_Y = 1, R_M = 2, R_D = 3, R_I = 4, R_MR = 5, R_E = 6, R_N = 7, R_TBD = 8 };

typedef struct
{
    int kind, rlen, plen, lflen, mlen, rflen, W, nst;
    const char* read;
    const char* model[3];
    float T[7][7];
    float e_match, e_mismatch;
    float* V[6];
    int32_t* U[6];
    float max_score;
    int32_t max_track;
    int32_t* disc; /* motif_discordance */
    hmmo_result* res;
} hctx;

#define IDX(h, i, j) ((size_t)(i) * (h)->W + (j))
#define track_get_base(h, t) ((h)->model[track_get_d(t)][(int)track_get_p(t) - 1])

/* LogTool::pl2log10_varp, log_tool.cpp:90-108 */
static double pl2log10_varp(uint32_t pl)
{
    if (pl > 3236) pl = 3236;
    double e = pow(10, -((double)pl) / 10.0);
    return log10(e * (1 - e)) / 2;
}

void hmmo_transitions(int kind, const hmmo_params* par, float* Tout)
{
    float delta = par->delta, epsilon = par->epsilon, tau = par->tau, eta = par->eta;
    float T[7][7];
    for (int i = 0; i < 7; ++i)
        for (int j = 0; j < 7; ++j) T[i][j] = -INFINITY;
    if (kind == HMMO_AHMM)
    { /* ahmm.cpp:172-182 */
        T[A_S][A_M] = log10(((1 - delta)) / ((1 - eta) * (1 - eta)));
        T[A_M][A_M] = log10(((1 - 2 * delta - tau)) / ((1 - eta) * (1 - eta)));
        T[A_D][A_M] = log10(((1 - epsilon - tau)) / ((1 - eta) * (1 - eta)));
        T[A_I][A_M] = T[A_D][A_M];
        T[A_S][A_D] = log10(delta / (1 - eta));
        T[A_M][A_D] = log10(delta / (1 - eta));
        T[A_D][A_D] = log10(epsilon / (1 - eta));
        T[A_M][A_I] = T[A_M][A_D];
        T[A_I][A_I] = T[A_D][A_D];
    }
    else if (kind == HMMO_LFHMM)
    { /* lfhmm.cpp:179-199 */
        T[L_S][L_ML] = 0;
        T[L_ML][L_ML] = 0;
        T[L_ML][L_M] = log10(((1 - 2 * delta - tau)) / (eta * (1 - eta) * (1 - eta)));
        T[L_M][L_M] = log10(((1 - 2 * delta - tau)) / ((1 - eta) * (1 - eta)));
        T[L_D][L_M] = log10(((1 - epsilon - tau)) / ((1 - eta) * (1 - eta)));
        T[L_I][L_M] = T[L_D][L_M];
        T[L_ML][L_D] = log10((tau * delta) / (eta * (1 - eta)));
        T[L_M][L_D] = log10(delta / (1 - eta));
        T[L_D][L_D] = log10(epsilon / (1 - eta));
        T[L_ML][L_I] = T[L_ML][L_D];
        T[L_M][L_I] = T[L_M][L_D];
        T[L_I][L_I] = T[L_D][L_D];
        T[L_ML][L_Z] = log10((tau * tau * (1 - eta)) / (eta * eta * eta * (1 - eta)));
        T[L_M][L_Z] = log10((tau * (1 - eta)) / (eta * (1 - eta)));
        T[L_D][L_Z] = log10((tau * (1 - eta)) / (eta * (1 - eta)));
        T[L_I][L_Z] = log10((tau * (1 - eta)) / (eta * (1 - eta)));
        T[L_Z][L_Z] = log10((1 - eta) / (1 - eta));
    }
    else
    { /* rfhmm.cpp:190-214; T[I][M] is assigned to itself at :197 and so stays -inf */
        T[R_S][R_Y] = 0;
        T[R_Y][R_Y] = 0;
        T[R_S][R_M] = log10((tau * (1 - 2 * delta - tau)) / (tau * eta * (1 - eta) * (1 - eta)));
        T[R_Y][R_M] = T[R_S][R_M];
        T[R_M][R_M] = log10(((1 - 2 * delta - tau)) / ((1 - eta) * (1 - eta)));
        T[R_D][R_M] = log10(((1 - epsilon - tau)) / ((1 - eta) * (1 - eta)));
        T[R_I][R_M] = T[R_I][R_M];
        T[R_S][R_D] = log10((tau * delta) / (eta * (1 - eta)));
        T[R_Y][R_D] = T[R_S][R_D];
        T[R_M][R_D] = log10(delta / (1 - eta));
        T[R_D][R_D] = log10(epsilon / (1 - eta));
        T[R_S][R_I] = log10((tau * delta) / (eta * (1 - eta)));
        T[R_Y][R_I] = T[R_S][R_I];
        T[R_M][R_I] = log10(delta / (1 - eta));
        T[R_I][R_I] = log10(epsilon / (1 - eta));
        T[R_S][R_MR] = log10((tau * tau * (1 - tau)) / (tau * eta * eta * eta * (1 - eta) * (1 - eta)));
        T[R_Y][R_MR] = log10((eta * tau) / (eta * eta * eta));
        T[R_M][R_MR] = log10((tau * (1 - tau)) / (eta * eta * (1 - eta) * (1 - eta)));
        T[R_D][R_MR] = T[R_M][R_MR];
        T[R_I][R_MR] = T[R_D][R_MR];
        T[R_MR][R_MR] = log10(((1 - 2 * delta - tau)) / ((1 - eta) * (1 - eta)));
    }
    memcpy(Tout, T, sizeof(T));
}

/* p==mlen wraps to the next motif copy (ahmm.h:433-440 and the identical bodies in lfhmm.h / rfhmm.h) */
static int32_t advance(const hctx* h, int32_t t, int u)
{
    int32_t p = track_get_p(t);
    if (p == h->mlen) return make_track(u, MOTIF, track_get_c(t) + 1, 1);
    return make_track(u, MOTIF, track_get_c(t), p + 1);
}

/* ahmm.h:418-556 */
static int32_t move_ahmm(const hctx* h, int A, int B, int32_t t, int32_t j)
{
    int32_t p = track_get_p(t);
    switch (A * 8 + B)
    {
    case A_S * 8 + A_M:
    case A_S * 8 + A_D:
    case A_S * 8 + A_I:
        return t == START_TRACK ? make_track(A_S, MOTIF, 1, 1) : NULL_TRACK;
    case A_M * 8 + A_M:
    case A_D * 8 + A_M:
        return (p <= h->mlen && j < h->rlen) ? advance(h, t, A) : NULL_TRACK;
    case A_I * 8 + A_M:
    case A_M * 8 + A_D:
    case A_D * 8 + A_D:
        return p <= h->mlen ? advance(h, t, A) : NULL_TRACK;
    case A_M * 8 + A_I:
    case A_I * 8 + A_I:
        return (t != NULL_TRACK && j < h->rlen) ? track_set_u(t, A) : NULL_TRACK;
    }
    return NULL_TRACK;
}

/* lfhmm.h:413-638 */
static int32_t move_lfhmm(const hctx* h, int A, int B, int32_t t, int32_t j)
{
    int32_t p = track_get_p(t);
    switch (A * 8 + B)
    {
    case L_S * 8 + L_ML:
        return t == START_TRACK ? make_track(L_S, LFLANK, 0, 1) : NULL_TRACK;
    case L_ML * 8 + L_ML:
        return (p < h->lflen && j < h->rlen) ? make_track(L_ML, LFLANK, 0, p + 1) : NULL_TRACK;
    case L_ML * 8 + L_M:
        return (p == h->lflen && j < h->rlen) ? make_track(L_ML, MOTIF, 1, 1) : NULL_TRACK;
    case L_M * 8 + L_M:
    case L_D * 8 + L_M:
        return (p <= h->mlen && j < h->rlen) ? advance(h, t, A) : NULL_TRACK;
    case L_I * 8 + L_M:
    case L_M * 8 + L_D:
    case L_D * 8 + L_D:
        return p <= h->mlen ? advance(h, t, A) : NULL_TRACK;
    case L_ML * 8 + L_D:
        return p == h->lflen ? make_track(L_ML, MOTIF, 1, 1) : NULL_TRACK;
    case L_ML * 8 + L_I:
        return j < h->rlen ? make_track(L_ML, MOTIF, 1, 0) : NULL_TRACK;
    case L_M * 8 + L_I:
    case L_I * 8 + L_I:
        return (t != NULL_TRACK && j < h->rlen) ? track_set_u(t, A) : NULL_TRACK;
    case L_M * 8 + L_Z:
        return (track_get_d(t) == MOTIF && j < h->rlen) ? track_set_u(t, L_M) : NULL_TRACK;
    case L_D * 8 + L_Z:
        return (track_get_d(t) == MOTIF && p == h->mlen && j < h->rlen) ? track_set_u(t, L_D) : NULL_TRACK;
    case L_Z * 8 + L_Z:
        return j < h->rlen ? track_set_u(t, L_Z) : NULL_TRACK;
    }
    return NULL_TRACK;
}

/* rfhmm.h:407-668 */
static int32_t move_rfhmm(const hctx* h, int A, int B, int32_t t, int32_t j)
{
    int32_t p = track_get_p(t);
    int32_t body = h->rlen - h->rflen;
    switch (A * 8 + B)
    {
    case R_S * 8 + R_M:
        return (t == START_TRACK && j == 0) ? make_track(R_S, MOTIF, 1, 1) : NULL_TRACK;
    case R_Y * 8 + R_M:
    case R_Y * 8 + R_D:
        return j < body ? make_track(R_Y, MOTIF, 1, 1) : NULL_TRACK;
    case R_M * 8 + R_M:
        return (p <= h->mlen && j < body) ? advance(h, t, R_M) : NULL_TRACK;
    case R_D * 8 + R_M:
        return (p <= h->mlen && j < h->rlen) ? advance(h, t, R_D) : NULL_TRACK;
    case R_I * 8 + R_M:
    case R_M * 8 + R_D:
    case R_D * 8 + R_D:
        return p <= h->mlen ? advance(h, t, A) : NULL_TRACK;
    case R_S * 8 + R_D:
        return t == START_TRACK ? make_track(R_S, MOTIF, 1, 1) : NULL_TRACK;
    case R_S * 8 + R_I:
        return t == START_TRACK ? make_track(R_S, MOTIF, 1, 0) : NULL_TRACK;
    case R_Y * 8 + R_I:
        return j < h->rlen ? make_track(R_Y, MOTIF, 1, 0) : NULL_TRACK;
    case R_M * 8 + R_I:
    case R_I * 8 + R_I:
        return (t != NULL_TRACK && j < h->rlen) ? track_set_u(t, A) : NULL_TRACK;
    case R_S * 8 + R_MR:
        return t == START_TRACK ? make_track(R_S, RFLANK, 1, 1) : NULL_TRACK;
    case R_Y * 8 + R_MR:
        return j == body ? make_track(R_Y, RFLANK, 1, 1) : NULL_TRACK;
    case R_M * 8 + R_MR:
        return (p == h->mlen && j == body) ? make_track(R_M, RFLANK, 1, 1) : NULL_TRACK;
    case R_D * 8 + R_MR:
        return (p == h->mlen && j == body) ? make_track(R_D, RFLANK, 0, 1) : NULL_TRACK;
    case R_I * 8 + R_MR:
        return (p == h->mlen && j == body) ? make_track(R_I, RFLANK, 0, 1) : NULL_TRACK;
    case R_MR * 8 + R_MR:
        return (p < h->rflen && j < h->rlen) ? make_track(R_MR, RFLANK, 0, p + 1) : NULL_TRACK;
    }
    return NULL_TRACK;
}

/* ahmm.cpp:427-449 (same body in lfhmm.cpp:448-470, rfhmm.cpp:475-497); emission ahmm.cpp:797-812 */
static void proc_comp(hctx* h, int A, int B, size_t index1, int32_t j, int match_type)
{
    float emission = 0, score = 0, valid = 0;
    int32_t t0 = h->U[A][index1];
    int32_t t = h->kind == HMMO_AHMM ? move_ahmm(h, A, B, t0, j)
                : h->kind == HMMO_LFHMM ? move_lfhmm(h, A, B, t0, j) : move_rfhmm(h, A, B, t0, j);
    if (t == NULL_TRACK)
        valid = -INFINITY;
    else if (match_type == MT_MATCH)
        emission = track_get_base(h, t) != h->read[j] ? h->e_mismatch : h->e_match;
    score = h->V[A][index1] + h->T[A][B] + emission + valid;
    if (score > h->max_score)
    {
        h->max_score = score;
        h->max_track = t;
    }
}

#define BEGIN_STATE(h) ((h)->max_score = -INFINITY, (h)->max_track = NULL_TRACK)
#define END_STATE(h, st, c) ((h)->V[st][c] = (h)->max_score, (h)->U[st][c] = (h)->max_track)

/* initialize_UV restricted to the cells align() can read: ahmm.cpp:188-235, lfhmm.cpp:205-276, rfhmm.cpp:220-302 */
static int init_uv(hctx* h)
{
    size_t n = (size_t)(h->plen + 1) * h->W;
    int tbd = h->kind == HMMO_AHMM ? A_TBD : L_TBD, nn = h->kind == HMMO_AHMM ? A_N : L_N;
    for (int s = 0; s < 6; ++s)
    {
        h->V[s] = (float*)malloc(n * sizeof(float));
        h->U[s] = (int32_t*)malloc(n * sizeof(int32_t));
        if (!h->V[s] || !h->U[s]) return -1;
    }
    for (int i = 0; i <= h->plen; ++i)
        for (int j = 0; j <= h->rlen; ++j)
        {
            size_t c = IDX(h, i, j);
            h->V[0][c] = -INFINITY;
            h->U[0][c] = NULL_TRACK;
            for (int s = 1; s < 6; ++s)
            {
                h->V[s][c] = -INFINITY;
                h->U[s][c] = (!i || !j) ? make_track(nn, UNMODELED, 0, 0) : make_track(tbd, UNCERTAIN, 0, 0);
            }
            if (h->kind == HMMO_RFHMM)
            { /* Y: row 0 is a free head, rfhmm.cpp:232-249 */
                if (i >= 1)
                {
                    h->V[R_Y][c] = -INFINITY;
                    h->U[R_Y][c] = NULL_TRACK;
                }
                else
                {
                    h->V[R_Y][c] = 0;
                    h->U[R_Y][c] = j == 1 ? make_track(R_S, LFLANK, 0, j) : make_track(R_Y, LFLANK, 0, j);
                }
            }
        }
    h->V[0][0] = 0;
    h->U[0][0] = START_TRACK;
    if (h->kind == HMMO_LFHMM) h->U[L_ML][0] = START_TRACK; /* lfhmm.cpp:271-272 */
    return 0;
}

static void fill_ahmm(hctx* h)
{ /* ahmm.cpp:490-539 */
    for (int i = 1; i <= h->plen; ++i)
        for (int j = 1; j <= h->rlen; ++j)
        {
            size_t c = IDX(h, i, j), d = IDX(h, i - 1, j - 1), u = IDX(h, i - 1, j), l = IDX(h, i, j - 1);
            BEGIN_STATE(h);
            proc_comp(h, A_S, A_M, d, j - 1, MT_MATCH);
            proc_comp(h, A_M, A_M, d, j - 1, MT_MATCH);
            proc_comp(h, A_D, A_M, d, j - 1, MT_MATCH);
            proc_comp(h, A_I, A_M, d, j - 1, MT_MATCH);
            END_STATE(h, A_M, c);
            BEGIN_STATE(h);
            proc_comp(h, A_S, A_D, u, j, MT_MODEL);
            proc_comp(h, A_M, A_D, u, j, MT_MODEL);
            proc_comp(h, A_D, A_D, u, j, MT_MODEL);
            END_STATE(h, A_D, c);
            BEGIN_STATE(h);
            proc_comp(h, A_S, A_I, l, j - 1, MT_READ);
            proc_comp(h, A_M, A_I, l, j - 1, MT_READ);
            proc_comp(h, A_I, A_I, l, j - 1, MT_READ);
            END_STATE(h, A_I, c);
        }
}

static void fill_lfhmm(hctx* h)
{ /* lfhmm.cpp:511-597 */
    for (int i = 1; i <= h->plen; ++i)
        for (int j = 1; j <= h->rlen; ++j)
        {
            size_t c = IDX(h, i, j), d = IDX(h, i - 1, j - 1), u = IDX(h, i - 1, j), l = IDX(h, i, j - 1);
            BEGIN_STATE(h);
            if (i <= h->lflen)
            {
                proc_comp(h, L_S, L_ML, d, j - 1, MT_MATCH);
                proc_comp(h, L_ML, L_ML, d, j - 1, MT_MATCH);
            }
            END_STATE(h, L_ML, c);
            BEGIN_STATE(h);
            if (i > h->lflen)
            {
                proc_comp(h, L_ML, L_M, d, j - 1, MT_MATCH);
                proc_comp(h, L_M, L_M, d, j - 1, MT_MATCH);
                proc_comp(h, L_D, L_M, d, j - 1, MT_MATCH);
                proc_comp(h, L_I, L_M, d, j - 1, MT_MATCH);
            }
            END_STATE(h, L_M, c);
            BEGIN_STATE(h);
            if (i > h->lflen)
            {
                proc_comp(h, L_ML, L_D, u, j, MT_MODEL);
                proc_comp(h, L_M, L_D, u, j, MT_MODEL);
                proc_comp(h, L_D, L_D, u, j, MT_MODEL);
            }
            END_STATE(h, L_D, c);
            BEGIN_STATE(h);
            if (i > h->lflen)
            {
                proc_comp(h, L_ML, L_I, l, j - 1, MT_READ);
                proc_comp(h, L_M, L_I, l, j - 1, MT_READ);
                proc_comp(h, L_I, L_I, l, j - 1, MT_READ);
            }
            END_STATE(h, L_I, c);
            BEGIN_STATE(h);
            if (i > h->lflen)
            {
                proc_comp(h, L_M, L_Z, l, j - 1, MT_READ);
                proc_comp(h, L_D, L_Z, l, j - 1, MT_READ);
                proc_comp(h, L_Z, L_Z, l, j - 1, MT_READ);
            }
            END_STATE(h, L_Z, c);
        }
}

static void fill_rfhmm(hctx* h)
{ /* rfhmm.cpp:535-611; Y is invariant */
    for (int i = 1; i <= h->plen; ++i)
        for (int j = 1; j <= h->rlen; ++j)
        {
            size_t c = IDX(h, i, j), d = IDX(h, i - 1, j - 1), u = IDX(h, i - 1, j), l = IDX(h, i, j - 1);
            BEGIN_STATE(h);
            proc_comp(h, R_S, R_M, d, j - 1, MT_MATCH);
            proc_comp(h, R_Y, R_M, d, j - 1, MT_MATCH);
            proc_comp(h, R_M, R_M, d, j - 1, MT_MATCH);
            proc_comp(h, R_D, R_M, d, j - 1, MT_MATCH);
            proc_comp(h, R_I, R_M, d, j - 1, MT_MATCH);
            END_STATE(h, R_M, c);
            BEGIN_STATE(h);
            proc_comp(h, R_S, R_D, u, j, MT_MODEL);
            proc_comp(h, R_Y, R_D, u, j, MT_MODEL);
            proc_comp(h, R_M, R_D, u, j, MT_MODEL);
            proc_comp(h, R_D, R_D, u, j, MT_MODEL);
            END_STATE(h, R_D, c);
            BEGIN_STATE(h);
            proc_comp(h, R_S, R_I, l, j - 1, MT_READ);
            proc_comp(h, R_Y, R_I, l, j - 1, MT_READ);
            proc_comp(h, R_M, R_I, l, j - 1, MT_READ);
            proc_comp(h, R_I, R_I, l, j - 1, MT_READ);
            END_STATE(h, R_I, c);
            BEGIN_STATE(h);
            proc_comp(h, R_S, R_MR, d, j - 1, MT_MATCH);
            proc_comp(h, R_Y, R_MR, d, j - 1, MT_MATCH);
            proc_comp(h, R_M, R_MR, d, j - 1, MT_MATCH);
            proc_comp(h, R_D, R_MR, d, j - 1, MT_MATCH);
            proc_comp(h, R_I, R_MR, d, j - 1, MT_MATCH);
            proc_comp(h, R_MR, R_MR, d, j - 1, MT_MATCH);
            END_STATE(h, R_MR, c);
        }
}

enum { LFS = 0, LFE = 2, MOS = 4, MOE = 6, RFS = 8, RFE = 10 }; /* + MT_MODEL / MT_READ */

static void init_disc(hctx* h)
{
    for (int k = 1; k <= h->res->motif_count; ++k) h->disc[k] = 0;
}

/* ahmm.cpp:676-730 */
static void stats_ahmm(hctx* h, int32_t src_t, int32_t des_t, int32_t j)
{
    hmmo_result* r = h->res;
    int src_u = track_get_u(src_t), des_u = track_get_u(des_t);
    if (src_u == A_E)
    {
        if (des_u == A_M || des_u == A_D || des_u == A_I)
        {
            r->pos[RFS + MT_MODEL] = r->pos[RFS + MT_READ] = r->pos[RFE + MT_MODEL] = r->pos[RFE + MT_READ] = -1;
            r->pos[MOE + MT_MODEL] = track_get_c(des_t);
            r->motif_count = track_get_c(des_t);
            r->last_motif_pos = track_get_p(des_t);
            r->pos[MOE + MT_READ] = j;
            init_disc(h);
            if (des_u == A_D || track_get_base(h, des_t) != h->read[j - 1]) ++h->disc[r->motif_count];
        }
    }
    if (des_u == A_M)
    {
        if (track_get_base(h, des_t) != h->read[j - 1])
        {
            r->trf_score -= 7;
            ++h->disc[track_get_c(des_t)];
        }
        else
            r->trf_score += 2;
    }
    if (des_u == A_D || des_u == A_I)
    {
        r->trf_score -= 7;
        ++h->disc[track_get_c(des_t)];
    }
    r->frac_no_repeats = r->motif_count - (h->mlen - r->last_motif_pos) / ((float)h->mlen);
}

/* lfhmm.cpp:750-861 */
static void stats_lfhmm(hctx* h, int32_t src_t, int32_t des_t, int32_t j)
{
    hmmo_result* r = h->res;
    int src_u = track_get_u(src_t), des_u = track_get_u(des_t);
    if (src_u == L_E)
    {
        if (des_u == L_Z)
        {
            r->pos[RFE + MT_MODEL] = -1;
            r->pos[RFE + MT_READ] = j + 1;
        }
        else if (des_u == L_M || des_u == L_D || des_u == L_I)
        {
            r->pos[RFS + MT_MODEL] = 0;
            r->pos[RFS + MT_READ] = j;
            r->pos[RFE + MT_MODEL] = 0;
            r->pos[RFE + MT_READ] = j;
            r->pos[MOE + MT_MODEL] = track_get_c(des_t);
            r->motif_count = track_get_c(des_t);
            r->pos[MOE + MT_READ] = j;
            r->trf_score = 0;
            init_disc(h);
            if (des_u == L_D || des_u == L_I || track_get_base(h, des_t) != h->read[j - 1]) ++h->disc[r->motif_count];
        }
    }
    else if (src_u == L_Z)
    {
        if (des_u == L_M || des_u == L_D)
        {
            r->pos[RFS + MT_MODEL] = track_get_p(src_t);
            r->pos[RFS + MT_READ] = j + 1;
            r->pos[MOE + MT_MODEL] = track_get_c(des_t);
            r->motif_count = track_get_c(des_t);
            r->last_motif_pos = track_get_p(des_t);
            r->pos[MOE + MT_READ] = j;
            r->trf_score = 0;
            init_disc(h);
            if (des_u == L_D || track_get_base(h, des_t) != h->read[j - 1]) ++h->disc[r->motif_count];
        }
    }
    else if (src_u == L_M)
    {
        if (des_u == L_ML)
        {
            r->pos[MOS + MT_MODEL] = track_get_c(src_t);
            r->pos[MOS + MT_READ] = j + 1;
            r->pos[LFE + MT_MODEL] = track_get_p(des_t);
            r->pos[LFE + MT_READ] = j;
        }
    }
    else if (src_u == L_ML)
    {
        if (des_u == L_S)
        {
            r->pos[LFS + MT_MODEL] = track_get_p(src_t);
            r->pos[LFS + MT_READ] = j + 1;
        }
    }
    if (des_u == L_M)
    {
        if (track_get_base(h, des_t) != h->read[j - 1])
        {
            r->trf_score -= 7;
            ++h->disc[track_get_c(des_t)];
        }
        else
            r->trf_score += 2;
    }
    if (des_u == L_D || des_u == L_I)
    {
        r->trf_score -= 7;
        ++h->disc[track_get_c(des_t)];
    }
    r->frac_no_repeats = r->motif_count - (h->mlen - r->last_motif_pos) / ((float)h->mlen);
}

/* rfhmm.cpp:727-840 */
static void stats_rfhmm(hctx* h, int32_t src_t, int32_t des_t, int32_t j)
{
    hmmo_result* r = h->res;
    int src_u = track_get_u(src_t), des_u = track_get_u(des_t);
    if (src_u == R_E)
    {
        if (des_u == R_MR)
        {
            r->pos[RFE + MT_MODEL] = track_get_p(des_t);
            r->pos[RFE + MT_READ] = j;
        }
    }
    else if (src_u == R_MR)
    {
        if (des_u == R_M || des_u == R_D)
        {
            r->pos[RFS + MT_MODEL] = track_get_p(src_t);
            r->pos[RFS + MT_READ] = j + 1;
            r->pos[MOE + MT_MODEL] = track_get_c(des_t);
            r->motif_count = track_get_c(des_t);
            r->pos[MOE + MT_READ] = j;
            init_disc(h);
            if (des_u == R_D || track_get_base(h, des_t) != h->read[j - 1]) ++h->disc[r->motif_count];
        }
        else if (des_u == R_Y)
        {
            r->pos[RFS + MT_MODEL] = track_get_p(src_t);
            r->pos[RFS + MT_READ] = j + 1;
            r->pos[LFE + MT_MODEL] = track_get_p(des_t);
            r->pos[LFE + MT_READ] = j;
        }
    }
    else if (src_u == R_M)
    {
        if (des_u == R_Y)
        {
            r->pos[MOS + MT_MODEL] = track_get_c(src_t);
            r->pos[MOS + MT_READ] = j + 1;
            r->pos[LFE + MT_MODEL] = track_get_p(des_t);
            r->pos[LFE + MT_READ] = j;
        }
    }
    else if (src_u == R_Y)
    {
        if (des_u == R_S)
        {
            r->pos[LFS + MT_MODEL] = track_get_p(des_t);
            r->pos[LFS + MT_READ] = j + 1;
        }
    }
    /* the per-copy discordance counts (rfhmm.cpp:823-835) feed nothing align() reports; kept for form */
    if (des_u == R_M && track_get_c(des_t) < 1100 && track_get_base(h, des_t) != h->read[j - 1]) ++h->disc[track_get_c(des_t)];
    if ((des_u == R_D || des_u == R_I) && track_get_c(des_t) < 1100) ++h->disc[track_get_c(des_t)];
}

static void stats(hctx* h, int32_t src_t, int32_t des_t, int32_t j)
{
    if (h->kind == HMMO_AHMM)
        stats_ahmm(h, src_t, des_t, j);
    else if (h->kind == HMMO_LFHMM)
        stats_lfhmm(h, src_t, des_t, j);
    else
        stats_rfhmm(h, src_t, des_t, j);
}

#define CONSIDER(st, op)                            \
    if (h->V[st][c] op r->optimal_score)            \
    {                                               \
        r->optimal_score = h->V[st][c];             \
        r->optimal_state = st;                      \
        r->optimal_probe_len = i;                   \
    }

/* ahmm.cpp:592-671, lfhmm.cpp:659-745, rfhmm.cpp:650-725 */
static int trace_path(hctx* h, int32_t* path)
{
    hmmo_result* r = h->res;
    int32_t rev[4096];
    r->optimal_score = -INFINITY;
    r->optimal_probe_len = 0;
    if (h->kind == HMMO_AHMM)
    {
        r->optimal_state = A_TBD;
        r->trf_score = 0;
        for (int i = h->lflen; i <= h->plen; ++i)
        {
            size_t c = IDX(h, i, h->rlen);
            CONSIDER(A_M, >=)
            CONSIDER(A_D, >=)
        }
    }
    else if (h->kind == HMMO_LFHMM)
    {
        r->optimal_state = L_TBD;
        for (int i = h->lflen; i <= h->plen; ++i)
        {
            size_t c = IDX(h, i, h->rlen);
            CONSIDER(L_Z, >=)
            CONSIDER(L_M, >=)
            CONSIDER(L_D, >=)
            CONSIDER(L_I, >=)
        }
    }
    else
    {
        r->optimal_state = R_TBD;
        for (int i = 0; i <= h->plen; ++i)
        {
            size_t c = IDX(h, i, h->rlen);
            CONSIDER(R_MR, >)
        }
    }
    if (!(r->optimal_score > -INFINITY)) return -1; /* the reference would index U[] out of range here */

    int32_t i = r->optimal_probe_len, j = h->rlen, u;
    int E = h->kind == HMMO_AHMM ? A_E : L_E;
    int32_t last_t = h->kind == HMMO_RFHMM ? make_track(r->optimal_state, RFLANK, 0, h->rflen + 1)
                                           : make_track(r->optimal_state, UNMODELED, 0, 0);
    int32_t src_t = h->kind == HMMO_RFHMM ? make_track(E, RFLANK, 0, h->rflen + 1) : make_track(E, UNMODELED, 0, 0);
    int32_t des_t;
    int n = 0;
    do
    {
        u = track_get_u(last_t);
        if (u > 5 || i < 0 || j < 0 || n >= 4096) return -1;
        last_t = h->U[u][IDX(h, i, j)];
        des_t = track_set_u(last_t, u);
        rev[n++] = des_t;
        stats(h, src_t, des_t, j);
        src_t = des_t;
        if (h->kind == HMMO_AHMM)
        {
            if (u == A_M) { --i; --j; }
            else if (u == A_D) --i;
            else if (u == A_I || u == 4) --j;
        }
        else if (h->kind == HMMO_LFHMM)
        {
            if (u == L_ML || u == L_M) { --i; --j; }
            else if (u == L_D) --i;
            else if (u == L_I || u == L_Z) --j;
        }
        else
        {
            if (u == R_M || u == R_MR) { --i; --j; }
            else if (u == R_D) --i;
            else if (u == R_Y || u == R_I) --j;
        }
    } while (track_get_u(last_t) != 0 /* S */);
    stats(h, src_t, last_t, j);
    r->path_len = n;
    if (path)
        for (int k = 0; k < n; ++k) path[k] = rev[n - 1 - k];
    return 0;
}

int hmmo_align(int kind, const hmmo_params* par, const char* motif, const char* flank, const char* read,
               hmmo_result* res, int32_t* path)
{
    hctx H, *h = &H;
    memset(h, 0, sizeof(H));
    h->kind = kind;
    h->res = res;
    h->read = read;
    h->rlen = (int)strlen(read);
    h->mlen = (int)strlen(motif);
    h->model[MOTIF] = motif;
    if (kind == HMMO_AHMM)
    {
        if (h->rlen >= 1024) h->rlen = 1023; /* ahmm.cpp:476-480 */
        h->plen = h->rlen;
    }
    else if (kind == HMMO_LFHMM)
    {
        h->model[LFLANK] = flank;
        h->lflen = (int)strlen(flank);
        h->plen = h->lflen + h->rlen;
        if (h->rlen >= 1024) return -1; /* lfhmm.cpp:498-502 exit(1) */
    }
    else
    {
        h->model[RFLANK] = flank;
        h->rflen = (int)strlen(flank);
        h->plen = h->rlen + h->rflen;
        if (h->rlen > 1024 || h->plen >= 1024) return -1; /* rfhmm.cpp:524-528; rows >= 1024 leave the arrays */
    }
    if (h->rlen < 1 || h->mlen < 1) return -1;
    h->W = h->rlen + 1;
    hmmo_transitions(kind, par, &h->T[0][0]);
    /* ahmm.cpp:797-812 with LogTool::pl2log10_varp */
    h->e_mismatch = (float)(pl2log10_varp((uint32_t)par->pl) - par->mismatch_penalty);
    h->e_match = (float)(-(pl2log10_varp((uint32_t)par->pl)));
    /* clear_statistics(): ahmm.cpp:735-755 (same in the other two) */
    for (int k = 0; k < 12; ++k) res->pos[k] = -1;
    res->motif_count = 0;
    res->exact_motif_count = 0;
    res->motif_concordance = 0;
    int rc = init_uv(h);
    h->disc = (int32_t*)calloc(1200, sizeof(int32_t));
    if (rc == 0 && h->disc)
    {
        if (kind == HMMO_AHMM)
            fill_ahmm(h);
        else if (kind == HMMO_LFHMM)
            fill_lfhmm(h);
        else
            fill_rfhmm(h);
        rc = trace_path(h, path);
        if (rc == 0 && kind != HMMO_RFHMM)
        { /* ahmm.cpp:569-583, lfhmm.cpp:637-653 */
            res->exact_motif_count = res->motif_count;
            res->motif_concordance = 0;
            for (int32_t k = 1; k <= res->motif_count; ++k)
            {
                if (h->disc[k]) --res->exact_motif_count;
                if (h->mlen >= h->disc[k]) res->motif_concordance += (float)(h->mlen - h->disc[k]) / h->mlen;
            }
            if (kind == HMMO_AHMM)
                res->motif_concordance /= res->motif_count;
            else
                res->motif_concordance *= 1.0 / res->motif_count;
        }
    }
    for (int s = 0; s < 6; ++s)
    {
        free(h->V[s]);
        free(h->U[s]);
    }
    free(h->disc);
    return rc;
}
