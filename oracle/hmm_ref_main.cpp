/* oracle/hmm_ref_main.cpp -- TEST INFRASTRUCTURE ONLY.
 * Known-answer generator for the HMM family: links the UNMODIFIED reference classes
 * AHMM (ahmm.h:94), LFHMM (lfhmm.h), RFHMM (rfhmm.h) compiled from /root/reference by oracle/Makefile
 * and dumps everything their align() leaves behind.  One object per model is reused across
 * alignments, exactly like FlankDetector does (flank_detector.cpp:29-80), so fields the reference
 * does not reset between calls are stale here as they are there (tests mask those).
 *
 * usage: hmm_ref delta epsilon tau eta mismatch_penalty < cases.txt > answers.txt
 * case line:   A <motif> - <read>   |   L <motif> <lflank> <read>   |   R <motif> <rflank> <read>
 * first three output lines: "T A|L|R" + the 7x7 (6x6 for A) transition table as float bit patterns.
 * answer line: kind score_bits state probe_len path_len 12 positions motif_count exact_motif_count
 *              concordance_bits trf_score frac_bits last_motif_pos path tracks (hex) ... */
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <stdint.h>
#include "ahmm.h"
#include "lfhmm.h"
#include "rfhmm.h"

static uint32_t fbits(float f)
{
    uint32_t u;
    memcpy(&u, &f, 4);
    return u;
}

/* RFHMM has no trf_score / frac_no_repeats / last_motif_pos members (rfhmm.h:112-126) */
template <class H>
static void dump_tail(H* h)
{
    printf(" %d %08x %d", h->trf_score, fbits(h->frac_no_repeats), h->last_motif_pos);
}
static void dump_tail(RFHMM*)
{
    printf(" 0 00000000 -1");
}

template <class H>
static void dump(const char* kind, H* h)
{
    printf("%s %08x %d %d %d", kind, fbits(h->optimal_score), h->optimal_state, h->optimal_probe_len,
           h->optimal_path_len);
    printf(" %d %d %d %d %d %d %d %d %d %d %d %d", h->lflank_start[0], h->lflank_start[1], h->lflank_end[0],
           h->lflank_end[1], h->motif_start[0], h->motif_start[1], h->motif_end[0], h->motif_end[1],
           h->rflank_start[0], h->rflank_start[1], h->rflank_end[0], h->rflank_end[1]);
    printf(" %d %d %08x", h->motif_count, h->exact_motif_count, fbits(h->motif_concordance));
    dump_tail(h);
    for (int k = 0; k < h->optimal_path_len; ++k) printf(" %x", (unsigned)h->optimal_path_ptr[k]);
    printf("\n");
}

template <class H>
static void dump_T(const char* kind, H* h, int n)
{
    printf("T %s", kind);
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < n; ++j) printf(" %08x", fbits(h->T[i][j]));
    printf("\n");
}

template <class H>
static void setup(H* h, char** argv)
{
    h->set_delta(atof(argv[1]));
    h->set_epsilon(atof(argv[2]));
    h->set_tau(atof(argv[3]));
    h->set_eta(atof(argv[4]));
    h->set_mismatch_penalty(atof(argv[5]));
    h->initialize_T();
}

int main(int argc, char** argv)
{
    if (argc < 6)
    {
        fprintf(stderr, "usage: hmm_ref delta epsilon tau eta mismatch_penalty < cases > answers\n");
        return 1;
    }
    AHMM* a = new AHMM(false);
    LFHMM* l = new LFHMM(false);
    RFHMM* r = new RFHMM(false);
    setup(a, argv);
    setup(l, argv);
    setup(r, argv);
    /* stale-able fields start from a known value so that runs are reproducible */
    a->last_motif_pos = l->last_motif_pos = -1;
    a->trf_score = l->trf_score = 0;
    a->frac_no_repeats = l->frac_no_repeats = 0;
    dump_T("A", a, 6);
    dump_T("L", l, 7);
    dump_T("R", r, 7);
    std::string qual(1100, 'K'); /* flank_detector.cpp:74 */
    static char kind[8], motif[2048], flank[2048], read[4096];
    while (scanf("%7s %2047s %2047s %4095s", kind, motif, flank, read) == 4)
    {
        if (kind[0] == 'A')
        {
            a->set_model(motif);
            a->align(read, qual.c_str());
            dump("A", a);
        }
        else if (kind[0] == 'L')
        {
            l->set_model(flank, motif);
            l->align(read, qual.c_str());
            dump("L", l);
        }
        else
        {
            r->set_model(motif, flank);
            r->align(read, qual.c_str());
            dump("R", r);
        }
    }
    return 0;
}
