/* oracle/fd_trace.cpp -- TEST INFRASTRUCTURE ONLY.
 * Link-time taps (GNU ld --wrap) on the UNMODIFIED reference classes: every AHMM / LFHMM / RFHMM set_model + align call that
 * FlankDetector issues inside `vt annotate_indels` (flank_detector.cpp:199-276 detect_flanks(FUZZY), :646-712
 * compute_purity_score) is written, with what the caller reads back, to the file named by $VT_HMM_TRACE:
 *     R <motif> <rflank> <read> <lflank_read_epos1>
 *     L <motif> <lflank> <read> <rflank_read_spos1>
 *     A <motif> - <read> <motif_concordance bits> <exact_motif_count> <motif_count> <frac_no_repeats bits> <trf_score>
 * No reference source is modified or copied: oracle/Makefile links the same objects as _ref/vt_ref plus this file with
 * -Wl,--wrap=<mangled name> for the six member functions below (the result: _ref/vt_ref_trace).  This pins the driver row
 * (SURVEY 8 a18): the (model, sequence) tuples and the accepted positions must be the ones vt_b200's driver produces. */
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <stdint.h>
#include "ahmm.h"
#include "lfhmm.h"
#include "rfhmm.h"

static FILE* trace_file()
{
    static FILE* f = NULL;
    static bool tried = false;
    if (!tried)
    {
        tried = true;
        const char* p = getenv("VT_HMM_TRACE");
        if (p && *p) f = fopen(p, "w");
    }
    return f;
}
static uint32_t fbits(float x)
{
    uint32_t u;
    memcpy(&u, &x, 4);
    return u;
}
static std::string g_a_motif, g_l_flank, g_l_motif, g_r_motif, g_r_flank;

extern "C"
{
    void __real__ZN4AHMM9set_modelEPKc(AHMM*, const char*);
    void __real__ZN4AHMM5alignEPKcS1_(AHMM*, const char*, const char*);
    void __real__ZN5LFHMM9set_modelEPKcS1_(LFHMM*, const char*, const char*);
    void __real__ZN5LFHMM5alignEPKcS1_(LFHMM*, const char*, const char*);
    void __real__ZN5RFHMM9set_modelEPKcS1_(RFHMM*, const char*, const char*);
    void __real__ZN5RFHMM5alignEPKcS1_(RFHMM*, const char*, const char*);

    void __wrap__ZN4AHMM9set_modelEPKc(AHMM* h, const char* motif)
    {
        g_a_motif = motif;
        __real__ZN4AHMM9set_modelEPKc(h, motif);
    }
    void __wrap__ZN4AHMM5alignEPKcS1_(AHMM* h, const char* read, const char* qual)
    {
        __real__ZN4AHMM5alignEPKcS1_(h, read, qual);
        if (FILE* f = trace_file())
            fprintf(f, "A %s - %s %08x %d %d %08x %d\n", g_a_motif.c_str(), read, fbits(h->motif_concordance), h->exact_motif_count, h->motif_count,
                    fbits(h->frac_no_repeats), h->trf_score);
    }
    void __wrap__ZN5LFHMM9set_modelEPKcS1_(LFHMM* h, const char* lflank, const char* motif)
    {
        g_l_flank = lflank;
        g_l_motif = motif;
        __real__ZN5LFHMM9set_modelEPKcS1_(h, lflank, motif);
    }
    void __wrap__ZN5LFHMM5alignEPKcS1_(LFHMM* h, const char* read, const char* qual)
    {
        __real__ZN5LFHMM5alignEPKcS1_(h, read, qual);
        if (FILE* f = trace_file()) fprintf(f, "L %s %s %s %d\n", g_l_motif.c_str(), g_l_flank.c_str(), read, h->get_rflank_read_spos1());
    }
    void __wrap__ZN5RFHMM9set_modelEPKcS1_(RFHMM* h, const char* motif, const char* rflank)
    {
        g_r_motif = motif;
        g_r_flank = rflank;
        __real__ZN5RFHMM9set_modelEPKcS1_(h, motif, rflank);
    }
    void __wrap__ZN5RFHMM5alignEPKcS1_(RFHMM* h, const char* read, const char* qual)
    {
        __real__ZN5RFHMM5alignEPKcS1_(h, read, qual);
        if (FILE* f = trace_file()) fprintf(f, "R %s %s %s %d\n", g_r_motif.c_str(), g_r_flank.c_str(), read, h->get_lflank_read_epos1());
    }
}
