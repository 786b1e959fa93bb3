/* oracle/hmm_oracle.h -- TEST INFRASTRUCTURE ONLY (CPU restatement of the reference's HMM family).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may use it.
 * Pinned against the unmodified reference classes by oracle/_ref/hmm_ref (oracle/hmm_ref_main.cpp) and
 * the committed vectors tests/golden/hmm_*.txt.  The reference has no tests of its own for this family
 * ("parity unpinned by reference tests"); the pin is the reference's own code run here. */
#ifndef HMM_ORACLE_H
#define HMM_ORACLE_H
#include <stdint.h>

#define HMMO_AHMM 0
#define HMMO_LFHMM 1
#define HMMO_RFHMM 2

typedef struct
{
    float delta, epsilon, tau, eta, mismatch_penalty; /* ahmm.h:74-92 */
    int32_t pl;                                       /* qual - 33; FlankDetector passes 'K' = 42 */
} hmmo_params;

/* What align() leaves behind (ahmm.h:109-132).  Fields the reference does not touch in a given call keep
 * the caller's previous value: pass the same struct to successive calls to model one reused object. */
typedef struct
{
    float optimal_score;
    int32_t optimal_state, optimal_probe_len, path_len;
    /* lflank_start[MODEL,READ], lflank_end[..], motif_start, motif_end, rflank_start, rflank_end */
    int32_t pos[12];
    int32_t motif_count, exact_motif_count;
    float motif_concordance;
    int32_t trf_score;
    float frac_no_repeats;
    int32_t last_motif_pos;
} hmmo_result;

/* transition table as the reference computes it (initialize_T); T is 7x7 row-major, unused entries -inf */
void hmmo_transitions(int kind, const hmmo_params* par, float* T);

/* One alignment.  flank = left flank (LFHMM), right flank (RFHMM), ignored (AHMM).
 * path receives the optimal path tracks in forward order (capacity 4096 as in the reference).
 * returns 0, or -1 where the reference would exit(1)/run off its arrays. */
int hmmo_align(int kind, const hmmo_params* par, const char* motif, const char* flank, const char* read,
               hmmo_result* res, int32_t* path);

#endif
