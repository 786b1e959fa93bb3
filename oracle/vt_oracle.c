/* oracle/vt_oracle.c -- TEST INFRASTRUCTURE ONLY.  See vt_oracle.h.
 * CPU restatement of the reference path; literal loops, one base at a time, as the reference does. */
#include "vt_oracle.h"
#include <ctype.h>
#include <stdlib.h>
#include <string.h>

#define VT_REF 0
#define VT_SNP 1
#define VT_MNP 2
#define VT_INDEL 4
#define VT_CLUMPED 8
#define VT_VNTR 16
#define VT_SV 32

static int has_char(const uint8_t* s, uint32_t n, uint8_t c)
{
    for (uint32_t i = 0; i < n; ++i)
        if (s[i] == c) return 1;
    return 0;
}

static int is_ts(uint8_t a, uint8_t b)
{
    return (a == 'G' && b == 'A') || (a == 'A' && b == 'G') || (a == 'C' && b == 'T') || (a == 'T' && b == 'C');
}

/* variant_manip.cpp:227-508 */
int32_t vto_classify(int32_t n_allele, const uint8_t* const* al, const uint32_t* len)
{
    int32_t vtype = VT_REF;
    int homogeneous_length = 1;
    int32_t rlen = (int32_t)len[0];

    for (int32_t i = 1; i < n_allele; ++i)
    {
        const uint8_t* a = al[i];
        uint32_t l = len[i];
        int32_t type = VT_REF;
        if (has_char(a, l, '<')) /* :262-322 */
        {
            if (l >= 5)
            {
                if (a[0] == '<' && a[1] == 'V' && a[2] == 'N' && a[l - 1] == '>')
                {
                    for (uint32_t j = 3; j < l - 1; ++j)
                        if (a[j] < '0' || a[j] > '9') type = VT_VNTR;
                }
                else if (l == 6 && memcmp(a, "<VNTR>", 6) == 0)
                    type = VT_VNTR;
                else if (l == 5 && memcmp(a, "<STR>", 5) == 0)
                    type = VT_VNTR;
                else if (a[0] == '<' && a[1] == 'S' && a[2] == 'T' && a[l - 1] == '>')
                {
                    type = VT_VNTR;
                    for (uint32_t j = 3; j < l - 1; ++j)
                        if ((a[j] < '0' || a[j] > '9') && a[j] != '.') type = VT_SV;
                }
            }
            vtype |= (type == VT_VNTR) ? VT_VNTR : VT_SV;
        }
        else if (has_char(a, l, '[') || has_char(a, l, ']')) /* :324-330 */
        {
            vtype |= VT_SV;
        }
        else if (a[0] == '.' || (l == len[0] && memcmp(a, al[0], l) == 0)) /* :332-335 */
        {
            /* VT_REF */
        }
        else /* :337-491 */
        {
            const uint8_t* ref = al[0];
            const uint8_t* alt = a;
            int32_t alen = (int32_t)l;
            if (rlen != alen) homogeneous_length = 0;
            int32_t rl = rlen, alx = alen;
            while (rl != 1 && alx != 1)
            {
                if (ref[rl - 1] == alt[alx - 1]) { --rl; --alx; }
                else break;
            }
            while (rl != 1 && alx != 1)
            {
                if (ref[0] == alt[0]) { ++ref; ++alt; --rl; --alx; }
                else break;
            }
            int32_t mlen = rl < alx ? rl : alx;
            int32_t dlen = alx - rl;
            int32_t diff = 0;
            if (mlen == 1 && dlen)
            {
                uint8_t ls, le, ss;
                if (rl > alx) { ls = ref[0]; le = ref[rl - 1]; ss = alt[0]; }
                else { ls = alt[0]; le = alt[alx - 1]; ss = ref[0]; }
                if (ls != ss && le != ss) ++diff;
            }
            else
            {
                for (int32_t j = 0; j < mlen; ++j)
                    if (ref[j] != alt[j]) ++diff;
            }
            (void)is_ts;
            if (mlen == diff) type |= (mlen == 1) ? VT_SNP : VT_MNP;
            if (dlen) type |= VT_INDEL;
            if (diff && diff < mlen) type |= VT_CLUMPED;
            vtype |= type;
        }
    }
    if (!(vtype & (VT_VNTR | VT_SV)) && vtype != VT_REF) /* :499-505 */
    {
        if (homogeneous_length && rlen > 1 && n_allele > 1) vtype |= VT_MNP;
    }
    return vtype;
}

/* faidx.c:832-860 position clipping, then raw byte */
static int64_t clip_pos(int64_t p, int64_t len)
{
    if (p < 0) return 0;
    if (p >= len) return len - 1;
    return p;
}

/* variant_manip.cpp:61-103 ; faidx_fetch_seq is raw case (faidx.c:692-736) */
int32_t vto_is_not_ref_consistent(const vto_ref* ref, int32_t rid, int32_t pos1, const uint8_t* a0, uint32_t len0)
{
    if (rid < 0 || rid >= ref->n_contig) return -1;
    int64_t clen = ref->len[rid];
    int64_t beg = (int64_t)pos1 - 1, end = beg + (int64_t)len0 - 1;
    if (end < beg) beg = end;
    beg = clip_pos(beg, clen);
    end = clip_pos(end, clen);
    int64_t n = end - beg + 1; /* fetched length; may be shorter than len0 at the contig end */
    const uint8_t* s = ref->seq[rid];
    int32_t code = 0;
    for (int64_t i = 0; i < n && i < (int64_t)len0; ++i)
    {
        int u = toupper(a0[i]);
        if (u != s[beg + i])
        {
            if (s[beg + i] == 'N' || u == 'N') code = 1;
            else { code = 2; break; }
        }
    }
    return code;
}

/* variant_manip.cpp:109-197 */
int32_t vto_is_normalized(int32_t n_allele, const uint8_t* const* al, const uint32_t* len)
{
    if (n_allele == 1) return 1;
    int exists_len_one = 0, first_same = 1, last_same = 1;
    int first_base = 0, last_base = 0;
    if (n_allele == 2)
    {
        uint32_t rlen = len[0], alen = len[1];
        if (rlen == 1 && alen == 1) return 1;
        if (rlen == alen && strncasecmp((const char*)al[0], (const char*)al[1], rlen) == 0) return 1;
        if (rlen == 1) exists_len_one = 1;
        first_base = toupper(al[0][0]);
        last_base = toupper(al[0][rlen - 1]);
        if (alen == 1) exists_len_one = 1;
        if (first_base != toupper(al[1][0])) first_same = 0;
        if (last_base != toupper(al[1][alen - 1])) last_same = 0;
        if (last_same || (!exists_len_one && first_same)) return 0;
        return 1;
    }
    int same = 1;
    for (int32_t i = 0; i < n_allele; ++i)
    {
        uint32_t l = len[i];
        if (i)
        {
            if (l == 1) exists_len_one = 1;
            if (first_base != toupper(al[i][0])) first_same = 0;
            if (last_base != toupper(al[i][l - 1])) last_same = 0;
            same = same && (l == len[0] && strncasecmp((const char*)al[i], (const char*)al[0], l) == 0);
        }
        else
        {
            if (l == 1) exists_len_one = 1;
            first_base = toupper(al[0][0]);
            last_base = toupper(al[0][l - 1]);
        }
    }
    if (same) return 1;
    if (last_same || (!exists_len_one && first_same)) return 0;
    return 1;
}

/* ---- literal string-vector machinery for right_trim_or_left_extend / left_trim ---- */
typedef struct
{
    uint8_t* buf; /* allocation */
    size_t cap;   /* capacity */
    size_t beg;   /* first live byte */
    size_t n;     /* live length */
} vstr;

static int vstr_init(vstr* s, const uint8_t* src, uint32_t n)
{
    s->cap = (size_t)n + 64;
    s->buf = (uint8_t*)malloc(s->cap);
    if (!s->buf) return -1;
    s->beg = 64;
    s->n = n;
    for (uint32_t i = 0; i < n; ++i) s->buf[s->beg + i] = (uint8_t)toupper(src[i]); /* normalize.cpp:223-233 */
    return 0;
}

static int vstr_prepend(vstr* s, uint8_t c) /* std::string::insert(0,1,c) */
{
    if (s->beg == 0)
    {
        size_t grow = s->cap + 64;
        uint8_t* nb = (uint8_t*)malloc(s->cap + grow);
        if (!nb) return -1;
        memcpy(nb + grow, s->buf, s->cap);
        free(s->buf);
        s->buf = nb;
        s->beg += grow;
        s->cap += grow;
    }
    s->buf[--s->beg] = c;
    s->n++;
    return 0;
}

/* hts_utils.cpp:41-51 over faidx.c:832-881: one upper-cased base at 0-based p0, clipped into the contig */
static uint8_t fetch_uc_base(const vto_ref* ref, int32_t rid, int64_t p0)
{
    int64_t p = clip_pos(p0, ref->len[rid]);
    uint8_t c = ref->seq[rid][p];
    if (isgraph(c)) c = (uint8_t)toupper(c);
    return c;
}

/* variant_manip.cpp:513-583 */
static int right_trim_or_left_extend(const vto_ref* ref, int32_t rid, vstr* A, int32_t n, int32_t* pos1,
                                     uint32_t* left_extended, uint32_t* right_trimmed)
{
    int to_right_trim = 1, to_left_extend = 0;
    if (n == 1) return 0;
    while (to_right_trim || to_left_extend)
    {
        to_right_trim = 1;
        to_left_extend = 0;
        for (int32_t i = 0; i < n; ++i)
        {
            if (A[i].n)
            {
                if (A[0].buf[A[0].beg + A[0].n - 1] != A[i].buf[A[i].beg + A[i].n - 1]) to_right_trim = 0;
                if (*pos1 == 1 && A[i].n == 1)
                {
                    to_right_trim = 0;
                    break;
                }
            }
            else
            {
                to_right_trim = 0;
                to_left_extend = 1;
                break;
            }
        }
        if (to_right_trim)
        {
            for (int32_t i = 0; i < n; ++i) A[i].n--;
            ++*right_trimmed;
        }
        if (to_left_extend)
        {
            --*pos1;
            uint8_t base = fetch_uc_base(ref, rid, (int64_t)*pos1 - 1);
            for (int32_t i = 0; i < n; ++i)
                if (vstr_prepend(&A[i], base)) return -1;
            ++*left_extended;
        }
    }
    return 0;
}

/* variant_manip.cpp:588-615 */
static void left_trim(vstr* A, int32_t n, int32_t* pos1, uint32_t* left_trimmed)
{
    int to_left_trim = 1;
    while (to_left_trim)
    {
        for (int32_t i = 0; i < n; ++i)
        {
            if (A[i].n == 1 || A[i].buf[A[i].beg] != A[0].buf[A[0].beg])
            {
                to_left_trim = 0;
                break;
            }
        }
        if (to_left_trim)
        {
            for (int32_t i = 0; i < n; ++i) { A[i].beg++; A[i].n--; }
            ++*pos1;
            ++*left_trimmed;
        }
    }
}

/* normalize.cpp:174-330 */
int32_t vto_normalize_batch(const vto_ref* ref, const vto_batch* in, vto_norm_out* out)
{
    memset(out->counters, 0, sizeof(out->counters));
    out->n_changed = 0;
    out->n_changed_bytes = 0;
    if (out->chg_off) out->chg_off[0] = 0;
    for (uint32_t r = 0; r < in->n_rec; ++r)
    {
        uint32_t ab = in->allele_beg[r];
        int32_t n = (int32_t)(in->allele_beg[r + 1] - ab);
        const uint8_t** al = (const uint8_t**)malloc(sizeof(uint8_t*) * (size_t)(n > 0 ? n : 1));
        uint32_t* len = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)(n > 0 ? n : 1));
        if (!al || !len) return -1;
        for (int32_t i = 0; i < n; ++i)
        {
            al[i] = in->bytes + in->allele_off[ab + i];
            len[i] = in->allele_off[ab + i + 1] - in->allele_off[ab + i];
        }
        int32_t pos1 = in->pos1[r], rid = in->rid[r];
        int32_t type = vto_classify(n, al, len);
        uint8_t st = 0;
        int32_t nrc = 0;
        if (type != VT_SNP) /* normalize.cpp:205 */
        {
            nrc = vto_is_not_ref_consistent(ref, rid, pos1, al[0], len[0]);
            if (nrc < 0) { st |= VTO_ST_FETCH_FAIL; nrc = 0; }
            st |= (uint8_t)nrc;
        }
        if (!(st & (VTO_ST_FETCH_FAIL | VTO_ST_REFCODE_MASK)) && !(type & VT_SV) && !(type & VT_VNTR) &&
            !vto_is_normalized(n, al, len)) /* :218 */
        {
            if (rid < 0 || rid >= ref->n_contig)
            {
                st |= VTO_ST_FETCH_FAIL; /* pure SNP-typed record on an unknown contig entering the path */
            }
            else
            {
                st |= VTO_ST_ENTERED;
                vstr* A = (vstr*)malloc(sizeof(vstr) * (size_t)n);
                if (!A) return -1;
                for (int32_t i = 0; i < n; ++i)
                    if (vstr_init(&A[i], al[i], len[i])) return -1;
                uint32_t le = 0, lt = 0, rt = 0;
                if (right_trim_or_left_extend(ref, rid, A, n, &pos1, &le, &rt)) return -1;
                left_trim(A, n, &pos1, &lt);
                if (lt || le || rt) /* :239 */
                {
                    st |= VTO_ST_CHANGED;
                    uint32_t cat = 0; /* :254-311 */
                    if (le) cat = (rt > le) ? 4 : 5;
                    else if (lt && rt > le) cat = 3;
                    else if (lt) cat = 1;
                    else if (rt > le) cat = 2;
                    st |= (uint8_t)(cat << VTO_ST_CAT_SHIFT);
                    if (cat) out->counters[(n == 2 ? 0 : 5) + (cat - 1)]++;
                    uint32_t k = out->n_changed++;
                    out->chg_rec[k] = r;
                    out->chg_pos1[k] = pos1;
                    out->chg_lt[k] = lt;
                    out->chg_le[k] = le;
                    out->chg_rt[k] = rt;
                    uint64_t o = out->chg_off[k];
                    for (int32_t i = 0; i < n; ++i)
                    {
                        memcpy(out->chg_bytes + o, A[i].buf + A[i].beg, A[i].n);
                        o += A[i].n;
                    }
                    out->chg_off[k + 1] = o;
                    out->n_changed_bytes = o;
                }
                for (int32_t i = 0; i < n; ++i) free(A[i].buf);
                free(A);
            }
        }
        if (type == VT_REF) out->counters[11]++;
        else out->counters[10]++;
        out->status[r] = st;
        out->vtype[r] = (uint8_t)type;
        free(al);
        free(len);
    }
    return 0;
}

/* decompose.cpp:194-966 (fan-out and counters only; INFO/FORMAT subsetting is host-side record work) */
int32_t vto_decompose_plan(const vto_batch* in, uint32_t* src_rec, uint16_t* alt_idx, uint64_t* n_out, uint32_t counters[4])
{
    uint64_t k = 0;
    counters[0] = counters[1] = counters[2] = counters[3] = 0;
    for (uint32_t r = 0; r < in->n_rec; ++r)
    {
        uint32_t n = in->allele_beg[r + 1] - in->allele_beg[r];
        if (n > 2)
        {
            counters[2]++;          /* no_multiallelic :214 */
            counters[3] += n - 2;   /* no_additional_biallelic :215 */
            for (uint32_t i = 1; i < n; ++i) { src_rec[k] = r; alt_idx[k] = (uint16_t)i; ++k; }
        }
        else
        {
            counters[1]++;          /* no_biallelic :962 */
            src_rec[k] = r;
            alt_idx[k] = (uint16_t)(n == 2 ? 1 : 0);
            ++k;
        }
        counters[0]++;              /* no_variants :965 */
    }
    *n_out = k;
    return 0;
}

/* ---------------------------------------------------------------------------------------------------------
 * vt uniq: uniq.cpp:110-156 over BCFSyncedReader(SYNC_BY_VAR) with ONE input file, sequential
 * (non-indexed) reading.  Restated literally: a record buffer that is topped up until it holds records of two
 * different (rid,pos) (bcf_synced_reader.cpp:497-640, the `else` branch :594-637), a priority queue ordered by
 * CompareBCFPtr (bcf_synced_reader.h:86-122) on (rid, pos1, strcmp of the sorted allele string
 * hts_utils.cpp:1136-1167), and read_next_position (:388-433) popping all records that compare equal to the top.
 * std::priority_queue is libstdc++'s: push = push_back + std::push_heap, pop = std::pop_heap + pop_back
 * (bits/stl_queue.h, bits/stl_heap.h __push_heap / __adjust_heap), restated below because WHICH of several equal
 * records surfaces first (= crecs[0], the record uniq keeps) depends on the heap's sift order, not on file order.
 * --------------------------------------------------------------------------------------------------------- */
typedef struct uq_rec
{
    int32_t rid, pos1;
    char* key; /* bcf_alleles2string_sorted: REF,ALT (n==2) or REF,sorted ALTs */
    uint32_t idx;
} uq_rec;

static int uq_cmpstr(const void* a, const void* b) { return strcmp(*(char* const*)a, *(char* const*)b); } /* hts_utils.cpp:1031-1037 */

static char* uq_key(const vto_batch* in, uint32_t r)
{
    uint32_t ab = in->allele_beg[r], n = in->allele_beg[r + 1] - ab;
    uint32_t total = in->allele_off[ab + n] - in->allele_off[ab];
    char* s = (char*)malloc(total + n + 1);
    char** al = (char**)malloc(sizeof(char*) * (n ? n : 1));
    if (!s || !al) return NULL;
    for (uint32_t i = 0; i < n; ++i)
    {
        uint32_t o = in->allele_off[ab + i], l = in->allele_off[ab + i + 1] - o;
        al[i] = (char*)malloc(l + 1);
        memcpy(al[i], in->bytes + o, l);
        al[i][l] = 0;
    }
    if (n > 2) qsort(al + 1, n - 1, sizeof(char*), uq_cmpstr); /* :1156 (n==2 needs no sort, :1141-1146) */
    size_t k = 0;
    for (uint32_t i = 0; i < n; ++i)
    {
        if (i) s[k++] = ',';
        size_t l = strlen(al[i]);
        memcpy(s + k, al[i], l);
        k += l;
        free(al[i]);
    }
    s[k] = 0;
    free(al);
    return s;
}

/* CompareBCFPtr::operator() (bcf_synced_reader.h:89-121), one file => file_index equal => false on ties */
static int uq_comp(const uq_rec* a, const uq_rec* b)
{
    if (a->rid == b->rid)
    {
        if (a->pos1 == b->pos1)
        {
            int d = strcmp(a->key, b->key);
            if (d == 0) return 0;
            return d > 0;
        }
        return a->pos1 > b->pos1;
    }
    return a->rid > b->rid;
}

/* BCFSyncedReader::bcfptr_cmp, sequential branch (bcf_synced_reader.cpp:361-383) */
static int uq_cmp3(const uq_rec* a, const uq_rec* b)
{
    if (a->rid == b->rid)
    {
        if (a->pos1 == b->pos1) return strcmp(a->key, b->key);
        return a->pos1 >= b->pos1 ? 1 : -1;
    }
    return a->rid >= b->rid ? 1 : -1;
}

/* libstdc++ bits/stl_heap.h */
static void uq_push_heap(uq_rec** first, long hole, long top, uq_rec* value)
{
    long parent = (hole - 1) / 2;
    while (hole > top && uq_comp(first[parent], value))
    {
        first[hole] = first[parent];
        hole = parent;
        parent = (hole - 1) / 2;
    }
    first[hole] = value;
}

static void uq_adjust_heap(uq_rec** first, long hole, long len, uq_rec* value)
{
    const long top = hole;
    long child = hole;
    while (child < (len - 1) / 2)
    {
        child = 2 * (child + 1);
        if (uq_comp(first[child], first[child - 1])) child--;
        first[hole] = first[child];
        hole = child;
    }
    if ((len & 1) == 0 && child == (len - 2) / 2)
    {
        child = 2 * (child + 1);
        first[hole] = first[child - 1];
        hole = child - 1;
    }
    uq_push_heap(first, hole, top, value);
}

typedef struct uq_state
{
    const vto_batch* in;
    uq_rec* recs;
    uint32_t next;       /* next record bcf_read would return */
    uint32_t* buf;       /* buffer[0] as an array list of record indices (front = buf[0]) */
    uint32_t nbuf;
    uq_rec** pq;
    long npq;
    int64_t bad;
} uq_state;

static void uq_pq_push(uq_state* S, uq_rec* v)
{
    S->pq[S->npq++] = v; /* c.push_back(x); std::push_heap(c.begin(), c.end(), comp) */
    uq_push_heap(S->pq, S->npq - 1, 0, S->pq[S->npq - 1]);
}

static void uq_pq_pop(uq_state* S)
{
    if (S->npq > 1)
    { /* std::pop_heap: --last; value = *last; *last = *first; __adjust_heap(first, 0, last-first, value) */
        uq_rec* value = S->pq[S->npq - 1];
        S->pq[S->npq - 1] = S->pq[0];
        uq_adjust_heap(S->pq, 0, S->npq - 1, value);
    }
    S->npq--;
}

/* BCFSyncedReader::fill_buffer, sequential branch (bcf_synced_reader.cpp:594-637) */
static void uq_fill_buffer(uq_state* S)
{
    if (S->nbuf >= 2 || S->bad >= 0) return;
    int32_t rid = S->nbuf == 0 ? -1 : S->recs[S->buf[0]].rid;
    int32_t pos1 = S->nbuf == 0 ? 0 : S->recs[S->buf[0]].pos1;
    while (S->next < S->in->n_rec)
    {
        uq_rec* v = &S->recs[S->next];
        if (S->nbuf)
        { /* bcf_is_in_order(buffer.back(), v), hts_utils.cpp:1042-1050 */
            const uq_rec* a = &S->recs[S->buf[S->nbuf - 1]];
            int ok = a->rid == v->rid ? a->pos1 <= v->pos1 : a->rid < v->rid;
            if (!ok)
            {
                S->bad = S->next; /* "VCF file not in order" + exit(1) */
                return;
            }
        }
        S->next++;
        S->buf[S->nbuf++] = v->idx;
        uq_pq_push(S, v);
        if (rid == -1)
        {
            rid = v->rid;
            pos1 = v->pos1;
        }
        if (v->rid != rid || v->pos1 != pos1) break;
    }
}

/* order[k] = input index of the k-th record popped; first[k] = 1 iff it is crecs[0] of a read_next_position
 * call (the record uniq writes; the following first[]==0 entries are the duplicates whose OLD_VARIANT values are
 * merged into it, uniq.cpp:114-143).  counters: no_total_variants, no_unique_variants (uniq.cpp:147-148).
 * *not_in_order = -1, or the index of the record at which the reference prints "VCF file not in order" and exits
 * (outputs up to that point are what was produced before).  rid must be >= 0. */
int32_t vto_uniq_plan(const vto_batch* in, uint32_t* order, uint8_t* first, uint32_t counters[2], int64_t* not_in_order)
{
    uq_state S;
    memset(&S, 0, sizeof S);
    S.in = in;
    S.bad = -1;
    S.recs = (uq_rec*)calloc(in->n_rec ? in->n_rec : 1, sizeof(uq_rec));
    S.buf = (uint32_t*)malloc(sizeof(uint32_t) * ((size_t)in->n_rec + 2));
    S.pq = (uq_rec**)malloc(sizeof(uq_rec*) * ((size_t)in->n_rec + 2));
    if (!S.recs || !S.buf || !S.pq) return -1;
    for (uint32_t r = 0; r < in->n_rec; ++r)
    { /* bcfptr ctor, bcf_synced_reader.h:57-70 */
        S.recs[r].rid = in->rid[r];
        S.recs[r].pos1 = in->pos1[r];
        S.recs[r].idx = r;
        S.recs[r].key = uq_key(in, r);
        if (!S.recs[r].key) return -1;
    }
    counters[0] = counters[1] = 0;
    uint32_t k = 0;
    for (;;)
    { /* read_next_position, :388-433 */
        if (S.npq == 0)
        {
            uq_fill_buffer(&S); /* initialize_next_interval, sequential branch :476-489 */
            if (S.npq == 0) break;
        }
        uq_rec* variant = S.pq[0];
        uq_rec* cvariant = variant;
        uint32_t ncrecs = 0;
        while (uq_cmp3(cvariant, variant) == 0)
        {
            uq_rec* b = S.pq[0];
            order[k] = b->idx;
            first[k] = ncrecs == 0;
            ++k;
            ++ncrecs;
            uint32_t j = 0; /* buffer.remove(b->v) */
            while (j < S.nbuf && S.buf[j] != b->idx) ++j;
            if (j < S.nbuf)
            {
                memmove(S.buf + j, S.buf + j + 1, sizeof(uint32_t) * (S.nbuf - j - 1));
                S.nbuf--;
            }
            uq_fill_buffer(&S);
            uq_pq_pop(&S);
            if (S.npq == 0) break;
            cvariant = S.pq[0];
        }
        counters[1]++;          /* ++no_unique_variants */
        counters[0] += ncrecs;  /* no_total_variants += crecs.size() */
        if (S.bad >= 0) break;
    }
    *not_in_order = S.bad;
    for (uint32_t r = 0; r < in->n_rec; ++r) free(S.recs[r].key);
    free(S.recs);
    free(S.buf);
    free(S.pq);
    return 0;
}

/* ---------------------------------------------------------------------------------------------------------
 * BCFOrderedWriter (bcf_ordered_writer.cpp:141-210 write, :252-300 flush), restated literally with the buffer as a
 * list whose front holds the largest position.  order[k] = arrival index of the k-th record written;
 * late_front[v] = position of the buffer's front when record v triggered the "Might not be sorted" warning (:163-179),
 * INT32_MIN otherwise.  window == 0 writes straight through (:193-203).
 * --------------------------------------------------------------------------------------------------------- */
int32_t vto_writer_order(uint32_t n, const int32_t* rid, const int32_t* pos1, int32_t window, uint32_t* order, int32_t* late_front)
{
    uint32_t* buf = (uint32_t*)malloc(sizeof(uint32_t) * ((size_t)n + 1)); /* buf[0] = front */
    uint32_t nb = 0, k = 0;
    if (!buf) return -1;
    for (uint32_t v = 0; v < n; ++v)
    {
        late_front[v] = INT32_MIN;
        if (!window)
        {
            order[k++] = v;
            continue;
        }
        if (nb == 0)
        {
            buf[nb++] = v; /* push_front on an empty list */
            continue;
        }
        if (rid[v] != rid[buf[nb - 1]])
        { /* flush(true) then push_front */
            while (nb) order[k++] = buf[--nb];
            buf[nb++] = v;
            continue;
        }
        uint32_t i = 0;
        while (i < nb && !(pos1[v] >= pos1[buf[i]])) ++i; /* the equal sign keeps records in original order */
        if (i == nb)
        {
            int32_t cutoff = pos1[buf[0]] - window;
            if (cutoff < 1) cutoff = 1;
            if (pos1[v] < cutoff) late_front[v] = pos1[buf[0]];
        }
        memmove(buf + i + 1, buf + i, sizeof(uint32_t) * (nb - i));
        buf[i] = v;
        ++nb;
        if (nb > 1)
        { /* flush(false) */
            int32_t cutoff = pos1[buf[0]] - window;
            if (cutoff < 1) cutoff = 1;
            while (nb > 1 && pos1[buf[nb - 1]] <= cutoff) order[k++] = buf[--nb];
        }
    }
    while (nb) order[k++] = buf[--nb]; /* close(): flush(true) */
    free(buf);
    return 0;
}

/* ---------------------------------------------------------------------------------------------------------
 * VCF text -> site columns: what the reference gets from htslib for every record line before the hot path
 * sees it: vcf_parse1 (lib/htslib/vcf.c) splits the line on tabs, CHROM = column 1, POS = column 2 as a decimal integer,
 * REF = column 4, ALT = column 5 split on ',' unless it is "." (then the record has the REF allele only; vcf.c fills
 * d.allele[] from "REF,ALT"), and bcf_unpack(BCF_UN_STR) exposes them as the allele strings the path reads
 * (bcf_get_allele, normalize.cpp:196, decompose.cpp:205).  Restated with memchr / strtol, one line at a time.
 * text = record lines only (no header), '\n'-terminated (a missing final '\n' is accepted; a '\r' before it is stripped).
 * status[r] = 0, or 1 when the line has fewer than 8 columns (vcf.c would report a truncated record).
 * Returns the number of records, or -1 when a capacity is too small.
 * --------------------------------------------------------------------------------------------------------- */
int64_t vto_vcf_tokenize(const char* text, uint64_t n_text, uint64_t cap_rec, uint64_t cap_allele, uint64_t cap_bytes,
                         uint64_t* line_off, uint32_t* chrom_len, int32_t* pos1, uint32_t* tail_off, uint8_t* status,
                         uint32_t* allele_beg, uint32_t* allele_off, uint8_t* bytes)
{
    uint64_t r = 0, na = 0, nb = 0, o = 0;
    allele_beg[0] = 0;
    allele_off[0] = 0;
    while (o < n_text)
    {
        const char* nl = (const char*)memchr(text + o, '\n', n_text - o);
        uint64_t next = nl ? (uint64_t)(nl - text) + 1 : n_text;
        uint64_t end = nl ? next - 1 : n_text; /* without the '\n' */
        if (end > o && text[end - 1] == '\r') --end;
        if (r >= cap_rec) return -1;
        line_off[r] = o;
        /* the first seven tabs */
        uint64_t tab[7];
        int nt = 0;
        for (uint64_t i = o; i < end && nt < 7; ++i)
            if (text[i] == '\t') tab[nt++] = i;
        status[r] = nt < 7 ? 1 : 0;
        chrom_len[r] = 0;
        pos1[r] = 0;
        tail_off[r] = 0;
        if (nt >= 7)
        {
            chrom_len[r] = (uint32_t)(tab[0] - o);
            pos1[r] = (int32_t)strtol(text + tab[0] + 1, NULL, 10);
            tail_off[r] = (uint32_t)(tab[4] + 1 - o); /* QUAL column: everything from here on passes through untouched */
            /* REF */
            uint64_t len = tab[3] - (tab[2] + 1);
            if (na + 1 > cap_allele || nb + len > cap_bytes) return -1;
            memcpy(bytes + nb, text + tab[2] + 1, len);
            nb += len;
            allele_off[++na] = (uint32_t)nb;
            /* ALT */
            uint64_t a = tab[3] + 1, e = tab[4];
            if (!(e - a == 1 && text[a] == '.'))
            {
                while (a <= e)
                {
                    uint64_t c = a;
                    while (c < e && text[c] != ',') ++c;
                    if (na + 1 > cap_allele || nb + (c - a) > cap_bytes) return -1;
                    memcpy(bytes + nb, text + a, c - a);
                    nb += c - a;
                    allele_off[++na] = (uint32_t)nb;
                    a = c + 1;
                }
            }
        }
        allele_beg[++r] = (uint32_t)na;
        o = next;
    }
    line_off[r] = n_text;
    return (int64_t)r;
}

/* ---------------------------------------------------------------------------------------------------------
 * vt decompose_blocksub, record loop of Igor::decompose_blocksub(), decompose_blocksub.cpp:185-421, without
 * the alignment mode (-a, :198-303, needs needle.cpp: not restated).  A biallelic record whose REF and ALT have the same
 * length > 1 (:304) is replaced by one SNP per differing position (:341-375) or, with keep_mnv (-m), by one record per run
 * of differences that are at most max_mnv_dist apart (:322-337, :382-401, :405-410; the kstring buffers of the reference
 * are restated as such, with the truncation at the last difference).  Everything else passes through (:413-418).
 * Output k: src_rec[k], off[k], len[k]: new POS = pos1 + off, alleles REF[off, off+len) / ALT[off, off+len);
 * len 0 = the record itself, unchanged.
 * counters: no_variants, no_biallelic_blocksub, no_additional_snps, new_no_variants (:157-160, :377-378, :412, :417, :420;
 * the last two count DIFFERENCES also in -m mode, like the reference).
 * --------------------------------------------------------------------------------------------------------- */
int32_t vto_blocksub_plan(const vto_batch* in, int32_t keep_mnv, int32_t max_mnv_dist, uint32_t* src_rec, uint32_t* off, uint32_t* len,
                          uint64_t* n_out, uint32_t counters[4])
{
    uint64_t k = 0;
    counters[0] = counters[1] = counters[2] = counters[3] = 0;
    for (uint32_t r = 0; r < in->n_rec; ++r)
    {
        const uint32_t ab = in->allele_beg[r], n_allele = in->allele_beg[r + 1] - ab;
        const uint32_t o0 = in->allele_off[ab], o1 = in->allele_off[ab + 1];
        const size_t ref_len = o1 - o0;
        const uint8_t* ref = in->bytes + o0;
        if (n_allele == 2 && ref_len != 1 && ref_len == (size_t)(in->allele_off[ab + 2] - o1))
        {
            const uint8_t* alt = in->bytes + o1;
            int prev_snv_pos = -1, mnv_start_pos = -1;
            char* ref_mnv = (char*)malloc(ref_len + 1);
            char* alt_mnv = (char*)malloc(ref_len + 1);
            size_t l_mnv = 0; /* ref_mnv.l == alt_mnv.l throughout */
            if (!ref_mnv || !alt_mnv) return -1;
            for (size_t i = 0; i < ref_len; ++i)
            {
                if (ref[i] != alt[i])
                {
                    if (keep_mnv)
                    {
                        if (prev_snv_pos == -1 || i - (size_t)prev_snv_pos <= (size_t)max_mnv_dist)
                        {
                            prev_snv_pos = (int)i;
                            if (mnv_start_pos == -1) mnv_start_pos = (int)i;
                            ref_mnv[l_mnv] = (char)ref[i];
                            alt_mnv[l_mnv] = (char)alt[i];
                            ++l_mnv;
                        }
                    }
                    else
                    {
                        src_rec[k] = r;
                        off[k] = (uint32_t)i;
                        len[k] = 1;
                        ++k;
                    }
                    counters[3]++; /* ++new_no_variants */
                    counters[2]++; /* ++no_additional_snps */
                }
                else if (keep_mnv)
                {
                    if (prev_snv_pos != -1 && i - (size_t)prev_snv_pos < (size_t)max_mnv_dist)
                    {
                        ref_mnv[l_mnv] = (char)ref[i];
                        alt_mnv[l_mnv] = (char)alt[i];
                        ++l_mnv;
                    }
                    else if (prev_snv_pos != -1)
                    { /* print_decomposed_var(v, pos1 + mnv_start_pos, ref_mnv.s, alt_mnv.s) after the truncation */
                        src_rec[k] = r;
                        off[k] = (uint32_t)mnv_start_pos;
                        len[k] = (uint32_t)(prev_snv_pos - mnv_start_pos + 1);
                        /* the buffers hold exactly REF/ALT[mnv_start_pos ...]: check the restated kstrings against the substring form */
                        if (memcmp(ref_mnv, ref + mnv_start_pos, len[k]) || memcmp(alt_mnv, alt + mnv_start_pos, len[k])) return -2;
                        ++k;
                        prev_snv_pos = -1;
                        mnv_start_pos = -1;
                        l_mnv = 0;
                    }
                }
            }
            if (keep_mnv && prev_snv_pos != -1)
            {
                src_rec[k] = r;
                off[k] = (uint32_t)mnv_start_pos;
                len[k] = (uint32_t)(prev_snv_pos - mnv_start_pos + 1);
                if (memcmp(ref_mnv, ref + mnv_start_pos, len[k]) || memcmp(alt_mnv, alt + mnv_start_pos, len[k])) return -2;
                ++k;
            }
            free(ref_mnv);
            free(alt_mnv);
            counters[1]++; /* ++no_biallelic_blocksub */
        }
        else
        {
            src_rec[k] = r;
            off[k] = 0;
            len[k] = 0;
            ++k;
            counters[3]++; /* ++new_no_variants */
        }
        counters[0]++; /* ++no_variants */
    }
    *n_out = k;
    return 0;
}

/* ---------------------------------------------------------------------------------------------------------
 * vt decompose_blocksub -a, the aggressive branch (decompose_blocksub.cpp:190-303) with NeedlemanWunsch::align and trace_path
 * (needle.cpp:36-133; NWParameters match 0, mismatch -1, gap -1, needle.h:37), restated in the reference's own order: fill the
 * traceback matrix row by row with one rolling score row, walk it back from the last cell, reverse, force-align the first bases,
 * then cut the trace FORWARD into chunks (:209-246).  For record r the mode applies when ref_len != alt_len, ref_len != 1 and
 * alt_len > 1, alt being ALT 1 only (:188-190); its chunks are written at chunk_beg[r] .. + n_chunk[r], chunk_beg laid out with
 * min(ref_len, alt_len) slots per such record.  Pinned to the reference's fixture test/decompose_blocksub/02_* and to
 * oracle/_ref/vt_ref runs with -a (tests/test_blocksub_oracle.py).
 * --------------------------------------------------------------------------------------------------------- */
int32_t vto_blocksub_align(const vto_batch* in, uint32_t* chunk_beg, uint32_t* n_chunk, uint32_t* pos_ref, uint32_t* pos_alt, uint32_t* len_ref,
                           uint32_t* len_alt)
{
    enum { EMPTY, CIGAR_D, CIGAR_M, CIGAR_X, CIGAR_I };
    uint32_t out = 0;
    for (uint32_t r = 0; r < in->n_rec; ++r)
    {
        const uint32_t ab = in->allele_beg[r], n_allele = in->allele_beg[r + 1] - ab;
        const size_t ref_len = in->allele_off[ab + 1] - in->allele_off[ab];
        const size_t alt_len = n_allele == 1 ? 0 : in->allele_off[ab + 2] - in->allele_off[ab + 1];
        chunk_beg[r] = out;
        n_chunk[r] = 0;
        if (!(ref_len != alt_len && ref_len != 1 && alt_len > 1)) continue;
        const uint8_t* a0 = in->bytes + in->allele_off[ab];
        const uint8_t* a1 = in->bytes + in->allele_off[ab + 1];
        const uint8_t *ref = a0 + 1, *read = a1 + 1;
        const int lr = (int)ref_len - 1, lq = (int)alt_len - 1;
        uint8_t* matrix = (uint8_t*)calloc((size_t)(lr + 1) * (lq + 1), 1);
        int* scores = (int*)malloc(sizeof(int) * (size_t)(lq + 1));
        uint8_t* trace = (uint8_t*)malloc((size_t)lr + lq + 2);
        for (int i = 0; i <= lr; ++i) matrix[(size_t)i * (lq + 1)] = CIGAR_D;
        for (int i = 0; i <= lq; ++i)
        {
            matrix[i] = CIGAR_I;
            scores[i] = i * -1;
        }
        matrix[0] = CIGAR_M;
        size_t offset = 0;
        int score_n = 0;
        for (int i = 1; i <= lr; ++i)
        {
            offset += (size_t)lq + 1;
            score_n += -1;
            int best_score = 0;
            for (int j = 1; j <= lq; ++j)
            {
                int score_nw = scores[j - 1], score_w = scores[j], score_diag = score_nw, best_dir;
                if (ref[i - 1] == read[j - 1])
                    best_dir = CIGAR_M;
                else
                {
                    best_dir = CIGAR_X;
                    score_diag += -1;
                }
                best_score = score_diag;
                if (score_w + -1 > best_score)
                {
                    best_score = score_w + -1;
                    best_dir = CIGAR_D;
                }
                else if (score_n + -1 > best_score)
                {
                    best_score = score_n + -1;
                    best_dir = CIGAR_I;
                }
                scores[j - 1] = score_n;
                score_n = best_score;
                matrix[offset + j] = (uint8_t)best_dir;
            }
            scores[lq] = best_score;
        }
        /* trace_path */
        int i = lr, j = lq;
        long k = (long)(lq + 1) * (lr + 1) - 1;
        size_t nt = 0;
        while (i > 0 || j > 0)
        {
            trace[nt++] = matrix[k];
            switch (matrix[k])
            {
            case CIGAR_X:
            case CIGAR_M: --i; --j; k -= lq + 2; break;
            case CIGAR_I: --j; --k; break;
            case CIGAR_D: --i; k -= lq + 1; break;
            }
        }
        trace[nt++] = a0[0] == a1[0] ? CIGAR_M : CIGAR_X; /* force-aligned first characters, inserted at the front */
        for (size_t a = 0, b = nt - 1; a < b; ++a, --b)
        {
            uint8_t t = trace[a];
            trace[a] = trace[b];
            trace[b] = t;
        }
        /* break apart the alignment */
        int has_error = 0, p_ref = 0, p_alt = 0;
        size_t kk = 0;
        uint32_t c_pr = 0, c_pa = 0, c_lr = 0, c_la = 0, n = 0;
        while (p_ref <= lr || p_alt <= lq)
        {
            switch (trace[kk++])
            {
            case CIGAR_M:
            case CIGAR_X:
                if (has_error)
                {
                    pos_ref[out + n] = c_pr, pos_alt[out + n] = c_pa, len_ref[out + n] = c_lr, len_alt[out + n] = c_la;
                    ++n;
                }
                c_pr = (uint32_t)p_ref++, c_pa = (uint32_t)p_alt++, c_lr = 1, c_la = 1;
                has_error = trace[kk - 1] == CIGAR_X;
                break;
            case CIGAR_D: c_lr++; p_ref++; has_error = 1; break;
            case CIGAR_I: c_la++; p_alt++; has_error = 1; break;
            }
        }
        if (has_error)
        {
            pos_ref[out + n] = c_pr, pos_alt[out + n] = c_pa, len_ref[out + n] = c_lr, len_alt[out + n] = c_la;
            ++n;
        }
        n_chunk[r] = n;
        out += (uint32_t)(ref_len < alt_len ? ref_len : alt_len);
        free(matrix);
        free(scores);
        free(trace);
    }
    chunk_beg[in->n_rec] = out;
    return 0;
}
