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
