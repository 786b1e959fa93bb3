// vtn_annot_host.h -- the host-side control logic of vtn_annotate_indels (see vtn_annot.cu for what runs where), written against
// an abstract backend so that the same code is checked on the CPU against the unmodified reference (tests/annot_host_main.cpp).
#pragma once
#include <math.h>
#include <stdio.h>
#include <string.h>
#include <algorithm>
#include <map>
#include <queue>
#include <string>
#include <vector>
#include "../../include/vtn.h"

#define VTA_KEEP 1u       // the reference annotates / writes the record (normalized, no N: annotate_indels.cpp:306-323)
#define VTA_INDEL 2u      // kept and vtype & VT_INDEL: VNTRAnnotator::annotate runs (annotate_indels.cpp:330-334)
#define VTA_NOCONTIG 4u   // an indel on a contig the reference does not have (the reference exits in fetch_base)
#define VTA_RUNAWAY 8u    // right_align did not stop inside the contig (REF == ALT after trimming; the reference loops forever)

namespace vta
{

// ---------------------------------------------------------------------------------------------------------
// host: VNTR string helpers (vntr.cpp)
// ---------------------------------------------------------------------------------------------------------
std::string shift_str(const std::string& s, size_t i) { return i ? s.substr(i) + s.substr(0, i) : s; } // vntr.cpp:283-292

char complement(char b) // vntr.h:163 "TGNCNNNNNA"[(b-65)>>1]
{
    const int k = ((int)b - 65) >> 1;
    return (k >= 0 && k < 10) ? "TGNCNNNNNA"[k] : 'N';
}

std::string reverse_complement(const std::string& s) // vntr.cpp:124-134
{
    std::string r;
    for (size_t i = s.size(); i-- > 0;) r.push_back(complement(s[i]));
    return r;
}

std::string canonicalize2(const std::string& motif) // vntr.cpp:140-161
{
    std::string c = motif;
    for (size_t i = 0; i < motif.size(); ++i)
    {
        const std::string sm = shift_str(motif, i), rc = reverse_complement(sm);
        if (sm < c) c = sm;
        if (rc < c) c = rc;
    }
    return c;
}

std::string get_basis(const std::string& motif) // vntr.cpp:238-278
{
    bool b[4] = {false, false, false, false};
    for (char ch : motif)
    {
        if (ch == 'A') b[0] = true;
        else if (ch == 'C') b[1] = true;
        else if (ch == 'G') b[2] = true;
        else if (ch == 'T') b[3] = true;
    }
    std::string s;
    for (int i = 0; i < 4; ++i)
        if (b[i]) s.push_back("ACGT"[i]);
    return s;
}

int is_periodic(const std::string& motif) // vntr.cpp:189-202
{
    for (size_t i = 1; i + 1 <= motif.size(); ++i)
        if (shift_str(motif, i) == motif) return (int)i;
    return 0;
}

// ---------------------------------------------------------------------------------------------------------
// host: MotifTree / MotifMap (motif_tree.cpp, motif_map.cpp).  The k-mer census of a tract: every k-mer (k <= cmax_len) at
// every position, consolidated by canonical rotation; candidates ordered by the reference's priority queue.
// ---------------------------------------------------------------------------------------------------------
struct CandidateMotif
{
    std::string motif;
    float score;
    uint32_t len;
    float fit;
};
struct CompareCandidateMotif // motif_tree.h:80-95
{
    bool operator()(const CandidateMotif& a, const CandidateMotif& b) const { return a.score != b.score ? a.score < b.score : a.len > b.len; }
};
typedef std::priority_queue<CandidateMotif, std::vector<CandidateMotif>, CompareCandidateMotif> MotifQueue;

struct MotifCensus
{
    static const uint32_t MAX_LEN = 10; // CandidateMotifPicker::max_mlen (candidate_motif_picker.cpp:31)
    uint32_t len_count[MAX_LEN + 1];
    std::vector<uint32_t> count;   // per node index; all zero between calls
    std::vector<uint32_t> touched;

    MotifCensus()
    {
        len_count[0] = 0;
        for (uint32_t i = 1; i <= MAX_LEN; ++i) len_count[i] = len_count[i - 1] + (1u << (2 * i)); // motif_map.cpp:47-52
        count.assign(len_count[MAX_LEN], 0);
    }
    // 2 bits per base, base i at bits 31-2i .. 30-2i (motif_map.h:31-34)
    static uint32_t get_i(uint32_t s, uint32_t i) { return (s >> ((15 - i) << 1)) & 3u; }
    static uint32_t set_i(uint32_t s, uint32_t i, uint32_t b) { return (b << ((15 - i) << 1)) | (s & ~(3u << ((15 - i) << 1))); }
    static uint32_t rot(uint32_t s, uint32_t l) { return ((s >> 30) << ((16 - l) << 1)) | (s << 2); }
    uint32_t seq2index(uint32_t s, uint32_t len) const // motif_map.cpp:148-157
    {
        uint32_t index = len_count[len - 1];
        for (uint32_t i = 0; i < len; ++i) index += get_i(s, i) << ((len - i - 1) << 1);
        return index;
    }
    uint32_t index2seq(uint32_t index, uint32_t& len) const // motif_map.cpp:123-143
    {
        for (len = 1; len <= MAX_LEN; ++len)
            if (index < len_count[len])
            {
                index -= len_count[len - 1];
                uint32_t s = 0;
                for (int32_t i = (int32_t)len; i > 0; --i) s = set_i(s, len - i, (index >> ((i - 1) << 1)) & 3u);
                return s;
            }
        return 0;
    }
    static uint32_t canonical(uint32_t seq, uint32_t len) // motif_map.cpp:86-101
    {
        if (len == 1) return seq;
        uint32_t c = seq;
        for (uint32_t i = 1; i < len; ++i)
        {
            seq = rot(seq, len);
            c = seq < c ? seq : c;
        }
        return c;
    }
    static bool is_aperiodic(uint32_t motif, uint32_t len) // motif_map.cpp:106-118
    {
        for (uint32_t i = 1; i <= (len >> 1); ++i)
            if (len % i == 0)
            {
                uint32_t n = 0;
                for (uint32_t j = 0; j < len / i; ++j)
                    for (uint32_t k = 0; k < i; ++k) n = set_i(n, i * j + k, get_i(motif, k));
                if (n == motif) return false;
            }
        return true;
    }
    static std::string seq2str(uint32_t seq, uint32_t len)
    {
        std::string s;
        for (uint32_t i = 0; i < len; ++i) s.push_back("ACGT"[get_i(seq, i)]);
        return s;
    }
    static uint32_t base2index(char b) { return b == 'A' ? 0u : b == 'C' ? 1u : b == 'G' ? 2u : b == 'T' ? 3u : 4u; } // motif_tree.cpp:372-392

    // MotifTree::detect_candidate_motifs (motif_tree.cpp:278-367)
    void detect(const std::string& seq, MotifQueue& pcm)
    {
        while (!pcm.empty()) pcm.pop();
        const uint32_t len0 = (uint32_t)seq.size();
        std::string s;
        if (seq.find('N') != std::string::npos)
        {
            for (char ch : seq)
                if (ch != 'N') s.push_back(ch);
        }
        else
            s = seq;
        const uint32_t len = (uint32_t)s.size();
        const uint32_t cmax = (len >> 1) < MAX_LEN ? (len >> 1) : MAX_LEN; // set_sequence, motif_tree.cpp:100-125
        touched.clear();
        uint32_t w = 0;
        for (uint32_t i = 0; i < len; ++i)
        {
            uint32_t l;
            if (i == 0) // get_sub_seq, motif_tree.cpp:64-95
            {
                w = 0;
                for (uint32_t j = 0; j < cmax; ++j) w = set_i(w, j, base2index(s[j]));
                l = cmax;
            }
            else if (len - i < cmax)
            {
                l = len - i;
                w = rot(w, l + 1);
                w = set_i(w, len - i, 0);
            }
            else
            {
                l = cmax;
                w = rot(w, cmax);
                w = set_i(w, cmax - 1, base2index(s[i + cmax - 1]));
            }
            for (uint32_t k = 1; k <= l; ++k) // insert_prefix, :130-137
            {
                const uint32_t idx = seq2index(w, k);
                if (count[idx]++ == 0) touched.push_back(idx);
            }
        }
        uint32_t base[4] = {count[0], count[1], count[2], count[3]}; // scontent, :304-310
        const uint32_t nb = base[0] + base[1] + base[2] + base[3];
        std::map<uint32_t, uint32_t> cm; // consolidate_motif_counts, :142-184
        uint32_t lc[MAX_LEN + 1] = {0};
        for (uint32_t idx : touched)
        {
            uint32_t l;
            const uint32_t q = index2seq(idx, l);
            cm[seq2index(canonical(q, l), l)] += count[idx];
            lc[l] += count[idx];
            count[idx] = 0;
        }
        auto fit_of = [&](uint32_t q, uint32_t l) { // compute_fit, :236-268
            uint32_t e[4] = {0, 0, 0, 0};
            for (uint32_t i = 0; i < l; ++i) ++e[get_i(q, i)];
            float fit = 0, t = 0;
            for (uint32_t i = 0; i < 4; ++i)
            {
                t = (float)base[i] / nb - (float)e[i] / l;
                fit += t * t;
            }
            return fit;
        };
        for (auto it = cm.begin(); it != cm.end(); ++it)
        {
            uint32_t l;
            const uint32_t q = index2seq(it->first, l);
            if (!is_aperiodic(q, l)) continue;
            const float p = (float)it->second / lc[l];
            const float f = fit_of(q, l);
            if (len0 < 10 || (l == 1 && p > 0.6) || l > 1) pcm.push(CandidateMotif{seq2str(q, l), p, l, f});
        }
        if (pcm.empty() && !cm.empty())
        {
            uint32_t l;
            const uint32_t q = index2seq(cm.begin()->first, l);
            float p = (float)cm.begin()->second / lc[l];
            const float f = fit_of(q, l);
            p -= f;
            pcm.push(CandidateMotif{seq2str(q, l), p, l, f});
        }
    }
};

// ---------------------------------------------------------------------------------------------------------
// host: per-indel state (the fields of VNTR that annotate_indels writes, vntr.h)
// ---------------------------------------------------------------------------------------------------------
struct Side
{
    std::string motif, ru, basis, tract;
    int32_t mlen = 0, blen = 0, beg1 = 0, end1 = 0, comp[4] = {0, 0, 0, 0}, rl = 0, ll = 0, no_perfect_ru = 0, no_ru = 0, trf_score = 0;
    float entropy = 0, entropy2 = 0, kl = 0, kl2 = 0, ref = 0, score = 0;
};
struct Indel
{
    uint32_t rec = 0;
    int32_t rid = 0, max_dlen = 0;
    std::vector<std::string> alleles;
    Side ex, fz;
    bool ambiguous = false, large = false;
    // fuzzy detection state (flank_detector.cpp:188-281)
    int32_t slen = 100, lflank_end1 = 0, rflank_beg1 = 0;
    std::string rflank, lflank, win; // the five flank bases and the widest window of the current direction
    bool open = false;
};

// CandidateMotifPicker::get_indel (candidate_motif_picker.cpp:153-203)
bool get_indel(std::string ref, std::string alt, std::string& indel, bool& fatal)
{
    while (!ref.empty() && !alt.empty() && ref.back() == alt.back()) { ref.pop_back(); alt.pop_back(); }
    while (!ref.empty() && !alt.empty() && ref[0] == alt[0]) { ref.erase(0, 1); alt.erase(0, 1); }
    if (ref.empty()) { indel = alt; return true; }
    if (alt.empty()) { indel = ref; return true; }
    if (ref.size() > alt.size()) indel = ref.substr(alt.size(), ref.size() - alt.size());
    else if (ref.size() < alt.size()) indel = alt.substr(ref.size(), alt.size() - ref.size());
    else fatal = true; // "Not an indel!!!!": the reference exits
    return false;
}

// CandidateMotifPicker::update_exact_repeat_unit (candidate_motif_picker.cpp:211-324)
bool update_exact_repeat_unit(Indel& d)
{
    std::string indel, ru;
    bool fatal = false;
    auto unit_of = [&](bool simple) {
        if (simple)
        {
            const int i = is_periodic(indel);
            return i ? indel.substr(0, i) : indel;
        }
        return indel;
    };
    if (d.alleles.size() == 2)
    {
        const bool simple = get_indel(d.alleles[0], d.alleles[1], indel, fatal);
        ru = unit_of(simple);
        d.ambiguous = !simple;
    }
    else
    {
        bool all_simple = true;
        std::map<std::string, int32_t> indels;
        for (size_t i = 1; i < d.alleles.size(); ++i)
        {
            if (d.alleles[0].size() == d.alleles[i].size()) continue;
            const bool simple = get_indel(d.alleles[0], d.alleles[i], indel, fatal);
            all_simple = all_simple && simple;
            ++indels[unit_of(simple)];
        }
        int32_t best = 0;
        for (auto& kv : indels)
            if (kv.second > best)
            {
                ru = kv.first;
                best = kv.second;
            }
        d.ambiguous = !all_simple;
    }
    d.ex.ru = ru;
    d.ex.motif = canonicalize2(ru);
    d.ex.basis = get_basis(d.ex.motif);
    d.ex.mlen = (int32_t)d.ex.motif.size();
    d.ex.blen = (int32_t)d.ex.basis.size();
    return !fatal;
}

// CandidateMotifPicker::is_in_indel_fragment (candidate_motif_picker.cpp:452-472)
bool in_indel_fragment(const Indel& d, const std::string& motif)
{
    for (size_t i = 0; i < motif.size(); ++i)
    {
        const std::string sm = motif.substr(i) + motif.substr(0, i);
        for (const std::string& a : d.alleles)
            if (a.find(sm) != std::string::npos) return true;
    }
    return false;
}

// FlankDetector::compute_score / choose_fuzzy_{3,5}prime_repeat_unit (flank_detector.cpp:303-332, :382-417, :449-484)
int32_t phase_score(size_t start, size_t len, const std::string& a, const std::string& b, bool& bad)
{
    if (len > b.size()) len = b.size();
    int32_t score = 0;
    for (size_t j = 0; j < len; ++j)
    {
        if (start + j >= a.size()) { bad = true; return score; } // std::string::at would throw in the reference
        score += a[start + j] == b[j] ? 2 : -7;
    }
    return score;
}
std::string choose_fuzzy_unit(const std::string& seq, const std::string& motif, bool three_prime, bool& bad)
{
    int32_t best_score = -10000;
    std::string best = motif;
    const int32_t mx = 2 * (int32_t)motif.size();
    const size_t mlen = motif.size();
    if (three_prime && seq.size() < mlen) { bad = true; return best; }
    const size_t start = three_prime ? seq.size() - mlen : 0;
    for (size_t i = 0; i < mlen; ++i)
    {
        const std::string sm = shift_str(motif, i);
        const int32_t s1 = phase_score(start, mlen, seq, sm, bad);
        if (bad) return best;
        if (s1 == mx) return sm;
        if (s1 > best_score) { best_score = s1; best = sm; }
        const std::string rc = reverse_complement(sm);
        const int32_t s2 = phase_score(start, mlen, seq, rc, bad);
        if (s2 == mx) return rc;
        if (s2 > best_score) { best_score = s2; best = rc; }
    }
    return best;
}
std::string choose_exact_unit(const std::string& seq, const std::string& motif) // flank_detector.cpp:491-507
{
    for (size_t i = 0; i < motif.size(); ++i)
    {
        const std::string sm = shift_str(motif, i);
        if (seq.compare(0, sm.size(), sm) == 0) return sm;
    }
    return "";
}

// FlankDetector::compute_composition_and_entropy (flank_detector.cpp:771-887); float arithmetic and libm's log2f as there
void composition_and_entropy(Side& v)
{
    const std::string& t = v.tract;
    int32_t c1[4] = {0, 0, 0, 0}, c2[16] = {0};
    static const int32_t b2i[10] = {0, 1, 0, 2, 0, 0, 0, 0, 0, 3};
    const int32_t n = (int32_t)t.size();
    auto code = [&](char ch) {
        const int k = ((int)ch - 65) >> 1;
        return (uint32_t)((k >= 0 && k < 10) ? b2i[k] : 0);
    };
    for (int32_t i = 0; i < n; ++i)
    {
        const uint32_t b1 = code(t[i]);
        ++c1[b1];
        if (i < n - 1) ++c2[(b1 << 2) + code(t[i + 1])];
    }
    float p[4] = {(float)c1[0] / n, (float)c1[1] / n, (float)c1[2] / n, (float)c1[3] / n};
    float entropy = 0;
    for (int i = 0; i < 4; ++i)
        if (p[i]) entropy += p[i] * std::log2(p[i]);
    float kl = p[0] + p[1] + p[2] + p[3];
    kl = entropy + 2 * kl;
    kl = std::round(100 * kl) / 100;
    kl = (kl == 0 ? 0 : kl);
    entropy = -entropy;
    entropy = std::round(100 * entropy) / 100;
    entropy = (entropy == 0 ? 0 : entropy);
    for (int i = 0; i < 4; ++i) v.comp[i] = (int32_t)std::round(p[i] * 100);
    float entropy2 = 0, kl2 = 0;
    if (n != 1)
    {
        float p2[16];
        for (int i = 0; i < 16; ++i) p2[i] = (float)c2[i] / (n - 1);
        for (int i = 0; i < 16; ++i)
            if (p2[i])
            {
                entropy2 += p2[i] * std::log2(p2[i]);
                kl2 += p2[i];
            }
        kl2 = entropy2 + 4 * kl2;
        kl2 = std::round(100 * kl2) / 100;
        kl2 = (kl2 == 0 ? 0 : kl2);
        entropy2 = -entropy2;
        entropy2 = std::round(100 * entropy2) / 100;
        entropy2 = (entropy2 == 0 ? 0 : entropy2);
    }
    v.entropy = entropy;
    v.entropy2 = entropy2;
    v.kl = kl;
    v.kl2 = kl2;
}

// faidx_adjust_position (lib/htslib/faidx.c:832-860) for 0-based inclusive [beg0, end0]; false = nothing to fetch
bool clip_window(int64_t clen, int64_t& beg0, int64_t& end0)
{
    if (clen <= 0) return false;
    if (end0 < beg0) beg0 = end0;
    beg0 = beg0 < 0 ? 0 : (beg0 >= clen ? clen - 1 : beg0);
    end0 = end0 < 0 ? 0 : (end0 >= clen ? clen - 1 : end0);
    return true;
}

struct FetchList
{
    std::vector<int32_t> rid;
    std::vector<int64_t> beg0;
    std::vector<uint64_t> off{0};
    // ReferenceSequence::fetch_seq(chrom, beg1, end1); clen = length of contig r, <= 0 when the reference does not have it
    void add(int32_t r, int64_t clen, int64_t beg1, int64_t end1)
    {
        int64_t b = beg1 - 1, e = end1 - 1;
        uint64_t len = 0;
        if (clip_window(clen, b, e)) len = (uint64_t)(e - b + 1);
        rid.push_back(len ? r : 0);
        beg0.push_back(len ? b : 0);
        off.push_back(off.back() + len);
    }
    size_t size() const { return rid.size(); }
};

std::string window_of(const FetchList& L, const std::vector<uint8_t>& buf, size_t k)
{
    return std::string((const char*)buf.data() + L.off[k], (size_t)(L.off[k + 1] - L.off[k]));
}

// ---- batched HMM calls ----
struct HmmJob
{
    std::vector<uint8_t> kind;
    std::vector<uint32_t> seq_off{0};
    std::string bytes;
    void add(int k, const std::string& motif, const std::string& flank, const std::string& read)
    {
        kind.push_back((uint8_t)k);
        bytes += motif; seq_off.push_back((uint32_t)bytes.size());
        bytes += flank; seq_off.push_back((uint32_t)bytes.size());
        bytes += read; seq_off.push_back((uint32_t)bytes.size());
    }
    size_t size() const { return kind.size(); }
};

const vtn_hmm_params FD_PARAMS = {1e-7f, 1e-7f, 0.01f, 0.01f, 5.0f, 42}; // flank_detector.cpp:44-74

// What the control logic needs from the machine underneath: the device in libvtn (vtn_annot.cu); the CPU oracles in
// tests/annot_host_main.cpp, which checks this logic against the unmodified reference without a GPU.
struct RegionOut
{
    std::vector<uint8_t> flags, vtype;
    std::vector<int32_t> reg_beg, reg_end, max_dlen;
    std::vector<uint32_t> a_lt, a_rt, a_shift;
};
struct AnnotBackend
{
    virtual ~AnnotBackend() {}
    virtual int64_t contig_len(int32_t rid) = 0; // <= 0: unknown
    virtual int regions(const vtn_batch* in, RegionOut& out) = 0;
    virtual int fetch(const FetchList& L, int upper, std::vector<uint8_t>& out) = 0; // out.size() > L.off.back()
    virtual int hmm(const HmmJob& J, std::vector<vtn_hmm_result>& res) = 0;
    virtual int fail(int code, const std::string& msg) = 0;
};

uint32_t fbits(float x)
{
    uint32_t u;
    memcpy(&u, &x, 4);
    return u;
}

// FlankDetector::compute_purity_score(tract, motif) (flank_detector.cpp:646-712): the exact branch here; false = needs AHMM
bool purity_exact(Side& v, const std::string& motif_in, std::string& ru)
{
    ru = choose_exact_unit(v.tract, motif_in);
    if (!ru.empty())
    {
        const size_t mlen = ru.size();
        size_t j = 0;
        bool exact = true;
        for (size_t i = 0; i < v.tract.size(); ++i)
        {
            if (ru[j] != v.tract[i]) { exact = false; break; }
            j = (j == mlen - 1) ? 0 : j + 1;
        }
        if (exact)
        {
            v.score = 1;
            v.no_perfect_ru = (int32_t)(v.tract.size() / motif_in.size());
            v.no_ru = v.no_perfect_ru;
            v.ref = (float)v.tract.size() / motif_in.size();
            v.rl = (int32_t)v.tract.size();
            v.trf_score = (int32_t)(v.tract.size() << 1);
            return true;
        }
    }
    ru = motif_in;
    return false;
}
void purity_from_ahmm(Side& v, const vtn_hmm_result& r)
{
    float score = r.motif_concordance;
    score = std::round(100 * score) / 100;
    v.score = score;
    v.no_perfect_ru = r.exact_motif_count;
    v.no_ru = r.motif_count;
    v.ref = r.frac_no_repeats;
    v.rl = (int32_t)v.tract.size();
    v.trf_score = r.trf_score;
}


// Igor::annotate_indels()'s record loop (annotate_indels.cpp:293-410) over a batch; see vtn_annot.cu for the stages
inline int annotate_core(AnnotBackend& be, const vtn_batch* in, vtn_annot_result* out, const std::string& trace_path)
{
    out->n_indel = 0;
    out->n_str = 0;
    out->n_alignments = 0;
    const uint32_t n = in->n_rec;
    if (!n) return VTN_OK;
    // ---- stage 1: classification, the two record filters, exact left/right alignment of every indel ----
    RegionOut RO;
    int rc = be.regions(in, RO);
    if (rc) return rc;
    const std::vector<uint8_t>&flags = RO.flags, &vtype = RO.vtype;
    const std::vector<int32_t>&reg_beg = RO.reg_beg, &reg_end = RO.reg_end, &max_dlen = RO.max_dlen;
    const std::vector<uint32_t>&a_lt = RO.a_lt, &a_rt = RO.a_rt, &a_shift = RO.a_shift;
    std::vector<Indel> D;
    for (uint32_t r = 0; r < n; ++r)
    {
        out->flags[r] = flags[r] & (VTA_KEEP | VTA_INDEL);
        out->vtype[r] = vtype[r];
        out->indel_idx[r] = 0xffffffffu;
        if (out->shift)
            for (uint32_t sl = in->allele_beg[r]; sl < in->allele_beg[r + 1]; ++sl)
            {
                out->trim_left[sl] = a_lt[sl];
                out->trim_right[sl] = a_rt[sl];
                out->shift[sl] = a_shift[sl];
            }
        if (!(flags[r] & VTA_INDEL)) continue;
        if (flags[r] & VTA_NOCONTIG) return be.fail(VTN_E_ARG, "vtn_annotate_indels: an indel lies on a contig the reference does not have (record " + std::to_string(r) + ")");
        if (flags[r] & VTA_RUNAWAY) return be.fail(VTN_E_ARG, "vtn_annotate_indels: right alignment does not end (REF equals an ALT after trimming; record " + std::to_string(r) + ")");
        Indel d;
        d.rec = r;
        d.rid = in->rid[r];
        d.max_dlen = max_dlen[r];
        for (uint32_t sl = in->allele_beg[r]; sl < in->allele_beg[r + 1]; ++sl)
            d.alleles.emplace_back((const char*)in->bytes + in->allele_off[sl], (size_t)(in->allele_off[sl + 1] - in->allele_off[sl]));
        d.ex.beg1 = reg_beg[r];
        d.ex.end1 = reg_end[r];
        out->indel_idx[r] = (uint32_t)D.size();
        D.push_back(std::move(d));
    }
    const size_t K = D.size();
    FILE* trace = trace_path.empty() ? nullptr : fopen(trace_path.c_str(), "w");
    auto trace_hmm = [&](size_t k, int stage, const char* kind, const std::string& motif, const std::string& flank, const std::string& read, const vtn_hmm_result& r) {
        if (!trace) return;
        if (kind[0] == 'A')
            fprintf(trace, "%zu %d A %s - %s %08x %d %d %08x %d\n", k, stage, motif.c_str(), read.c_str(), fbits(r.motif_concordance), r.exact_motif_count, r.motif_count, fbits(r.frac_no_repeats), r.trf_score);
        else
            fprintf(trace, "%zu %d %s %s %s %s %d\n", k, stage, kind, motif.c_str(), flank.c_str(), read.c_str(), kind[0] == 'R' ? r.pos[3] : r.pos[9]);
    };
    uint64_t n_align = 0;
    std::vector<uint8_t> buf;
    std::vector<vtn_hmm_result> hres;
    // ---- stage 2: exact tract, exact repeat unit, candidate motifs -> fuzzy motif (vntr_annotator.cpp:103-123) ----
    {
        FetchList L;
        for (Indel& d : D) L.add(d.rid, be.contig_len(d.rid), d.ex.beg1, d.ex.end1);
        rc = be.fetch(L, 1, buf);
        if (rc) { if (trace) fclose(trace); return rc; }
        MotifCensus census;
        MotifQueue pcm;
        for (size_t k = 0; k < K; ++k)
        {
            Indel& d = D[k];
            d.ex.tract = window_of(L, buf, k);
            if (!update_exact_repeat_unit(d))
            {
                if (trace) fclose(trace);
                return be.fail(VTN_E_ARG, "vtn_annotate_indels: not an indel (record " + std::to_string(d.rec) + "); the reference exits here");
            }
            // generate_candidate_motifs (candidate_motif_picker.cpp:54-130): the tract with the longest allele spliced in
            size_t longest = 0, longest_len = d.alleles[0].size();
            for (size_t i = 1; i < d.alleles.size(); ++i)
                if (d.alleles[i].size() > longest_len) { longest = i; longest_len = d.alleles[i].size(); }
            std::string seq = d.ex.tract;
            if (longest)
            {
                const int32_t pos1 = in->pos1[d.rec];
                const size_t offset = d.ex.beg1 < pos1 ? (size_t)(pos1 - d.ex.beg1) : 0;
                if (offset <= seq.size()) seq.replace(offset, d.alleles[0].size(), d.alleles[longest]);
            }
            census.detect(seq, pcm);
            // next_motif(CHECK_MOTIF_PRESENCE_IN_ALLELE) (:332-381), else fall back on the exact motif (vntr_annotator.cpp:112-123)
            bool found = false;
            while (!pcm.empty())
            {
                const CandidateMotif cm = pcm.top();
                pcm.pop();
                if (in_indel_fragment(d, cm.motif))
                {
                    d.fz.motif = canonicalize2(cm.motif);
                    d.fz.basis = get_basis(cm.motif);
                    d.fz.mlen = (int32_t)d.fz.motif.size();
                    d.fz.blen = (int32_t)d.fz.basis.size();
                    found = true;
                    break;
                }
            }
            if (!found)
            {
                d.fz.motif = d.ex.motif;
                d.fz.ru = d.ex.ru;
                d.fz.basis = d.ex.basis;
                d.fz.mlen = d.ex.mlen;
                d.fz.blen = d.ex.blen;
            }
            // detect_flanks(EXACT) (flank_detector.cpp:103-170): VNTR::mlen is 0 on this path, so the anchors go when the tract has >= 3 bases
            if (d.ex.tract.size() > 2)
            {
                d.ex.tract = d.ex.tract.substr(1, d.ex.tract.size() - 2);
                ++d.ex.beg1;
                --d.ex.end1;
            }
        }
    }
    // ---- stage 3: purity of the exact tract (AHMM where it is not perfectly periodic), composition / entropy ----
    {
        HmmJob J;
        std::vector<size_t> owner;
        std::vector<std::string> rus(K);
        for (size_t k = 0; k < K; ++k)
        {
            Indel& d = D[k];
            if (!purity_exact(d.ex, d.ex.ru, rus[k]))
            {
                J.add(VTN_HMM_AHMM, rus[k], "", d.ex.tract);
                owner.push_back(k);
            }
        }
        rc = be.hmm(J, hres);
        if (rc) { if (trace) fclose(trace); return rc; }
        n_align += J.size();
        for (size_t j = 0; j < owner.size(); ++j)
        {
            purity_from_ahmm(D[owner[j]].ex, hres[j]);
            trace_hmm(owner[j], 0, "A", rus[owner[j]], "", D[owner[j]].ex.tract, hres[j]);
        }
        for (size_t k = 0; k < K; ++k)
        {
            Indel& d = D[k];
            d.ex.ru = rus[k];
            d.ex.ll = d.ex.rl + d.max_dlen;
            composition_and_entropy(d.ex);
        }
    }
    // ---- stage 4: detect_flanks(FUZZY) (flank_detector.cpp:171-281), every indel in lock step ----
    {
        // RFHMM: rflank = [end1+1, end1+5]; seq = [end1-slen, end1], slen = 100, 200 .. 1000: suffixes of the widest window
        FetchList L;
        for (Indel& d : D)
        {
            L.add(d.rid, be.contig_len(d.rid), (int64_t)d.ex.end1 + 1, (int64_t)d.ex.end1 + 5);
            L.add(d.rid, be.contig_len(d.rid), (int64_t)d.ex.end1 - 1000, (int64_t)d.ex.end1);
        }
        rc = be.fetch(L, 1, buf);
        if (rc) { if (trace) fclose(trace); return rc; }
        std::vector<size_t> todo;
        for (size_t k = 0; k < K; ++k)
        {
            D[k].rflank = window_of(L, buf, 2 * k);
            D[k].win = window_of(L, buf, 2 * k + 1);
            D[k].slen = 100;
            todo.push_back(k);
        }
        while (!todo.empty())
        {
            HmmJob J;
            std::vector<std::string> reads(todo.size());
            for (size_t j = 0; j < todo.size(); ++j)
            {
                Indel& d = D[todo[j]];
                // fetch_seq(end1 - slen, end1): the last slen+1 bases of the window unless the contig start clips it
                int64_t b = (int64_t)d.ex.end1 - d.slen - 1, e = (int64_t)d.ex.end1 - 1;
                std::string seq;
                if (clip_window(be.contig_len(d.rid), b, e))
                {
                    const size_t len = (size_t)(e - b + 1);
                    seq = len <= d.win.size() ? d.win.substr(d.win.size() - len) : d.win;
                }
                bool bad = false;
                d.fz.ru = choose_fuzzy_unit(seq, d.fz.motif, true, bad);
                if (bad)
                {
                    if (trace) fclose(trace);
                    return be.fail(VTN_E_ARG, "vtn_annotate_indels: repeat tract shorter than its motif at a contig start (record " + std::to_string(d.rec) + "); the reference throws here");
                }
                reads[j] = seq + d.rflank;
                J.add(VTN_HMM_RFHMM, d.fz.ru, d.rflank, reads[j]);
            }
            rc = be.hmm(J, hres);
            if (rc) { if (trace) fclose(trace); return rc; }
            n_align += J.size();
            std::vector<size_t> next;
            for (size_t j = 0; j < todo.size(); ++j)
            {
                Indel& d = D[todo[j]];
                trace_hmm(todo[j], 1, "R", d.fz.ru, d.rflank, reads[j], hres[j]);
                if (hres[j].status != VTN_HMM_ST_OK)
                {
                    if (trace) fclose(trace);
                    return be.fail(VTN_E_ARG, "vtn_annotate_indels: RFHMM rejected a model (record " + std::to_string(d.rec) + ", status " + std::to_string(hres[j].status) + ")");
                }
                const int32_t epos1 = hres[j].pos[3]; // get_lflank_read_epos1
                if (epos1 > std::min((int32_t)(10 * d.fz.ru.size()), 50))
                    d.lflank_end1 = d.ex.end1 - d.slen - 1 + 1 + epos1 - 1;
                else if (d.slen == 1000)
                {
                    d.lflank_end1 = d.ex.end1 - 1000 - 1;
                    d.large = true;
                }
                else
                {
                    d.slen += 100;
                    next.push_back(todo[j]);
                }
            }
            todo.swap(next);
        }
        // LFHMM: lflank = [lflank_end1-4, lflank_end1]; seq = [lflank_end1+1, lflank_end1+slen], slen carried over
        FetchList L2;
        for (Indel& d : D)
        {
            L2.add(d.rid, be.contig_len(d.rid), (int64_t)d.lflank_end1 - 5 + 1, (int64_t)d.lflank_end1);
            L2.add(d.rid, be.contig_len(d.rid), (int64_t)d.lflank_end1 + 1, (int64_t)d.lflank_end1 + 1000);
        }
        rc = be.fetch(L2, 1, buf);
        if (rc) { if (trace) fclose(trace); return rc; }
        for (size_t k = 0; k < K; ++k)
        {
            D[k].lflank = window_of(L2, buf, 2 * k);
            D[k].win = window_of(L2, buf, 2 * k + 1);
            todo.push_back(k);
        }
        while (!todo.empty())
        {
            HmmJob J;
            std::vector<std::string> reads(todo.size());
            for (size_t j = 0; j < todo.size(); ++j)
            {
                Indel& d = D[todo[j]];
                int64_t b = (int64_t)d.lflank_end1, e = (int64_t)d.lflank_end1 + d.slen - 1;
                std::string seq;
                if (clip_window(be.contig_len(d.rid), b, e))
                {
                    const size_t len = (size_t)(e - b + 1);
                    seq = len <= d.win.size() ? d.win.substr(0, len) : d.win;
                }
                bool bad = false;
                d.fz.ru = choose_fuzzy_unit(seq, d.fz.motif, false, bad);
                if (bad)
                {
                    if (trace) fclose(trace);
                    return be.fail(VTN_E_ARG, "vtn_annotate_indels: repeat tract shorter than its motif at a contig end (record " + std::to_string(d.rec) + "); the reference throws here");
                }
                reads[j] = d.lflank + seq;
                J.add(VTN_HMM_LFHMM, d.fz.ru, d.lflank, reads[j]);
            }
            rc = be.hmm(J, hres);
            if (rc) { if (trace) fclose(trace); return rc; }
            n_align += J.size();
            std::vector<size_t> next;
            for (size_t j = 0; j < todo.size(); ++j)
            {
                Indel& d = D[todo[j]];
                trace_hmm(todo[j], 2, "L", d.fz.ru, d.lflank, reads[j], hres[j]);
                if (hres[j].status != VTN_HMM_ST_OK)
                {
                    if (trace) fclose(trace);
                    return be.fail(VTN_E_ARG, "vtn_annotate_indels: LFHMM rejected a model (record " + std::to_string(d.rec) + ", status " + std::to_string(hres[j].status) + ")");
                }
                const int32_t spos1 = hres[j].pos[9]; // get_rflank_read_spos1
                if (spos1 < d.slen - std::min((int32_t)(10 * d.fz.ru.size()), 50))
                    d.rflank_beg1 = d.lflank_end1 - 5 + spos1;
                else if (d.slen == 1000)
                {
                    d.rflank_beg1 = d.lflank_end1 + 1000;
                    d.large = true;
                }
                else
                {
                    d.slen += 100;
                    next.push_back(todo[j]);
                }
            }
            todo.swap(next);
        }
    }
    // ---- stage 5: the fuzzy tract, its purity (AHMM) and composition ----
    {
        FetchList L;
        for (Indel& d : D)
        {
            d.fz.beg1 = d.lflank_end1 + 1;
            d.fz.end1 = d.rflank_beg1 - 1;
            L.add(d.rid, be.contig_len(d.rid), d.fz.beg1, d.fz.end1);
        }
        rc = be.fetch(L, 1, buf);
        if (rc) { if (trace) fclose(trace); return rc; }
        HmmJob J;
        std::vector<size_t> owner;
        std::vector<std::string> rus(K);
        for (size_t k = 0; k < K; ++k)
        {
            Indel& d = D[k];
            d.fz.tract = window_of(L, buf, k);
            if (!purity_exact(d.fz, d.fz.ru, rus[k]))
            {
                J.add(VTN_HMM_AHMM, rus[k], "", d.fz.tract);
                owner.push_back(k);
            }
        }
        rc = be.hmm(J, hres);
        if (rc) { if (trace) fclose(trace); return rc; }
        n_align += J.size();
        for (size_t j = 0; j < owner.size(); ++j)
        {
            purity_from_ahmm(D[owner[j]].fz, hres[j]);
            trace_hmm(owner[j], 3, "A", rus[owner[j]], "", D[owner[j]].fz.tract, hres[j]);
        }
        for (size_t k = 0; k < K; ++k)
        {
            Indel& d = D[k];
            d.fz.ru = rus[k];
            d.fz.ll = d.fz.rl + d.max_dlen;
            composition_and_entropy(d.fz);
        }
    }
    if (trace) fclose(trace);
    // ---- results ----
    out->n_indel = (uint32_t)K;
    out->n_alignments = n_align;
    uint64_t need = 0;
    for (const Indel& d : D) need += d.ex.motif.size() + d.ex.ru.size() + d.ex.basis.size() + d.fz.motif.size() + d.fz.ru.size() + d.fz.basis.size();
    out->n_str = need;
    if (K > out->cap_indel || need > out->cap_str) return be.fail(VTN_E_CAPACITY, "vtn_annotate_indels: result capacity too small (n_indel / n_str are set)");
    uint64_t so = 0;
    for (size_t k = 0; k < K; ++k)
    {
        const Indel& d = D[k];
        vtn_vntr_annot& a = out->vntr[k];
        memset(&a, 0, sizeof a);
        const Side* sd[2] = {&d.ex, &d.fz};
        for (int x = 0; x < 2; ++x)
        {
            vtn_vntr_side& o = x ? a.fuzzy : a.exact;
            const Side& v = *sd[x];
            o.beg1 = v.beg1; o.end1 = v.end1; o.mlen = v.mlen; o.blen = v.blen;
            for (int i = 0; i < 4; ++i) o.comp[i] = v.comp[i];
            o.rl = v.rl; o.ll = v.ll; o.no_perfect_ru = v.no_perfect_ru; o.no_ru = v.no_ru; o.trf_score = v.trf_score;
            o.entropy = v.entropy; o.entropy2 = v.entropy2; o.kl_divergence = v.kl; o.kl_divergence2 = v.kl2; o.ref = v.ref; o.score = v.score;
            const std::string* st[3] = {&v.motif, &v.ru, &v.basis};
            for (int i = 0; i < 3; ++i)
            {
                o.str_off[i] = so;
                o.str_len[i] = (uint32_t)st[i]->size();
                memcpy(out->strings + so, st[i]->data(), st[i]->size());
                so += st[i]->size();
            }
        }
        a.rec = d.rec;
        a.exact_ru_ambiguous = d.ambiguous ? 1 : 0;
        a.is_large_repeat_tract = d.large ? 1 : 0;
    }
    return VTN_OK;
}

} // namespace vta
