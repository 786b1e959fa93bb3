/* tests/annot_host_main.cpp -- TEST INFRASTRUCTURE ONLY.
 * Runs libvtn's annotate_indels control logic (vt_b200/csrc/vtn_annot_host.h, the product code) on the CPU, with the CPU oracles
 * in place of the device: record classification / normalization from oracle/vt_oracle.c, the three aligners from
 * oracle/hmm_oracle.c, reference windows from the FASTA in memory, and a literal std::string restatement of
 * CandidateRegionExtractor's trim / right_align / left_align (candidate_region_extractor.cpp:239-267, :472-502) below.
 * tests/test_annot_host_cpu.py compares its output with what the unmodified reference (`vt_ref annotate_indels`) wrote for the
 * same input, and its alignment trace with oracle/_ref/vt_ref_trace's.  No GPU is involved; nothing here ships.
 *
 * usage: annot_host_main ref.fa in.vcf [trace.txt]    -> one line per record on stdout: rec, flags, vtype[, INFO tags] */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <fstream>
#include <iostream>
#include <sstream>
#include "../vt_b200/csrc/vtn_annot_host.h"
extern "C"
{
#include "../oracle/hmm_oracle.h"
#include "../oracle/vt_oracle.h"
}

struct Fasta
{
    std::vector<std::string> names, seqs;
    int find(const std::string& n) const
    {
        for (size_t i = 0; i < names.size(); ++i)
            if (names[i] == n) return (int)i;
        return -1;
    }
};

static Fasta read_fasta(const char* path)
{
    Fasta f;
    std::ifstream in(path);
    std::string line;
    while (std::getline(in, line))
    {
        if (!line.empty() && line[0] == '>')
        {
            f.names.push_back(line.substr(1, line.find_first_of(" \t") == std::string::npos ? std::string::npos : line.find_first_of(" \t") - 1));
            f.seqs.emplace_back();
        }
        else if (!f.seqs.empty())
            f.seqs.back() += line;
    }
    return f;
}

struct CpuBackend : vta::AnnotBackend
{
    const Fasta& fa;
    explicit CpuBackend(const Fasta& f) : fa(f) {}
    int64_t contig_len(int32_t rid) override { return (rid >= 0 && rid < (int32_t)fa.seqs.size()) ? (int64_t)fa.seqs[rid].size() : -1; }
    int fail(int code, const std::string& msg) override
    {
        fprintf(stderr, "annot_host_main: %s\n", msg.c_str());
        return code;
    }
    /* ReferenceSequence::fetch_base (reference_sequence.cpp:76-106) */
    char base1(int32_t rid, int64_t p1) const
    {
        const std::string& s = fa.seqs[rid];
        int64_t p0 = p1 - 1;
        p0 = p0 < 0 ? 0 : (p0 >= (int64_t)s.size() ? (int64_t)s.size() - 1 : p0);
        return (char)toupper((unsigned char)s[p0]);
    }
    int regions(const vtn_batch* in, vta::RegionOut& RO) override
    {
        const uint32_t n = in->n_rec;
        RO.flags.assign(n, 0); RO.vtype.assign(n, 0); RO.reg_beg.assign(n, 0); RO.reg_end.assign(n, 0); RO.max_dlen.assign(n, 0);
        RO.a_lt.assign(in->n_allele, 0); RO.a_rt.assign(in->n_allele, 0); RO.a_shift.assign(in->n_allele, 0);
        for (uint32_t r = 0; r < n; ++r)
        {
            const uint32_t ab = in->allele_beg[r], na = in->allele_beg[r + 1] - ab;
            std::vector<const uint8_t*> al(na);
            std::vector<uint32_t> len(na);
            std::vector<std::string> A(na);
            for (uint32_t i = 0; i < na; ++i)
            {
                al[i] = in->bytes + in->allele_off[ab + i];
                len[i] = in->allele_off[ab + i + 1] - in->allele_off[ab + i];
                A[i].assign((const char*)al[i], len[i]);
            }
            const int32_t vtype = vto_classify((int32_t)na, al.data(), len.data());
            RO.vtype[r] = (uint8_t)vtype;
            bool has_n = false; /* contains_N, variant_manip.cpp:202-222 */
            for (uint32_t i = 0; i < na; ++i)
                if (A[i].find('N') != std::string::npos && !(i && A[i][0] == '<')) has_n = true;
            int32_t max_dlen = 0; /* variant_manip.cpp:486, explicit-bases branch only */
            for (uint32_t i = 1; i < na; ++i)
            {
                if (A[i].find('<') != std::string::npos || A[i].find('[') != std::string::npos || A[i].find(']') != std::string::npos) continue;
                if (A[i][0] == '.' || A[i] == A[0]) continue;
                max_dlen = std::max(max_dlen, (int32_t)A[i].size() - (int32_t)A[0].size());
            }
            RO.max_dlen[r] = max_dlen;
            uint32_t flags = (vto_is_normalized((int32_t)na, al.data(), len.data()) && !has_n) ? VTA_KEEP : 0;
            int32_t min_beg1 = in->pos1[r], max_end1 = in->pos1[r];
            if (flags && (vtype & VTN_VT_INDEL))
            {
                flags |= VTA_INDEL;
                const int32_t rid = in->rid[r];
                if (contig_len(rid) <= 0)
                    flags |= VTA_NOCONTIG;
                else
                    for (uint32_t i = 1; i < na; ++i)
                    {
                        std::string ref = A[0], alt = A[i];
                        int32_t pos1 = in->pos1[r];
                        uint32_t lt = 0;
                        for (;;) /* trim, candidate_region_extractor.cpp:472-502 */
                        {
                            if (ref.size() == 1 || alt.size() == 1) break;
                            if (ref[0] != alt[0] && ref.back() != alt.back()) break;
                            if (ref.back() == alt.back()) { ref.pop_back(); alt.pop_back(); }
                            else if (ref[0] == alt[0]) { ref.erase(0, 1); alt.erase(0, 1); ++pos1; ++lt; }
                        }
                        RO.a_lt[ab + i] = lt;
                        RO.a_rt[ab + i] = (uint32_t)(A[0].size() - lt - ref.size());
                        int32_t end1 = pos1 + (int32_t)ref.size() - 1;
                        uint32_t s = 0;
                        while (ref[0] == alt[0]) /* right_align, :256-267 */
                        {
                            const char b = base1(rid, (int64_t)end1 + 1);
                            ref.erase(0, 1); alt.erase(0, 1);
                            ref.push_back(b); alt.push_back(b);
                            ++end1;
                            if (++s > fa.seqs[rid].size() + A[0].size() + A[i].size()) { flags |= VTA_RUNAWAY; break; }
                        }
                        RO.a_shift[ab + i] = s;
                        int32_t beg1 = end1 - (int32_t)ref.size() + 1;
                        while (ref.back() == alt.back() && beg1 > 1) /* left_align, :239-251 (fetches pos1-2) */
                        {
                            const char b = base1(rid, (int64_t)beg1 - 2);
                            ref.pop_back(); alt.pop_back();
                            ref.insert(0, 1, b); alt.insert(0, 1, b);
                            --beg1;
                        }
                        min_beg1 = std::min(min_beg1, beg1);
                        max_end1 = std::max(max_end1, end1);
                    }
            }
            RO.flags[r] = (uint8_t)flags;
            RO.reg_beg[r] = min_beg1;
            RO.reg_end[r] = max_end1;
        }
        return VTN_OK;
    }
    int fetch(const vta::FetchList& L, int upper, std::vector<uint8_t>& out) override
    {
        out.assign((size_t)L.off.back() + 1, 0);
        for (size_t k = 0; k < L.size(); ++k)
            for (uint64_t i = 0; i < L.off[k + 1] - L.off[k]; ++i)
            {
                const char c = fa.seqs[L.rid[k]][(size_t)L.beg0[k] + i];
                out[L.off[k] + i] = (uint8_t)(upper ? toupper((unsigned char)c) : c);
            }
        return VTN_OK;
    }
    int hmm(const vta::HmmJob& J, std::vector<vtn_hmm_result>& res) override
    {
        res.assign(J.size(), vtn_hmm_result{});
        const hmmo_params par = {vta::FD_PARAMS.delta, vta::FD_PARAMS.epsilon, vta::FD_PARAMS.tau, vta::FD_PARAMS.eta, vta::FD_PARAMS.mismatch_penalty, vta::FD_PARAMS.pl};
        std::vector<int32_t> path(8192);
        for (size_t a = 0; a < J.size(); ++a)
        {
            const std::string motif = J.bytes.substr(J.seq_off[3 * a], J.seq_off[3 * a + 1] - J.seq_off[3 * a]);
            const std::string flank = J.bytes.substr(J.seq_off[3 * a + 1], J.seq_off[3 * a + 2] - J.seq_off[3 * a + 1]);
            std::string read = J.bytes.substr(J.seq_off[3 * a + 2], J.seq_off[3 * a + 3] - J.seq_off[3 * a + 2]);
            hmmo_result h;
            memset(&h, 0, sizeof h);
            const int rc = hmmo_align(J.kind[a], &par, motif.c_str(), flank.c_str(), read.c_str(), &h, path.data());
            vtn_hmm_result& o = res[a];
            o.status = rc ? VTN_HMM_ST_TOO_LONG : VTN_HMM_ST_OK;
            o.optimal_score = h.optimal_score;
            for (int i = 0; i < 12; ++i) o.pos[i] = h.pos[i];
            o.motif_count = h.motif_count;
            o.exact_motif_count = h.exact_motif_count;
            o.motif_concordance = h.motif_concordance;
            o.trf_score = h.trf_score;
            o.frac_no_repeats = h.frac_no_repeats;
        }
        return VTN_OK;
    }
};

int main(int argc, char** argv)
{
    if (argc < 3)
    {
        fprintf(stderr, "usage: annot_host_main ref.fa in.vcf [trace.txt]\n");
        return 2;
    }
    const Fasta fa = read_fasta(argv[1]);
    std::vector<int32_t> pos1, rid;
    std::vector<uint32_t> abeg{0}, aoff{0};
    std::string bytes;
    std::ifstream in(argv[2]);
    std::string line;
    while (std::getline(in, line))
    {
        if (line.empty() || line[0] == '#') continue;
        std::vector<std::string> f;
        std::stringstream ss(line);
        std::string tok;
        while (std::getline(ss, tok, '\t')) f.push_back(tok);
        if (f.size() < 5) continue;
        rid.push_back(fa.find(f[0]));
        pos1.push_back(atoi(f[1].c_str()));
        std::vector<std::string> al{f[3]};
        if (f[4] != ".")
        {
            std::stringstream sa(f[4]);
            while (std::getline(sa, tok, ',')) al.push_back(tok);
        }
        for (auto& a : al)
        {
            bytes += a;
            aoff.push_back((uint32_t)bytes.size());
        }
        abeg.push_back((uint32_t)aoff.size() - 1);
    }
    const uint32_t n = (uint32_t)pos1.size();
    vtn_batch b{n, abeg.back(), (uint64_t)bytes.size(), pos1.data(), rid.data(), abeg.data(), aoff.data(), (const uint8_t*)bytes.data()};
    std::vector<uint8_t> flags(n + 1), vtype(n + 1);
    std::vector<uint32_t> idx(n + 1), tl(abeg.back() + 1), tr(abeg.back() + 1), sh(abeg.back() + 1);
    std::vector<vtn_vntr_annot> vn(n + 1);
    std::vector<char> strs(4 * bytes.size() + 4096 * (size_t)n + 4096);
    vtn_annot_result out{};
    out.flags = flags.data(); out.vtype = vtype.data(); out.indel_idx = idx.data();
    out.trim_left = tl.data(); out.trim_right = tr.data(); out.shift = sh.data();
    out.cap_indel = n; out.cap_str = strs.size(); out.vntr = vn.data(); out.strings = strs.data();
    CpuBackend be(fa);
    const int rc = vta::annotate_core(be, &b, &out, argc > 3 ? argv[3] : "");
    if (rc)
    {
        fprintf(stderr, "annotate_core failed: %d\n", rc);
        return 1;
    }
    for (uint32_t r = 0; r < n; ++r)
    {
        printf("%u\t%u\t%u", r, flags[r], vtype[r]);
        if (flags[r] & VTN_ANNOT_INDEL)
        {
            const vtn_vntr_annot& a = vn[idx[r]];
            const vtn_vntr_side* sd[2] = {&a.exact, &a.fuzzy};
            const char* pre[2] = {"EX", "FZ"};
            printf("\t");
            for (int x = 0; x < 2; ++x)
            {
                const vtn_vntr_side& v = *sd[x];
                auto S = [&](int i) { return std::string(out.strings + v.str_off[i], v.str_len[i]); };
                printf("%s_MOTIF=%s;%s_MLEN=%d;%s_RU=%s;%s_BASIS=%s;%s_BLEN=%d;%s_REPEAT_TRACT=%d,%d;%s_COMP=%d,%d,%d,%d;", pre[x], S(0).c_str(), pre[x], v.mlen,
                       pre[x], S(1).c_str(), pre[x], S(2).c_str(), pre[x], v.blen, pre[x], v.beg1, v.end1, pre[x], v.comp[0], v.comp[1], v.comp[2], v.comp[3]);
                printf("%s_ENTROPY=%.9g;%s_ENTROPY2=%.9g;%s_KL_DIVERGENCE=%.9g;%s_KL_DIVERGENCE2=%.9g;%s_REF=%.9g;%s_RL=%d;%s_LL=%d;%s_RU_COUNTS=%d,%d;%s_SCORE=%.9g;%s_TRF_SCORE=%d",
                       pre[x], v.entropy, pre[x], v.entropy2, pre[x], v.kl_divergence, pre[x], v.kl_divergence2, pre[x], v.ref, pre[x], v.rl, pre[x], v.ll, pre[x],
                       v.no_perfect_ru, v.no_ru, pre[x], v.score, pre[x], v.trf_score);
                if (x == 0) printf("%s;", a.exact_ru_ambiguous ? ";EXACT_RU_AMBIGUOUS" : "");
            }
            printf("\tlarge=%d", a.is_large_repeat_tract);
            for (uint32_t sl = abeg[r] + 1; sl < abeg[r + 1]; ++sl) printf("\talt:%u,%u,%u", tl[sl], tr[sl], sh[sl]);
        }
        printf("\n");
    }
    fprintf(stderr, "annot_host_main: %u records, %u indels, %llu alignments\n", n, out.n_indel, (unsigned long long)out.n_alignments);
    return 0;
}
