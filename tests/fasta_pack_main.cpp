// CPU check of vt_b200/csrc/vtn_fasta_pack.h (the host half of the resident reference): pack a FASTA file the parallel way and
// compare every group and exception run with a byte-by-byte restatement of what faidx streams (isgraph bytes from the contig's
// offset, lib/htslib/faidx.c fai_retrieve).  usage: fasta_pack_main <fasta> [threads]   (reads <fasta>.fai when present)
// Prints "ok <contigs> <bases> <runs>" or the first difference.
#include "../vt_b200/csrc/vtn_fasta_pack.h"

int main(int argc, char** argv)
{
    if (argc < 2) return 2;
    const unsigned nthr = argc > 2 ? (unsigned)atoi(argv[2]) : 4;
    vtf::FastaText T;
    if (!T.open(argv[1]))
    {
        printf("cannot open\n");
        return 1;
    }
    std::vector<vtf::FaiEntry> fai;
    if (!vtf::read_fai(std::string(argv[1]) + ".fai", fai)) vtf::build_fai(T.data, T.size, fai);
    const size_t nc = fai.size();
    std::vector<uint64_t> goff(nc), hoff(nc);
    uint64_t ng = 0;
    for (size_t c = 0; c < nc; ++c)
    {
        goff[c] = ng + 7; // global and host offsets differ on purpose
        hoff[c] = ng;
        ng += (uint64_t)((fai[c].len + 31) / 32) + 2;
    }
    std::vector<vtf::Group> got(ng + 2), exp(ng + 2);
    memset(got.data(), 0, got.size() * sizeof(vtf::Group));
    memset(exp.data(), 0, exp.size() * sizeof(vtf::Group));
    std::vector<std::vector<vtf::ExcRun>> exc(nc);
    std::vector<int> which, order;
    for (size_t c = 0; c < nc; ++c) which.push_back((int)c);
    int bad = vtf::pack_contigs(T.data, T.size, fai, which, goff, hoff, got.data(), exc, nthr, [&](int c) { order.push_back(c); });
    if (bad >= 0)
    {
        printf("short %s\n", fai[bad].name.c_str());
        return 0;
    }
    uint64_t bases = 0, runs = 0;
    for (size_t c = 0; c < nc; ++c)
    {
        if (order[c] != (int)c)
        {
            printf("order\n");
            return 1;
        }
        // restatement: the isgraph stream, one byte at a time
        std::vector<uint8_t> seq;
        size_t p = (size_t)fai[c].offset;
        while ((int64_t)seq.size() < fai[c].len && p < T.size)
        {
            const uint8_t ch = T.data[p++];
            if (isgraph(ch)) seq.push_back(ch);
        }
        std::vector<vtf::ExcRun> er;
        for (int64_t b = 0; b < fai[c].len; ++b)
        {
            const uint8_t ch = seq[(size_t)b];
            vtf::Group& g = exp[hoff[c] + (uint64_t)(b >> 5)];
            const int j = (int)(b & 31);
            int k = -1;
            switch (ch)
            {
            case 'A': case 'a': k = 0; break;
            case 'C': case 'c': k = 1; break;
            case 'T': case 't': k = 2; break;
            case 'G': case 'g': k = 3; break;
            }
            if (k >= 0)
            {
                if (j < 16) g.code_lo |= (uint32_t)k << (2 * j); else g.code_hi |= (uint32_t)k << (2 * (j - 16));
                if (ch >= 'a') g.nonstd |= 1u << j, g.lower |= 1u << j;
            }
            else
            {
                g.nonstd |= 1u << j;
                const uint64_t gp = (goff[c] << 5) + (uint64_t)b;
                if (!er.empty() && er.back().byte == ch && er.back().start + er.back().len == gp) er.back().len++;
                else er.push_back(vtf::ExcRun{gp, 1u, ch});
            }
        }
        const uint64_t n = (uint64_t)((fai[c].len + 31) / 32);
        for (uint64_t g = 0; g < n + 2; ++g)
            if (memcmp(&got[hoff[c] + g], &exp[hoff[c] + g], sizeof(vtf::Group)))
            {
                printf("group differs: contig %zu group %llu\n", c, (unsigned long long)g);
                return 1;
            }
        if (er.size() != exc[c].size())
        {
            printf("run count differs: contig %zu %zu vs %zu\n", c, exc[c].size(), er.size());
            return 1;
        }
        for (size_t k = 0; k < er.size(); ++k)
            if (er[k].start != exc[c][k].start || er[k].len != exc[c][k].len || er[k].byte != exc[c][k].byte)
            {
                printf("run differs: contig %zu run %zu\n", c, k);
                return 1;
            }
        bases += (uint64_t)fai[c].len;
        runs += er.size();
    }
    printf("ok %zu %llu %llu\n", nc, (unsigned long long)bases, (unsigned long long)runs);
    return 0;
}
