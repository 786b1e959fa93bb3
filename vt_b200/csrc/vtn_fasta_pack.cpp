// vt_b200 -- the FASTA packing loops of vtn_fasta_pack.h compiled by the host compiler (AVX2 path with run-time dispatch); vtn_ref.cu
// calls these two.
#include "vtn_fasta_pack.h"

namespace vtf
{

bool pack_bases_ol(const uint8_t* seq, int64_t b0, int64_t b1, uint64_t gpos0, Group* groups, std::vector<ExcRun>& exc)
{
    return pack_bases(seq, b0, b1, gpos0, groups, exc);
}

int pack_contigs_ol(const uint8_t* text, size_t n_text, const std::vector<FaiEntry>& fai, const std::vector<int>& which, const std::vector<uint64_t>& goff,
                    const std::vector<uint64_t>& hoff, Group* groups, std::vector<std::vector<ExcRun>>& exc, unsigned nthr, const std::function<void(int)>& on_done)
{
    return pack_contigs(text, n_text, fai, which, goff, hoff, groups, exc, nthr, on_done);
}

} // namespace vtf
