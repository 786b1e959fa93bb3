// vtn_variant_manip.hpp -- a header-only, VariantManip-shaped facade over the batch C ABI (vtn.h).
//
// For callers written against the reference's class (variant_manip.h:35-114): the same method names, the same in/out references that
// ACCUMULATE (the caller zeroes left_extended / left_trimmed / right_trimmed, normalize.cpp:234), the same stderr text and exit(1) on a
// failed FASTA fetch (variant_manip.cpp:70-75, 94-98).  Every method forwards ONE record to the device through libvtn (a one-record
// batch): the trimming, extension, classification and the reference bases all come from the device (the only host arithmetic is the byte
// comparison of is_not_ref_consistent over bases fetched from the device copy); without a CUDA device the constructor throws.  It is a compatibility layer, not the
// fast path: a caller that has many records should fill a vtn_batch and call vtn_normalize once.
//
// Differences from the reference class (also listed in DESIGN.md):
//   * records come in as vtn::Record {chrom, pos1, alleles} instead of (bcf_hdr_t*, bcf1_t*): this library does not link htslib;
//   * `fai` is replaced by `ctx` (the resident reference); a destructor releases it (the reference never frees its faidx_t);
//   * right_trim_or_left_extend upper-cases the alleles it is given, as its only caller does before the call (normalize.cpp:223-233),
//     and returns them untouched when REF disagrees with the FASTA (normalize.cpp:203-216 never calls it for such a record);
//   * left_trim is a member, not static, and must follow right_trim_or_left_extend on the same record (its only call site,
//     normalize.cpp:236-237): the device computes both steps in one pass and the facade hands the result out in the two steps;
//   * generate_probes (variant_manip.h:91-112) is not on the decompose / normalize path and is not provided.
#ifndef VTN_VARIANT_MANIP_HPP
#define VTN_VARIANT_MANIP_HPP

#include <cctype>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "vtn.h"

namespace vtn
{

// the type mask of classify_variant (variant.h VT_*)
enum
{
    VT_REF = VTN_VT_REF,
    VT_SNP = VTN_VT_SNP,
    VT_MNP = VTN_VT_MNP,
    VT_INDEL = VTN_VT_INDEL,
    VT_CLUMPED = VTN_VT_CLUMPED,
    VT_VNTR = VTN_VT_VNTR,
    VT_SV = VTN_VT_SV
};

struct Variant // what the path reads of the reference's Variant after classify_variant
{
    int32_t type = VT_REF;
};

struct Record // what the reference reads of (bcf_hdr_t*, bcf1_t*): bcf_get_chrom, bcf_get_pos1, bcf_get_allele
{
    std::string chrom;
    int32_t pos1 = 0;
    std::vector<std::string> alleles; // REF first
};

class VariantManip
{
  public:
    vtn_ctx* ctx = nullptr; // in place of faidx_t* fai
    bool reference_present = false;

    explicit VariantManip(std::string ref_fasta_file, int device = 0)
    {
        open(device);
        // fai_load reads only the index (variant_manip.cpp:35); contigs become resident on first use
        if (vtn_ref_open_fasta(ctx, ref_fasta_file.c_str()))
        {
            fprintf(stderr, "[%s:%d %s] Cannot load genome index: %s\n", "variant_manip.cpp", 38, "VariantManip", ref_fasta_file.c_str());
            exit(1);
        }
        reference_present = true;
    }
    explicit VariantManip(int device = 0) { open(device); }
    VariantManip(const VariantManip&) = delete;
    VariantManip& operator=(const VariantManip&) = delete;
    ~VariantManip()
    {
        if (ctx) vtn_ctx_destroy(ctx);
    }

    // Classifies variants (variant_manip.cpp:227-505): returns the type mask and stores it in variant.type
    int32_t classify_variant(const Record& v, Variant& variant)
    {
        uint8_t vt = 0, pr = 0;
        props(v, vt, pr);
        variant.type = vt;
        return vt;
    }

    // Checks if a variant is normalized (variant_manip.cpp:109-197)
    bool is_normalized(const Record& v)
    {
        uint8_t vt = 0, pr = 0;
        props(v, vt, pr);
        return (pr & VTN_PROP_NORMALIZED) != 0;
    }

    // Checks if a variant contains N bases (variant_manip.cpp:202-222)
    bool contains_N(const Record& v)
    {
        uint8_t vt = 0, pr = 0;
        props(v, vt, pr);
        return (pr & VTN_PROP_CONTAINS_N) != 0;
    }

    // Checks if the REF sequence of a VCF entry is consistent (variant_manip.cpp:61-103): 0 consistent, 1 differs only where an N is
    // involved, 2 differs; prints the reference's message for 1 and 2; a contig the FASTA lacks ends the process like the reference.
    // The bases come from the device copy (vtn_ref_fetch = faidx_fetch_seq); for every record the normalize kernel checks itself
    // (everything but SNPs, normalize.cpp:204) the value equals the VTN_ST_REFCODE bits of vtn_normalize's status.
    int32_t is_not_ref_consistent(const Record& v)
    {
        const std::string& vcf_ref = v.alleles[0];
        const int rl = (int)vcf_ref.size();
        const int pos0 = v.pos1 - 1;
        const int rid = vtn_ref_contig_id(ctx, v.chrom.c_str());
        std::string ref((size_t)rl + 1, '\0');
        const int64_t got = rid < 0 ? -1 : vtn_ref_fetch(ctx, rid, pos0, (int64_t)pos0 + rl - 1, 0, (uint8_t*)&ref[0], rl);
        if (got < 0)
        {
            if (rid < 0) fprintf(stderr, "[E::faidx_adjust_position] The sequence \"%s\" was not found\n", v.chrom.c_str());
            fprintf(stderr, "[variant_manip.cpp:72 is_not_ref_consistent] failure to extract base from fasta file: %s:%d-%d\n", v.chrom.c_str(), pos0, pos0 + rl - 1);
            fprintf(stderr, "FAQ: http://genome.sph.umich.edu/wiki/Vt#1._vt_cannot_retrieve_sequences_from_my_reference_sequence_file\n");
            exit(1);
        }
        ref.resize((size_t)got);
        int32_t is_not_consistent = 0;
        for (int64_t i = 0; i < got; ++i)
            if (toupper((unsigned char)vcf_ref[(size_t)i]) != (unsigned char)ref[(size_t)i])
            {
                if (ref[(size_t)i] == 'N' || toupper((unsigned char)vcf_ref[(size_t)i]) == 'N')
                    is_not_consistent = 1;
                else
                {
                    is_not_consistent = 2;
                    break;
                }
            }
        if (is_not_consistent)
            fprintf(stderr, "[variant_manip.cpp:96 is_not_ref_consistent] reference bases not consistent: %s:%d-%d  %s(REF) vs %s(FASTA)\n", v.chrom.c_str(), pos0,
                    pos0 + rl - 1, vcf_ref.c_str(), ref.c_str());
        return is_not_consistent;
    }

    // Right trims or left extends a variant (variant_manip.cpp:513-583).  pos1 / left_extended / right_trimmed are in/out and accumulate.
    void right_trim_or_left_extend(std::vector<std::string>& alleles, int32_t& pos1, const char* chrom, int32_t& left_extended, int32_t& right_trimmed)
    {
        for (std::string& a : alleles)
            for (char& c : a) c = (char)toupper((unsigned char)c);
        Record v;
        v.chrom = chrom ? chrom : "";
        v.pos1 = pos1;
        v.alleles = alleles;
        One o;
        run(v, o);
        pending_ = false;
        if (!(o.status & VTN_ST_CHANGED))
        { // nothing to trim or extend (or a record normalize.cpp would not have passed in): left_trim has nothing to do either
            pending_ = true;
            p_lt_ = 0;
            p_pos1_ = pos1;
            p_inter_ = alleles;
            p_final_ = alleles;
            return;
        }
        // the state between the two steps: every allele is (le bases to the left of POS) + (the allele without its last rt bases)
        std::string prefix;
        if (o.le)
        {
            prefix.assign((size_t)o.le, '\0');
            const int64_t got = vtn_ref_fetch(ctx, o.rid, (int64_t)pos1 - 1 - o.le, (int64_t)pos1 - 2, 1, (uint8_t*)&prefix[0], o.le);
            if (got != (int64_t)o.le) fail("vtn_ref_fetch");
        }
        std::vector<std::string> inter;
        for (const std::string& a : alleles) inter.push_back(prefix + a.substr(0, a.size() - o.rt));
        pos1 -= (int32_t)o.le;
        left_extended += (int32_t)o.le;
        right_trimmed += (int32_t)o.rt;
        alleles = inter;
        pending_ = true;
        p_lt_ = o.lt;
        p_pos1_ = pos1;
        p_inter_ = inter;
        p_final_ = o.alleles;
    }

    // Left trims a variant of unnecessary nucleotides (variant_manip.cpp:588-616).  Must follow right_trim_or_left_extend on the same record.
    void left_trim(std::vector<std::string>& alleles, int32_t& pos1, int32_t& left_trimmed)
    {
        if (!pending_ || pos1 != p_pos1_ || alleles != p_inter_)
            throw std::logic_error("vtn::VariantManip::left_trim must follow right_trim_or_left_extend on the same record (normalize.cpp:236-237)");
        pending_ = false;
        alleles = p_final_;
        pos1 += (int32_t)p_lt_;
        left_trimmed += (int32_t)p_lt_;
    }

  private:
    struct One // vtn_normalize's answer for one record
    {
        int rid = -1;
        uint8_t status = 0, vtype = 0;
        uint32_t lt = 0, le = 0, rt = 0;
        int32_t pos1 = 0;
        std::vector<std::string> alleles;
    };
    bool pending_ = false;
    uint32_t p_lt_ = 0;
    int32_t p_pos1_ = 0;
    std::vector<std::string> p_inter_, p_final_;

    void open(int device)
    {
        if (vtn_ctx_create(&ctx, device)) throw std::runtime_error(std::string("vtn_ctx_create: ") + vtn_last_error(nullptr));
    }
    [[noreturn]] void fail(const char* what) { throw std::runtime_error(std::string(what) + ": " + vtn_last_error(ctx)); }

    struct Soa // a one-record vtn_batch
    {
        int32_t pos1 = 0, rid = -1;
        std::vector<uint32_t> beg, off;
        std::vector<uint8_t> bytes;
        vtn_batch b;
        Soa(const Record& v, int rid_)
        {
            pos1 = v.pos1, rid = rid_;
            beg = {0u, (uint32_t)v.alleles.size()};
            off.push_back(0);
            for (const std::string& a : v.alleles)
            {
                bytes.insert(bytes.end(), a.begin(), a.end());
                off.push_back((uint32_t)bytes.size());
            }
            b.n_rec = 1, b.n_allele = (uint32_t)v.alleles.size(), b.n_bytes = bytes.size();
            if (bytes.empty()) bytes.push_back(0);
            b.pos1 = &pos1, b.rid = &rid, b.allele_beg = beg.data(), b.allele_off = off.data(), b.bytes = bytes.data();
        }
    };

    void props(const Record& v, uint8_t& vt, uint8_t& pr)
    {
        if (v.alleles.empty()) throw std::invalid_argument("vtn::VariantManip: a record without alleles");
        Soa s(v, -1);
        if (vtn_variant_props(ctx, &s.b, &vt, &pr)) fail("vtn_variant_props");
    }

    void run(const Record& v, One& o)
    {
        if (!reference_present) throw std::logic_error("vtn::VariantManip: no reference (use the constructor that takes the FASTA file)");
        if (v.alleles.empty()) throw std::invalid_argument("vtn::VariantManip: a record without alleles");
        o.rid = vtn_ref_contig_id(ctx, v.chrom.c_str());
        Soa s(v, o.rid);
        uint8_t status[2] = {0, 0}, vtype[2] = {0, 0};
        uint32_t rec[2], lt[2], le[2], rt[2];
        int32_t pos[2];
        uint64_t off[3] = {0, 0, 0};
        std::vector<uint8_t> bytes(s.b.n_bytes + 16);
        vtn_norm_result r;
        memset(&r, 0, sizeof r);
        r.status = status, r.vtype = vtype;
        r.cap_changed = 1, r.cap_changed_bytes = s.b.n_bytes;
        r.chg_rec = rec, r.chg_pos1 = pos, r.chg_lt = lt, r.chg_le = le, r.chg_rt = rt, r.chg_off = off, r.chg_bytes = bytes.data();
        if (vtn_normalize(ctx, &s.b, &r)) fail("vtn_normalize");
        o.status = status[0], o.vtype = vtype[0], o.pos1 = v.pos1;
        if ((o.status & VTN_ST_CHANGED) && r.n_changed == 1)
        {
            o.lt = lt[0], o.le = le[0], o.rt = rt[0], o.pos1 = pos[0];
            uint64_t q = off[0];
            for (const std::string& a : v.alleles)
            {
                const size_t nl = a.size() - o.rt + o.le - o.lt;
                o.alleles.emplace_back((const char*)bytes.data() + q, nl);
                q += nl;
            }
        }
    }
};

} // namespace vtn
#endif
