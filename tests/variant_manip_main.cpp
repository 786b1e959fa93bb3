// Replays a sites file through the VariantManip-shaped facade (include/vtn_variant_manip.hpp) record by record, with the control flow
// of the reference's normalize loop (normalize.cpp:188-250), and prints one line per record:
//   chrom pos1 ref alts | type refcode normalized containsN | new_pos1 new_ref new_alts lt le rt
// usage: variant_manip_main <ref.fa> <sites.tsv>     sites.tsv: chrom \t pos1 \t ref \t alt[,alt..]
#include <fstream>
#include <iostream>
#include <sstream>
#include "../include/vtn_variant_manip.hpp"

int main(int argc, char** argv)
{
    if (argc < 3) return 2;
    vtn::VariantManip vm{std::string(argv[1])};
    vtn::VariantManip* vm0 = new vtn::VariantManip(); // the constructor without a reference: allele-only methods work
    std::ifstream in(argv[2]);
    std::string line;
    while (std::getline(in, line))
    {
        std::istringstream ls(line);
        vtn::Record v;
        std::string ref, alts;
        ls >> v.chrom >> v.pos1 >> ref >> alts;
        v.alleles.push_back(ref);
        if (alts != ".")
        {
            std::istringstream as(alts);
            std::string a;
            while (std::getline(as, a, ',')) v.alleles.push_back(a);
        }
        vtn::Variant variant;
        const int32_t type = vm.classify_variant(v, variant);
        if (vm0->classify_variant(v, variant) != type) return 3;
        int32_t code = 0;
        if (type != vtn::VT_SNP) code = vm.is_not_ref_consistent(v);
        const bool norm = vm.is_normalized(v), has_n = vm.contains_N(v);
        int32_t pos1 = v.pos1, left_extended = 0, left_trimmed = 0, right_trimmed = 0;
        std::vector<std::string> alleles = v.alleles;
        if (!code && !(type & vtn::VT_SV) && !(type & vtn::VT_VNTR) && !norm)
        {
            vm.right_trim_or_left_extend(alleles, pos1, v.chrom.c_str(), left_extended, right_trimmed);
            vm.left_trim(alleles, pos1, left_trimmed);
        }
        std::cout << v.chrom << ' ' << v.pos1 << ' ' << ref << ' ' << alts << " | " << type << ' ' << code << ' ' << norm << ' ' << has_n << " | " << pos1;
        for (size_t i = 0; i < alleles.size(); ++i) std::cout << (i < 2 ? ' ' : ',') << alleles[i];
        if (alleles.size() == 1) std::cout << " .";
        std::cout << ' ' << left_trimmed << ' ' << left_extended << ' ' << right_trimmed << '\n';
    }
    delete vm0;
    return 0;
}
