// CPU check of vt_b200/csrc/vtn_bcf.h:  bcf_main c <in.vcf> <out.bcf or out.ubcf>  text -> BCF2 ;  bcf_main d <in: BCF2 or text, any compression> <out.vcf>  -> text
#include "../vt_b200/csrc/vtn_bcf.h"
int main(int argc, char** argv)
{
    if (argv[1][0] == 'c')
    {
        vtz::ByteSource* s = vtb2::open_any(argv[2]);
        FILE* out = fopen(argv[3], "wb");
        vtb2::BcfSink k; vtz::BgzfWriter w;
        k.f = out;
        std::string o = argv[3];
        if (o.size() > 4 && o.substr(o.size() - 4) == ".bcf" ) { w.open(out, -1); k.bz = &w; }
        std::vector<char> b(100000); long n;
        while ((n = s->read(b.data(), b.size())) > 0) k.write(b.data(), n);
        bool ok = k.close(); fclose(out); return ok ? 0 : 1;
    }
    vtz::ByteSource* s = vtb2::open_any(argv[2]);
    FILE* out = fopen(argv[3], "wb");
    std::vector<char> b(100000); long n;
    while ((n = s->read(b.data(), b.size())) > 0) fwrite(b.data(), 1, n, out);
    fclose(out); return 0;
}
