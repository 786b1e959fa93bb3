// CPU check of vt_b200/csrc/vtn_bgzf.h: usage  bgzf_main c <in> <out.gz> [level]   compress a file into BGZF
//                                              bgzf_main d <in.gz> <out>           read through open_source (BGZF, gzip or plain)
#include "../vt_b200/csrc/vtn_bgzf.h"

int main(int argc, char** argv)
{
    if (argc < 4) return 2;
    if (argv[1][0] == 'c')
    {
        FILE* in = fopen(argv[2], "rb");
        FILE* out = fopen(argv[3], "wb");
        if (!in || !out) return 1;
        vtz::BgzfWriter w;
        w.open(out, argc > 4 ? atoi(argv[4]) : -1);
        std::vector<char> b(777777);
        size_t n;
        while ((n = fread(b.data(), 1, b.size(), in)) > 0) w.write(b.data(), n);
        const bool ok = w.close();
        fclose(out);
        return ok ? 0 : 1;
    }
    vtz::ByteSource* s = vtz::open_source(argv[2]);
    if (!s) return 1;
    FILE* out = fopen(argv[3], "wb");
    std::vector<char> b(1234567);
    long n;
    while ((n = s->read(b.data(), b.size())) > 0) fwrite(b.data(), 1, (size_t)n, out);
    fclose(out);
    delete s;
    return n < 0 ? 1 : 0;
}
