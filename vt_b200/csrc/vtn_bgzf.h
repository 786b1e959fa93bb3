// vtn_bgzf.h -- file I/O either side of the tools: byte sources for plain / gzip / BGZF input and a BGZF writer.
// Host-only C++ (zlib); used by vtn_tools.cpp and checked on the CPU by tests/bgzf_main.cpp.
//
// What it stands in for: htslib's hts_open / bgzf.c under BCFOrderedReader / BCFOrderedWriter (bcf_ordered_writer.cpp:35-83 picks the
// format from the file name: .vcf plain, .vcf.gz BGZF, .bcf BGZF-compressed BCF2, .ubcf / "+" uncompressed BCF2, "-" plain VCF on
// stdout).  BGZF (the SAM specification, section 4.1): a series of gzip members of at most 64 KiB, each with an extra subfield
// 'B','C' that holds the member's total size - 1, and an empty member as the end-of-file marker.  Blocks are independent, so a batch
// of them is inflated (reader) or deflated (writer) by all host threads at once.
#pragma once
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include <zlib.h>

namespace vtz
{

inline unsigned io_threads()
{
    unsigned n = std::max(1u, std::min(16u, std::thread::hardware_concurrency()));
    if (const char* e = getenv("VTN_HOST_THREADS"))
        if (atoi(e) > 0) n = (unsigned)atoi(e);
    return n;
}

template <class F>
void parallel_items(size_t n, unsigned nthr, F fn) // fn(i) for i in [0, n), contiguous ranges per thread
{
    nthr = (unsigned)std::max<size_t>(1, std::min<size_t>(nthr, n));
    if (nthr <= 1)
    {
        for (size_t i = 0; i < n; ++i) fn(i);
        return;
    }
    std::vector<std::thread> th;
    for (unsigned t = 0; t < nthr; ++t)
        th.emplace_back([=]() {
            for (size_t i = n * t / nthr; i < n * (t + 1) / nthr; ++i) fn(i);
        });
    for (auto& x : th) x.join();
}

struct ByteSource
{
    virtual ~ByteSource() {}
    virtual long read(char* dst, size_t n) = 0; // bytes read (possibly fewer than n), 0 at the end, < 0 on error
};

// plain or gzip (any number of members) through zlib; also stdin
struct GzSource : ByteSource
{
    gzFile f = nullptr;
    bool open(const std::string& path)
    {
        f = path == "-" ? gzdopen(0, "rb") : gzopen(path.c_str(), "rb");
        if (f) gzbuffer(f, 1 << 20);
        return f != nullptr;
    }
    long read(char* dst, size_t n) override
    {
        const int got = gzread(f, dst, (unsigned)std::min<size_t>(n, 1u << 30));
        return got;
    }
    ~GzSource() override
    {
        if (f) gzclose(f);
    }
};

inline bool is_bgzf_header(const unsigned char* h, size_t n)
{ // gzip magic, deflate, FEXTRA, and the BC subfield first (what bgzf.c writes and checks)
    return n >= 18 && h[0] == 0x1f && h[1] == 0x8b && h[2] == 8 && (h[3] & 4) && h[10] + 256 * h[11] >= 6 && h[12] == 'B' && h[13] == 'C' && h[14] == 2 && h[15] == 0;
}

// BGZF input: compressed blocks are read in batches and inflated by all threads straight into their place in the output buffer
struct BgzfSource : ByteSource
{
    FILE* f = nullptr;
    unsigned nthr = io_threads();
    std::vector<unsigned char> raw;
    std::vector<char> out;
    size_t out_pos = 0, out_end = 0;
    bool eof = false, bad = false;
    static bool sniff(const std::string& path)
    {
        FILE* g = fopen(path.c_str(), "rb");
        if (!g) return false;
        unsigned char h[18];
        const size_t n = fread(h, 1, sizeof h, g);
        fclose(g);
        return is_bgzf_header(h, n);
    }
    bool open(const std::string& path)
    {
        f = fopen(path.c_str(), "rb");
        return f != nullptr;
    }
    ~BgzfSource() override
    {
        if (f) fclose(f);
    }
    bool fill()
    {
        struct Blk
        {
            size_t raw_off, raw_len, out_off, out_len;
        };
        std::vector<Blk> blks;
        const size_t max_blocks = 512; // up to 32 MiB of text per batch
        raw.clear();
        size_t out_total = 0;
        while (blks.size() < max_blocks)
        {
            unsigned char h[18];
            const size_t got = fread(h, 1, 18, f);
            if (got == 0)
            {
                eof = true;
                break;
            }
            if (!is_bgzf_header(h, got)) return false;
            // the BC subfield is the first one in files written by htslib and by this library; walk the extra field to be safe
            const size_t xlen = h[10] + 256u * h[11];
            size_t bsize = 0;
            const size_t r0 = raw.size();
            raw.insert(raw.end(), h, h + 18);
            if (xlen > 6)
            {
                raw.resize(raw.size() + (xlen - 6));
                if (fread(raw.data() + r0 + 18, 1, xlen - 6, f) != xlen - 6) return false;
            }
            for (size_t p = 12; p + 4 <= 12 + xlen;)
            {
                const unsigned char* s = raw.data() + r0 + p;
                const size_t sl = s[2] + 256u * s[3];
                if (s[0] == 'B' && s[1] == 'C' && sl == 2) bsize = s[4] + 256u * s[5] + 1;
                p += 4 + sl;
            }
            const size_t hdr = 12 + xlen;
            if (bsize < hdr + 8) return false;
            raw.resize(r0 + bsize);
            if (fread(raw.data() + r0 + hdr, 1, bsize - hdr, f) != bsize - hdr) return false;
            const unsigned char* t = raw.data() + r0 + bsize - 4;
            const size_t isize = t[0] + 256u * t[1] + 65536u * t[2] + 16777216u * (size_t)t[3];
            blks.push_back(Blk{r0 + hdr, bsize - hdr - 8, out_total, isize});
            out_total += isize;
        }
        if (out.size() < out_total) out.resize(out_total);
        std::vector<char> ok(blks.size(), 1);
        parallel_items(blks.size(), nthr, [&](size_t i) {
            const Blk& b = blks[i];
            if (!b.out_len) return;
            z_stream zs;
            memset(&zs, 0, sizeof zs);
            if (inflateInit2(&zs, -15) != Z_OK)
            {
                ok[i] = 0;
                return;
            }
            zs.next_in = raw.data() + b.raw_off, zs.avail_in = (uInt)b.raw_len;
            zs.next_out = (Bytef*)out.data() + b.out_off, zs.avail_out = (uInt)b.out_len;
            const int rc = inflate(&zs, Z_FINISH);
            if (rc != Z_STREAM_END || zs.avail_out != 0) ok[i] = 0;
            inflateEnd(&zs);
            const unsigned char* t = raw.data() + b.raw_off + b.raw_len;
            const uint32_t crc = t[0] | (t[1] << 8) | (t[2] << 16) | ((uint32_t)t[3] << 24);
            if (ok[i] && crc32(crc32(0L, Z_NULL, 0), (const Bytef*)out.data() + b.out_off, (uInt)b.out_len) != crc) ok[i] = 0;
        });
        for (char c : ok)
            if (!c) return false;
        out_pos = 0, out_end = out_total;
        return true;
    }
    long read(char* dst, size_t n) override
    {
        while (out_pos == out_end)
        {
            if (eof) return 0;
            if (bad || !fill())
            {
                bad = true;
                return -1;
            }
        }
        const size_t k = std::min(n, out_end - out_pos);
        memcpy(dst, out.data() + out_pos, k);
        out_pos += k;
        return (long)k;
    }
};

inline ByteSource* open_source(const std::string& path)
{
    if (path != "-" && BgzfSource::sniff(path))
    {
        BgzfSource* s = new BgzfSource();
        if (s->open(path)) return s;
        delete s;
        return nullptr;
    }
    GzSource* s = new GzSource();
    if (s->open(path)) return s;
    delete s;
    return nullptr;
}

// BGZF output: text is cut into blocks of at most 0xff00 bytes (bgzf.h BGZF_BLOCK_SIZE), a batch of blocks is deflated by all threads,
// the members are written in order, and close() appends the 28-byte end-of-file member.  level < 0 = zlib's default (6), 0 = stored.
struct BgzfWriter
{
    enum { BLOCK = 0xff00, BATCH = 256 };
    FILE* f = nullptr;
    int level = -1;
    unsigned nthr = io_threads();
    std::vector<char> buf;
    std::vector<std::vector<unsigned char>> comp;
    bool failed = false;
    void open(FILE* file, int lvl)
    {
        f = file, level = lvl;
        buf.reserve((size_t)BLOCK * BATCH);
    }
    static bool deflate_block(const char* src, size_t n, int level, std::vector<unsigned char>& dst)
    {
        dst.resize(18 + compressBound((uLong)n) + 16 + 8);
        static const unsigned char head[16] = {0x1f, 0x8b, 8, 4, 0, 0, 0, 0, 0, 0xff, 6, 0, 'B', 'C', 2, 0};
        memcpy(dst.data(), head, 16);
        z_stream zs;
        memset(&zs, 0, sizeof zs);
        if (deflateInit2(&zs, level, Z_DEFLATED, -15, 8, Z_DEFAULT_STRATEGY) != Z_OK) return false;
        zs.next_in = (Bytef*)src, zs.avail_in = (uInt)n;
        zs.next_out = dst.data() + 18, zs.avail_out = (uInt)(dst.size() - 18 - 8);
        const int rc = deflate(&zs, Z_FINISH);
        const size_t clen = zs.total_out;
        deflateEnd(&zs);
        if (rc != Z_STREAM_END) return false;
        const size_t total = 18 + clen + 8;
        if (total > 65536) return false; // cannot happen for n <= 0xff00 (stored blocks add 5 bytes per 65535)
        dst[16] = (unsigned char)((total - 1) & 0xff), dst[17] = (unsigned char)((total - 1) >> 8);
        const uint32_t crc = (uint32_t)crc32(crc32(0L, Z_NULL, 0), (const Bytef*)src, (uInt)n);
        unsigned char* t = dst.data() + 18 + clen;
        t[0] = crc & 0xff, t[1] = (crc >> 8) & 0xff, t[2] = (crc >> 16) & 0xff, t[3] = (crc >> 24) & 0xff;
        t[4] = n & 0xff, t[5] = (n >> 8) & 0xff, t[6] = (n >> 16) & 0xff, t[7] = (n >> 24) & 0xff;
        dst.resize(total);
        return true;
    }
    void flush_blocks(bool all)
    { // whole blocks only, unless all
        size_t nb = buf.size() / BLOCK;
        if (all && buf.size() % BLOCK) ++nb;
        if (!nb) return;
        if (comp.size() < nb) comp.resize(nb);
        std::vector<char> ok(nb, 1);
        parallel_items(nb, nthr, [&](size_t i) {
            const size_t o = i * BLOCK, n = std::min<size_t>(BLOCK, buf.size() - o);
            ok[i] = deflate_block(buf.data() + o, n, level, comp[i]) ? 1 : 0;
        });
        for (size_t i = 0; i < nb; ++i)
            if (!ok[i] || fwrite(comp[i].data(), 1, comp[i].size(), f) != comp[i].size()) failed = true;
        const size_t used = std::min(buf.size(), nb * BLOCK);
        buf.erase(buf.begin(), buf.begin() + used);
    }
    bool write(const char* p, size_t n)
    {
        while (n)
        {
            const size_t room = (size_t)BLOCK * BATCH - buf.size(), k = std::min(room, n);
            buf.insert(buf.end(), p, p + k);
            p += k, n -= k;
            if (buf.size() == (size_t)BLOCK * BATCH) flush_blocks(false);
        }
        return !failed;
    }
    bool close()
    {
        flush_blocks(true);
        static const unsigned char eof_block[28] = {0x1f, 0x8b, 8, 4, 0, 0, 0, 0, 0, 0xff, 6, 0, 'B', 'C', 2, 0, 0x1b, 0, 3, 0, 0, 0, 0, 0, 0, 0, 0, 0};
        if (fwrite(eof_block, 1, 28, f) != 28) failed = true;
        if (fflush(f) != 0) failed = true;
        return !failed;
    }
};

} // namespace vtz
