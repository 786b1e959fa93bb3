// vt_b200 -- host side of the resident reference: FASTA text -> 2-bit groups + exception runs.  Host-only C++ (no CUDA), so that
// tests/fasta_pack_main.cpp can check it on the CPU against a byte-by-byte restatement of what faidx streams.
//
// What the reference does on this path: fai_load (lib/htslib/faidx.c) reads <fasta>.fai (building it when absent), and every
// faidx_fetch_seq seeks to offset + beg / line_blen * line_len + beg % line_blen and keeps the isgraph() bytes it reads until it
// has end - beg + 1 of them (faidx.c fai_retrieve).  Here a contig is packed once, in parallel, straight from the mapped file:
// the .fai geometry gives every 2^20-base chunk its byte offset, each worker gathers its lines, checks that the line ends are
// where the index says, and packs.  A contig whose text does not have the regular geometry falls back to the sequential isgraph stream.
#pragma once
#include <algorithm>
#include <atomic>
#include <cctype>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <functional>
#include <string>
#include <thread>
#include <vector>

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <zlib.h>
#if defined(__x86_64__) && !defined(__CUDACC__)
#include <immintrin.h>
#endif

namespace vtf
{

struct ExcRun
{
    uint64_t start;
    uint32_t len;
    uint8_t byte;
};

struct Group // one 32-base group as the kernels read it (a uint4): 2-bit codes, "not plain ACGT" mask, lower-case mask
{
    uint32_t code_lo, code_hi, nonstd, lower;
};

struct FaiEntry
{
    std::string name;
    int64_t len = 0, offset = 0, line_blen = 0, line_len = 0;
};

// byte classes: bits 0-1 = 2-bit code ((ascii >> 1) & 3 for ACGT), bit 2 = lower-case acgt, bit 3 = any other printing byte (an
// exception run keeps it verbatim), bit 4 = not isgraph (cannot be a base: the geometry was wrong)
struct ByteClass
{
    uint8_t v[256];
    ByteClass()
    {
        for (int c = 0; c < 256; ++c) v[c] = isgraph(c) ? 8 : 16;
        v['A'] = 0, v['C'] = 1, v['T'] = 2, v['G'] = 3;
        v['a'] = 4, v['c'] = 5, v['t'] = 6, v['g'] = 7;
    }
};
inline const ByteClass& byte_class()
{
    static const ByteClass k;
    return k;
}

#ifndef __CUDACC__ // the packing loops are compiled by g++ (vtn_fasta_pack.cpp, tests/fasta_pack_main.cpp); nvcc sees the declarations below
#if defined(__x86_64__)
// 32 bases that are all [ACGTacgt] -> 64 code bits + the lower-case mask, or false.  (ascii >> 1) & 3 is the code for either case.
__attribute__((target("avx2"))) inline bool pack32_avx2(const uint8_t* s, uint64_t& code, uint32_t& lower)
{
    const __m256i v = _mm256_loadu_si256((const __m256i*)s);
    const __m256i u = _mm256_and_si256(v, _mm256_set1_epi8((char)0xDF));
    const __m256i ok = _mm256_or_si256(_mm256_or_si256(_mm256_cmpeq_epi8(u, _mm256_set1_epi8('A')), _mm256_cmpeq_epi8(u, _mm256_set1_epi8('C'))),
                                       _mm256_or_si256(_mm256_cmpeq_epi8(u, _mm256_set1_epi8('G')), _mm256_cmpeq_epi8(u, _mm256_set1_epi8('T'))));
    if ((uint32_t)_mm256_movemask_epi8(ok) != 0xffffffffu) return false;
    lower = (uint32_t)_mm256_movemask_epi8(_mm256_slli_epi16(v, 2)); // bit 5 (0x20) of every byte
    const __m256i c2 = _mm256_and_si256(_mm256_srli_epi16(v, 1), _mm256_set1_epi8(3));
    const __m256i n4 = _mm256_maddubs_epi16(c2, _mm256_set1_epi16(0x0401));         // c0 + 4 c1 per 16 bits
    const __m256i n8 = _mm256_madd_epi16(n4, _mm256_set1_epi32(0x00100001));        // n0 + 16 n1 per 32 bits: 4 bases in one byte
    const __m256i g = _mm256_shuffle_epi8(n8, _mm256_setr_epi8(0, 4, 8, 12, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, 0, 4, 8, 12, -1, -1, -1, -1, -1, -1,
                                                               -1, -1, -1, -1, -1, -1));
    code = (uint64_t)(uint32_t)_mm256_extract_epi32(g, 0) | ((uint64_t)(uint32_t)_mm256_extract_epi32(g, 4) << 32);
    return true;
}
inline bool have_avx2()
{
    static const bool k = __builtin_cpu_supports("avx2");
    return k;
}
#endif

// pack bases [b0, b1) of a contig given as seq[0 .. b1 - b0) (b0 a multiple of 32) into groups[b0 / 32 ..]; exception runs are
// appended in order with global positions gpos0 + base index.  Returns false when a byte that is not isgraph was among the bases.
inline bool pack_bases(const uint8_t* seq, int64_t b0, int64_t b1, uint64_t gpos0, Group* groups, std::vector<ExcRun>& exc)
{
    const uint8_t* cls = byte_class().v;
    unsigned bad = 0;
#if defined(__x86_64__)
    const bool simd = have_avx2();
#endif
    for (int64_t g0 = b0; g0 < b1; g0 += 32)
    {
        const int cnt = (int)std::min<int64_t>(32, b1 - g0);
        const uint8_t* s = seq + (g0 - b0);
        uint64_t code = 0;
        uint32_t lower = 0, other = 0;
#if defined(__x86_64__)
        if (simd && cnt == 32 && pack32_avx2(s, code, lower))
        {
            groups[g0 >> 5] = Group{(uint32_t)code, (uint32_t)(code >> 32), lower, lower};
            continue;
        }
        code = 0, lower = 0;
#endif
        for (int j = 0; j < cnt; ++j)
        {
            const unsigned v = cls[s[j]];
            code |= (uint64_t)(v & 3u) << (2 * j);
            lower |= ((v >> 2) & 1u) << j;
            other |= ((v >> 3) & 1u) << j;
            bad |= v;
        }
        if (other)
        {
            for (int j = 0; j < cnt; ++j)
                if ((other >> j) & 1u)
                {
                    const uint64_t gp = gpos0 + (uint64_t)(g0 + j);
                    const uint8_t c = s[j];
                    if (!exc.empty() && exc.back().byte == c && exc.back().start + exc.back().len == gp && exc.back().len < 0xffffffffu)
                        exc.back().len++;
                    else
                        exc.push_back(ExcRun{gp, 1u, c});
                }
        }
        groups[g0 >> 5] = Group{(uint32_t)code, (uint32_t)(code >> 32), lower | other, lower};
    }
    return !(bad & 16u);
}
#endif // !__CUDACC__

inline void append_runs(std::vector<ExcRun>& dst, const std::vector<ExcRun>& src)
{ // fuse runs that were split at a chunk border
    for (const ExcRun& e : src)
    {
        if (!dst.empty() && dst.back().byte == e.byte && dst.back().start + dst.back().len == e.start && (uint64_t)dst.back().len + e.len <= 0xffffffffull)
            dst.back().len += e.len;
        else
            dst.push_back(e);
    }
}

inline bool read_fai(const std::string& path, std::vector<FaiEntry>& out)
{
    FILE* f = fopen(path.c_str(), "r");
    if (!f) return false;
    char line[65536];
    while (fgets(line, sizeof line, f))
    {
        FaiEntry e;
        char* tab = strchr(line, '\t');
        if (!tab) continue;
        e.name.assign(line, tab - line);
        long long a, b, c, d;
        if (sscanf(tab + 1, "%lld\t%lld\t%lld\t%lld", &a, &b, &c, &d) != 4) continue;
        e.len = a, e.offset = b, e.line_blen = c, e.line_len = d;
        out.push_back(e);
    }
    fclose(f);
    return true;
}

// the index built from the FASTA text itself (what fai_load does when no .fai exists; faidx.c fai_build_core)
inline void build_fai(const uint8_t* t, size_t n, std::vector<FaiEntry>& out)
{
    size_t i = 0;
    while (i < n)
    {
        if (t[i] != '>')
        {
            ++i;
            continue;
        }
        size_t j = i + 1;
        while (j < n && !isspace(t[j])) ++j;
        FaiEntry e;
        e.name.assign((const char*)&t[i + 1], j - i - 1);
        const uint8_t* nl = (const uint8_t*)memchr(t + j, '\n', n - j);
        size_t p = nl ? (size_t)(nl - t) + 1 : n;
        e.offset = (int64_t)p;
        bool first = true;
        while (p < n && t[p] != '>')
        {
            const uint8_t* le = (const uint8_t*)memchr(t + p, '\n', n - p);
            const size_t q = le ? (size_t)(le - t) : n;
            int64_t bl = 0;
            for (size_t k = p; k < q; ++k) bl += isgraph(t[k]) ? 1 : 0;
            const int64_t ll = (int64_t)(q - p) + (q < n ? 1 : 0);
            if (first && bl)
            {
                e.line_blen = bl, e.line_len = ll;
                first = false;
            }
            e.len += bl;
            p = q + 1;
        }
        out.push_back(e);
        i = p;
    }
}

// the FASTA bytes: a read-only mapping of a plain file, or the inflated stream of a gzip / bgzip file
struct FastaText
{
    const uint8_t* data = nullptr;
    size_t size = 0;
    bool mapped = false;
    std::vector<uint8_t> owned;
    FastaText() = default;
    FastaText(const FastaText&) = delete;
    FastaText& operator=(const FastaText&) = delete;
    ~FastaText() { close(); }
    void close()
    {
        if (mapped && data) munmap((void*)data, size);
        data = nullptr, size = 0, mapped = false;
        owned.clear();
        owned.shrink_to_fit();
    }
    bool open(const char* path)
    {
        close();
        int fd = ::open(path, O_RDONLY);
        if (fd < 0) return false;
        struct stat st;
        unsigned char magic[2] = {0, 0};
        const bool ok = fstat(fd, &st) == 0 && S_ISREG(st.st_mode);
        const ssize_t got = ok ? pread(fd, magic, 2, 0) : 0;
        if (ok && !(got == 2 && magic[0] == 0x1f && magic[1] == 0x8b))
        {
            if (st.st_size == 0)
            {
                ::close(fd);
                return true;
            }
            void* p = mmap(nullptr, (size_t)st.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
            ::close(fd);
            if (p == MAP_FAILED) return false;
            data = (const uint8_t*)p, size = (size_t)st.st_size, mapped = true;
            return true;
        }
        ::close(fd);
        gzFile f = gzopen(path, "rb");
        if (!f) return false;
        gzbuffer(f, 1 << 20);
        size_t cap = 1 << 26, n = 0;
        owned.resize(cap);
        for (;;)
        {
            if (n == cap) owned.resize(cap *= 2);
            const int g = gzread(f, owned.data() + n, (unsigned)std::min<size_t>(cap - n, 1u << 30));
            if (g < 0)
            {
                gzclose(f);
                return false;
            }
            if (g == 0) break;
            n += (size_t)g;
        }
        gzclose(f);
        owned.resize(n);
        data = owned.data(), size = n;
        return true;
    }
};

#ifndef __CUDACC__
enum { PACK_CHUNK = 1 << 20 }; // bases per job (a multiple of 32): the gathered bases stay in the core's L2

// bases [b0, b1) of a contig with the regular geometry (every line line_blen bases + line_len - line_blen line-end bytes, the last
// line possibly shorter).  false = the text does not look like that here.
inline bool pack_chunk_regular(const uint8_t* text, size_t n_text, const FaiEntry& e, int64_t b0, int64_t b1, uint64_t gpos0, Group* groups,
                               std::vector<ExcRun>& exc, std::vector<uint8_t>& buf)
{
    buf.resize((size_t)(b1 - b0));
    const int64_t blen = e.line_blen, llen = e.line_len;
    int64_t col = b0 % blen;
    uint64_t p = (uint64_t)e.offset + (uint64_t)(b0 / blen) * (uint64_t)llen + (uint64_t)col;
    int64_t b = b0;
    while (b < b1)
    {
        const int64_t take = std::min(blen - col, b1 - b);
        if (p + (uint64_t)take > n_text) return false;
        memcpy(buf.data() + (b - b0), text + p, (size_t)take);
        b += take;
        if (col + take == blen)
        { // a full line: its end bytes must not be bases (a longer line would shift everything after it)
            const uint64_t le = p + (uint64_t)take, ln = le + (uint64_t)(llen - blen);
            for (uint64_t q = le; q < ln && q < n_text; ++q)
                if (isgraph(text[q])) return false;
            p = ln;
            col = 0;
        }
        else
            p += (uint64_t)take;
    }
    return pack_bases(buf.data(), b0, b1, gpos0, groups, exc);
}

// the sequential stream fai_retrieve reads: isgraph bytes from the contig's offset
inline bool pack_contig_stream(const uint8_t* text, size_t n_text, const FaiEntry& e, uint64_t gpos0, Group* groups, std::vector<ExcRun>& exc)
{
    std::vector<uint8_t> buf((size_t)PACK_CHUNK);
    size_t p = (size_t)e.offset;
    for (int64_t b0 = 0; b0 < e.len; b0 += PACK_CHUNK)
    {
        const int64_t b1 = std::min<int64_t>(e.len, b0 + PACK_CHUNK);
        size_t k = 0;
        while (k < (size_t)(b1 - b0) && p < n_text)
        {
            const uint8_t ch = text[p++];
            if (isgraph(ch)) buf[k++] = ch;
        }
        if (k < (size_t)(b1 - b0)) return false;
        pack_bases(buf.data(), b0, b1, gpos0, groups, exc);
    }
    return true;
}

// Parallel packer over a list of contigs.  groups: the host array (contig c starts at hoff[c] groups there; its global base position is
// goff[c] * 32); exc[c]: the contig's exception runs in order.  on_done(c) is called from the calling thread, in list order, as soon as
// contig c is complete -- the caller uploads it while the workers pack the next ones.  Returns the index of a contig shorter than its
// index entry, or -1.
template <class Done>
int pack_contigs(const uint8_t* text, size_t n_text, const std::vector<FaiEntry>& fai, const std::vector<int>& which, const std::vector<uint64_t>& goff,
                 const std::vector<uint64_t>& hoff, Group* groups, std::vector<std::vector<ExcRun>>& exc, unsigned nthr, Done on_done)
{
    struct Job
    {
        int w; // index into which
        int64_t b0, b1;
        std::vector<ExcRun> exc;
        bool ok = true;
    };
    std::vector<Job> jobs;
    std::vector<size_t> first_job(which.size() + 1, 0);
    std::vector<char> regular(which.size(), 0);
    for (size_t w = 0; w < which.size(); ++w)
    {
        const FaiEntry& e = fai[which[w]];
        first_job[w] = jobs.size();
        regular[w] = e.line_blen > 0 && e.line_len >= e.line_blen && e.offset >= 0;
        if (regular[w])
            for (int64_t b = 0; b < e.len; b += PACK_CHUNK) jobs.push_back(Job{(int)w, b, std::min<int64_t>(e.len, b + PACK_CHUNK), {}, true});
        else if (e.len > 0)
            jobs.push_back(Job{(int)w, 0, e.len, {}, true}); // one sequential job
    }
    first_job[which.size()] = jobs.size();
    std::vector<std::atomic<uint32_t>> left(which.size());
    for (size_t w = 0; w < which.size(); ++w) left[w].store((uint32_t)(first_job[w + 1] - first_job[w]));
    std::atomic<size_t> next{0};
    nthr = std::max(1u, std::min<unsigned>(nthr, (unsigned)std::max<size_t>(1, jobs.size())));
    auto work = [&]() {
        std::vector<uint8_t> buf;
        for (;;)
        {
            const size_t j = next.fetch_add(1);
            if (j >= jobs.size()) return;
            Job& J = jobs[j];
            const FaiEntry& e = fai[which[J.w]];
            Group* g = groups + hoff[which[J.w]];
            const uint64_t gpos0 = goff[which[J.w]] << 5;
            if (regular[J.w])
                J.ok = pack_chunk_regular(text, n_text, e, J.b0, J.b1, gpos0, g, J.exc, buf);
            else
                J.ok = pack_contig_stream(text, n_text, e, gpos0, g, J.exc);
            left[J.w].fetch_sub(1, std::memory_order_release);
        }
    };
    std::vector<std::thread> th;
    for (unsigned t = 1; t < nthr; ++t) th.emplace_back(work);
    int bad = -1;
    if (nthr == 1) work();
    for (size_t w = 0; w < which.size() && bad < 0; ++w)
    {
        while (left[w].load(std::memory_order_acquire) != 0)
        { // the calling thread helps instead of spinning
            const size_t j = next.fetch_add(1);
            if (j < jobs.size())
            {
                std::vector<uint8_t> buf;
                Job& J = jobs[j];
                const FaiEntry& e = fai[which[J.w]];
                Group* g = groups + hoff[which[J.w]];
                const uint64_t gpos0 = goff[which[J.w]] << 5;
                J.ok = regular[J.w] ? pack_chunk_regular(text, n_text, e, J.b0, J.b1, gpos0, g, J.exc, buf) : pack_contig_stream(text, n_text, e, gpos0, g, J.exc);
                left[J.w].fetch_sub(1, std::memory_order_release);
            }
            else
                std::this_thread::yield();
        }
        const int c = which[w];
        bool ok = true;
        for (size_t j = first_job[w]; j < first_job[w + 1]; ++j) ok = ok && jobs[j].ok;
        exc[c].clear();
        if (!ok && regular[w])
        { // irregular text: redo this contig the sequential way
            ok = pack_contig_stream(text, n_text, fai[c], goff[c] << 5, groups + hoff[c], exc[c]);
        }
        else
            for (size_t j = first_job[w]; j < first_job[w + 1]; ++j) append_runs(exc[c], jobs[j].exc);
        if (!ok)
            bad = c;
        else
            on_done(c);
    }
    for (auto& t : th) t.join();
    return bad;
}

#endif // !__CUDACC__

// out of line, in vtn_fasta_pack.cpp (what vtn_ref.cu calls)
bool pack_bases_ol(const uint8_t* seq, int64_t b0, int64_t b1, uint64_t gpos0, Group* groups, std::vector<ExcRun>& exc);
int pack_contigs_ol(const uint8_t* text, size_t n_text, const std::vector<FaiEntry>& fai, const std::vector<int>& which, const std::vector<uint64_t>& goff,
                    const std::vector<uint64_t>& hoff, Group* groups, std::vector<std::vector<ExcRun>>& exc, unsigned nthr, const std::function<void(int)>& on_done);

} // namespace vtf
