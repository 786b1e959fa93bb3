// vtn_tools.cpp -- libvtn_tools.so: the host shim that keeps vt's tool entry points and command surface for the path (include/vtn_tools.h):  `vtn normalize`, `vtn decompose`, `vtn uniq`,
// `vtn decompose_blocksub`, `vtn annotate_indels`.
//
// Mirrors the reference tools' flags, stderr text and VCF output (normalize.cpp:84-171, :332-382; decompose.cpp:
// 80-150, :960-995) while the per-record arithmetic runs in libvtn's CUDA kernels through the C ABI (include/vtn.h):
//   normalize : record lines -> struct-of-arrays batches on the device (vtn_vcf_tokenize) -> vtn_normalize -> changed records
//               get the new POS/alleles and INFO/OLD_VARIANT (bcf_variant2string, hts_utils.cpp:1013-1026) -> written in the
//               order of BCFOrderedWriter's insertion/flush rules (bcf_ordered_writer.cpp:141-300), computed on the device
//               (vtn_writer_order).
//   uniq      : vtn_uniq decides which duplicate survives and the reader's output order (uniq.cpp:110-156); the host merges
//               the OLD_VARIANT values of a group.
//   decompose : vtn_decompose gives the (source record, ALT index) fan-out plan and the counters; the INFO/FORMAT
//               sub-setting of each emitted record is host-side text work (decompose.cpp:210-392 plain, :393-957 -s).
//   decompose_blocksub : vtn_blocksub gives the (source record, offset, length) plan and the counters (decompose_blocksub.cpp:
//               185-421 without -a); the host writes POS/REF/ALT, OLD_CLUMPED and with -p the PS field (:352-372).
// VCF text only (plain or gzip/bgzip in, plain out); BCF and the -f filter / -i intervals are not handled here
// (SURVEY.md 8f).  Values of Type=Float fields and QUAL are re-serialised the way htslib does when it formats a
// record it has parsed (float32 storage, 6 significant digits), because the reference's output goes through that.
// No CPU fallback: without a CUDA device vtn_ctx_create fails and the tool exits 1.
#include <unistd.h>
#include <zlib.h>
#include <algorithm>
#include <atomic>
#include <cctype>
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <condition_variable>
#include <cstring>
#include <ctime>
#include <deque>
#include <list>
#include <map>
#include <mutex>
#include <string>
#include <string_view>
#include <thread>
#include <vector>
#include "../../include/vtn.h"
#include "vtn_bcf.h"
#include "vtn_bgzf.h"

namespace
{
const char* VERSION = "0.5"; // normalize.cpp:88 / decompose.cpp:84

// VTN_CLI_TIMING=1: wall-clock seconds per phase on stderr at the end of `vtn normalize` (where the command's time goes)
struct PhaseTimer
{
    bool on = getenv("VTN_CLI_TIMING") != nullptr;
    std::map<std::string, double> t;
    std::vector<std::string> order;
    double t0 = 0;
    static double now()
    {
        timespec ts;
        clock_gettime(CLOCK_MONOTONIC, &ts);
        return ts.tv_sec + 1e-9 * ts.tv_nsec;
    }
    void start() { t0 = now(); }
    void stop(const char* name)
    {
        const double x = now();
        if (!t.count(name)) order.push_back(name);
        t[name] += x - t0;
        t0 = x;
    }
    void report()
    {
        if (!on) return;
        for (const std::string& k : order) fprintf(stderr, "[vtn timing] %-22s %8.3f s\n", k.c_str(), t[k]);
    }
};

typedef std::pair<const char*, size_t> Tok; // a slice of the current line

struct Tag
{
    char len_kind = '.'; // 'A','R','G','.' or 'N' (fixed count)
    int count = 0;
    char type = 'S'; // I(nteger) F(loat) S(tring) C(haracter) f(L)ag
};

typedef std::map<std::string, Tag, std::less<>> TagMap;

struct Header
{
    std::vector<std::string> meta; // ## lines
    std::string columns;           // #CHROM line
    TagMap info, fmt; // looked up by string_view: no temporary strings on the per-record paths
    int n_samples = 0;
    bool float_info() const
    {
        for (const auto& kv : info)
            if (kv.second.type == 'F') return true;
        return false;
    }
    bool float_fmt() const
    {
        for (const auto& kv : fmt)
            if (kv.second.type == 'F') return true;
        return false;
    }
};

std::string attr(const std::string& line, const char* key)
{
    std::string k = std::string(key) + "=";
    size_t p = line.find(k);
    while (p != std::string::npos && p > 0 && line[p - 1] != '<' && line[p - 1] != ',') p = line.find(k, p + 1);
    if (p == std::string::npos) return "";
    p += k.size();
    size_t e = p;
    while (e < line.size() && line[e] != ',' && line[e] != '>') ++e;
    return line.substr(p, e - p);
}

void add_tag(TagMap& m, const std::string& line)
{
    Tag t;
    std::string id = attr(line, "ID"), num = attr(line, "Number"), ty = attr(line, "Type");
    if (num == "A" || num == "R" || num == "G" || num == ".")
        t.len_kind = num[0];
    else
    {
        t.len_kind = 'N';
        t.count = atoi(num.c_str());
    }
    t.type = ty == "Integer" ? 'I' : ty == "Float" ? 'F' : ty == "Flag" ? 'L' : ty == "Character" ? 'C' : 'S';
    m[id] = t;
}

// -i / -I: genomic intervals (genome_interval.cpp:50-103: "X", "X:2000", "X:2000-4000").  The reference reads by interval only
// through an index (bcf_ordered_reader.cpp:45-130): a plain VCF, or a compressed file without .tbi / .csi, is read from the start with a
// message and the intervals are ignored.  With an index it returns, interval after interval, the records whose span [POS, POS +
// len(REF) - 1] (INFO/END when given) overlaps.  This tool does not walk the index: when one exists it makes one linear pass over the
// file per interval and keeps the overlapping records, which gives the same records in the same order.
struct Interval
{
    std::string seq;
    int32_t start1 = 1, end1 = (1 << 29) - 1;
    std::string to_string() const // genome_interval.cpp:123-137: a whole contig prints as its name
    {
        return (start1 != 1 || end1 != (1 << 29) - 1) ? seq + ":" + std::to_string(start1) + "-" + std::to_string(end1) : seq;
    }
};
void add_interval(std::vector<Interval>& v, std::map<std::string, int>& seen, const std::string& text)
{
    if (text.empty() || seen.count(text)) return;
    seen[text] = 1;
    std::vector<std::string> part;
    size_t b = 0;
    for (size_t k = 0; k <= text.size(); ++k)
        if (k == text.size() || text[k] == ':' || text[k] == '-' || text[k] == '\t')
        {
            if (k > b) part.push_back(text.substr(b, k - b));
            b = k + 1;
        }
    Interval iv;
    auto num = [&](const std::string& x, int32_t& out) {
        char* e = nullptr;
        const long val = strtol(x.c_str(), &e, 10);
        if (x.empty() || *e)
        {
            fprintf(stderr, "[genome_interval.cpp:95 set] Invalid genomic interval: %s\n", text.c_str());
            exit(1);
        }
        out = (int32_t)val;
    };
    if (part.empty() || part.size() > 3)
    {
        fprintf(stderr, "[genome_interval.cpp:101 set] Invalid genomic interval: %s\n", text.c_str());
        exit(1);
    }
    iv.seq = part[0];
    if (part.size() == 2)
    {
        num(part[1], iv.start1);
        iv.end1 = iv.start1;
    }
    else if (part.size() == 3)
    {
        num(part[1], iv.start1);
        num(part[2], iv.end1);
    }
    v.push_back(iv);
}
// Program::parse_intervals (program.cpp:152-190): the list file first, then the comma separated string, duplicates dropped
std::vector<Interval> parse_intervals(const std::string& list_file, const std::string& str)
{
    std::vector<Interval> v;
    std::map<std::string, int> seen;
    if (!list_file.empty())
    {
        FILE* f = fopen(list_file.c_str(), "r");
        if (f)
        {
            char line[65536];
            while (fgets(line, sizeof line, f))
            {
                std::string t(line);
                while (!t.empty() && (t.back() == '\n' || t.back() == '\r')) t.pop_back();
                add_interval(v, seen, t);
            }
            fclose(f);
        }
    }
    size_t b = 0;
    for (size_t k = 0; k <= str.size(); ++k)
        if (k == str.size() || str[k] == ',')
        {
            if (k > b) add_interval(v, seen, str.substr(b, k - b));
            b = k + 1;
        }
    return v;
}
void print_intervals(const char* option_line, const std::vector<Interval>& iv)
{ // Program::print_int_op, program.cpp:315-334
    if (iv.empty()) return;
    fputs(option_line, stderr);
    for (size_t i = 0; i < std::min<size_t>(iv.size(), 5); ++i) fprintf(stderr, "%s%s", i ? "," : "", iv[i].to_string().c_str());
    if (iv.size() > 5) fprintf(stderr, " and %zu other intervals\n", iv.size() - 5);
    else fputc('\n', stderr);
}

// text of the records that overlap the intervals: one pass over the file per interval, the header from the first pass
struct IntervalSource : vtz::ByteSource
{
    std::string path;
    std::vector<Interval> iv;
    size_t cur = 0;
    vtz::ByteSource* in = nullptr;
    std::string pend, out;
    size_t out_pos = 0;
    bool in_eof = false;
    IntervalSource(const std::string& p, const std::vector<Interval>& v, vtz::ByteSource* first) : path(p), iv(v), in(first) {}
    ~IntervalSource() override { delete in; }
    void take_line(const char* l, size_t n)
    { // n without the line end
        if (n && l[0] == '#')
        {
            if (cur == 0) out.append(l, n), out.push_back('\n');
            return;
        }
        const Interval& I = iv[cur];
        const char* t1 = (const char*)memchr(l, '\t', n);
        if (!t1 || (size_t)(t1 - l) != I.seq.size() || memcmp(l, I.seq.data(), I.seq.size())) return;
        const char* e = l + n;
        const char* t2 = (const char*)memchr(t1 + 1, '\t', (size_t)(e - t1 - 1));
        if (!t2) return;
        const long pos1 = strtol(std::string(t1 + 1, t2).c_str(), nullptr, 10);
        const char* t3 = (const char*)memchr(t2 + 1, '\t', (size_t)(e - t2 - 1));
        const char* t4 = t3 ? (const char*)memchr(t3 + 1, '\t', (size_t)(e - t3 - 1)) : nullptr;
        if (!t4) return;
        long end1 = pos1 + (t4 - t3 - 1) - 1;
        // INFO/END (tbx_parse1 for the VCF preset; vcf_parse for BCF)
        const char* c = t4;
        for (int k = 0; k < 3 && c; ++k) c = (const char*)memchr(c + 1, '\t', (size_t)(e - c - 1));
        if (c)
        {
            const char* ie = (const char*)memchr(c + 1, '\t', (size_t)(e - c - 1));
            if (!ie) ie = e;
            const std::string info(c + 1, ie);
            size_t q = info.compare(0, 4, "END=") == 0 ? 0 : info.find(";END=");
            if (q != std::string::npos)
            {
                const long v = strtol(info.c_str() + q + (q ? 5 : 4), nullptr, 10);
                if (v >= pos1) end1 = v;
            }
        }
        if (pos1 <= I.end1 && end1 >= I.start1) out.append(l, n), out.push_back('\n');
    }
    long read(char* dst, size_t n) override
    {
        while (out_pos == out.size())
        {
            out.clear();
            out_pos = 0;
            if (in_eof)
            {
                if (++cur >= iv.size()) return 0;
                delete in;
                in = vtb2::open_any(path);
                if (!in) return -1;
                in_eof = false;
                pend.clear();
            }
            char buf[1 << 16];
            const long got = in->read(buf, sizeof buf);
            if (got < 0) return -1;
            if (got == 0)
            {
                in_eof = true;
                if (!pend.empty()) take_line(pend.data(), pend.size());
                pend.clear();
                continue;
            }
            pend.append(buf, (size_t)got);
            size_t b = 0;
            for (;;)
            {
                const char* nl = (const char*)memchr(pend.data() + b, '\n', pend.size() - b);
                if (!nl) break;
                size_t e = (size_t)(nl - pend.data());
                take_line(pend.data() + b, (e > b && pend[e - 1] == '\r') ? e - b - 1 : e - b);
                b = e + 1;
            }
            pend.erase(0, b);
        }
        const size_t k = std::min(n, out.size() - out_pos);
        memcpy(dst, out.data() + out_pos, k);
        out_pos += k;
        return (long)k;
    }
};

// line reader over a byte source: plain, gzip, or BGZF inflated block-parallel (vtn_bgzf.h); BCF2 arrives as text (vtn_bcf.h)
struct Reader
{
    vtz::ByteSource* src = nullptr;
    std::vector<char> buf;
    size_t pos = 0, end = 0;
    bool at_eof = false;
    bool open(const std::string& path, const std::vector<Interval>& intervals = std::vector<Interval>())
    {
        src = vtb2::open_any(path); // plain / gzip / BGZF text, or BCF2 (raw or BGZF) turned into text
        buf.resize(1 << 20);
        if (src && !intervals.empty())
        { // BCFOrderedReader's constructor, bcf_ordered_reader.cpp:84-128
            const bool bgzf = path != "-" && vtz::BgzfSource::sniff(path);
            auto exists = [](const std::string& p) {
                FILE* f = fopen(p.c_str(), "rb");
                if (f) fclose(f);
                return f != nullptr;
            };
            if (!bgzf && !dynamic_cast<vtb2::BcfTextSource*>(src))
                fprintf(stderr, "[E:BCFOrderedReader] no random access support for VCF file: %s\n", path.c_str());
            else if (!exists(path + ".tbi") && !exists(path + ".csi"))
                fprintf(stderr, "[E::idx_find_and_load] Could not retrieve index file for '%s'\n", path.c_str()), // htslib's own line comes first
                    fprintf(stderr, "[E:BCFOrderedReader] index cannot be loaded for %s for random access, ignoring specified intervals and reading from start.\n", path.c_str());
            else
                src = new IntervalSource(path, intervals, src);
        }
        return src != nullptr;
    }
    bool refill()
    { // keeps the unread bytes, appends more; false at the end of the input
        if (at_eof) return false;
        if (pos && pos == end) pos = end = 0;
        if (end == buf.size())
        {
            if (pos)
            {
                memmove(buf.data(), buf.data() + pos, end - pos);
                end -= pos, pos = 0;
            }
            else
                buf.resize(buf.size() * 2);
        }
        const long got = src->read(buf.data() + end, buf.size() - end);
        if (got < 0)
        {
            fprintf(stderr, "[vtn] read error (damaged compressed input?)\n");
            exit(1);
        }
        if (got == 0)
        {
            at_eof = true;
            return false;
        }
        end += (size_t)got;
        return true;
    }
    bool getline(std::string& out)
    {
        out.clear();
        size_t scanned = pos;
        for (;;)
        {
            const char* nl = (const char*)memchr(buf.data() + scanned, '\n', end - scanned);
            if (nl)
            {
                out.assign((const char*)buf.data() + pos, nl);
                pos = (size_t)(nl - buf.data()) + 1;
                if (!out.empty() && out.back() == '\r') out.pop_back();
                return true;
            }
            const size_t have = end - pos;
            if (!refill())
            {
                if (!have) return false;
                out.assign(buf.data() + pos, buf.data() + end);
                pos = end;
                return true;
            }
            scanned = pos + have; // refill may have moved the unread bytes to the front
        }
    }
    // raw bytes for the block reader: what is buffered first, then straight from the source
    long read(char* dst, size_t n)
    {
        if (pos < end)
        {
            const size_t k = std::min(n, end - pos);
            memcpy(dst, buf.data() + pos, k);
            pos += k;
            return (long)k;
        }
        if (at_eof) return 0;
        const long got = src->read(dst, n);
        if (got < 0)
        {
            fprintf(stderr, "[vtn] read error (damaged compressed input?)\n");
            exit(1);
        }
        if (got == 0) at_eof = true;
        return got;
    }
    ~Reader() { delete src; }
};

bool read_header(Reader& rd, Header& h, std::string& first_record)
{
    std::string line;
    while (rd.getline(line))
    {
        if (line.compare(0, 2, "##") == 0)
        {
            if (line.compare(0, 7, "##INFO=") == 0) add_tag(h.info, line);
            if (line.compare(0, 9, "##FORMAT=") == 0) add_tag(h.fmt, line);
            h.meta.push_back(line);
        }
        else if (!line.empty() && line[0] == '#')
        {
            h.columns = line;
            int tabs = 0;
            for (char c : line) tabs += c == '\t';
            h.n_samples = tabs > 8 ? tabs - 8 : 0;
        }
        else
        {
            first_record = line;
            return true;
        }
    }
    first_record.clear();
    return !h.columns.empty();
}

// htslib formats a parsed Float (float32) with 6 significant digits: fixed notation inside [1e-4, 999999], rounding
// half-up on the decimal expansion truncated at 1e-10, trailing zeros removed; "%g" outside (kstring.c:38-147, kputd).
// Restated here: scale to an integer count of 1e-10 units, add half a unit of the 6th significant digit, cut.
void put_float(std::string& out, const char* s, size_t n)
{
    if (n == 1 && s[0] == '.')
    {
        out.push_back('.');
        return;
    }
    std::string tmp(s, n);
    char* end = nullptr;
    double d = (double)strtof(tmp.c_str(), &end);
    if (end == tmp.c_str() || std::isnan(d) || std::isinf(d))
    { // nan / inf / not a number: as written
        out.append(s, n);
        return;
    }
    if (d == 0)
    {
        out += std::signbit(d) ? "-0" : "0";
        return;
    }
    if (d < 0)
    {
        out.push_back('-');
        d = -d;
    }
    if (!(d >= 0.0001 && d <= 999999))
    {
        char b[64];
        snprintf(b, sizeof b, "%g", d);
        out += b;
        return;
    }
    uint64_t units = (uint64_t)(d * 10000000000LL);
    uint64_t half = 5; // half of the 6th significant digit, in 1e-10 units, for d in [1e-4, 1e-3)
    for (double lim = 0.001; d >= lim && lim < 1e6; lim *= 10) half *= 10;
    units += half;
    char dig[32];
    int p = snprintf(dig, sizeof dig, "%llu", (unsigned long long)units);
    std::string r;
    if (p <= 10)
    { // d < 1
        r = "0." + std::string(10 - p, '0') + std::string(dig, dig + (p < 6 ? p : 6));
    }
    else
    {
        int ni = p - 10;
        r.assign(dig, dig + ni);
        if (ni < 6)
        {
            r.push_back('.');
            r.append(dig + ni, dig + 6);
        }
    }
    if (r.find('.') != std::string::npos)
    {
        while (r.back() == '0') r.pop_back();
        if (r.back() == '.') r.pop_back();
    }
    out += r;
}

void put_float_list(std::string& out, const char* s, size_t n)
{
    size_t b = 0;
    for (size_t k = 0; k <= n; ++k)
        if (k == n || s[k] == ',')
        {
            if (b) out.push_back(',');
            put_float(out, s + b, k - b);
            b = k + 1;
        }
}

std::vector<Tok> split(const char* s, size_t n, char sep)
{
    std::vector<Tok> v;
    size_t b = 0;
    for (size_t k = 0; k <= n; ++k)
        if (k == n || s[k] == sep)
        {
            v.push_back(Tok(s + b, k - b));
            b = k + 1;
        }
    return v;
}

// true when htslib would print this Float value exactly as it is written: "." or a plain decimal -?(0|[1-9]d*)(.d*[1-9])? with at most
// six significant digits inside [0.0001, 999999] (float32 keeps such a decimal and kputd prints its six-digit rounding without
// trailing zeros: the same characters).  Anything else (exponents, more digits, trailing zeros, "+", "nan") goes through put_float.
bool canon_float(const char* p, size_t n)
{
    if (n == 1 && p[0] == '.') return true;
    size_t i = 0;
    if (i < n && p[i] == '-') ++i;
    if (i >= n) return false;
    size_t int_digits = 0, sig = 0;
    bool started = false; // a non-zero digit has been seen
    if (p[i] == '0')
    {
        ++i;
        int_digits = 1;
        if (i < n && p[i] != '.') return false; // "00", "01"
    }
    else
    {
        while (i < n && p[i] >= '0' && p[i] <= '9') ++i, ++int_digits, ++sig, started = true;
        if (!int_digits || int_digits > 6) return false;
    }
    if (i == n) return true; // an integer of at most six digits, or "0"
    if (p[i] != '.') return false;
    ++i;
    if (i == n) return false; // "1."
    size_t lead_zeros = 0;
    for (; i < n; ++i)
    {
        if (p[i] < '0' || p[i] > '9') return false;
        if (!started && p[i] == '0')
        {
            ++lead_zeros;
            continue;
        }
        started = true;
        ++sig;
    }
    if (p[n - 1] == '0') return false; // trailing zero ("0.0", "1.50")
    return sig <= 6 && lead_zeros <= 3;
}
bool canon_float_list(const char* p, size_t n)
{
    size_t b = 0;
    for (size_t k = 0; k <= n; ++k)
        if (k == n || p[k] == ',')
        {
            if (!canon_float(p + b, k - b)) return false;
            b = k + 1;
        }
    return true;
}

// true when htslib would print the columns from QUAL on exactly as they were read (what put_tail would produce is the input): QUAL and
// every Float-typed INFO / FORMAT value already in htslib's form, no empty INFO field to drop, every sample with as many values as
// FORMAT has keys.  p[0..n) = the line from QUAL to its end, without the line end.
bool tail_is_plain(const Header& h, bool float_info, bool float_fmt, const char* p, size_t n)
{
    size_t i = 0;
    while (i < n && p[i] != '\t') ++i;
    if (!canon_float(p, i)) return false;
    if (i >= n) return false;
    ++i;
    while (i < n && p[i] != '\t') ++i; // FILTER
    if (i >= n) return false;
    const size_t i0 = ++i;
    size_t f0 = i0; // start of the current INFO field
    while (i <= n)
    {
        if (i == n || p[i] == '\t' || p[i] == ';')
        {
            if (i == f0) return false; // an empty field (";;", leading or trailing ';') or an empty column
            if (float_info)
            {
                const char* eq = (const char*)memchr(p + f0, '=', i - f0);
                if (eq)
                {
                    auto it = h.info.find(std::string_view(p + f0, (size_t)(eq - (p + f0))));
                    if (it != h.info.end() && it->second.type == 'F' && !canon_float_list(eq + 1, (size_t)(p + i - eq - 1))) return false;
                }
            }
            if (i == n || p[i] == '\t') break;
            f0 = i + 1;
        }
        ++i;
    }
    if (i >= n) return true; // sites only
    // FORMAT
    ++i;
    const size_t k0 = i;
    size_t keys = 1;
    while (i < n && p[i] != '\t') keys += p[i++] == ':';
    uint64_t fl = 0; // bit k: key k is Float-typed
    if (float_fmt)
    {
        if (keys > 64) return false;
        size_t b = k0, k = 0;
        for (size_t q = k0; q <= i; ++q)
            if (q == i || p[q] == ':')
            {
                auto it = h.fmt.find(std::string_view(p + b, q - b));
                if (it != h.fmt.end() && it->second.type == 'F') fl |= 1ull << k;
                b = q + 1;
                ++k;
            }
    }
    while (i < n)
    {
        ++i;
        size_t vals = 1, b = i;
        for (;; ++i)
        {
            if (i == n || p[i] == '\t' || p[i] == ':')
            {
                if (fl && ((fl >> (vals - 1)) & 1) && !canon_float_list(p + b, i - b)) return false;
                if (i == n || p[i] == '\t') break;
                ++vals;
                b = i + 1;
            }
        }
        if (vals != keys) return false;
    }
    return true;
}

// INFO column as htslib re-emits it: Float-typed values canonicalised, everything else as written
void put_info(std::string& out, const Header& h, const char* s, size_t n)
{
    if (n == 1 && s[0] == '.')
    {
        out.push_back('.');
        return;
    }
    bool first = true;
    for (const Tok& kv : split(s, n, ';'))
    {
        if (!kv.second) continue;
        if (!first) out.push_back(';');
        first = false;
        const char* eq = (const char*)memchr(kv.first, '=', kv.second);
        if (!eq)
        {
            out.append(kv.first, kv.second);
            continue;
        }
        auto it = h.info.find(std::string_view(kv.first, (size_t)(eq - kv.first)));
        out.append(kv.first, eq - kv.first + 1);
        if (it != h.info.end() && it->second.type == 'F')
            put_float_list(out, eq + 1, kv.second - (eq - kv.first) - 1);
        else
            out.append(eq + 1, kv.second - (eq - kv.first) - 1);
    }
    if (first) out.push_back('.');
}

// one sample column, FORMAT keys given; Float-typed values canonicalised
void put_sample(std::string& out, const Header& h, const std::vector<Tok>& keys, const char* s, size_t n)
{
    std::vector<Tok> vals = split(s, n, ':');
    for (size_t k = 0; k < keys.size(); ++k)
    {
        if (k) out.push_back(':');
        if (k >= vals.size())
        {
            out.push_back('.');
            continue;
        }
        auto it = h.fmt.find(std::string_view(keys[k].first, keys[k].second));
        if (it != h.fmt.end() && it->second.type == 'F')
            put_float_list(out, vals[k].first, vals[k].second);
        else
            out.append(vals[k].first, vals[k].second);
    }
}

// columns 6.. of a record (QUAL FILTER INFO [FORMAT samples]) re-emitted; extra_info ("KEY=VALUE") goes into INFO the way
// bcf_update_info_string puts it there: an existing KEY keeps its place and gets the new value, a new KEY goes last
void put_tail(std::string& out, const Header& h, const std::vector<Tok>& col, const std::string& extra_info)
{
    if (extra_info.empty() && col.size() >= 8)
    { // nothing to add and nothing htslib would print differently: the columns go out as they came in
        const char* t0 = col[5].first;
        const size_t tn = (size_t)(col.back().first + col.back().second - t0);
        bool contiguous = true; // the columns are slices of one line (a caller may have swapped one for a rebuilt string)
        for (size_t c = 6; c < col.size() && contiguous; ++c) contiguous = col[c].first == col[c - 1].first + col[c - 1].second + 1;
        if (contiguous && tail_is_plain(h, true, true, t0, tn))
        {
            out.push_back('\t');
            out.append(t0, tn);
            return;
        }
    }
    out.push_back('\t');
    put_float(out, col[5].first, col[5].second);
    out.push_back('\t');
    out.append(col[6].first, col[6].second);
    out.push_back('\t');
    if (extra_info.empty())
        put_info(out, h, col[7].first, col[7].second);
    else if (col[7].second == 1 && col[7].first[0] == '.')
        out += extra_info;
    else
    {
        const size_t kl = extra_info.find('=') == std::string::npos ? extra_info.size() : extra_info.find('=');
        std::string info;
        bool replaced = false;
        for (const Tok& kv : split(col[7].first, col[7].second, ';'))
        {
            if (!kv.second) continue;
            if (!info.empty()) info.push_back(';');
            if (kv.second >= kl && !memcmp(kv.first, extra_info.data(), kl) && (kv.second == kl || kv.first[kl] == '='))
            {
                info += extra_info;
                replaced = true;
            }
            else
                info.append(kv.first, kv.second);
        }
        if (!replaced) info += (info.empty() ? "" : ";") + extra_info;
        put_info(out, h, info.data(), info.size());
    }
    if (col.size() > 8)
    {
        out.push_back('\t');
        out.append(col[8].first, col[8].second);
        std::vector<Tok> keys = split(col[8].first, col[8].second, ':');
        for (size_t c = 9; c < col.size(); ++c)
        {
            out.push_back('\t');
            put_sample(out, h, keys, col[c].first, col[c].second);
        }
    }
}

// BCFOrderedWriter (bcf_ordered_writer.cpp:141-300) with the ORDER computed on the device (vtn_writer_order): records are
// collected per batch, the device returns the order in which BCFOrderedWriter would have written them and how many are still inside its buffer; those are
// carried into the next batch (replaying them rebuilds the buffer, include/vtn.h).  stderr text of a record (its own
// messages, then the writer's warning for it) is printed in arrival order, like the sequential reference.
struct DeviceWriter
{
    struct Rec
    {
        int rid;
        int32_t pos1;
        std::string chrom, line;
        std::string msg; // stderr text of this record (printed when the record is handed to the writer)
    };
    FILE* out = stdout;
    int window = 0;
    vtn_ctx* ctx = nullptr;
    std::vector<Rec> pend;
    size_t n_carry = 0;
    std::vector<int32_t> rid, pos, late;
    std::vector<uint32_t> order;
    void emit(const Rec& r)
    {
        fwrite(r.line.data(), 1, r.line.size(), out);
        fputc('\n', out);
    }
    void write(Rec&& v)
    {
        if (!window)
        {
            if (!v.msg.empty()) fputs(v.msg.c_str(), stderr);
            emit(v);
            return;
        }
        pend.push_back(std::move(v));
    }
    void drain(bool final)
    {
        const size_t n = pend.size();
        if (!n) return;
        rid.resize(n), pos.resize(n), late.resize(n), order.resize(n);
        for (size_t k = 0; k < n; ++k) rid[k] = pend[k].rid, pos[k] = pend[k].pos1;
        uint32_t n_pending = 0;
        if (vtn_writer_order(ctx, (uint32_t)n, rid.data(), pos.data(), window, order.data(), late.data(), &n_pending))
        {
            fprintf(stderr, "[vtn] vtn_writer_order: %s\n", vtn_last_error(ctx));
            exit(1);
        }
        for (size_t k = n_carry; k < n; ++k)
        {
            if (!pend[k].msg.empty()) fputs(pend[k].msg.c_str(), stderr);
            if (late[k] != INT32_MIN)
            { // bcf_ordered_writer.cpp:163-179
                const int32_t front = late[k], cutoff = std::max(front - window, 1);
                fprintf(stderr, "[bcf_ordered_writer.cpp:170 write] Might not be sorted for window size %d at current record %s:%d < %d (%d [last record] - %d), please increase window size to at least %d.\n",
                        window, pend[k].chrom.c_str(), pend[k].pos1, cutoff, front, window, front - pend[k].pos1 + 1);
            }
        }
        const size_t done = final ? n : n - n_pending;
        for (size_t k = 0; k < done; ++k) emit(pend[order[k]]);
        std::vector<uint32_t> keep(order.begin() + done, order.end());
        std::sort(keep.begin(), keep.end()); // arrival order
        std::vector<Rec> next;
        next.reserve(keep.size());
        for (uint32_t k : keep) next.push_back(std::move(pend[k]));
        pend.swap(next);
        n_carry = pend.size();
    }
};

struct Args
{
    std::string in = "-", out = "-", ref, fexp, interval_list, interval_str;
    std::vector<Interval> intervals;
    int window = 10000;
    bool quiet = false, debug = false, smart = false, opt_n = false, opt_m = false;
    int device = 0;
};

void usage_and_exit(const char* cmd, const char* msg)
{
    fprintf(stderr, "%s\nusage: vtn %s [options] <in.vcf>\n", msg, cmd);
    exit(1);
}

Args parse_args(const char* cmd, int argc, char** argv)
{
    Args a;
    if (!strcmp(cmd, "decompose")) a.window = 0;
    bool have_in = false;
    for (int k = 1; k < argc; ++k)
    {
        std::string s = argv[k];
        auto need = [&](const char* what) -> std::string {
            if (k + 1 >= argc) usage_and_exit(cmd, (std::string("missing value for ") + what).c_str());
            return argv[++k];
        };
        if (s == "-o")
            a.out = need("-o");
        else if (s == "-r")
            a.ref = need("-r");
        else if (s == "-w")
            a.window = atoi(need("-w").c_str());
        else if (s == "-q")
            a.quiet = true;
        else if (s == "-d")
            a.debug = true;
        else if (s == "-s")
            a.smart = true;
        else if (s == "-n")
            a.opt_n = true;
        else if (s == "-m")
            a.opt_m = true;
        else if (s == "--device")
            a.device = atoi(need("--device").c_str());
        else if (s == "-f")
            a.fexp = need("-f");
        else if (s == "-i")
            a.interval_str = need("-i");
        else if (s == "-I")
            a.interval_list = need("-I");
        else if (s == "-" || s[0] != '-')
        {
            a.in = s;
            have_in = true;
        }
        else
            usage_and_exit(cmd, ("unknown option " + s).c_str());
    }
    (void)have_in;
    a.intervals = parse_intervals(a.interval_list, a.interval_str);
    return a;
}

void die(vtn_ctx* ctx, const char* what)
{
    fprintf(stderr, "[vtn] %s: %s\n", what, vtn_last_error(ctx));
    exit(1);
}

// The output by file name, like BCFOrderedWriter (bcf_ordered_writer.cpp:35-83): "-" and .vcf plain text, .vcf.gz BGZF (deflated by all
// host threads, vtn_bgzf.h), .bcf BGZF-compressed BCF2, .ubcf and "+" (stdout) uncompressed BCF2 (vtn_bcf.h).  The tools write text into
// the FILE* they get; for the other formats that FILE* is a stream whose write function feeds the encoder / block writer (fopencookie),
// and fclose finishes the file (last blocks + the end-of-file member).
struct OutCookie
{
    vtz::BgzfWriter bz;
    vtb2::BcfSink bcf;
    bool use_bz = false, use_bcf = false;
    FILE* under = nullptr;
};
ssize_t out_cookie_write(void* c, const char* p, size_t n)
{
    OutCookie* k = (OutCookie*)c;
    const bool ok = k->use_bcf ? k->bcf.write(p, n) : k->bz.write(p, n);
    return ok ? (ssize_t)n : 0;
}
int out_cookie_close(void* c)
{
    OutCookie* k = (OutCookie*)c;
    const bool ok = k->use_bcf ? k->bcf.close() : k->bz.close();
    const int rc = k->under == stdout ? fflush(stdout) : fclose(k->under);
    delete k;
    if (!ok || rc)
    {
        fprintf(stderr, "[vtn] write error\n");
        _exit(1);
    }
    return 0;
}
bool ends_with(const std::string& s, const char* suf)
{
    const size_t n = strlen(suf);
    return s.size() >= n && !s.compare(s.size() - n, n, suf);
}

FILE* open_out(const std::string& path)
{
    if (path == "-") return stdout;
    const bool gz = ends_with(path, ".vcf.gz"), bcf = ends_with(path, ".bcf"), ubcf = path == "+" || ends_with(path, ".ubcf");
    if (!gz && !bcf && !ubcf && !ends_with(path, ".vcf"))
    {
        fprintf(stderr, "[bcf_ordered_writer.cpp:78 BCFOrderedWriter] Not a VCF/BCF file: %s\n", path.c_str());
        exit(1);
    }
    FILE* f = path == "+" ? stdout : fopen(path.c_str(), "w");
    if (!f)
    {
        fprintf(stderr, "[bcf_ordered_writer.cpp:86 BCFOrderedWriter] Cannot open VCF/BCF file for writing: %s\n", path.c_str());
        exit(1);
    }
    if (!gz && !bcf && !ubcf) return f;
    OutCookie* k = new OutCookie();
    k->under = f;
    k->use_bz = gz || bcf, k->use_bcf = bcf || ubcf;
    if (k->use_bz) k->bz.open(f, -1);
    k->bcf.f = f;
    k->bcf.bz = bcf ? &k->bz : nullptr;
    cookie_io_functions_t io = {nullptr, out_cookie_write, nullptr, out_cookie_close};
    FILE* s = fopencookie(k, "w", io);
    if (!s)
    {
        fprintf(stderr, "[vtn] cannot open %s for writing\n", path.c_str());
        exit(1);
    }
    setvbuf(s, nullptr, _IOFBF, 1 << 20);
    return s;
}

void write_header(FILE* out, const Header& h, const char* extra_line)
{
    // htslib parses the first line, registers FILTER PASS, then the rest (a repeated PASS line is dropped): vcf.c:809-829
    static const char* pass_line = "##FILTER=<ID=PASS,Description=\"All filters passed\">";
    bool pass_done = false;
    for (size_t i = 0; i < h.meta.size(); i++)
    {
        const std::string& l = h.meta[i];
        if (i > 0 && l.compare(0, 18, "##FILTER=<ID=PASS,") == 0) continue;
        fprintf(out, "%s\n", l.c_str());
        if (i == 0)
        {
            fprintf(out, "%s\n", pass_line);
            pass_done = true;
        }
    }
    if (!pass_done) fprintf(out, "%s\n", pass_line);
    if (extra_line)
    { // bcf_hdr_append of an ID the dictionary already holds leaves the header as it is (re-normalizing vt's own output)
        const std::string all(extra_line);
        size_t b = 0;
        while (b <= all.size())
        {
            size_t e = all.find('\n', b);
            if (e == std::string::npos) e = all.size();
            const std::string l = all.substr(b, e - b);
            bool dup = false;
            if (l.compare(0, 7, "##INFO=") == 0 || l.compare(0, 9, "##FORMAT=") == 0)
            {
                const std::string id = attr(l, "ID"), cls = l.substr(0, l.find('=') + 1);
                for (const std::string& mline : h.meta)
                    if (mline.compare(0, cls.size(), cls) == 0 && attr(mline, "ID") == id) dup = true;
            }
            if (!l.empty() && !dup) fprintf(out, "%s\n", l.c_str());
            b = e + 1;
        }
    }
    fprintf(out, "%s\n", h.columns.c_str());
}

// ------------------------------------------------------------------------------------------------------------
// -f: the reference's filter expression (filter.cpp) for the fields this path knows on the host side of the boundary: VTYPE (from
// the device's classify_variant mask, vtn_variant_props), N_ALLELE, QUAL, PASS / FILTER.<id> / N_FILTER, integer and float literals and
// the type names SNP MNP INDEL CLUMPED VNTR SV REF, joined by && || ~ == != < <= > >= & and brackets.  The expression tree is built
// the way Filter::parse builds it (filter.cpp:1563-1650: the binary operator with the LOWEST code outside brackets, first occurrence,
// becomes the root, so && splits before || before the comparisons before &) and evaluated with Node::evaluate's rules (:56-478,
// :509-545, :1129-1165, :1322-1332; missing QUAL makes a comparison false and a ~ true), including the state it keeps between
// records: a <= between a float and an integer marks its node on the first evaluation and keeps that answer afterwards (:307).
// A record the filter rejects is skipped entirely, like `continue` in normalize.cpp:196-202.  INFO.*, strings, regular expressions,
// arithmetic, DLEN / LEN and | are not handled: the tool exits with a message that says so.
// ------------------------------------------------------------------------------------------------------------
struct Filter
{
    enum
    {
        F_BOOL = 64, F_INT = 128, F_FLT = 256, F_LOGIC = 2048, F_CMP = 4096, F_MATH = 8192, F_BCF = 16384,
        F_NOT = 0 | F_LOGIC | F_BOOL, F_AND = 1 | F_LOGIC | F_BOOL, F_OR = 2 | F_LOGIC | F_BOOL,
        F_EQ = 3 | F_CMP | F_BOOL, F_NE = 4 | F_CMP | F_BOOL, F_LT = 5 | F_CMP | F_BOOL, F_LE = 6 | F_CMP | F_BOOL, F_GT = 7 | F_CMP | F_BOOL,
        F_GE = 8 | F_CMP | F_BOOL, F_MATCH = 9 | F_CMP | F_BOOL, F_NO_MATCH = 10 | F_CMP | F_BOOL,
        F_ADD = 11 | F_MATH | F_FLT, F_SUB = 12 | F_MATH | F_FLT, F_MUL = 13 | F_MATH | F_FLT, F_DIV = 14 | F_MATH | F_FLT,
        F_BIT_AND = 15 | F_INT | F_MATH | F_BOOL, F_BIT_OR = 16 | F_INT | F_MATH | F_BOOL,
        F_VTYPE = 33 | F_INT | F_BCF | F_BOOL, F_N_ALLELE = 37 | F_INT | F_BCF, F_QUAL = 38 | F_BCF | F_FLT, F_FILTER = 39 | F_BCF | F_BOOL,
        F_N_FILTER = 40 | F_INT | F_BCF
    };
    struct Node
    {
        int type = -1;
        Node *left = nullptr, *right = nullptr;
        bool value_exists = false, b = false;
        int32_t i = 0;
        float f = 0;
        std::string tag;
    };
    struct Rec // what a record offers the filter
    {
        int32_t vtype = 0, n_allele = 0;
        const char* qual = nullptr; // QUAL and FILTER columns as written
        size_t qual_len = 0;
        const char* filt = nullptr;
        size_t filt_len = 0;
    };
    Node* tree = nullptr;
    std::vector<std::string> header_filters; // IDs the header declares (bcf_has_filter answers -1, which is true, for any other)

    [[noreturn]] static void unsupported(const std::string& what)
    {
        fprintf(stderr, "[vtn] filter expression: %s is not handled by this tool (VTYPE, N_ALLELE, QUAL, PASS, FILTER.<id>, N_FILTER, numbers, type names, && || ~ == != < <= > >= & and brackets are)\n",
                what.c_str());
        exit(1);
    }
    static bool is_literal(const char* e, int len)
    {
        if (len > 1 && e[0] == '\'' && memchr(e + 1, '\'', len - 1) == e + len - 1) return true;
        for (int k = 0; k < len; ++k)
        {
            const char c = e[k];
            if (c == '=' || c == '&' || c == '|' || c == '>' || c == '<' || c == '!' || c == '+' || c == '*' || c == '/' || c == '~' || (c == '-' && k)) return false;
        }
        return true;
    }
    static bool is_bracketed(const char* e, int len)
    {
        if (len < 2 || e[0] != '(' || e[len - 1] != ')') return false;
        int open = 1;
        for (int j = 1; j < len - 1; ++j)
        {
            if (e[j] == '(') ++open;
            else if (e[j] == ')')
            {
                if (open <= 0) unsupported("an expression with unbalanced brackets");
                --open;
            }
        }
        return open == 1;
    }
    static void trim_brackets(const char*& e, int& len)
    { // only a pair that encloses the whole expression is removed (filter.cpp:2069-2133)
        while (len >= 2 && e[0] == '(' && e[len - 1] == ')')
        {
            int open = 1;
            bool nested = true;
            for (int j = 1; j < len - 1; ++j)
            {
                if (e[j] == '(') ++open;
                else if (e[j] == ')') --open;
                if (open == 0)
                {
                    nested = false;
                    break;
                }
            }
            if (!nested || open != 1) return;
            ++e;
            len -= 2;
        }
    }
    static int peek_op(const char* s, int left, int& oplen)
    {
        const bool two = left > 1;
        oplen = 2;
        if (s[0] == '&' && two && s[1] == '&') return F_AND;
        if (s[0] == '|' && two && s[1] == '|') return F_OR;
        if (s[0] == '=' && two && s[1] == '=') return F_EQ;
        if (s[0] == '=' && two && s[1] == '~') return F_MATCH;
        if (s[0] == '~' && two && s[1] == '~') return F_NO_MATCH;
        if (s[0] == '>' && two && s[1] == '=') return F_GE;
        if (s[0] == '<' && two && s[1] == '=') return F_LE;
        if (s[0] == '!' && two && s[1] == '=') return F_NE;
        oplen = 1;
        switch (s[0])
        {
        case '+': return F_ADD;
        case '-': return F_SUB;
        case '*': return F_MUL;
        case '/': return F_DIV;
        case '&': return F_BIT_AND;
        case '|': return F_BIT_OR;
        case '>': return F_GT;
        case '<': return F_LT;
        }
        oplen = -1;
        return -1;
    }
    void parse_literal(const char* e, int len, Node* n)
    { // filter.cpp:1818-2010, in its order (strncmp with the LITERAL's length, so a prefix of a keyword matches it, like there)
        auto is = [&](const char* kw) { return strncmp(e, kw, (size_t)len) == 0; };
        const std::string lit(e, (size_t)len);
        if (is("PASS"))
        {
            n->type = F_FILTER;
            n->tag = lit;
            return;
        }
        if (!strncmp(e, "REF", 3) && len >= 3) unsupported("REF"); // REF as a column wins over REF as a type name (:1828)
        if (!strncmp(e, "ALT", 3) && len >= 3) unsupported("ALT");
        if (!strncmp(e, "QUAL", 4) && len >= 4)
        {
            n->type = F_QUAL;
            return;
        }
        if (!strncmp(e, "FILTER.", 7) && len >= 7)
        {
            n->type = F_FILTER;
            n->tag.assign(e + 7, (size_t)len - 7);
            return;
        }
        if (is("N_FILTER"))
        {
            n->type = F_N_FILTER;
            return;
        }
        if (!strncmp(e, "INFO.", 5) && len >= 5) unsupported("INFO." + lit.substr(5));
        if (!strncmp(e, "VTYPE", 5) && len >= 5)
        {
            n->type = F_VTYPE;
            return;
        }
        if (is("N_ALLELE"))
        {
            n->type = F_N_ALLELE;
            return;
        }
        static const struct { const char* kw; int v; } names[] = {{"INDEL", VTN_VT_INDEL}, {"SNP", VTN_VT_SNP}, {"MNP", VTN_VT_MNP}, {"CLUMPED", VTN_VT_CLUMPED},
                                                                   {"VNTR", VTN_VT_VNTR}, {"SV", VTN_VT_SV}, {"REF", VTN_VT_REF}};
        for (const auto& k : names)
            if (is(k.kw))
            {
                n->type = F_INT;
                n->i = k.v;
                n->f = (float)k.v;
                return;
            }
        if (is("DLEN") || is("LEN") || is("VARIANT_CONTAINS_N")) unsupported(lit);
        char* end = nullptr;
        const std::string z(lit);
        const int32_t iv = (int32_t)strtoul(z.c_str(), &end, 10);
        if (end == z.c_str() + len)
        {
            n->type = F_INT;
            n->i = iv;
            n->f = (float)iv;
            return;
        }
        const float fv = (float)strtod(z.c_str(), &end);
        if (end == z.c_str() + len)
        {
            n->type = F_FLT;
            n->f = fv;
            return;
        }
        unsupported("'" + lit + "'");
    }
    void parse(const char* e, int len, Node* n)
    {
        while (len && *e == ' ') ++e, --len;
        while (len && e[len - 1] == ' ') --len;
        if (len <= 0) unsupported("an empty sub-expression");
        trim_brackets(e, len);
        if (is_literal(e, len)) return parse_literal(e, len, n);
        if (e[0] == '~' && (is_literal(e + 1, len - 1) || is_bracketed(e + 1, len - 1)))
        {
            n->type = F_NOT;
            n->left = new Node();
            return parse(e + 1, len - 1, n->left);
        }
        int type = INT_MAX, p = 0, q = 0;
        for (int r = 0; r < len; ++r)
        {
            if (e[r] == '(')
            { // bypass a bracketed expression
                int open = 0, s = r;
                for (; s < len; ++s)
                {
                    if (e[s] == '(') ++open;
                    else if (e[s] == ')' && --open == 0) break;
                }
                if (s >= len) unsupported("an expression with unbalanced brackets");
                r = s;
            }
            int oplen = -1;
            const int c = peek_op(e + r, len, oplen);
            if (c != -1)
            {
                if (c < type)
                {
                    type = c;
                    p = r - 1;
                    q = r + oplen;
                }
                r += oplen - 1;
            }
        }
        if (type == INT_MAX)
        {
            fprintf(stderr, "[filter.cpp:1636 parse] binary operator not found in \"%.*s\". Valid operators are  ==,!=,=~,~~,&&,||,&,|,+,-,*,/\n", len, e);
            exit(1);
        }
        if (type == F_MATCH || type == F_NO_MATCH || type == F_ADD || type == F_SUB || type == F_MUL || type == F_DIV || type == F_BIT_OR)
            unsupported(std::string("the operator in \"") + std::string(e, (size_t)len) + "\"");
        n->type = type;
        n->left = new Node();
        parse(e, p + 1, n->left);
        n->right = new Node();
        parse(e + q, len - q, n->right);
    }
    bool set(const std::string& exp, const std::vector<std::string>& header_meta)
    {
        if (exp.empty()) return false;
        tree = new Node();
        parse(exp.c_str(), (int)exp.size(), tree);
        header_filters.push_back("PASS");
        for (const std::string& m : header_meta)
            if (m.compare(0, 13, "##FILTER=<ID=") == 0)
            {
                const size_t e = m.find_first_of(",>", 13);
                header_filters.push_back(m.substr(13, e == std::string::npos ? std::string::npos : e - 13));
            }
        return true;
    }
    void evaluate(Node* n, const Rec& v)
    { // Node::evaluate
        n->value_exists = true;
        Node *L = n->left, *R = n->right;
        if (n->type & F_LOGIC)
        {
            if (!L->value_exists && n->type == F_NOT)
            {
                n->b = true;
                return;
            }
            if (!L->value_exists)
            {
                n->b = false;
                n->value_exists = false;
                return;
            }
            if (n->type == F_NOT)
            {
                n->b = !L->b;
                return;
            }
            if (!R->value_exists)
            {
                n->b = false;
                n->value_exists = false;
                return;
            }
            n->b = n->type == F_AND ? (L->b && R->b) : (L->b || R->b);
        }
        else if (n->type & F_CMP)
        {
            if (!L->value_exists && !R->value_exists)
            {
                n->value_exists = false;
                return;
            }
            const int op = n->type;
            if (op != F_EQ && op != F_NE && op != F_LE && op != F_GE && op != F_GT && op != F_LT) return; // a <= that was marked (:307): b stays
            if (!L->value_exists || !R->value_exists)
            {
                n->b = false;
                return;
            }
            const bool li = (L->type & F_INT) != 0, lf = (L->type & F_FLT) != 0, ri = (R->type & F_INT) != 0, rf = (R->type & F_FLT) != 0;
            // int with int as integers; a mixed pair as floats (the integer side converted), like the reference's C++ expressions
            int c = 2; // -1 less, 0 equal, 1 greater, 2 unordered
            float x = 0, y = 0;
            bool as_int = false;
            if (li && ri)
                as_int = true;
            else if (li && rf)
                x = (float)L->i, y = R->f;
            else if (!li && lf && ri)
                x = L->f, y = (float)R->i;
            else if (!li && lf && rf)
                x = L->f, y = R->f;
            else
                unsupported("a comparison between these operands");
            if (as_int)
                c = L->i < R->i ? -1 : (L->i > R->i ? 1 : 0);
            else
                c = x < y ? -1 : (x > y ? 1 : (x == y ? 0 : 2));
            switch (op)
            {
            case F_EQ: n->b = c == 0; break;
            case F_NE: n->b = c != 0; break;
            case F_LE:
                if (!li && ri) n->type |= F_INT; // filter.cpp:307
                n->b = c == -1 || c == 0;
                break;
            case F_GE: n->b = c == 1 || c == 0; break;
            case F_GT: n->b = c == 1; break;
            case F_LT: n->b = c == -1; break;
            }
        }
        else if (n->type & F_BCF)
        {
            if (n->type == F_QUAL)
            {
                if (v.qual_len == 1 && v.qual[0] == '.')
                {
                    n->f = -1;
                    n->value_exists = false;
                }
                else
                    n->f = strtof(std::string(v.qual, v.qual_len).c_str(), nullptr);
            }
            else if (n->type == F_FILTER)
            {
                const bool none = v.filt_len == 1 && v.filt[0] == '.';
                bool has = false, declared = false;
                for (const std::string& id : header_filters) declared = declared || id == n->tag;
                if (!none)
                {
                    if (!declared)
                        has = true; // bcf_has_filter returns -1 for an id the header lacks, and -1 is true (:524)
                    else
                        for (const Tok& t : split(v.filt, v.filt_len, ';')) has = has || (t.second == n->tag.size() && !memcmp(t.first, n->tag.data(), t.second));
                }
                n->b = has;
            }
            else if (n->type == F_N_FILTER)
            {
                n->i = (v.filt_len == 1 && v.filt[0] == '.') ? 0 : (int32_t)split(v.filt, v.filt_len, ';').size();
                n->f = (float)n->i;
                n->b = true;
            }
            else if (n->type == F_VTYPE)
            {
                n->i = v.vtype;
                n->b = n->i != 0;
            }
            else if (n->type == F_N_ALLELE)
            {
                n->i = v.n_allele;
                n->f = (float)n->i;
            }
        }
        else if (n->type & F_MATH)
        { // only & reaches here
            if (!L->value_exists && !R->value_exists)
            {
                n->value_exists = false;
                return;
            }
            if (!((L->type & F_INT) && (R->type & F_INT))) unsupported("& between operands that are not integers");
            n->i = L->i & R->i;
            n->b = n->i != 0;
        }
    }
    void apply(Node* n, const Rec& v)
    { // Filter::apply(Node*), filter.cpp:2308-2358: && and || do not look at the right side when the left one decides
        n->value_exists = false;
        if (n->left) apply(n->left, v);
        if (n->type & F_LOGIC)
        {
            if (n->type == F_AND)
            {
                if (n->left->value_exists && !n->left->b)
                {
                    n->b = false;
                    n->value_exists = true;
                    return;
                }
                apply(n->right, v);
            }
            else if (n->type == F_OR)
            {
                if (n->left->value_exists && n->left->b)
                {
                    n->b = true;
                    n->value_exists = true;
                    return;
                }
                apply(n->right, v);
            }
        }
        else if (n->right)
            apply(n->right, v);
        evaluate(n, v);
    }
    bool pass(const Rec& v)
    {
        if (!tree) return true;
        apply(tree, v);
        return tree->value_exists ? tree->b : false;
    }
};

// SoA batch under construction (vtn_batch)
// the record lines of a batch in one buffer, each followed by '\n' (exactly the text the device tokenizer takes); a line is a view
struct LineRef
{
    const char* p;
    size_t n;
    const char* data() const { return p; }
    size_t size() const { return n; }
    std::string str() const { return std::string(p, n); }
};
struct Lines
{
    std::string text;
    std::vector<size_t> off{0};
    size_t size() const { return off.size() - 1; }
    bool empty() const { return off.size() == 1; }
    void clear()
    {
        text.clear();
        off.assign(1, 0);
    }
    void push_back(const char* p, size_t n)
    {
        text.append(p, n);
        text.push_back('\n');
        off.push_back(text.size());
    }
    void push_back(const std::string& l) { push_back(l.data(), l.size()); }
    LineRef operator[](size_t r) const { return LineRef{text.data() + off[r], off[r + 1] - off[r] - 1}; }
    LineRef back() const { return (*this)[size() - 1]; }
};

struct BatchBuf
{
    Lines lines;
    std::vector<int32_t> pos1, rid;
    std::vector<uint32_t> allele_beg{0}, allele_off{0};
    std::vector<uint8_t> bytes;
    void clear()
    {
        lines.clear(), pos1.clear(), rid.clear(), bytes.clear();
        allele_beg.assign(1, 0);
        allele_off.assign(1, 0);
    }
    void add_allele(const char* s, size_t n)
    {
        bytes.insert(bytes.end(), s, s + n);
        allele_off.push_back((uint32_t)bytes.size());
    }
    vtn_batch view()
    {
        vtn_batch b;
        b.n_rec = (uint32_t)pos1.size();
        b.n_allele = (uint32_t)allele_off.size() - 1;
        b.n_bytes = bytes.size();
        if (bytes.empty()) bytes.push_back(0);
        b.pos1 = pos1.data(), b.rid = rid.data(), b.allele_beg = allele_beg.data(), b.allele_off = allele_off.data();
        b.bytes = bytes.data();
        return b;
    }
};

// -f over a tokenized batch: keep[r] = the filter accepts record r; VTYPE comes from the device (vtn_variant_props), QUAL / FILTER from
// the text.  Returns true when at least one record is rejected.
bool filter_records(Filter& flt, vtn_ctx* ctx, vtn_batch b, const char* text, const uint64_t* line_off, const uint32_t* tail_off, std::vector<uint8_t>& keep)
{
    const size_t n = b.n_rec;
    std::vector<uint8_t> vtype(n + 1), props(n + 1);
    if (vtn_variant_props(ctx, &b, vtype.data(), props.data())) die(ctx, "vtn_variant_props");
    keep.assign(n, 1);
    bool dropped = false;
    for (size_t r = 0; r < n; ++r)
    {
        Filter::Rec v;
        v.vtype = vtype[r];
        v.n_allele = (int32_t)(b.allele_beg[r + 1] - b.allele_beg[r]);
        const char* q = text + line_off[r] + tail_off[r];
        const char* e = text + line_off[r + 1];
        const char* t1 = (const char*)memchr(q, '\t', (size_t)(e - q));
        v.qual = q, v.qual_len = t1 ? (size_t)(t1 - q) : 0;
        const char* f0 = t1 ? t1 + 1 : q;
        const char* t2 = (const char*)memchr(f0, '\t', (size_t)(e - f0));
        v.filt = f0, v.filt_len = t2 ? (size_t)(t2 - f0) : 0;
        if (!flt.pass(v))
        {
            keep[r] = 0;
            dropped = true;
        }
    }
    return dropped;
}

// The site columns of a batch of record lines, parsed on the DEVICE (vtn_vcf_tokenize) instead of line by line on the host:
// reads up to max_lines lines (the first one is already in `line`), fills the SoA batch; rid comes from the reference
// (ref_ids) or from the order of appearance / the header dictionary in rid_cache, looked up once per run of equal CHROM.
struct LineParser
{
    std::vector<uint64_t> line_off;
    std::vector<uint32_t> chrom_len, tail_off;
    std::vector<uint8_t> status;
    Filter* filter = nullptr; // -f: records it rejects are dropped right after the parse, before the tool sees them
    bool keep_groups = false; // uniq: a batch never ends inside a run of records at one CHROM / POS
    // uniq: which duplicate leaves the reference's reader first depends on whether a record comes before / after the group, so a batch
    // carries one record of context on either side: lead_line goes in front, and with peek_next a copy of the next unread line is
    // appended (n_lead / n_peek say how many of the batch's records are such context)
    std::string lead_line;
    bool peek_next = false;
    size_t n_lead = 0, n_peek = 0;
    void tokenize(BatchBuf& bb, vtn_ctx* ctx)
    {
        const std::string& text = bb.lines.text; // already the tokenizer's input: no second copy
        const size_t commas = (size_t)std::count(text.begin(), text.end(), ',');
        const size_t n = bb.lines.size();
        line_off.resize(n + 1), chrom_len.resize(n), tail_off.resize(n), status.resize(n);
        bb.pos1.resize(n), bb.allele_beg.resize(n + 1), bb.allele_off.resize(commas + 2 * n + 1), bb.bytes.resize(text.size() + 1);
        vtn_vcf_fields f;
        memset(&f, 0, sizeof f);
        f.cap_rec = (uint32_t)n, f.cap_allele = (uint32_t)(commas + 2 * n), f.cap_bytes = text.size();
        f.line_off = line_off.data(), f.chrom_len = chrom_len.data(), f.pos1 = bb.pos1.data(), f.tail_off = tail_off.data();
        f.status = status.data(), f.allele_beg = bb.allele_beg.data(), f.allele_off = bb.allele_off.data(), f.bytes = bb.bytes.data();
        if (vtn_vcf_tokenize(ctx, text.data(), text.size(), &f)) die(ctx, "vtn_vcf_tokenize");
        bb.allele_off.resize((size_t)f.n_allele + 1);
        bb.bytes.resize((size_t)f.n_bytes);
        for (size_t r = 0; r < n; ++r)
            if (status[r])
            {
                fprintf(stderr, "[vtn] malformed VCF record (fewer than 8 columns): %.*s\n", (int)bb.lines[r].size(), bb.lines[r].data());
                exit(1);
            }
    }
    bool lines_ready = false; // read_lines has filled the batch's lines already (done while the CUDA context was still coming up)
    // host only: the lines of the next batch
    void read_lines(Reader& rd, std::string& line, bool& more, size_t max_lines, BatchBuf& bb)
    {
        bb.clear();
        n_lead = n_peek = 0;
        if (!lead_line.empty())
        {
            bb.lines.push_back(lead_line);
            n_lead = 1;
        }
        size_t bytes = 0;
        const size_t max_bytes = (size_t)1 << 30; // a batch is also cut by bytes (the tokenizer takes less than 4 GiB per call)
        auto site_len = [](const char* p, size_t n) { // length of "CHROM\tPOS"
            const char* t1 = (const char*)memchr(p, '\t', n);
            const char* t2 = t1 ? (const char*)memchr(t1 + 1, '\t', n - (size_t)(t1 + 1 - p)) : nullptr;
            return t2 ? (size_t)(t2 - p) : n;
        };
        while (more)
        {
            if (bb.lines.size() >= max_lines + n_lead || bytes >= max_bytes)
            { // full; with keep_groups the records at the last position stay together
                if (!keep_groups) break;
                const LineRef last = bb.lines.back();
                const size_t a = site_len(last.data(), last.size()), b = site_len(line.data(), line.size());
                if (a != b || memcmp(last.data(), line.data(), a)) break;
            }
            bytes += line.size() + 1;
            bb.lines.push_back(line);
            more = rd.getline(line);
        }
        if (bb.lines.size() > n_lead && peek_next && more)
        {
            bb.lines.push_back(line);
            n_peek = 1;
        }
        lines_ready = bb.lines.size() > n_lead;
    }
    bool read_batch(Reader& rd, std::string& line, bool& more, size_t max_lines, BatchBuf& bb, vtn_ctx* ctx, bool ref_ids,
                    std::map<std::string, int>& rid_cache)
    {
        if (!lines_ready) read_lines(rd, line, more, max_lines, bb);
        lines_ready = false;
        if (bb.lines.size() == n_lead) return false;
        tokenize(bb, ctx);
        if (filter)
        {
            std::vector<uint8_t> keep;
            if (filter_records(*filter, ctx, bb.view(), bb.lines.text.data(), line_off.data(), tail_off.data(), keep))
            {
                Lines kept;
                for (size_t r = 0; r < bb.lines.size(); ++r)
                    if (keep[r]) kept.push_back(bb.lines[r].data(), bb.lines[r].size());
                bb.lines.text.swap(kept.text);
                bb.lines.off.swap(kept.off);
                if (!bb.lines.empty()) tokenize(bb, ctx);
            }
        }
        const size_t n = bb.lines.size();
        bb.rid.resize(n);
        if (!n)
        { // everything filtered out: an empty batch
            bb.pos1.clear(), bb.bytes.clear();
            bb.allele_beg.assign(1, 0), bb.allele_off.assign(1, 0);
            return true;
        }
        const char* prev = nullptr;
        uint32_t prev_len = 0;
        int prev_rid = -1;
        for (size_t r = 0; r < n; ++r)
        {
            const char* c = bb.lines.text.data() + line_off[r];
            if (!(prev && prev_len == chrom_len[r] && !memcmp(prev, c, prev_len)))
            {
                std::string chrom(c, chrom_len[r]);
                auto it = rid_cache.find(chrom);
                if (it == rid_cache.end()) it = rid_cache.emplace(chrom, ref_ids ? vtn_ref_contig_id(ctx, chrom.c_str()) : (int)rid_cache.size()).first;
                prev = c, prev_len = chrom_len[r], prev_rid = it->second;
            }
            bb.rid[r] = prev_rid;
        }
        return true;
    }
};

// ------------------------------------------------------------------------------------------------------------
// Host pipeline of `normalize`: the file is read in blocks of whole lines by a reader thread, a block is tokenized and
// normalized on the device, its records are formatted by all host threads (a record the run does not touch is passed through as the bytes
// it came in with), ordered on the device and written with one write per block.
// ------------------------------------------------------------------------------------------------------------
unsigned host_threads()
{
    unsigned n = std::max(1u, std::min(16u, std::thread::hardware_concurrency()));
    if (const char* e = getenv("VTN_HOST_THREADS"))
        if (atoi(e) > 0) n = (unsigned)atoi(e);
    return n;
}

// fn(t, lo, hi) over [0, n) split into one contiguous range per thread (ranges in thread order)
template <class F>
void parallel_ranges(size_t n, unsigned nthr, F fn)
{
    nthr = (unsigned)std::max<size_t>(1, std::min<size_t>(nthr, (n + 4095) / 4096));
    if (nthr == 1)
    {
        fn(0u, (size_t)0, n);
        return;
    }
    std::vector<std::thread> th;
    for (unsigned t = 0; t < nthr; ++t) th.emplace_back([=]() { fn(t, n * t / nthr, n * (t + 1) / nthr); });
    for (auto& x : th) x.join();
}

bool g_allow_quick_exit = false; // vtn_tools_end_process_after_output(1): the command-line program ends the process after the output is complete
bool g_quick_exit = false;       // set by a tool once its output is closed, when allowed: nothing is torn down piece by piece
struct CtxStarter // vtn_ctx_create on its own thread; get() joins and ends the process like die() when there is no device
{
    std::thread th;
    vtn_ctx* ctx = nullptr;
    int rc = 0;
    std::string err;
    std::atomic<bool> done{false};
    void start(int device)
    {
        th = std::thread([this, device]() {
            rc = vtn_ctx_create(&ctx, device);
            if (rc) err = vtn_last_error(nullptr); // the text lives in the creating thread
            done.store(true);
        });
    }
    bool ready() const { return done.load(); }
    vtn_ctx* get()
    {
        if (th.joinable()) th.join();
        if (rc)
        {
            fprintf(stderr, "[vtn] vtn_ctx_create: %s\n", err.c_str());
            exit(1);
        }
        return ctx;
    }
    ~CtxStarter()
    {
        if (th.joinable()) th.join();
    }
};
struct CtxGuard // destroys the context after every buffer allocated through it (declare it right after vtn_ctx_create)
{
    vtn_ctx* c;
    ~CtxGuard()
    {
        if (c && !g_quick_exit) vtn_ctx_destroy(c);
    }
};

// page-locked host array (vtn_host_alloc): what the device calls copy from / into without a staging pass; grows, never shrinks
template <class T>
struct PinnedBuf
{
    vtn_ctx* ctx = nullptr;
    T* p = nullptr;
    size_t cap = 0;
    PinnedBuf() = default;
    PinnedBuf(const PinnedBuf&) = delete;
    PinnedBuf& operator=(const PinnedBuf&) = delete;
    ~PinnedBuf()
    {
        if (p && !g_quick_exit) vtn_host_free(ctx, p);
    }
    void resize(size_t n) // contents are not kept
    {
        if (n <= cap) return;
        if (p) vtn_host_free(ctx, p);
        p = nullptr;
        cap = n + n / 4 + 4096;
        void* q = nullptr;
        if (vtn_host_alloc(ctx, cap * sizeof(T), &q))
        {
            fprintf(stderr, "[vtn] vtn_host_alloc: %s\n", vtn_last_error(ctx));
            exit(1);
        }
        p = (T*)q;
    }
    void resize_keep(size_t n, size_t keep) // the first `keep` elements survive
    {
        if (n <= cap) return;
        T* old = p;
        p = nullptr;
        const size_t oc = cap;
        cap = 0;
        resize(n);
        if (old && keep) memcpy(p, old, std::min(keep, oc) * sizeof(T));
        if (old) vtn_host_free(ctx, old);
    }
    T* data() { return p; }
    const T* data() const { return p; }
    T& operator[](size_t i) { return p[i]; }
    const T& operator[](size_t i) const { return p[i]; }
    T* begin() { return p; }
};

// output on its own thread: the formatter fills one buffer while the other one is being written
struct AsyncOut
{
    FILE* out = stdout;
    std::vector<char> buf[2];
    size_t fill[2] = {0, 0};
    int state[2] = {0, 0}; // 0 free, 1 being filled, 2 queued for writing
    int cur = 0, next_write = 0;
    std::mutex m;
    std::condition_variable cv;
    std::thread th;
    bool stop = false;
    void start(FILE* f)
    {
        out = f;
        th = std::thread([this]() {
            for (;;)
            {
                int k;
                {
                    std::unique_lock<std::mutex> g(m);
                    cv.wait(g, [&]() { return state[next_write] == 2 || stop; });
                    if (state[next_write] != 2) return;
                    k = next_write;
                }
                if (fill[k] && fwrite(buf[k].data(), 1, fill[k], out) != fill[k])
                {
                    fprintf(stderr, "[vtn] write error\n");
                    _exit(1);
                }
                {
                    std::lock_guard<std::mutex> g(m);
                    state[k] = 0;
                    next_write ^= 1;
                }
                cv.notify_all();
            }
        });
    }
    char* acquire(size_t bytes)
    {
        std::unique_lock<std::mutex> g(m);
        cv.wait(g, [&]() { return state[cur] == 0; });
        state[cur] = 1;
        g.unlock();
        if (buf[cur].size() < bytes) buf[cur].resize(bytes + bytes / 4);
        return buf[cur].data();
    }
    void submit(size_t bytes)
    {
        {
            std::lock_guard<std::mutex> g(m);
            fill[cur] = bytes;
            state[cur] = 2;
            cur ^= 1;
        }
        cv.notify_all();
    }
    void finish()
    {
        {
            std::unique_lock<std::mutex> g(m);
            cv.wait(g, [&]() { return state[0] == 0 && state[1] == 0; });
            stop = true;
        }
        cv.notify_all();
        if (th.joinable()) th.join();
        fflush(out);
    }
};

struct TextBlock
{
    PinnedBuf<char> pinned; // page-locked storage (the H2D copy of the text needs no staging pass)
    std::vector<char> big;  // a block that outgrew it (one very long line)
    char* text = nullptr;   // whole lines, each ended by '\n' ("\r\n" already reduced to '\n')
    size_t cap = 0, n = 0, n_lines = 0, n_commas = 0;
    void ensure(size_t bytes, size_t keep)
    {
        if (bytes <= cap) return;
        if (!keep && pinned.ctx)
        {
            pinned.resize(bytes);
            text = pinned.data(), cap = pinned.cap;
            return;
        }
        std::vector<char> nb(bytes + bytes / 2);
        if (keep) memcpy(nb.data(), text, keep);
        big.swap(nb);
        text = big.data(), cap = big.size();
    }
};

// Reads the record lines of a VCF in blocks (target_bytes of whole lines, at most max_lines lines) on its own thread, two blocks ahead.
struct BlockReader
{
    Reader* f = nullptr;
    size_t target_bytes = 16u << 20, max_lines = 0;
    std::vector<char> pend; // read but not handed out yet
    TextBlock blocks[3];
    std::deque<TextBlock*> free_q, ready_q;
    std::mutex m;
    std::condition_variable cv;
    std::thread th;
    bool done = false, stop = false;

    void start(vtn_ctx* ctx, Reader* file, const std::string& first_line, bool have_first)
    {
        f = file;
        for (auto& b : blocks) b.pinned.ctx = ctx;
        if (have_first)
        {
            pend.assign(first_line.begin(), first_line.end());
            pend.push_back('\n');
        }
        else
            done = true;
        for (auto& b : blocks) free_q.push_back(&b);
        if (!done) th = std::thread([this]() { run(); });
    }
    ~BlockReader()
    {
        {
            std::lock_guard<std::mutex> g(m);
            stop = true;
        }
        cv.notify_all();
        if (th.joinable()) th.join();
    }
    TextBlock* next() // nullptr at the end of the file
    {
        std::unique_lock<std::mutex> g(m);
        cv.wait(g, [&]() { return !ready_q.empty() || done; });
        if (ready_q.empty()) return nullptr;
        TextBlock* b = ready_q.front();
        ready_q.pop_front();
        return b;
    }
    void release(TextBlock* b)
    {
        {
            std::lock_guard<std::mutex> g(m);
            free_q.push_back(b);
        }
        cv.notify_all();
    }
    void run()
    {
        bool eof = false;
        for (;;)
        {
            TextBlock* b = nullptr;
            {
                std::unique_lock<std::mutex> g(m);
                cv.wait(g, [&]() { return !free_q.empty() || stop; });
                if (stop) return;
                b = free_q.front();
                free_q.pop_front();
            }
            size_t want = std::max(target_bytes, pend.size());
            b->ensure(want + 2, 0);
            char* t = b->text;
            memcpy(t, pend.data(), pend.size());
            size_t n = pend.size(), cut = 0;
            pend.clear();
            for (;;)
            {
                while (!eof && n < want)
                {
                    const long got = f->read(t + n, want - n);
                    if (got <= 0)
                        eof = true;
                    else
                        n += (size_t)got;
                }
                if (eof)
                {
                    if (n && t[n - 1] != '\n') t[n++] = '\n';
                    cut = n;
                    break;
                }
                const void* nl = memrchr(t, '\n', n);
                if (nl)
                {
                    cut = (size_t)((const char*)nl - t) + 1;
                    break;
                }
                want *= 2; // one line longer than the block: keep reading
                b->ensure(want + 2, n);
                t = b->text;
            }
            if (max_lines)
            { // VTN_CLI_BATCH: at most that many lines per block
                size_t lines = 0, p = 0;
                while (p < cut && lines < max_lines)
                {
                    const void* nl = memchr(t + p, '\n', cut - p);
                    p = (size_t)((const char*)nl - t) + 1;
                    ++lines;
                }
                cut = p;
            }
            pend.assign(t + cut, t + n);
            n = cut;
            if (memchr(t, '\r', n))
            { // "\r\n" -> "\n" (the reference's reader drops the '\r' with the line end)
                size_t w = 0;
                for (size_t r = 0; r < n; ++r)
                    if (!(t[r] == '\r' && r + 1 < n && t[r + 1] == '\n')) t[w++] = t[r];
                n = w;
            }
            b->n = n;
            b->n_lines = (size_t)std::count(t, t + n, '\n');
            b->n_commas = (size_t)std::count(t, t + n, ',');
            const bool last = eof && pend.empty();
            {
                std::lock_guard<std::mutex> g(m);
                if (n) ready_q.push_back(b);
                if (last) done = true;
            }
            cv.notify_all();
            if (last) return;
        }
    }
};

// BCFOrderedWriter (bcf_ordered_writer.cpp:141-300) over record VIEWS: a record is (contig id of the output header, POS, pointer, length
// with its '\n'); the order comes from the device (vtn_writer_order) once per block, the bytes are gathered by all host threads into one
// buffer and written with one call.  Records still inside the reference writer's buffer after a block are copied and replayed in
// front of the next block, as DeviceWriter does.
struct FastWriter
{
    FILE* out = stdout;
    int window = 0;
    vtn_ctx* ctx = nullptr;
    unsigned nthr = 1;
    PinnedBuf<int32_t> rid, pos, late; // what vtn_writer_order copies: page-locked
    PinnedBuf<uint32_t> order;
    std::vector<const char*> ptr;
    std::vector<uint32_t> len;
    std::vector<uint64_t> ooff;
    std::vector<char> carry_text;
    AsyncOut aout;
    size_t n_carry = 0;
    // room for n_new records after the carried ones; the caller fills rid/pos/ptr/len at [n_carry, n_carry + n_new)
    size_t begin(size_t n_new)
    {
        const size_t n = n_carry + n_new;
        rid.ctx = pos.ctx = late.ctx = order.ctx = ctx;
        rid.resize_keep(n, n_carry), pos.resize_keep(n, n_carry), ptr.resize(n), len.resize(n);
        return n_carry;
    }
    // msgs: (index among the new records, stderr text) ascending
    void drain(size_t n_new, const std::vector<std::pair<uint32_t, std::string>>& msgs, bool final)
    {
        const size_t n = begin(n_new) + n_new;
        if (!n) return;
        size_t done = n;
        uint32_t n_pending = 0;
        const bool ordered = window > 0;
        if (ordered)
        {
            late.resize(n), order.resize(n);
            if (vtn_writer_order(ctx, (uint32_t)n, rid.data(), pos.data(), window, order.data(), late.data(), &n_pending))
            {
                fprintf(stderr, "[vtn] vtn_writer_order: %s\n", vtn_last_error(ctx));
                exit(1);
            }
            if (!final) done = n - n_pending;
        }
        size_t mi = 0;
        for (size_t k = n_carry; k < n; ++k)
        { // stderr in arrival order: the record's own messages, then the writer's warning for it
            const bool has_msg = mi < msgs.size() && msgs[mi].first == k - n_carry;
            const bool is_late = ordered && late[k] != INT32_MIN;
            if (has_msg) fputs(msgs[mi++].second.c_str(), stderr);
            if (is_late)
            { // bcf_ordered_writer.cpp:163-179
                const int32_t front = late[k], cutoff = std::max(front - window, 1);
                const char* tab = (const char*)memchr(ptr[k], '\t', len[k]);
                std::string chrom(ptr[k], tab ? (size_t)(tab - ptr[k]) : 0);
                fprintf(stderr, "[bcf_ordered_writer.cpp:170 write] Might not be sorted for window size %d at current record %s:%d < %d (%d [last record] - %d), please increase window size to at least %d.\n",
                        window, chrom.c_str(), pos[k], cutoff, front, window, front - pos[k] + 1);
            }
        }
        // gather the first `done` records of the order
        ooff.resize(done + 1);
        uint64_t tot = 0;
        for (size_t k = 0; k < done; ++k)
        {
            ooff[k] = tot;
            tot += len[ordered ? order[k] : k];
        }
        ooff[done] = tot;
        char* obuf = aout.acquire(tot);
        parallel_ranges(done, nthr, [&](unsigned, size_t lo, size_t hi) {
            size_t k = lo;
            while (k < hi)
            { // runs of records that are neighbours in memory (an untouched stretch of the input) move with one copy
                const size_t i = ordered ? order[k] : k;
                const char* p = ptr[i];
                size_t l = len[i], e = k + 1;
                while (e < hi)
                {
                    const size_t j = ordered ? order[e] : e;
                    if (ptr[j] != p + l) break;
                    l += len[j];
                    ++e;
                }
                memcpy(obuf + ooff[k], p, l);
                k = e;
            }
        });
        aout.submit(tot);
        // what stays inside the buffer: copied, in arrival order, in front of the next block
        std::vector<uint32_t> keep;
        if (ordered) keep.assign(order.data() + done, order.data() + n);
        std::sort(keep.begin(), keep.end());
        std::vector<char> nt;
        std::vector<uint64_t> noff;
        std::vector<int32_t> nr, np;
        std::vector<uint32_t> nl;
        for (uint32_t k : keep)
        {
            noff.push_back(nt.size());
            nt.insert(nt.end(), ptr[k], ptr[k] + len[k]);
            nr.push_back(rid[k]), np.push_back(pos[k]), nl.push_back(len[k]);
        }
        carry_text.swap(nt);
        n_carry = keep.size();
        for (size_t k = 0; k < n_carry; ++k) rid[k] = nr[k], pos[k] = np[k]; // n_carry <= n: the arrays are large enough
        len.assign(nl.begin(), nl.end());
        ptr.resize(n_carry);
        for (size_t k = 0; k < n_carry; ++k) ptr[k] = carry_text.data() + noff[k];
    }
};

// ------------------------------------------------------------------------------------------------------------
// normalize  (Igor::normalize, normalize.cpp:174-330)
// ------------------------------------------------------------------------------------------------------------
int cmd_normalize(int argc, char** argv)
{
    Args a = parse_args("normalize", argc, argv);
    const int strict_level = a.opt_n ? 0 : a.opt_m ? 1 : 2; // normalize.cpp:120-121
    if (a.ref.empty()) usage_and_exit("normalize", "required: -r <reference FASTA>");
    const bool print = !a.quiet;
    if (print)
    { // print_options, normalize.cpp:332-349
        fprintf(stderr, "normalize v%s\n\n", VERSION);
        fprintf(stderr, "options:     input VCF file                                  %s\n", a.in.c_str());
        fprintf(stderr, "         [o] output VCF file                                 %s\n", a.out.c_str());
        fprintf(stderr, "         [w] sorting window size                             %d\n", a.window);
        if (!a.fexp.empty()) fprintf(stderr, "         [f] filter                                          %s\n", a.fexp.c_str());
        fprintf(stderr, "         [m] no fail on masked reference inconsistency       %s\n", strict_level == 1 ? "true" : "false");
        fprintf(stderr, "         [n] no fail on reference inconsistency              %s\n", strict_level == 0 ? "true" : "false");
        fprintf(stderr, "         [q] quiet                                           %s\n", "false");
        fprintf(stderr, "         [d] debug                                           %s\n", a.debug ? "true" : "false");
        fprintf(stderr, "         [r] reference FASTA file                            %s\n", a.ref.c_str());
        print_intervals("         [i] intervals                                       ", a.intervals);
        fputc('\n', stderr);
    }
    PhaseTimer pt;
    pt.start();
    // The CUDA context comes up on its own thread (0.25-2 s); meanwhile the host opens the FASTA (mapping + index), opens the input,
    // reads its header and packs the contig of the first record.
    CtxGuard ctx_guard{nullptr};
    CtxStarter starter;
    starter.start(a.device);
    vtn_fasta* fa = nullptr;
    const int fa_rc = vtn_fasta_open(a.ref.c_str(), &fa);
    Reader rd;
    if (!rd.open(a.in, a.intervals))
    {
        fprintf(stderr, "[vtn] cannot open %s\n", a.in.c_str());
        return 1;
    }
    Header h;
    std::string line;
    if (!read_header(rd, h, line))
    {
        fprintf(stderr, "[vtn] %s has no VCF header\n", a.in.c_str());
        return 1;
    }
    if (fa && !line.empty())
    {
        const size_t t = line.find('\t');
        const int c = vtn_fasta_contig_id(fa, line.substr(0, t == std::string::npos ? line.size() : t).c_str());
        if (c >= 0)
        {
            vtn_fasta_prepack(fa, c); // a short FASTA is reported when the contig is required
            // a sorted input walks on to the next contigs: the host has nothing else to do until the context is there
            for (int k = c + 1; k < vtn_fasta_n_contig(fa) && !starter.ready(); ++k) vtn_fasta_prepack(fa, k);
        }
    }
    pt.stop("host: FASTA index, header, first contig");
    vtn_ctx* ctx = starter.get();
    ctx_guard.c = ctx;
    pt.stop("context (what was left of it)");
    if (fa_rc)
    { // the same call on the context gives the same failure and leaves its text where die() prints it from
        if (vtn_ref_open_fasta(ctx, a.ref.c_str())) die(ctx, "loading the reference");
    }
    else if (vtn_ref_attach(ctx, fa))
        die(ctx, "loading the reference");
    pt.stop("reference attach");
    const unsigned nthr = host_threads();
    FastWriter ow;
    ow.out = open_out(a.out);
    ow.window = a.window;
    ow.ctx = ctx;
    ow.nthr = nthr;
    write_header(ow.out, h, "##INFO=<ID=OLD_VARIANT,Number=.,Type=String,Description=\"Original chr:pos:ref:alt encoding\">");
    ow.aout.start(ow.out);
    bool float_info = false, float_fmt = false;
    for (const auto& kv : h.info) float_info = float_info || kv.second.type == 'F';
    for (const auto& kv : h.fmt) float_fmt = float_fmt || kv.second.type == 'F';

    BlockReader br;
    if (const char* e = getenv("VTN_CLI_BATCH")) // records per device batch (tests use a small one to exercise the carry-over)
        if (atol(e) > 0) br.max_lines = (size_t)atol(e);
    if (const char* e = getenv("VTN_CLI_BLOCK_MB"))
        if (atol(e) > 0) br.target_bytes = (size_t)atol(e) << 20;
    br.start(ctx, &rd, line, !line.empty());
    Filter flt;
    const bool have_filter = flt.set(a.fexp, h.meta);
    std::vector<char> ftext; // the lines of a block that pass -f

    std::map<std::string, int> rid_cache, writer_ids;
    // every array a device call reads or fills is page-locked
    PinnedBuf<uint64_t> line_off, chg_off;
    PinnedBuf<uint32_t> chrom_len, tail_off, allele_beg, allele_off, chg_rec, chg_lt, chg_le, chg_rt;
    PinnedBuf<int32_t> pos1, rid, chg_pos1;
    PinnedBuf<uint8_t> tstatus, bytes, status, vtype, chg_bytes;
    line_off.ctx = chg_off.ctx = chrom_len.ctx = tail_off.ctx = allele_beg.ctx = allele_off.ctx = chg_rec.ctx = chg_lt.ctx = chg_le.ctx = chg_rt.ctx = ctx;
    pos1.ctx = rid.ctx = chg_pos1.ctx = tstatus.ctx = bytes.ctx = status.ctx = vtype.ctx = chg_bytes.ctx = ctx;
    std::vector<int32_t> wid;
    std::vector<std::string> arena(nthr);
    std::vector<std::vector<uint32_t>> arena_recs(nthr);
    std::vector<std::pair<uint32_t, std::string>> msgs;
    vtn_norm_counters tot;
    memset(&tot, 0, sizeof tot);
    std::string last_chrom;
    int last_rid = -1, last_wid = -1;
    std::vector<int> new_contigs;
    pt.stop("open + header");
    while (TextBlock* blk = br.next())
    {
        pt.stop("wait for reader");
        const char* text = blk->text;
        size_t n = blk->n_lines, text_n = blk->n;
        vtn_vcf_fields f;
        auto tokenize = [&]() {
            line_off.resize(n + 1), chrom_len.resize(n), tail_off.resize(n), tstatus.resize(n), pos1.resize(n), allele_beg.resize(n + 1);
            allele_off.resize(blk->n_commas + 2 * n + 1), bytes.resize(text_n + 1);
            memset(&f, 0, sizeof f);
            f.cap_rec = (uint32_t)n, f.cap_allele = (uint32_t)(blk->n_commas + 2 * n), f.cap_bytes = text_n;
            f.line_off = line_off.data(), f.chrom_len = chrom_len.data(), f.pos1 = pos1.data(), f.tail_off = tail_off.data();
            f.status = tstatus.data(), f.allele_beg = allele_beg.data(), f.allele_off = allele_off.data(), f.bytes = bytes.data();
            if (vtn_vcf_tokenize(ctx, text, text_n, &f)) die(ctx, "vtn_vcf_tokenize");
            for (size_t r = 0; r < n; ++r)
                if (tstatus[r])
                {
                    fprintf(stderr, "[vtn] malformed VCF record (fewer than 8 columns): %.*s\n", (int)(line_off[r + 1] - line_off[r] - 1), text + line_off[r]);
                    exit(1);
                }
        };
        tokenize();
        if (have_filter)
        { // -f: the records the filter rejects leave the block before anything else sees them (normalize.cpp:196-202)
            vtn_batch fb;
            fb.n_rec = (uint32_t)n, fb.n_allele = f.n_allele, fb.n_bytes = f.n_bytes;
            fb.pos1 = pos1.data(), fb.rid = nullptr, fb.allele_beg = allele_beg.data(), fb.allele_off = allele_off.data(), fb.bytes = bytes.data();
            std::vector<uint8_t> keep;
            if (filter_records(flt, ctx, fb, text, line_off.data(), tail_off.data(), keep))
            {
                ftext.clear();
                size_t kept = 0;
                for (size_t r = 0; r < n; ++r)
                    if (keep[r])
                    {
                        ftext.insert(ftext.end(), text + line_off[r], text + line_off[r + 1]);
                        ++kept;
                    }
                text = ftext.data(), text_n = ftext.size(), n = kept;
                if (n) tokenize();
            }
        }
        pt.stop("tokenize");
        if (!n)
        {
            br.release(blk);
            continue;
        }
        // contig ids, once per run of equal CHROM: the FASTA's for the kernel, the order of appearance for the writer (the ordered writer
        // flushes when the record's header-dictionary id changes, bcf_ordered_writer.cpp:185-189; the FASTA id cannot stand in for it:
        // contigs the FASTA lacks all have id -1 there, and two of them in a row must still flush)
        rid.resize(n), wid.resize(n);
        for (size_t r = 0; r < n; ++r)
        {
            const char* c = text + line_off[r];
            if (!(last_chrom.size() == chrom_len[r] && !memcmp(last_chrom.data(), c, chrom_len[r])) || last_wid < 0)
            {
                last_chrom.assign(c, chrom_len[r]);
                auto it = rid_cache.find(last_chrom);
                if (it == rid_cache.end())
                {
                    it = rid_cache.emplace(last_chrom, vtn_ref_contig_id(ctx, last_chrom.c_str())).first;
                    if (it->second >= 0) new_contigs.push_back(it->second);
                }
                last_rid = it->second;
                last_wid = writer_ids.emplace(last_chrom, (int)writer_ids.size()).first->second;
            }
            rid[r] = last_rid, wid[r] = last_wid;
        }
        pt.stop("contig ids");
        for (int c : new_contigs) // first record on this contig: its bases become resident now; the next contig is packed in the background
        {
            if (vtn_ref_require(ctx, c)) die(ctx, "loading the reference");
            vtn_ref_prefetch(ctx, c + 1);
        }
        new_contigs.clear();
        pt.stop("reference load");
        vtn_batch in;
        in.n_rec = (uint32_t)n, in.n_allele = f.n_allele, in.n_bytes = f.n_bytes;
        in.pos1 = pos1.data(), in.rid = rid.data(), in.allele_beg = allele_beg.data(), in.allele_off = allele_off.data(), in.bytes = bytes.data();
        status.resize(n + 1), vtype.resize(n + 1);
        chg_rec.resize(n + 1), chg_lt.resize(n + 1), chg_le.resize(n + 1), chg_rt.resize(n + 1), chg_pos1.resize(n + 1);
        chg_off.resize(n + 2), chg_bytes.resize(in.n_bytes + 16);
        vtn_norm_result out;
        memset(&out, 0, sizeof out);
        out.status = status.data(), out.vtype = vtype.data();
        out.cap_changed = (uint32_t)n, out.cap_changed_bytes = in.n_bytes;
        out.chg_rec = chg_rec.data(), out.chg_pos1 = chg_pos1.data(), out.chg_lt = chg_lt.data(), out.chg_le = chg_le.data();
        out.chg_rt = chg_rt.data(), out.chg_off = chg_off.data(), out.chg_bytes = chg_bytes.data();
        if (vtn_normalize(ctx, &in, &out)) die(ctx, "vtn_normalize");
        pt.stop("normalize");

        // records whose REF disagrees with the FASTA (rare): their stderr text, and the first one that ends the run
        msgs.clear();
        size_t n_fmt = n;
        std::string fatal;
        for (size_t r = 0; r < n; ++r)
        {
            const uint8_t st = status[r];
            if (!(st & (VTN_ST_FETCH_FAIL | VTN_ST_REFCODE_MASK))) continue;
            const char* L = text + line_off[r];
            const std::string chrom(L, chrom_len[r]);
            const uint32_t a0 = allele_beg[r];
            const int64_t rl = (int64_t)(allele_off[a0 + 1] - allele_off[a0]);
            const char* ref = (const char*)bytes.data() + allele_off[a0];
            if (st & VTN_ST_FETCH_FAIL)
            { // faidx_fetch_seq returned NULL: variant_manip.cpp:70-75
                char b[512];
                fatal = "[E::faidx_adjust_position] The sequence \"" + chrom + "\" was not found\n";
                snprintf(b, sizeof b, "[variant_manip.cpp:72 is_not_ref_consistent] failure to extract base from fasta file: %s:%d-%d\n", chrom.c_str(),
                         pos1[r] - 1, pos1[r] - 1 + (int)rl - 1);
                fatal += b;
                fatal += "FAQ: http://genome.sph.umich.edu/wiki/Vt#1._vt_cannot_retrieve_sequences_from_my_reference_sequence_file\n";
                n_fmt = r;
                break;
            }
            const int code = st & VTN_ST_REFCODE_MASK;
            // variant_manip.cpp:94-98 then normalize.cpp:205-215
            const int64_t pos0 = pos1[r] - 1;
            std::vector<uint8_t> fa((size_t)rl + 1, 0);
            const int64_t got = vtn_ref_fetch(ctx, rid[r], pos0, pos0 + rl - 1, 0, fa.data(), rl);
            char hdrbuf[256];
            snprintf(hdrbuf, sizeof hdrbuf, "[variant_manip.cpp:96 is_not_ref_consistent] reference bases not consistent: %s:%d-%d  ", chrom.c_str(), (int)pos0, (int)(pos0 + rl - 1));
            std::string msg = hdrbuf;
            msg.append(ref, (size_t)rl);
            msg += "(REF) vs ";
            msg.append((const char*)fa.data(), (size_t)(got > 0 ? got : 0));
            msg += "(FASTA)\n";
            if (code > 2 - strict_level)
            {
                fatal = msg + "[normalize.cpp:209 normalize] Normalization not performed due to inconsistent reference sequences. (use -n or -m option to relax this)\n";
                n_fmt = r;
                break;
            }
            msg += "[normalize.cpp:214 normalize] Normalization skipped due to inconsistent reference sequences\n";
            msgs.emplace_back((uint32_t)r, std::move(msg));
        }
        pt.stop("reference messages");

        const size_t base = ow.begin(n_fmt);
        int32_t* o_rid = ow.rid.data() + base;
        int32_t* o_pos = ow.pos.data() + base;
        const char** o_ptr = ow.ptr.data() + base;
        uint32_t* o_len = ow.len.data() + base;
        const uint32_t n_changed = out.n_changed;
        parallel_ranges(n_fmt, nthr, [&](unsigned t, size_t lo, size_t hi) {
            std::string& o = arena[t];
            std::vector<uint32_t>& mine = arena_recs[t];
            o.clear(), mine.clear();
            size_t k = (size_t)(std::lower_bound(chg_rec.begin(), chg_rec.begin() + n_changed, (uint32_t)lo) - chg_rec.begin());
            for (size_t r = lo; r < hi; ++r)
            {
                const char* L = text + line_off[r];
                const size_t ll = (size_t)(line_off[r + 1] - line_off[r]); // with the '\n'
                const uint8_t st = status[r];
                const bool changed = (st & VTN_ST_CHANGED) && k < n_changed && chg_rec[k] == r;
                o_rid[r] = wid[r];
                o_pos[r] = pos1[r];
                if (!changed && !(st & VTN_ST_ENTERED) && tail_is_plain(h, float_info, float_fmt, L + tail_off[r], ll - 1 - tail_off[r]))
                {
                    o_ptr[r] = L;
                    o_len[r] = (uint32_t)ll;
                    continue;
                }
                const size_t o0 = o.size();
                const std::vector<Tok> col = split(L, ll - 1, '\t');
                const uint32_t a0 = allele_beg[r], a1 = allele_beg[r + 1];
                std::string extra;
                if (changed)
                {
                    // OLD_VARIANT = CHROM:POS:REF/ALT.. of the record after the in-place upper-casing (normalize.cpp:241-243)
                    extra = "OLD_VARIANT=";
                    extra.append(col[0].first, col[0].second);
                    extra += ":" + std::to_string(pos1[r]) + ":";
                    for (uint32_t s = a0; s < a1; ++s)
                    {
                        if (s > a0) extra.push_back('/');
                        for (uint32_t q = allele_off[s]; q < allele_off[s + 1]; ++q) extra.push_back((char)toupper(bytes[q]));
                    }
                    o_pos[r] = chg_pos1[k];
                    o.append(col[0].first, col[0].second);
                    o.push_back('\t');
                    o += std::to_string(chg_pos1[k]);
                    o.push_back('\t');
                    o.append(col[2].first, col[2].second);
                    uint64_t q = chg_off[k];
                    for (uint32_t s = a0; s < a1; ++s)
                    {
                        const uint32_t old_len = allele_off[s + 1] - allele_off[s];
                        const uint32_t new_len = old_len - chg_rt[k] + chg_le[k] - chg_lt[k];
                        o.push_back(s == a0 ? '\t' : (s == a0 + 1 ? '\t' : ','));
                        o.append((const char*)chg_bytes.data() + q, new_len);
                        q += new_len;
                    }
                    ++k;
                }
                else
                {
                    o.append(col[0].first, col[0].second);
                    o.push_back('\t');
                    o.append(col[1].first, col[1].second);
                    o.push_back('\t');
                    o.append(col[2].first, col[2].second);
                    o.push_back('\t');
                    const size_t b0 = o.size();
                    o.append(col[3].first, col[3].second);
                    o.push_back('\t');
                    o.append(col[4].first, col[4].second);
                    // entered the trim path but came out unchanged: the alleles were upper-cased in place and the VCF text
                    // writer prints them that way
                    if (st & VTN_ST_ENTERED)
                        for (size_t q = b0; q < o.size(); ++q) o[q] = (char)toupper(o[q]);
                }
                put_tail(o, h, col, extra);
                o.push_back('\n');
                o_ptr[r] = (const char*)(uintptr_t)o0; // an offset until the arena stops growing
                o_len[r] = (uint32_t)(o.size() - o0);
                mine.push_back((uint32_t)r);
            }
            for (uint32_t r : mine) o_ptr[r] = o.data() + (uintptr_t)o_ptr[r];
        });
        pt.stop("format records");
        ow.drain(n_fmt, msgs, false);
        pt.stop("order + write");
        if (!fatal.empty())
        {
            ow.aout.finish();
            fputs(fatal.c_str(), stderr);
            exit(1);
        }
        br.release(blk);
        for (int c = 0; c < 5; ++c) tot.bi[c] += out.counters.bi[c], tot.multi[c] += out.counters.multi[c];
        tot.no_variants += out.counters.no_variants;
        tot.no_refs += out.counters.no_refs;
    }
    pt.start();
    ow.drain(0, {}, true);
    ow.aout.finish();
    if (ow.out != stdout) fclose(ow.out);
    pt.stop("order + write");
    pt.report();
    if (print)
    { // print_stats, normalize.cpp:351-382
        const int32_t nb = tot.bi[0] + tot.bi[1] + tot.bi[2] + tot.bi[3] + tot.bi[4];
        const int32_t nm = tot.multi[0] + tot.multi[1] + tot.multi[2] + tot.multi[3] + tot.multi[4];
        static const char* names[5] = {"left trimmed                      ", "right trimmed                     ", "left and right trimmed            ",
                                       "right trimmed and left aligned    ", "left aligned                      "};
        fprintf(stderr, "\nstats: biallelic\n");
        for (int c = 0; c < 5; ++c) fprintf(stderr, "          no. %s: %u\n", names[c], tot.bi[c]);
        fprintf(stderr, "\n       total no. biallelic normalized           : %d\n\n       multiallelic\n", nb);
        for (int c = 0; c < 5; ++c) fprintf(stderr, "          no. %s: %u\n", names[c], tot.multi[c]);
        fprintf(stderr, "\n       total no. multiallelic normalized        : %d\n\n", nm);
        fprintf(stderr, "       total no. variants normalized            : %d\n", nb + nm);
        fprintf(stderr, "       total no. variants observed              : %u\n", tot.no_variants);
        fprintf(stderr, "       total no. reference observed             : %u\n\n", tot.no_refs);
    }
    g_quick_exit = g_allow_quick_exit; // the output file is closed: vtn_tool_main may end the process without unmapping the pinned buffers and the context one by one
    return print ? 1 : 0; // normalize() returns `print` so that main reports the elapsed time (normalize.cpp:391-400)
}

// ------------------------------------------------------------------------------------------------------------
// decompose  (Igor::decompose, decompose.cpp:189-958)
// ------------------------------------------------------------------------------------------------------------
uint32_t choose(uint32_t n, uint32_t r) // n choose r with the reference's integer arithmetic (decompose.cpp:154-186)
{
    if (r > n) return 0;
    if (r == n || r == 0) return 1;
    if (r > (n >> 1)) r = n - r;
    uint32_t num = n, den = 1;
    for (uint32_t i = 1; i < r; ++i)
    {
        num *= n - i;
        den *= i + 1;
    }
    return num / den;
}
uint32_t ap2g(uint32_t na, uint32_t pl) // hts_utils.cpp:838-856
{
    if (!pl || !na) return 0;
    if (pl == 1) return na;
    if (pl == 2) return ((na + 1) * na) >> 1;
    return choose(pl + na - 1, na - 1);
}
uint32_t ag2p(uint32_t na, uint32_t ng) // hts_utils.cpp:871-909
{
    if (na == ng) return 1;
    for (uint32_t pl = 1;; ++pl)
    {
        uint32_t k = choose(pl + na - 1, na - 1);
        if (k == ng) return pl;
        if (k > ng) return 0;
    }
}

void join(std::string& out, const std::vector<Tok>& v, bool is_float)
{
    // trailing holes (indices past the end of a short vector) are dropped; an empty vector prints as "."
    size_t n = v.size();
    while (n && !v[n - 1].first) --n;
    if (!n)
    {
        out.push_back('.');
        return;
    }
    for (size_t k = 0; k < n; ++k)
    {
        if (k) out.push_back(',');
        if (!v[k].first)
            out.push_back('.');
        else if (is_float)
            put_float(out, v[k].first, v[k].second);
        else
            out.append(v[k].first, v[k].second);
    }
}

// genotype-indexed vector -> the entries of the biallelic genotypes {0, i}: entry 0 plus the cumulative-choose indices
// (decompose.cpp:323-343, :451-511)
std::vector<Tok> pick_g(const std::vector<Tok>& g, uint32_t ploidy, uint32_t i)
{
    std::vector<Tok> o;
    const uint32_t ng2 = ap2g(2, ploidy);
    if (!ng2) return o;
    auto at = [&](uint32_t x) { return x < g.size() ? g[x] : Tok(nullptr, 0); };
    o.push_back(at(0));
    uint32_t index = 0;
    for (uint32_t k = 1; k < ng2; ++k)
    {
        index += choose(ploidy - (k - 1) + i - 1, i - 1);
        o.push_back(at(index));
    }
    return o;
}

// GT recode: allele 0 -> 0, allele i -> 1, any other -> missing; phase kept (decompose.cpp:300-321)
void put_gt(std::string& out, const char* s, size_t n, uint32_t i)
{
    size_t b = 0;
    for (size_t k = 0; k <= n; ++k)
        if (k == n || s[k] == '/' || s[k] == '|')
        {
            if (k - b == 1 && s[b] == '.')
                out.push_back('.');
            else
            {
                uint32_t a = (uint32_t)strtoul(std::string(s + b, k - b).c_str(), nullptr, 10);
                out.push_back(a == 0 ? '0' : a == i ? '1' : '.');
            }
            if (k < n) out.push_back(s[k]);
            b = k + 1;
        }
}

uint32_t gt_ploidy(const char* s, size_t n)
{
    uint32_t p = 1;
    for (size_t k = 0; k < n; ++k) p += s[k] == '/' || s[k] == '|';
    return p;
}

int cmd_decompose(int argc, char** argv)
{
    Args a = parse_args("decompose", argc, argv);
    const bool print = !a.quiet;
    if (print)
    { // print_options, decompose.cpp:960-975
        fprintf(stderr, "decompose v%s\n\n", VERSION);
        fprintf(stderr, "options:     input VCF file        %s\n", a.in.c_str());
        fprintf(stderr, "         [s] smart decomposition   %s (experimental)\n", a.smart ? "true" : "false");
        fprintf(stderr, "         [o] output VCF file       %s\n", a.out.c_str());
        print_intervals("         [i] intervals             ", a.intervals);
        fputc('\n', stderr);
    }
    PhaseTimer pt;
    pt.start();
    CtxStarter starter; // the context comes up on its own thread while the header and the first batch's lines are read
    starter.start(a.device);
    Reader rd;
    if (!rd.open(a.in, a.intervals))
    {
        fprintf(stderr, "[vtn] cannot open %s\n", a.in.c_str());
        return 1;
    }
    Header h;
    std::string line;
    if (!read_header(rd, h, line))
    {
        fprintf(stderr, "[vtn] %s has no VCF header\n", a.in.c_str());
        return 1;
    }
    FILE* out = open_out(a.out);
    write_header(out, h, "##INFO=<ID=OLD_MULTIALLELIC,Number=1,Type=String,Description=\"Original chr:pos:ref:alt encoding\">");
    const size_t BATCH = 1u << 20;
    BatchBuf bb;
    std::map<std::string, int> rid_cache;
    std::vector<uint32_t> src_rec;
    std::vector<uint16_t> alt_idx;
    vtn_decomp_counters tot;
    memset(&tot, 0, sizeof tot);
    bool more = !line.empty();
    const unsigned nthr = host_threads();
    std::vector<std::string> obuf(nthr);
    LineParser lp;
    Filter flt;
    if (flt.set(a.fexp, h.meta)) lp.filter = &flt;
    lp.read_lines(rd, line, more, BATCH, bb);
    pt.stop("header + first lines");
    vtn_ctx* ctx = starter.get();
    pt.stop("context (what was left of it)");
    while (more || lp.lines_ready)
    {
        pt.start();
        lp.read_batch(rd, line, more, BATCH, bb, ctx, false, rid_cache);
        pt.stop("read + tokenize");
        vtn_batch in = bb.view();
        const uint64_t bound = vtn_decompose_out_bound(&in);
        src_rec.resize(bound + 1), alt_idx.resize(bound + 1);
        uint64_t n_out = 0;
        vtn_decomp_counters c;
        if (vtn_decompose(ctx, &in, src_rec.data(), alt_idx.data(), &n_out, &c)) die(ctx, "vtn_decompose");
        pt.stop("decompose plan");
        tot.no_variants += c.no_variants, tot.no_biallelic += c.no_biallelic, tot.no_multiallelic += c.no_multiallelic;
        tot.no_additional_biallelic += c.no_additional_biallelic;
        // output q of the plan appended to o as one line; outputs are independent, so ranges of them are formatted by all host threads
        auto format_output = [&](uint64_t q, std::string& o) {
            const uint32_t r = src_rec[q], i = alt_idx[q];
            const LineRef ln = bb.lines[r];
            std::vector<Tok> col = split(ln.data(), ln.size(), '\t');
            const uint32_t n_allele = in.allele_beg[r + 1] - in.allele_beg[r];
            if (n_allele <= 2)
            { // passes through (decompose.cpp:946-951); still re-serialised by the writer
                o.append(col[0].first, col[4].first + col[4].second - col[0].first);
                put_tail(o, h, col, "");
                o.push_back('\n');
                return;
            }
            std::vector<Tok> alts = split(col[4].first, col[4].second, ',');
            o.append(col[0].first, col[3].first + col[3].second - col[0].first);
            o.push_back('\t');
            o.append(alts[i - 1].first, alts[i - 1].second);
            o.push_back('\t');
            put_float(o, col[5].first, col[5].second);
            o.push_back('\t');
            o.append(col[6].first, col[6].second);
            o.push_back('\t');
            // ---- INFO
            std::string old = "OLD_MULTIALLELIC=" + std::string(col[0].first, col[0].second) + ":" + std::string(col[1].first, col[1].second) + ":" +
                              std::string(col[3].first, col[3].second);
            for (const Tok& t : alts) old += "/" + std::string(t.first, t.second);
            bool first = true, old_placed = false;
            if (!(col[7].second == 1 && col[7].first[0] == '.'))
                for (const Tok& kv : split(col[7].first, col[7].second, ';'))
                {
                    if (!kv.second) continue;
                    const char* eq = (const char*)memchr(kv.first, '=', kv.second);
                    std::string key(kv.first, eq ? (size_t)(eq - kv.first) : kv.second);
                    auto it = h.info.find(key);
                    const Tag* tg = it == h.info.end() ? nullptr : &it->second;
                    std::string item;
                    if (key == "OLD_MULTIALLELIC")
                    { // bcf_update_info_string on a key the record already carries: new value, same place
                        item = old;
                        old_placed = true;
                    }
                    else if (!eq)
                        item = key;
                    else
                    {
                        const char* vs = eq + 1;
                        const size_t vn = kv.second - (eq - kv.first) - 1;
                        const bool fl = tg && tg->type == 'F';
                        const bool numeric = tg && (tg->type == 'I' || tg->type == 'F');
                        item = key + "=";
                        if (a.smart && tg && tg->len_kind == 'G' && numeric)
                        { // decompose.cpp:451-511
                            std::vector<Tok> g = split(vs, vn, ',');
                            std::vector<Tok> sel = pick_g(g, ag2p(n_allele, (uint32_t)g.size()), i);
                            if (sel.empty()) continue; // update with n=0 removes the tag
                            join(item, sel, fl);
                        }
                        else if (a.smart && tg && tg->len_kind == 'A' && (numeric || tg->type == 'S'))
                        { // decompose.cpp:517-586
                            std::vector<Tok> g = split(vs, vn, ',');
                            join(item, std::vector<Tok>{i - 1 < g.size() ? g[i - 1] : Tok(nullptr, 0)}, fl);
                        }
                        else if (a.smart && tg && tg->len_kind == 'R' && (numeric || tg->type == 'S'))
                        { // decompose.cpp:587-663
                            std::vector<Tok> g = split(vs, vn, ',');
                            join(item, std::vector<Tok>{g[0], i < g.size() ? g[i] : Tok(nullptr, 0)}, fl);
                        }
                        else if (fl)
                            put_float_list(item, vs, vn);
                        else
                            item.append(vs, vn);
                    }
                    if (!first) o.push_back(';');
                    first = false;
                    o += item;
                }
            if (!old_placed)
            {
                if (!first) o.push_back(';');
                o += old;
            }
            // ---- FORMAT / samples
            if (col.size() > 9)
            {
                std::vector<Tok> keys = split(col[8].first, col[8].second, ':');
                std::vector<int> keep;
                for (size_t k = 0; k < keys.size(); ++k)
                {
                    std::string key(keys[k].first, keys[k].second);
                    if (a.smart || key == "GT" || key == "PL" || key == "GL" || key == "DP") keep.push_back((int)k);
                }
                // ploidy = the longest GT of the record (bcf_get_genotypes pads to the maximum, decompose.cpp:247-251)
                uint32_t ploidy = 0;
                size_t n_g_max[64] = {0};
                for (size_t cidx = 9; cidx < col.size(); ++cidx)
                {
                    std::vector<Tok> vals = split(col[cidx].first, col[cidx].second, ':');
                    for (size_t k = 0; k < keys.size() && k < vals.size(); ++k)
                    {
                        if (keys[k].second == 2 && !memcmp(keys[k].first, "GT", 2)) ploidy = std::max(ploidy, gt_ploidy(vals[k].first, vals[k].second));
                        if (k < 64)
                        {
                            size_t cnt = 1;
                            for (size_t z = 0; z < vals[k].second; ++z) cnt += vals[k].first[z] == ',';
                            n_g_max[k] = std::max(n_g_max[k], cnt);
                        }
                    }
                }
                o.push_back('\t');
                if (keep.empty())
                    o.push_back('.');
                for (size_t z = 0; z < keep.size(); ++z)
                {
                    if (z) o.push_back(':');
                    o.append(keys[keep[z]].first, keys[keep[z]].second);
                }
                for (size_t cidx = 9; cidx < col.size(); ++cidx)
                {
                    o.push_back('\t');
                    if (keep.empty())
                    {
                        o.push_back('.');
                        continue;
                    }
                    std::vector<Tok> vals = split(col[cidx].first, col[cidx].second, ':');
                    for (size_t z = 0; z < keep.size(); ++z)
                    {
                        if (z) o.push_back(':');
                        const size_t k = (size_t)keep[z];
                        if (k >= vals.size())
                        {
                            o.push_back('.');
                            continue;
                        }
                        std::string key(keys[k].first, keys[k].second);
                        auto it = h.fmt.find(key);
                        const Tag* tg = it == h.fmt.end() ? nullptr : &it->second;
                        const bool fl = tg && tg->type == 'F';
                        const bool numeric = tg && (tg->type == 'I' || tg->type == 'F');
                        const bool missing = vals[k].second == 1 && vals[k].first[0] == '.';
                        if (key == "GT")
                            put_gt(o, vals[k].first, vals[k].second, i);
                        else if (missing)
                            o.push_back('.');
                        else if (!a.smart && (key == "PL" || key == "GL"))
                            join(o, pick_g(split(vals[k].first, vals[k].second, ','), ploidy, i), key == "GL");
                        else if (a.smart && tg && tg->len_kind == 'G' && numeric)
                            join(o, pick_g(split(vals[k].first, vals[k].second, ','), ag2p(n_allele, (uint32_t)(k < 64 ? n_g_max[k] : 0)), i), fl);
                        else if (a.smart && tg && tg->len_kind == 'A' && (numeric || tg->type == 'S'))
                        {
                            std::vector<Tok> g = split(vals[k].first, vals[k].second, ',');
                            join(o, std::vector<Tok>{i - 1 < g.size() ? g[i - 1] : Tok(nullptr, 0)}, fl);
                        }
                        else if (a.smart && tg && tg->len_kind == 'R' && (numeric || tg->type == 'S'))
                        {
                            std::vector<Tok> g = split(vals[k].first, vals[k].second, ',');
                            join(o, std::vector<Tok>{g[0], i < g.size() ? g[i] : Tok(nullptr, 0)}, fl);
                        }
                        else if (fl)
                            put_float_list(o, vals[k].first, vals[k].second);
                        else
                            o.append(vals[k].first, vals[k].second);
                    }
                }
            }
            o.push_back('\n');
        };
        for (auto& b : obuf) b.clear();
        parallel_ranges((size_t)n_out, nthr, [&](unsigned t, size_t lo, size_t hi) {
            for (size_t q = lo; q < hi; ++q) format_output(q, obuf[t]);
        });
        for (const std::string& b : obuf) // ranges are in thread order = output order
            if (!b.empty() && fwrite(b.data(), 1, b.size(), out) != b.size())
            {
                fprintf(stderr, "[vtn] write error\n");
                exit(1);
            }
    }
    if (out != stdout) fclose(out);
    pt.stop("format + write");
    pt.report();
    if (print)
    { // print_stats, decompose.cpp:977-993
        fprintf(stderr, "\nstats: no. variants                 : %u\n", tot.no_variants);
        fprintf(stderr, "       no. biallelic variants       : %u\n", tot.no_biallelic);
        fprintf(stderr, "       no. multiallelic variants    : %u\n\n", tot.no_multiallelic);
        fprintf(stderr, "       no. additional biallelics    : %u\n", tot.no_additional_biallelic);
        fprintf(stderr, "       total no. of biallelics      : %u\n\n", tot.no_additional_biallelic + tot.no_variants);
    }
    vtn_ctx_destroy(ctx);
    return 0;
}
} // namespace

// ------------------------------------------------------------------------------------------------------------
// uniq  (Igor::uniq, uniq.cpp:110-156): the device decides which duplicate survives and in which order records leave
// the synced reader (vtn_uniq); the host merges the OLD_VARIANT values of a group into its first record (:114-143).
// ------------------------------------------------------------------------------------------------------------
std::string info_value(const Tok& info, const char* key)
{
    const size_t kl = strlen(key);
    for (const Tok& kv : split(info.first, info.second, ';'))
        if (kv.second > kl && !memcmp(kv.first, key, kl) && kv.first[kl] == '=') return std::string(kv.first + kl + 1, kv.second - kl - 1);
    return "";
}

int cmd_uniq(int argc, char** argv)
{
    Args a = parse_args("uniq", argc, argv);
    fprintf(stderr, "uniq v0.57\n\n"); // print_options, uniq.cpp:158-165
    fprintf(stderr, "options:     input VCF file        %s\n", a.in.c_str());
    fprintf(stderr, "         [o] output VCF file       %s\n", a.out.c_str());
    CtxStarter starter;
    starter.start(a.device);
    Reader rd;
    if (!rd.open(a.in, a.intervals))
    {
        fprintf(stderr, "[vtn] cannot open %s\n", a.in.c_str());
        return 1;
    }
    Header h;
    std::string line;
    if (!read_header(rd, h, line))
    {
        fprintf(stderr, "[vtn] %s has no VCF header\n", a.in.c_str());
        return 1;
    }
    FILE* out = open_out(a.out);
    write_header(out, h, nullptr);
    // rid = index in the header's contig dictionary (contigs missing from the header are appended as met, like htslib)
    std::map<std::string, int> rid_cache;
    for (const std::string& m : h.meta)
        if (m.compare(0, 9, "##contig=") == 0)
        {
            std::string id = attr(m, "ID");
            if (!rid_cache.count(id))
            {
                const int k = (int)rid_cache.size();
                rid_cache[id] = k;
            }
        }
    // The file goes through in batches that end between position groups: duplicates share CHROM and POS, and inside a group the
    // device reproduces the reader's order (vtn_uniq.cu), so groups are independent and the outputs concatenate.
    size_t BATCH = 1u << 20;
    if (const char* e = getenv("VTN_CLI_BATCH"))
        if (atol(e) > 0) BATCH = (size_t)atol(e);
    BatchBuf bb;
    bool more = !line.empty();
    LineParser lp;
    lp.keep_groups = true;
    lp.peek_next = true;
    std::vector<uint32_t> order;
    std::vector<uint8_t> first;
    vtn_uniq_counters cnt = {0, 0};
    const unsigned nthr = host_threads();
    std::vector<std::string> obuf(nthr);
    std::vector<uint32_t> gstart;
    lp.read_lines(rd, line, more, BATCH, bb);
    vtn_ctx* ctx = starter.get();
    while (more || lp.lines_ready)
    {
    lp.read_batch(rd, line, more, BATCH, bb, ctx, false, rid_cache);
    vtn_batch in = bb.view();
    const size_t n = in.n_rec;
    if (!n) continue;
    order.resize(n + 1), first.resize(n + 1);
    vtn_uniq_counters c1 = {0, 0};
    int64_t bad = -1;
    if (vtn_uniq(ctx, &in, order.data(), first.data(), &c1, &bad)) die(ctx, "vtn_uniq");
    if (bad >= 0)
    { // fill_buffer, bcf_synced_reader.cpp:597-601
        fprintf(stderr, "[E:bcf_synced_reader.cpp:599 fill_buffer] VCF file not in order: %s\n", a.in.c_str());
        exit(1);
    }
    // the context records are position groups of their own: the first and the last one to come out; they are neither written nor counted
    const uint32_t ctx_recs = (uint32_t)(lp.n_lead + lp.n_peek);
    cnt.no_total_variants += c1.no_total_variants - ctx_recs, cnt.no_unique_variants += c1.no_unique_variants - ctx_recs;
    lp.lead_line = bb.lines[n - 1 - lp.n_peek].str();
    // the groups of this batch (the context records' groups left out), formatted by all host threads, written in order
    gstart.clear();
    for (size_t k = 0; k < n;)
    {
        size_t e = k + 1;
        while (e < n && !first[e]) ++e;
        if (!((lp.n_lead && order[k] == 0) || (lp.n_peek && order[k] == n - 1))) gstart.push_back((uint32_t)k);
        k = e;
    }
    auto format_group = [&](size_t k, std::string& out_buf) {
        size_t e = k + 1;
        while (e < n && !first[e]) ++e;
        std::string o;
        const LineRef ln = bb.lines[order[k]];
        std::vector<Tok> col = split(ln.data(), ln.size(), '\t');
        std::string merged;
        if (e - k > 1)
        { // uniq.cpp:114-143: distinct OLD_VARIANT values of the group in pop order
            std::vector<std::string> seen;
            for (size_t j = k; j < e; ++j)
            {
                const LineRef lj = bb.lines[order[j]];
                std::vector<Tok> cj = split(lj.data(), lj.size(), '\t');
                std::string v = info_value(cj[7], "OLD_VARIANT");
                if (v.empty() || std::find(seen.begin(), seen.end(), v) != seen.end()) continue;
                seen.push_back(v);
                if (!merged.empty()) merged.push_back(',');
                merged += v;
            }
        }
        o.clear();
        for (int c = 0; c < 5; ++c)
        {
            if (c) o.push_back('\t');
            o.append(col[c].first, col[c].second);
        }
        if (merged.empty())
            put_tail(o, h, col, "");
        else
        { // bcf_update_info_string: an existing OLD_VARIANT keeps its place, a new one goes last
            std::string info;
            bool replaced = false;
            for (const Tok& kv : split(col[7].first, col[7].second, ';'))
            {
                if (!kv.second || (kv.second == 1 && kv.first[0] == '.')) continue;
                if (!info.empty()) info.push_back(';');
                if (kv.second > 12 && !memcmp(kv.first, "OLD_VARIANT=", 12))
                {
                    info += "OLD_VARIANT=" + merged;
                    replaced = true;
                }
                else
                    info.append(kv.first, kv.second);
            }
            if (!replaced) info += (info.empty() ? "" : ";") + ("OLD_VARIANT=" + merged);
            std::vector<Tok> c2 = col;
            c2[7] = Tok(info.data(), info.size());
            put_tail(o, h, c2, "");
        }
        out_buf += o;
        out_buf.push_back('\n');
    };
    for (auto& b : obuf) b.clear();
    parallel_ranges(gstart.size(), nthr, [&](unsigned t, size_t lo, size_t hi) {
        for (size_t g = lo; g < hi; ++g) format_group(gstart[g], obuf[t]);
    });
    for (const std::string& b : obuf)
        if (!b.empty() && fwrite(b.data(), 1, b.size(), out) != b.size())
        {
            fprintf(stderr, "[vtn] write error\n");
            exit(1);
        }
    }
    if (out != stdout) fclose(out);
    fprintf(stderr, "\nstats: Total number of observed variants   %u\n", cnt.no_total_variants); // print_stats, uniq.cpp:167-171
    fprintf(stderr, "       Total number of unique variants     %u\n\n", cnt.no_unique_variants);
    vtn_ctx_destroy(ctx);
    return 0;
}

// ------------------------------------------------------------------------------------------------------------
// decompose_blocksub  (Igor::decompose_blocksub, decompose_blocksub.cpp:185-421; -a alignment mode is not available):
// vtn_blocksub gives the fan-out plan; each planned record keeps the source record's columns with new POS / REF / ALT and
// INFO/OLD_CLUMPED (print_decomposed_var :166-183, :341-372), -p adds FORMAT/PS and phased GT; 1000-bp ordered writer (:133).
// ------------------------------------------------------------------------------------------------------------
int cmd_blocksub(int argc, char** argv)
{
    std::string in = "-", outp = "-", interval_str, interval_list;
    bool keep_mnv = false, phased = false, aggressive = false;
    int max_dist = 2, device = 0;
    for (int k = 1; k < argc; ++k)
    {
        std::string a = argv[k];
        if (a == "-o" && k + 1 < argc) outp = argv[++k];
        else if (a == "-d" && k + 1 < argc) max_dist = atoi(argv[++k]);
        else if (a == "--device" && k + 1 < argc) device = atoi(argv[++k]);
        else if (a == "-m") keep_mnv = true;
        else if (a == "-p") phased = true;
        else if (a == "-a") aggressive = true;
        else if (a == "-i" && k + 1 < argc) interval_str = argv[++k];
        else if (a == "-I" && k + 1 < argc) interval_list = argv[++k];
        else if (a == "-" || a[0] != '-') in = a;
        else usage_and_exit("decompose_blocksub", ("unknown option " + a).c_str());
    }
    fprintf(stderr, "decompose_blocksub v%s\n\n", VERSION); // print_options, decompose_blocksub.cpp:423-436
    fprintf(stderr, "options:     input VCF file          %s\n", in.c_str());
    fprintf(stderr, "         [o] output VCF file         %s\n", outp.c_str());
    const std::vector<Interval> intervals = parse_intervals(interval_list, interval_str);
    print_intervals("         [i] intervals               ", intervals);
    fprintf(stderr, "         [a] align/aggressive mode   %s\n", aggressive ? "true" : "false");
    fprintf(stderr, "         [m] keep MNVs (Multi-Nucleotide Variants)   %s\n", keep_mnv ? "true" : "false");
    fprintf(stderr, "         [d] MNVs max distance (when -m option is used)   %s\n", max_dist ? "true" : "false");
    fprintf(stderr, "         [p] output phased genotypes %s\n\n", phased ? "true" : "false");
    vtn_ctx* ctx = nullptr;
    if (vtn_ctx_create(&ctx, device)) die(nullptr, "vtn_ctx_create");
    Reader rd;
    if (!rd.open(in, intervals))
    {
        fprintf(stderr, "[vtn] cannot open %s\n", in.c_str());
        return 1;
    }
    Header h;
    std::string line;
    if (!read_header(rd, h, line))
    {
        fprintf(stderr, "[vtn] %s has no VCF header\n", in.c_str());
        return 1;
    }
    DeviceWriter ow;
    ow.out = open_out(outp);
    ow.window = 1000;
    ow.ctx = ctx;
    std::string extra_hdr = "##INFO=<ID=OLD_CLUMPED,Number=1,Type=String,Description=\"Original chr:pos:ref:alt encoding\">";
    if (phased) extra_hdr += "\n##FORMAT=<ID=PS,Number=1,Type=Integer,Description=\"ID for set of phased genotypes\">";
    write_header(ow.out, h, extra_hdr.c_str());
    size_t BATCH = 1u << 20;
    if (const char* e = getenv("VTN_CLI_BATCH"))
        if (atol(e) > 0) BATCH = (size_t)atol(e);
    BatchBuf bb;
    std::map<std::string, int> rid_cache;
    LineParser lp;
    std::vector<uint32_t> src, off, len, a_beg, a_cnt, a_pr, a_pa, a_lr, a_la;
    vtn_blocksub_counters tot = {0, 0, 0, 0};
    bool more = !line.empty();
    while (more)
    {
        lp.read_batch(rd, line, more, BATCH, bb, ctx, false, rid_cache);
        vtn_batch b = bb.view();
        const uint64_t bound = vtn_blocksub_out_bound(&b);
        src.resize(bound + 1), off.resize(bound + 1), len.resize(bound + 1);
        uint64_t n_out = 0;
        vtn_blocksub_counters c;
        if (vtn_blocksub(ctx, &b, keep_mnv ? 1 : 0, max_dist, src.data(), off.data(), len.data(), bound, &n_out, &c)) die(ctx, "vtn_blocksub");
        if (aggressive)
        { // -a: records with REF / ALT of different lengths are cut along an alignment (decompose_blocksub.cpp:190-303)
            const uint64_t cap = vtn_blocksub_align_bound(&b);
            a_beg.resize((size_t)b.n_rec + 1), a_cnt.resize((size_t)b.n_rec + 1);
            a_pr.resize(cap + 1), a_pa.resize(cap + 1), a_lr.resize(cap + 1), a_la.resize(cap + 1);
            vtn_blocksub_align_plan ap;
            memset(&ap, 0, sizeof ap);
            ap.cap_chunk = cap;
            ap.chunk_beg = a_beg.data(), ap.n_chunk = a_cnt.data();
            ap.pos_ref = a_pr.data(), ap.pos_alt = a_pa.data(), ap.len_ref = a_lr.data(), ap.len_alt = a_la.data();
            if (vtn_blocksub_align(ctx, &b, &ap)) die(ctx, "vtn_blocksub_align");
        }
        uint32_t cur = 0xffffffffu;
        int32_t ps = 0;
        struct OutDesc
        {
            uint32_t r, p_ref, p_alt, l_ref, l_alt;
            int32_t ps;
            bool aligned;
        };
        std::vector<OutDesc> descs;
        descs.reserve((size_t)n_out);
        // one output of the plan; a record in alignment mode turns its single pass-through entry into its chunks
        uint64_t k = 0;
        uint32_t chunk = 0;
        while (k < n_out)
        {
            const uint32_t r = src[k];
            const bool aligned = aggressive && a_cnt[r] > 0;
            uint32_t p_ref = off[k], p_alt = off[k], l_ref = len[k], l_alt = len[k];
            if (aligned)
            {
                const uint32_t c = a_beg[r] + chunk;
                p_ref = a_pr[c], p_alt = a_pa[c], l_ref = a_lr[c], l_alt = a_la[c];
            }
            if (r != cur)
            {
                cur = r;
                ps = b.pos1[r] + (int32_t)p_ref; // the PS vector is sized at the record's first SNP and keeps that value (:343, resize; :263)
                if (aligned)
                { // counted as a pass-through by vtn_blocksub; the reference counts it as a block substitution and its chunks as SNPs (:288-299)
                    c.new_no_variants += a_cnt[r] - 1;
                    c.no_additional_snps += a_cnt[r];
                    c.no_biallelic_blocksub += 1;
                }
            }
            if (aligned)
            {
                if (++chunk == a_cnt[r])
                {
                    chunk = 0;
                    ++k;
                }
            }
            else
                ++k;
            descs.push_back(OutDesc{r, p_ref, p_alt, l_ref, l_alt, ps, aligned});
        }
        // the outputs are independent: formatted by all host threads, handed to the ordered writer in plan order
        std::vector<DeviceWriter::Rec> recs(descs.size());
        parallel_ranges(descs.size(), host_threads(), [&](unsigned, size_t lo, size_t hi) {
          uint32_t last_r = 0xffffffffu;
          std::vector<Tok> col;
          for (size_t d = lo; d < hi; ++d)
          {
            const OutDesc& D = descs[d];
            const uint32_t r = D.r, p_ref = D.p_ref, p_alt = D.p_alt, l_ref = D.l_ref, l_alt = D.l_alt;
            const int32_t ps = D.ps;
            const bool aligned = D.aligned;
            if (r != last_r)
            {
                const LineRef ln = bb.lines[r];
                col = split(ln.data(), ln.size(), '\t');
                last_r = r;
            }
            DeviceWriter::Rec& rec = recs[d];
            rec.rid = b.rid[r];
            rec.chrom.assign(col[0].first, col[0].second);
            std::string& o = rec.line;
            if (!l_ref && !aligned)
            {
                rec.pos1 = b.pos1[r];
                for (int q = 0; q < 5; ++q)
                {
                    if (q) o.push_back('\t');
                    o.append(col[q].first, col[q].second);
                }
                put_tail(o, h, col, "");
            }
            else
            {
                rec.pos1 = b.pos1[r] + (int32_t)p_ref;
                o = rec.chrom + "\t" + std::to_string(rec.pos1) + "\t";
                o.append(col[2].first, col[2].second);
                o.push_back('\t');
                o.append(col[3].first + p_ref, l_ref);
                o.push_back('\t');
                o.append(col[4].first + p_alt, l_alt);
                std::string extra = "OLD_CLUMPED=" + rec.chrom + ":" + std::to_string(b.pos1[r]) + ":";
                extra.append(col[3].first, col[3].second);
                extra.push_back('/');
                for (size_t q = 0; q < col[4].second; ++q) extra.push_back(col[4].first[q] == ',' ? '/' : col[4].first[q]); // bcf_variant2string
                if (phased && (!keep_mnv || aligned) && col.size() > 9)
                { // FORMAT gains PS, every GT separator becomes '|' (:346-366)
                    std::vector<Tok> c2(col.begin(), col.begin() + 8);
                    put_tail(o, h, c2, extra);
                    std::vector<Tok> keys = split(col[8].first, col[8].second, ':');
                    size_t gt_idx = keys.size();
                    for (size_t q = 0; q < keys.size(); ++q)
                        if (keys[q].second == 2 && !memcmp(keys[q].first, "GT", 2)) gt_idx = q;
                    o.push_back('\t');
                    o.append(col[8].first, col[8].second);
                    o += ":PS";
                    for (size_t cidx = 9; cidx < col.size(); ++cidx)
                    {
                        o.push_back('\t');
                        std::string smp;
                        put_sample(smp, h, keys, col[cidx].first, col[cidx].second);
                        if (gt_idx < keys.size())
                        {
                            size_t f0 = 0;
                            for (size_t q = 0; q < gt_idx; ++q) f0 = smp.find(':', f0) + 1;
                            size_t f1 = smp.find(':', f0);
                            if (f1 == std::string::npos) f1 = smp.size();
                            for (size_t q = f0; q < f1; ++q)
                                if (smp[q] == '/') smp[q] = '|';
                        }
                        o += smp;
                        o += ":" + std::to_string(ps);
                    }
                }
                else
                    put_tail(o, h, col, extra);
            }
          }
        });
        for (DeviceWriter::Rec& rec : recs) ow.write(std::move(rec));
        ow.drain(false);
        tot.no_variants += c.no_variants, tot.no_biallelic_blocksub += c.no_biallelic_blocksub;
        tot.no_additional_snps += c.no_additional_snps, tot.new_no_variants += c.new_no_variants;
    }
    ow.drain(true);
    if (ow.out != stdout) fclose(ow.out);
    fprintf(stderr, "\nstats: no. variants                       : %u\n", tot.no_variants); // print_stats, :438-447
    fprintf(stderr, "       no. biallelic block substitutions  : %u\n\n", tot.no_biallelic_blocksub);
    fprintf(stderr, "       no. additional SNPs                : %u\n", tot.no_additional_snps);
    fprintf(stderr, "       no. variants after decomposition   : %u\n\n", tot.new_no_variants);
    vtn_ctx_destroy(ctx);
    return 0;
}

// ------------------------------------------------------------------------------------------------------------
// annotate_indels  (Igor::annotate_indels, annotate_indels.cpp:155-410): vtn_annotate_indels decides which records the reference
// keeps and computes the 34 VNTR values of every indel (regions, motifs and the HMM rounds on the device); the host writes the
// INFO tags in the reference's bcf_update_info_* order (an existing key keeps its place), FLANKSEQ from batched reference
// fetches (:277-291), the two stderr lines per ALT that CandidateRegionExtractor prints unconditionally, 10000-bp ordered writer.
// ------------------------------------------------------------------------------------------------------------
struct InfoList
{
    std::vector<std::pair<std::string, std::string>> kv; // value "\x01" = flag
    void parse(const char* s, size_t n)
    {
        kv.clear();
        if (n == 1 && s[0] == '.') return;
        for (const Tok& t : split(s, n, ';'))
        {
            if (!t.second) continue;
            const char* eq = (const char*)memchr(t.first, '=', t.second);
            if (eq) kv.emplace_back(std::string(t.first, eq - t.first), std::string(eq + 1, t.second - (eq - t.first) - 1));
            else kv.emplace_back(std::string(t.first, t.second), std::string("\x01"));
        }
    }
    void set(const std::string& k, const std::string& v) // bcf_update_info: replace in place, else append
    {
        for (auto& e : kv)
            if (e.first == k)
            {
                e.second = v;
                return;
            }
        kv.emplace_back(k, v);
    }
    std::string str() const
    {
        if (kv.empty()) return ".";
        std::string o;
        for (size_t i = 0; i < kv.size(); ++i)
        {
            if (i) o.push_back(';');
            o += kv[i].first;
            if (kv[i].second != "\x01")
            {
                o.push_back('=');
                o += kv[i].second;
            }
        }
        return o;
    }
};

std::string fmt_float(float x)
{
    char b[64];
    snprintf(b, sizeof b, "%.9g", (double)x);
    std::string o;
    put_float(o, b, strlen(b));
    return o;
}

int cmd_annotate_indels(int argc, char** argv)
{
    Args a = parse_args("annotate_indels", argc, argv);
    if (a.ref.empty()) usage_and_exit("annotate_indels", "required: -r <reference FASTA>");
    // print_options, annotate_indels.cpp:241-252 (add_flank_annotation is never set there; it prints as false)
    fprintf(stderr, "annotate_indels v%s\n\n", VERSION);
    fprintf(stderr, "options:     input VCF file(s)        %s\n", a.in.c_str());
    fprintf(stderr, "         [o] output VCF file          %s\n", a.out.c_str());
    fprintf(stderr, "         [k] add_flank_annotation     false\n");
    fprintf(stderr, "         [d] debug                    %s\n", a.debug ? "true" : "false");
    fprintf(stderr, "         [r] ref FASTA file           %s\n", a.ref.c_str());
    print_intervals("         [i] intervals                ", a.intervals);
    fputc('\n', stderr);
    vtn_ctx* ctx = nullptr;
    if (vtn_ctx_create(&ctx, a.device)) die(nullptr, "vtn_ctx_create");
    if (vtn_ref_open_fasta(ctx, a.ref.c_str())) die(ctx, "loading the reference");
    Reader rd;
    if (!rd.open(a.in, a.intervals))
    {
        fprintf(stderr, "[vtn] cannot open %s\n", a.in.c_str());
        return 1;
    }
    Header h;
    std::string line;
    if (!read_header(rd, h, line))
    {
        fprintf(stderr, "[vtn] %s has no VCF header\n", a.in.c_str());
        return 1;
    }
    // the 36 INFO lines of initialize() (:168-229), each only if the ID is not in the header dictionary yet
    static const char* TAGS[][4] = {
        {"EX_MOTIF", "1", "String", "Canonical motif in a VNTR."}, {"EX_RU", "1", "String", "Repeat unit in the reference sequence."},
        {"EX_BASIS", "1", "String", "Basis nucleotides in the motif."}, {"EX_MLEN", "1", "Integer", "Motif length."}, {"EX_BLEN", "1", "Integer", "Basis length."},
        {"EX_REPEAT_TRACT", "2", "Integer", "Boundary of the repeat tract detected by exact alignment."},
        {"EX_COMP", "4", "Integer", "Composition(%) of bases in an exact repeat tract."}, {"EX_ENTROPY", "1", "Float", "Entropy measure of an exact repeat tract [0,2]."},
        {"EX_ENTROPY2", "1", "Float", "Dinucleotide entropy measure of an exact repeat tract [0,4]."},
        {"EX_KL_DIVERGENCE", "1", "Float", "Kullback-Leibler Divergence of an exact repeat tract."},
        {"EX_KL_DIVERGENCE2", "1", "Float", "Dinucleotide Kullback-Leibler Divergence of an exact repeat tract."},
        {"EX_REF", ".", "Float", "Allele lengths in repeat units from exact alignment."}, {"EX_RL", "1", "Integer", "Reference exact repeat tract length in bases."},
        {"EX_LL", "1", "Integer", "Longest exact repeat tract length in bases."},
        {"EX_RU_COUNTS", "2", "Integer", "Number of exact repeat units and total number of repeat units in exact repeat tract."},
        {"EX_SCORE", "1", "Float", "Score of repeat unit in exact repeat tract."}, {"EX_TRF_SCORE", "1", "Integer", "TRF Score for M/I/D as 2/-7/-7 in exact repeat tract."},
        {"FZ_MOTIF", "1", "String", "Canonical motif in a VNTR."}, {"FZ_RU", "1", "String", "Repeat unit in the reference sequence."},
        {"FZ_BASIS", "1", "String", "Basis nucleotides in the motif."}, {"FZ_MLEN", "1", "Integer", "Motif length."}, {"FZ_BLEN", "1", "Integer", "Basis length."},
        {"FZ_REPEAT_TRACT", "2", "Integer", "Boundary of the repeat tract detected by fuzzy alignment."},
        {"FZ_COMP", "4", "Integer", "Composition(%) of bases in a fuzzy repeat tract."}, {"FZ_ENTROPY", "1", "Float", "Entropy measure of a fuzzy repeat tract (0-2)."},
        {"FZ_ENTROPY2", "1", "Float", "Dinucleotide entropy measure of a fuzzy repeat tract (0-2)."},
        {"FZ_KL_DIVERGENCE", "1", "Float", "Kullback-Leibler Divergence of a fuzzyt repeat tract."},
        {"FZ_KL_DIVERGENCE2", "1", "Float", "Dinucleotide Kullback-Leibler Divergence of a fuzzy repeat tract."},
        {"FZ_REF", ".", "Float", "Allele lengths in repeat units from fuzzy alignment."}, {"FZ_RL", "1", "Integer", "Reference fuzzy repeat tract length in bases."},
        {"FZ_LL", "1", "Integer", "Longest fuzzy repeat tract length in bases."},
        {"FZ_RU_COUNTS", "2", "Integer", "Number of exact repeat units and total number of repeat units in fuzzy repeat tract."},
        {"FZ_SCORE", "1", "Float", "Score of repeat unit in fuzzy repeat tract."}, {"FZ_TRF_SCORE", "1", "Integer", "TRF Score for M/I/D as 2/-7/-7 in fuzzy repeat tract."},
        {"FLANKSEQ", "1", "String", "Flanking sequence 10bp on either side of REF."}, {"EXACT_RU_AMBIGUOUS", "0", "Flag", "Exact motif is ambiguous."}};
    std::string extra;
    for (auto& t : TAGS)
    {
        bool present = h.info.count(t[0]) || h.fmt.count(t[0]);
        for (const std::string& m : h.meta)
            if (m.compare(0, 9, "##FILTER=") == 0 && attr(m, "ID") == t[0]) present = true;
        if (present) continue;
        if (!extra.empty()) extra.push_back('\n');
        extra += std::string("##INFO=<ID=") + t[0] + ",Number=" + t[1] + ",Type=" + t[2] + ",Description=\"" + t[3] + "\">";
        Tag tg;
        tg.type = t[2][0] == 'I' ? 'I' : t[2][0] == 'F' ? (t[2][1] == 'l' && t[2][2] == 'a' ? 'L' : 'F') : 'S';
        h.info[t[0]] = tg;
    }
    DeviceWriter ow;
    ow.out = open_out(a.out);
    ow.window = 10000; // annotate_indels.cpp:161
    ow.ctx = ctx;
    write_header(ow.out, h, extra.empty() ? nullptr : extra.c_str());

    size_t BATCH = 1u << 18;
    if (const char* e = getenv("VTN_CLI_BATCH"))
        if (atol(e) > 0) BATCH = (size_t)atol(e);
    BatchBuf bb;
    std::map<std::string, int> rid_cache;
    LineParser lp;
    Filter flt;
    if (flt.set(a.fexp, h.meta)) lp.filter = &flt;
    bool more = !line.empty();
    uint32_t no_indels_annotated = 0;
    while (more)
    {
        lp.read_batch(rd, line, more, BATCH, bb, ctx, true, rid_cache);
        vtn_batch in = bb.view();
        const size_t n = in.n_rec;
        std::vector<uint8_t> flags(n + 1), vtype(n + 1);
        std::vector<uint32_t> idx(n + 1), tl(in.n_allele + 1), tr(in.n_allele + 1), sh(in.n_allele + 1);
        std::vector<vtn_vntr_annot> vn(n + 1);
        std::vector<char> strs(4 * (size_t)in.n_bytes + 64 * n + 4096);
        vtn_annot_result out;
        memset(&out, 0, sizeof out);
        out.flags = flags.data(), out.vtype = vtype.data(), out.indel_idx = idx.data();
        out.trim_left = tl.data(), out.trim_right = tr.data(), out.shift = sh.data();
        out.cap_indel = (uint32_t)n, out.cap_str = strs.size(), out.vntr = vn.data(), out.strings = strs.data();
        if (vtn_annotate_indels(ctx, &in, &out)) die(ctx, "vtn_annotate_indels");
        // reference windows: FLANKSEQ (three per record that gets one) and the bases right_align walked over (one per ALT)
        std::vector<int32_t> w_rid;
        std::vector<int64_t> w_beg, w_end;
        std::vector<size_t> flank_at(n, (size_t)-1), shift_at(in.n_allele + 1, (size_t)-1);
        for (size_t r = 0; r < n; ++r)
        {
            if (!(flags[r] & VTN_ANNOT_KEEP)) continue;
            const uint32_t a0 = in.allele_beg[r];
            const int64_t beg1 = in.pos1[r], end1 = beg1 + (int64_t)(in.allele_off[a0 + 1] - in.allele_off[a0]) - 1;
            const bool indel = flags[r] & VTN_ANNOT_INDEL;
            if (indel || vtype[r] == VTN_VT_VNTR || vtype[r] == VTN_VT_SNP || vtype[r] == VTN_VT_MNP)
            {
                if (in.rid[r] < 0)
                {
                    fprintf(stderr, "[vtn] contig of record %s:%d is not in the reference\n", std::string(bb.lines[r].data(), (size_t)((const char*)memchr(bb.lines[r].data(), '\t', bb.lines[r].size()) - bb.lines[r].data())).c_str(), in.pos1[r]);
                    exit(1);
                }
                flank_at[r] = w_rid.size();
                const int64_t wb[3] = {beg1 - 10, beg1, end1 + 1}, we[3] = {beg1 - 1, end1, end1 + 10};
                for (int q = 0; q < 3; ++q) w_rid.push_back(in.rid[r]), w_beg.push_back(wb[q] - 1), w_end.push_back(we[q] - 1);
            }
            if (indel)
                for (uint32_t s = a0 + 1; s < in.allele_beg[r + 1]; ++s)
                    if (sh[s])
                    {
                        const int64_t e1 = beg1 + tl[s] + (int64_t)(in.allele_off[a0 + 1] - in.allele_off[a0] - tl[s] - tr[s]) - 1;
                        shift_at[s] = w_rid.size();
                        w_rid.push_back(in.rid[r]), w_beg.push_back(e1), w_end.push_back(e1 + sh[s] - 1);
                    }
        }
        std::vector<uint64_t> w_off(w_rid.size() + 1, 0);
        std::vector<uint8_t> w_bytes(1);
        if (!w_rid.empty())
        {
            uint64_t cap = 0;
            for (size_t k = 0; k < w_rid.size(); ++k) cap += (uint64_t)std::max<int64_t>(w_end[k] - w_beg[k] + 1, 1);
            w_bytes.resize(cap + 1);
            if (vtn_ref_fetch_batch(ctx, (uint32_t)w_rid.size(), w_rid.data(), w_beg.data(), w_end.data(), 1, w_off.data(), w_bytes.data(), cap)) die(ctx, "vtn_ref_fetch_batch");
        }
        auto window = [&](size_t k) { return std::string((const char*)w_bytes.data() + w_off[k], (size_t)(w_off[k + 1] - w_off[k])); };
        for (size_t r = 0; r < n; ++r)
        {
            if (!(flags[r] & VTN_ANNOT_KEEP)) continue; // annotate_indels.cpp:306-323: dropped, not passed through
            const LineRef ln = bb.lines[r];
            std::vector<Tok> col = split(ln.data(), ln.size(), '\t');
            const uint32_t a0 = in.allele_beg[r];
            DeviceWriter::Rec rec;
            rec.rid = in.rid[r];
            rec.chrom.assign(col[0].first, col[0].second);
            rec.pos1 = in.pos1[r];
            InfoList info;
            info.parse(col[7].first, col[7].second);
            if (flags[r] & VTN_ANNOT_INDEL)
            {
                // candidate_region_extractor.cpp:195, :199: "<end1> <ref>  <alt>" after trim and after right_align, per ALT
                const std::string REF((const char*)in.bytes + in.allele_off[a0], in.allele_off[a0 + 1] - in.allele_off[a0]);
                for (uint32_t s = a0 + 1; s < in.allele_beg[r + 1]; ++s)
                {
                    const std::string ALT((const char*)in.bytes + in.allele_off[s], in.allele_off[s + 1] - in.allele_off[s]);
                    std::string rt = REF.substr(tl[s], REF.size() - tl[s] - tr[s]), at = ALT.substr(tl[s], ALT.size() - tl[s] - tr[s]);
                    const int64_t e1 = (int64_t)in.pos1[r] + tl[s] + (int64_t)rt.size() - 1;
                    rec.msg += std::to_string(e1) + " " + rt + "  " + at + "\n";
                    if (sh[s])
                    {
                        const std::string walked = window(shift_at[s]);
                        rt = (rt + walked).substr(sh[s]);
                        at = (at + walked).substr(sh[s]);
                    }
                    rec.msg += std::to_string(e1 + sh[s]) + " " + rt + "  " + at + "\n";
                }
                const vtn_vntr_annot& v = vn[idx[r]];
                const vtn_vntr_side* sd[2] = {&v.exact, &v.fuzzy};
                const char* pre[2] = {"EX_", "FZ_"};
                for (int x = 0; x < 2; ++x)
                {
                    const vtn_vntr_side& q = *sd[x];
                    const std::string P = pre[x];
                    auto S = [&](int i) { return std::string(out.strings + q.str_off[i], q.str_len[i]); };
                    info.set(P + "MOTIF", S(0));
                    info.set(P + "MLEN", std::to_string(q.mlen));
                    info.set(P + "RU", S(1));
                    info.set(P + "BASIS", S(2));
                    info.set(P + "BLEN", std::to_string(q.blen));
                    info.set(P + "REPEAT_TRACT", std::to_string(q.beg1) + "," + std::to_string(q.end1));
                    info.set(P + "COMP", std::to_string(q.comp[0]) + "," + std::to_string(q.comp[1]) + "," + std::to_string(q.comp[2]) + "," + std::to_string(q.comp[3]));
                    info.set(P + "ENTROPY", fmt_float(q.entropy));
                    info.set(P + "ENTROPY2", fmt_float(q.entropy2));
                    info.set(P + "KL_DIVERGENCE", fmt_float(q.kl_divergence));
                    info.set(P + "KL_DIVERGENCE2", fmt_float(q.kl_divergence2));
                    info.set(P + "REF", fmt_float(q.ref));
                    info.set(P + "RL", std::to_string(q.rl));
                    info.set(P + "LL", std::to_string(q.ll));
                    info.set(P + "RU_COUNTS", std::to_string(q.no_perfect_ru) + "," + std::to_string(q.no_ru));
                    info.set(P + "SCORE", fmt_float(q.score));
                    info.set(P + "TRF_SCORE", std::to_string(q.trf_score));
                    if (x == 0 && v.exact_ru_ambiguous) info.set("EXACT_RU_AMBIGUOUS", "\x01");
                }
                ++no_indels_annotated;
            }
            if (flank_at[r] != (size_t)-1) info.set("FLANKSEQ", window(flank_at[r]) + "[" + window(flank_at[r] + 1) + "]" + window(flank_at[r] + 2));
            std::string& o = rec.line;
            for (int c = 0; c < 5; ++c)
            {
                if (c) o.push_back('\t');
                o.append(col[c].first, col[c].second);
            }
            const std::string istr = info.str();
            std::vector<Tok> c2 = col;
            c2[7] = Tok(istr.data(), istr.size());
            put_tail(o, h, c2, "");
            ow.write(std::move(rec));
        }
        ow.drain(false);
    }
    ow.drain(true);
    if (ow.out != stdout) fclose(ow.out);
    fprintf(stderr, "\nstats: no. of indels annotated   %u\n\n", no_indels_annotated); // print_stats, :254-259
    vtn_ctx_destroy(ctx);
    return 0;
}

// ------------------------------------------------------------------------------------------------------------
// exports (include/vtn_tools.h): the reference's tool entry points, argv[0] = sub-command (main.cpp:193-196, 221-224, 393-396)
// ------------------------------------------------------------------------------------------------------------
extern "C" {

int vtn_tool_normalize(int argc, char** argv) { return cmd_normalize(argc, argv); }

void vtn_tool_decompose(int argc, char** argv) { cmd_decompose(argc, argv); }

void vtn_tool_annotate_indels(int argc, char** argv) { cmd_annotate_indels(argc, argv); }

void vtn_tool_uniq(int argc, char** argv) { cmd_uniq(argc, argv); }

void vtn_tool_decompose_blocksub(int argc, char** argv) { cmd_blocksub(argc, argv); }

void vtn_tools_end_process_after_output(int on) { g_allow_quick_exit = on != 0; }

int vtn_tool_main(int argc, char** argv)
{
    clock_t t0 = clock();
    if (argc < 2)
    {
        fprintf(stderr, "usage: vtn normalize|decompose|decompose_blocksub|uniq|annotate_indels [options] <in.vcf>   (B200 drop-in for vt's decompose/normalize path)\n");
        return 1;
    }
    std::string cmd = argv[1];
    bool print = false;
    if (cmd == "normalize")
        print = vtn_tool_normalize(argc - 1, argv + 1) != 0;
    else if (cmd == "decompose")
    {
        vtn_tool_decompose(argc - 1, argv + 1);
        print = true;
        for (int k = 2; k < argc; ++k) print = print && strcmp(argv[k], "-q");
    }
    else if (cmd == "uniq")
    {
        vtn_tool_uniq(argc - 1, argv + 1);
        print = true;
    }
    else if (cmd == "decompose_blocksub")
    {
        vtn_tool_decompose_blocksub(argc - 1, argv + 1);
        print = true;
    }
    else if (cmd == "annotate_indels")
    {
        vtn_tool_annotate_indels(argc - 1, argv + 1);
        print = true;
    }
    else if (cmd == "_filter" && argc >= 3)
    { // self-test hook (no GPU needed): vtn _filter <expression> [header line ...] ; stdin lines "vtype<TAB>n_allele<TAB>QUAL<TAB>FILTER" -> 1 / 0
        Filter flt;
        std::vector<std::string> meta;
        for (int k = 3; k < argc; ++k) meta.push_back(argv[k]);
        flt.set(argv[2], meta);
        char buf[4096];
        while (fgets(buf, sizeof buf, stdin))
        {
            size_t n = strlen(buf);
            while (n && (buf[n - 1] == '\n' || buf[n - 1] == '\r')) --n;
            std::vector<Tok> c = split(buf, n, '\t');
            if (c.size() < 4) continue;
            Filter::Rec v;
            v.vtype = atoi(std::string(c[0].first, c[0].second).c_str());
            v.n_allele = atoi(std::string(c[1].first, c[1].second).c_str());
            v.qual = c[2].first, v.qual_len = c[2].second;
            v.filt = c[3].first, v.filt_len = c[3].second;
            printf("%d\n", flt.pass(v) ? 1 : 0);
        }
        return 0;
    }
    else if (cmd == "_plain" && argc >= 3)
    { // self-test hook (no GPU needed): vtn _plain <header.vcf> ; stdin lines = the columns of a record from QUAL on -> 1 when they
      // would be written as they are (tail_is_plain), else 0 followed by what put_tail writes instead
        Reader rd;
        Header h;
        std::string first;
        if (!rd.open(argv[2]) || !read_header(rd, h, first)) return 1;
        char buf[65536];
        while (fgets(buf, sizeof buf, stdin))
        {
            size_t n = strlen(buf);
            while (n && (buf[n - 1] == '\n' || buf[n - 1] == '\r')) --n;
            const std::string line = std::string("1\t1\t.\tA\tC\t") + std::string(buf, n);
            std::vector<Tok> col = split(line.data(), line.size(), '\t');
            if (col.size() < 8) continue;
            const bool plain = tail_is_plain(h, true, true, col[5].first, (size_t)(line.data() + line.size() - col[5].first));
            std::string o;
            put_tail(o, h, col, "");
            printf("%d%s\n", plain ? 1 : 0, o.c_str());
        }
        return 0;
    }
    else if (cmd == "_intervals")
    { // self-test hook: vtn _intervals <comma separated intervals> [list file] -> one parsed interval per line (GenomeInterval::to_string)
        for (const Interval& iv : parse_intervals(argc > 3 ? argv[3] : "", argc > 2 ? argv[2] : "")) printf("%s\t%d\t%d\n", iv.to_string().c_str(), iv.start1, iv.end1);
        return 0;
    }
    else if (cmd == "_floats")
    { // self-test hook (no GPU needed): every stdin line -> "<1 if canon_float says htslib prints it as written>\t<put_float's text>"
        char buf[4096];
        while (fgets(buf, sizeof buf, stdin))
        {
            size_t n = strlen(buf);
            while (n && (buf[n - 1] == '\n' || buf[n - 1] == '\r')) --n;
            std::string o;
            put_float(o, buf, n);
            printf("%d\t%s\n", canon_float(buf, n) ? 1 : 0, o.c_str());
        }
        return 0;
    }
    else
    {
        fprintf(stderr, "Command not found: %s\n", argv[1]);
        return 1;
    }
    if (print) fprintf(stderr, "Time elapsed: %.2fs\n\n", (double)(clock() - t0) / CLOCKS_PER_SEC); // main.cpp:428-433
    if (g_quick_exit)
    { // only after vtn_tools_end_process_after_output(1): the output is on disk, the OS reclaims the context and the pinned pages at once
        fflush(nullptr);
        _exit(0);
    }
    return 0;
}

} // extern "C"
