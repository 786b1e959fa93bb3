// vtn_bcf.h -- BCF2 either side of the tools: a byte source that turns a BCF2 stream into VCF text, and a sink that
// turns VCF text into BCF2.  Host-only C++; the tools stay text pipelines, so the site columns of a BCF record reach the same
// struct-of-arrays batch as those of a VCF line (through vtn_vcf_tokenize).
//
// What it stands in for: htslib's bcf_read / vcf_format (BCF in) and vcf_parse / bcf_write (BCF out) under BCFOrderedReader /
// BCFOrderedWriter (lib/htslib/vcf.c).  The format (VCF specification, section 6): "BCF\2\2", the header text, then records of two
// length-prefixed blocks: shared (CHROM, POS, rlen, QUAL, counts, ID, alleles, FILTER, INFO) and per-sample (FORMAT).  Values are
// typed: a descriptor byte (count << 4 | type; count 15 = a typed integer with the real count follows), types 1/2/3 = int8/16/32,
// 5 = float, 7 = characters; integers carry a "missing" and an "end of vector" value (0x80 / 0x81 and their wider forms), floats
// 0x7F800001 / 0x7F800002.  Strings (FILTER / INFO / FORMAT ids, contigs) are dictionary indices taken from the header lines in
// order, PASS first, an IDX= attribute overriding the order.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "vtn_bgzf.h"

namespace vtb2
{

[[noreturn]] inline void bcf_die(const std::string& msg)
{
    fprintf(stderr, "[vtn] BCF: %s\n", msg.c_str());
    exit(1);
}

struct Tag
{
    std::string id;
    char type = 'S';   // I F S C L(flag)
    char number = '.'; // A R G . or N
};

// the dictionaries of a header
struct Dict
{
    std::vector<std::string> contigs;
    std::vector<std::string> ids; // FILTER / INFO / FORMAT ids share one index space
    std::map<std::string, int> contig_idx, id_idx;
    std::map<std::string, Tag> info, fmt;

    static std::string attr(const std::string& line, const char* key)
    { // value of key= inside <...>, quotes respected
        const size_t lt = line.find('<');
        if (lt == std::string::npos) return "";
        size_t p = lt + 1;
        const std::string k = std::string(key) + "=";
        while (p < line.size())
        {
            size_t e = p;
            bool q = false;
            while (e < line.size() && (q || (line[e] != ',' && line[e] != '>')))
            {
                if (line[e] == '"' && (e == 0 || line[e - 1] != '\\')) q = !q;
                ++e;
            }
            if (line.compare(p, k.size(), k) == 0) return line.substr(p + k.size(), e - p - k.size());
            p = e + 1;
        }
        return "";
    }
    void set(std::vector<std::string>& v, std::map<std::string, int>& m, const std::string& id, const std::string& idx)
    {
        if (m.count(id)) return;
        int k = idx.empty() ? (int)v.size() : atoi(idx.c_str());
        if (!idx.empty())
        { // an explicit IDX: the slot may lie beyond what has been seen
            if ((int)v.size() <= k) v.resize((size_t)k + 1);
        }
        else
        {
            while (k < (int)v.size() && !v[k].empty()) ++k; // next free slot
            if ((int)v.size() <= k) v.resize((size_t)k + 1);
        }
        v[(size_t)k] = id;
        m[id] = k;
    }
    void build(const std::string& text)
    {
        set(ids, id_idx, "PASS", "");
        size_t p = 0;
        while (p < text.size())
        {
            size_t e = text.find('\n', p);
            if (e == std::string::npos) e = text.size();
            const std::string line = text.substr(p, e - p);
            p = e + 1;
            if (line.compare(0, 2, "##")) continue;
            const bool is_info = !line.compare(0, 7, "##INFO="), is_fmt = !line.compare(0, 9, "##FORMAT="), is_flt = !line.compare(0, 9, "##FILTER=");
            if (!line.compare(0, 9, "##contig="))
                set(contigs, contig_idx, attr(line, "ID"), attr(line, "IDX"));
            else if (is_info || is_fmt || is_flt)
            {
                const std::string id = attr(line, "ID");
                if (id.empty()) continue;
                set(ids, id_idx, id, attr(line, "IDX"));
                if (!is_flt)
                {
                    Tag t;
                    t.id = id;
                    const std::string ty = attr(line, "Type"), nu = attr(line, "Number");
                    t.type = ty == "Integer" ? 'I' : ty == "Float" ? 'F' : ty == "Flag" ? 'L' : ty == "Character" ? 'C' : 'S';
                    t.number = (nu == "A" || nu == "R" || nu == "G" || nu == ".") ? nu[0] : 'N';
                    (is_info ? info : fmt)[id] = t;
                }
            }
        }
    }
};

// htslib's number formatting for Float values and QUAL (kputd, kstring.c: 6 significant digits, "%g" outside [1e-4, 999999])
inline void put_float_bits(std::string& out, float f)
{
    if (std::isnan(f))
    {
        out += "nan";
        return;
    }
    if (std::isinf(f))
    {
        out += f < 0 ? "-inf" : "inf";
        return;
    }
    double d = f;
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
    uint64_t half = 5;
    for (double lim = 0.001; d >= lim && lim < 1e6; lim *= 10) half *= 10;
    units += half;
    char dig[32];
    const int p = snprintf(dig, sizeof dig, "%llu", (unsigned long long)units);
    std::string r;
    if (p <= 10)
        r = "0." + std::string((size_t)(10 - p), '0') + std::string(dig, dig + (p < 6 ? p : 6));
    else
    {
        const int ni = p - 10;
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

enum { BT_NULL = 0, BT_INT8 = 1, BT_INT16 = 2, BT_INT32 = 3, BT_FLOAT = 5, BT_CHAR = 7 };
const uint32_t FLOAT_MISSING = 0x7F800001u, FLOAT_VEND = 0x7F800002u;

// ---------------------------------------------------------------------------------------------------------
// BCF2 in: a byte source of VCF text made from a byte source of BCF2 (after BGZF, if any)
// ---------------------------------------------------------------------------------------------------------
struct BcfTextSource : vtz::ByteSource
{
    vtz::ByteSource* in = nullptr; // owned
    Dict dict;
    std::string out; // text produced and not handed out yet
    size_t out_pos = 0;
    std::vector<unsigned char> buf; // raw bytes read ahead
    size_t pos = 0, end = 0;
    bool started = false, at_eof = false;
    explicit BcfTextSource(vtz::ByteSource* src, const unsigned char* already, size_t n_already) : in(src)
    {
        buf.resize(1 << 20);
        memcpy(buf.data(), already, n_already);
        end = n_already;
    }
    ~BcfTextSource() override { delete in; }
    bool need(size_t n)
    { // make n bytes available at buf[pos..]
        while (end - pos < n)
        {
            if (at_eof) return false;
            if (pos)
            {
                memmove(buf.data(), buf.data() + pos, end - pos);
                end -= pos, pos = 0;
            }
            if (buf.size() < n) buf.resize(n + (n >> 1));
            if (end == buf.size()) buf.resize(buf.size() * 2);
            const long got = in->read((char*)buf.data() + end, buf.size() - end);
            if (got < 0) bcf_die("read error");
            if (got == 0) at_eof = true;
            else end += (size_t)got;
        }
        return true;
    }
    static uint32_t u32(const unsigned char* p) { return p[0] | (p[1] << 8) | (p[2] << 16) | ((uint32_t)p[3] << 24); }
    // one typed value header: type and count; p advances
    static void typed_size(const unsigned char*& p, const unsigned char* e, int& type, int& count)
    {
        if (p >= e) bcf_die("truncated record");
        const unsigned char d = *p++;
        type = d & 15, count = d >> 4;
        if (count == 15)
        {
            int t2, c2;
            typed_size(p, e, t2, c2);
            count = read_int(p, e, t2);
        }
    }
    static int32_t read_int(const unsigned char*& p, const unsigned char* e, int type)
    { // raw value (missing / end-of-vector codes are mapped to INT32 codes)
        int32_t v = 0;
        if (type == BT_INT8)
        {
            if (p + 1 > e) bcf_die("truncated record");
            const int8_t x = (int8_t)p[0];
            v = x == INT8_MIN ? INT32_MIN : (x == INT8_MIN + 1 ? INT32_MIN + 1 : x);
            p += 1;
        }
        else if (type == BT_INT16)
        {
            if (p + 2 > e) bcf_die("truncated record");
            const int16_t x = (int16_t)(p[0] | (p[1] << 8));
            v = x == INT16_MIN ? INT32_MIN : (x == INT16_MIN + 1 ? INT32_MIN + 1 : x);
            p += 2;
        }
        else if (type == BT_INT32)
        {
            if (p + 4 > e) bcf_die("truncated record");
            v = (int32_t)u32(p);
            p += 4;
        }
        else
            bcf_die("an integer was expected");
        return v;
    }
    static int type_width(int type) { return type == BT_INT8 || type == BT_CHAR ? 1 : type == BT_INT16 ? 2 : type == BT_NULL ? 0 : 4; }
    // a vector of count values of the given type as VCF text (vcf.c bcf_fmt_array)
    static void fmt_array(std::string& o, const unsigned char* p, int type, int count)
    {
        if (type == BT_CHAR)
        {
            int n = 0;
            while (n < count && p[n]) ++n;
            if (!n) o.push_back('.');
            else o.append((const char*)p, (size_t)n);
            return;
        }
        int printed = 0;
        for (int k = 0; k < count; ++k)
        {
            if (type == BT_FLOAT)
            {
                const uint32_t b = u32(p + 4 * k);
                if (b == FLOAT_VEND) break;
                if (k) o.push_back(',');
                if (b == FLOAT_MISSING) o.push_back('.');
                else
                {
                    float f;
                    memcpy(&f, &b, 4);
                    put_float_bits(o, f);
                }
            }
            else
            {
                const unsigned char* q = p + (size_t)type_width(type) * k;
                const int32_t v = read_int(q, q + 4, type);
                if (v == INT32_MIN + 1) break;
                if (k) o.push_back(',');
                if (v == INT32_MIN) o.push_back('.');
                else o += std::to_string(v);
            }
            ++printed;
        }
        if (!printed) o.push_back('.');
    }
    const std::string& id_name(int k)
    {
        if (k < 0 || k >= (int)dict.ids.size() || dict.ids[(size_t)k].empty()) bcf_die("a dictionary index the header does not define");
        return dict.ids[(size_t)k];
    }
    bool produce()
    { // header first, then one record per call; false at the end
        if (!started)
        {
            started = true;
            if (!need(9) || memcmp(buf.data() + pos, "BCF\2\2", 5)) bcf_die("not a BCF2.2 stream");
            const uint32_t l_text = u32(buf.data() + pos + 5);
            pos += 9;
            if (!need(l_text)) bcf_die("truncated header");
            std::string text((const char*)buf.data() + pos, l_text);
            pos += l_text;
            while (!text.empty() && text.back() == '\0') text.pop_back();
            if (!text.empty() && text.back() != '\n') text.push_back('\n');
            dict.build(text);
            // the IDX attributes belong to the BCF form of the header; the VCF form htslib prints has none (bcf_hdr_format, is_bcf = 0)
            size_t q = 0;
            while ((q = text.find(",IDX=", q)) != std::string::npos)
            {
                size_t r = q + 5;
                while (r < text.size() && text[r] >= '0' && text[r] <= '9') ++r;
                if (r < text.size() && text[r] == '>') text.erase(q, r - q);
                else q = r;
            }
            out += text;
            return true;
        }
        if (!need(8)) return false;
        const uint32_t l_shared = u32(buf.data() + pos), l_indiv = u32(buf.data() + pos + 4);
        if (!need(8 + (size_t)l_shared + l_indiv)) bcf_die("truncated record");
        const unsigned char* p = buf.data() + pos + 8;
        const unsigned char* e = p + l_shared;
        const unsigned char* ie = e + l_indiv;
        pos += 8 + (size_t)l_shared + l_indiv;
        if (l_shared < 24) bcf_die("truncated record");
        const int32_t chrom = (int32_t)u32(p), pos0 = (int32_t)u32(p + 4);
        const uint32_t qbits = u32(p + 12), n_ai = u32(p + 16), n_fs = u32(p + 20);
        const int n_info = n_ai & 0xffff, n_allele = n_ai >> 16, n_fmt = n_fs >> 24, n_sample = n_fs & 0xffffff;
        p += 24;
        std::string& o = out;
        if (chrom < 0 || chrom >= (int)dict.contigs.size()) bcf_die("a contig index the header does not define");
        o += dict.contigs[(size_t)chrom];
        o.push_back('\t');
        o += std::to_string((int64_t)pos0 + 1);
        o.push_back('\t');
        int type, count;
        typed_size(p, e, type, count); // ID
        if (count && type == BT_CHAR) o.append((const char*)p, (size_t)count);
        else o.push_back('.');
        p += count;
        for (int a = 0; a < n_allele; ++a)
        {
            typed_size(p, e, type, count);
            o.push_back(a < 2 ? '\t' : ',');
            o.append((const char*)p, (size_t)count);
            p += count;
        }
        if (n_allele == 0) o += "\t.";
        if (n_allele < 2) o += "\t.";
        o.push_back('\t');
        if (qbits == FLOAT_MISSING) o.push_back('.');
        else
        {
            float q;
            memcpy(&q, &qbits, 4);
            put_float_bits(o, q);
        }
        o.push_back('\t');
        typed_size(p, e, type, count); // FILTER
        if (!count) o.push_back('.');
        for (int k = 0; k < count; ++k)
        {
            if (k) o.push_back(';');
            o += id_name(read_int(p, e, type));
        }
        o.push_back('\t');
        if (!n_info) o.push_back('.');
        for (int k = 0; k < n_info; ++k)
        {
            int kt, kc;
            typed_size(p, e, kt, kc);
            const int key = read_int(p, e, kt);
            if (k) o.push_back(';');
            o += id_name(key);
            typed_size(p, e, type, count);
            if (count)
            {
                o.push_back('=');
                fmt_array(o, p, type, count);
                p += (size_t)type_width(type) * count;
            }
        }
        if (n_sample && n_fmt)
        {
            struct Fmt
            {
                int key, type, count;
                const unsigned char* data;
            };
            std::vector<Fmt> F;
            const unsigned char* q = e;
            for (int k = 0; k < n_fmt; ++k)
            {
                int kt, kc;
                typed_size(q, ie, kt, kc);
                Fmt f;
                f.key = read_int(q, ie, kt);
                typed_size(q, ie, f.type, f.count);
                f.data = q;
                q += (size_t)type_width(f.type) * f.count * n_sample;
                if (q > ie) bcf_die("truncated record");
                F.push_back(f);
            }
            o.push_back('\t');
            for (size_t k = 0; k < F.size(); ++k)
            {
                if (k) o.push_back(':');
                o += id_name(F[k].key);
            }
            for (int s = 0; s < n_sample; ++s)
            {
                o.push_back('\t');
                for (size_t k = 0; k < F.size(); ++k)
                {
                    if (k) o.push_back(':');
                    const Fmt& f = F[k];
                    const unsigned char* d = f.data + (size_t)type_width(f.type) * f.count * s;
                    if (id_name(f.key) == "GT" && f.type != BT_CHAR && f.type != BT_FLOAT)
                    { // bcf_format_gt
                        int n = 0;
                        for (int a = 0; a < f.count; ++a)
                        {
                            const unsigned char* dd = d + (size_t)type_width(f.type) * a;
                            const int32_t v = read_int(dd, dd + 4, f.type);
                            if (v == INT32_MIN + 1) break;
                            if (a) o.push_back((v & 1) ? '|' : '/');
                            if (v == INT32_MIN || (v >> 1) == 0) o.push_back('.');
                            else o += std::to_string((v >> 1) - 1);
                            ++n;
                        }
                        if (!n) o.push_back('.');
                    }
                    else
                        fmt_array(o, d, f.type, f.count);
                }
            }
        }
        o.push_back('\n');
        return true;
    }
    long read(char* dst, size_t n) override
    {
        while (out_pos == out.size())
        {
            out.clear();
            out_pos = 0;
            if (!produce()) return 0;
        }
        const size_t k = std::min(n, out.size() - out_pos);
        memcpy(dst, out.data() + out_pos, k);
        out_pos += k;
        return (long)k;
    }
};

// a source for any input file: plain / gzip / BGZF text, or BCF2 (raw or BGZF) turned into text
inline vtz::ByteSource* open_any(const std::string& path)
{
    vtz::ByteSource* s = vtz::open_source(path);
    if (!s) return nullptr;
    unsigned char head[5];
    size_t n = 0;
    while (n < 5)
    {
        const long got = s->read((char*)head + n, 5 - n);
        if (got <= 0) break;
        n += (size_t)got;
    }
    if (n == 5 && !memcmp(head, "BCF\2\2", 5)) return new BcfTextSource(s, head, n);
    struct Pushback : vtz::ByteSource
    { // hands the sniffed bytes out first
        vtz::ByteSource* in;
        unsigned char h[5];
        size_t n, p = 0;
        Pushback(vtz::ByteSource* s, const unsigned char* b, size_t k) : in(s), n(k) { memcpy(h, b, k); }
        ~Pushback() override { delete in; }
        long read(char* dst, size_t want) override
        {
            if (p < n)
            {
                const size_t k = std::min(want, n - p);
                memcpy(dst, h + p, k);
                p += k;
                return (long)k;
            }
            return in->read(dst, want);
        }
    };
    return new Pushback(s, head, n);
}

// ---------------------------------------------------------------------------------------------------------
// BCF2 out: VCF text in, BCF2 bytes out (vcf.c vcf_parse + bcf_write restated for what a VCF line can hold)
// ---------------------------------------------------------------------------------------------------------
struct BcfSink
{
    FILE* f = nullptr;       // where the bytes go
    vtz::BgzfWriter* bz = nullptr; // through BGZF when set (.bcf); raw otherwise (.ubcf, "+")
    Dict dict;
    std::string pend, header;
    bool header_done = false, failed = false;
    std::string rec; // one encoded record

    void emit(const void* p, size_t n)
    {
        if (bz)
        {
            if (!bz->write((const char*)p, n)) failed = true;
        }
        else if (fwrite(p, 1, n, f) != n)
            failed = true;
    }
    static void put_u32(std::string& o, uint32_t v)
    {
        const char b[4] = {(char)(v & 0xff), (char)((v >> 8) & 0xff), (char)((v >> 16) & 0xff), (char)((v >> 24) & 0xff)};
        o.append(b, 4);
    }
    static void enc_size(std::string& o, int count, int type)
    { // bcf_enc_size
        if (count >= 15)
        {
            o.push_back((char)(15 << 4 | type));
            if (count >= 128)
            {
                if (count >= 32768)
                {
                    o.push_back((char)(1 << 4 | BT_INT32));
                    put_u32(o, (uint32_t)count);
                }
                else
                {
                    o.push_back((char)(1 << 4 | BT_INT16));
                    o.push_back((char)(count & 0xff));
                    o.push_back((char)(count >> 8));
                }
            }
            else
            {
                o.push_back((char)(1 << 4 | BT_INT8));
                o.push_back((char)count);
            }
        }
        else
            o.push_back((char)(count << 4 | type));
    }
    static int int_type(int32_t mn, int32_t mx)
    { // the narrowest type that holds [mn, mx] beside its reserved values (vcf.h BCF_MIN_BT_INT8 = -120, BCF_MIN_BT_INT16 = -32760)
        if (mx <= INT8_MAX && mn >= -120) return BT_INT8;
        if (mx <= INT16_MAX && mn >= -32760) return BT_INT16;
        return BT_INT32;
    }
    static void put_int(std::string& o, int32_t v, int type)
    { // INT32_MIN = missing, INT32_MIN + 1 = end of vector
        if (type == BT_INT8)
            o.push_back(v == INT32_MIN ? (char)0x80 : v == INT32_MIN + 1 ? (char)0x81 : (char)v);
        else if (type == BT_INT16)
        {
            const uint16_t x = v == INT32_MIN ? 0x8000 : v == INT32_MIN + 1 ? 0x8001 : (uint16_t)(int16_t)v;
            o.push_back((char)(x & 0xff));
            o.push_back((char)(x >> 8));
        }
        else
            put_u32(o, (uint32_t)v);
    }
    static void enc_vint(std::string& o, const std::vector<int32_t>& v)
    { // bcf_enc_vint: one type for the whole vector
        if (v.empty())
        {
            enc_size(o, 0, BT_NULL);
            return;
        }
        int32_t mn = 0, mx = 0;
        bool any = false; // a value that is neither missing nor end-of-vector
        for (int32_t x : v)
            if (x != INT32_MIN && x != INT32_MIN + 1)
            {
                mn = any ? std::min(mn, x) : x, mx = any ? std::max(mx, x) : x;
                any = true;
            }
        const int t = int_type(mn, mx);
        enc_size(o, (int)v.size(), t);
        for (int32_t x : v) put_int(o, x, t);
    }
    static void enc_str(std::string& o, const char* p, size_t n)
    {
        enc_size(o, (int)n, BT_CHAR);
        o.append(p, n);
    }
    static uint32_t float_bits(const char* p, size_t n)
    {
        if (n == 1 && p[0] == '.') return FLOAT_MISSING;
        const float f = strtof(std::string(p, n).c_str(), nullptr);
        uint32_t b;
        memcpy(&b, &f, 4);
        return b;
    }
    static int32_t int_of(const char* p, size_t n)
    {
        if (n == 1 && p[0] == '.') return INT32_MIN;
        return (int32_t)strtol(std::string(p, n).c_str(), nullptr, 10);
    }
    typedef std::pair<const char*, size_t> Tok;
    static void split(std::vector<Tok>& v, const char* s, size_t n, char sep)
    {
        v.clear();
        size_t b = 0;
        for (size_t k = 0; k <= n; ++k)
            if (k == n || s[k] == sep)
            {
                v.push_back(Tok(s + b, k - b));
                b = k + 1;
            }
    }
    int id_of(const Tok& t, const char* what)
    {
        auto it = dict.id_idx.find(std::string(t.first, t.second));
        if (it == dict.id_idx.end()) bcf_die(std::string(what) + " '" + std::string(t.first, t.second) + "' is not defined in the header");
        return it->second;
    }
    void write_header()
    {
        header_done = true;
        dict.build(header);
        std::string h = "BCF\2\2";
        put_u32(h, (uint32_t)header.size() + 1);
        h += header;
        h.push_back('\0');
        emit(h.data(), h.size());
    }
    void encode(const char* line, size_t n)
    {
        std::vector<Tok> col, part, vals;
        split(col, line, n, '\t');
        if (col.size() < 8) bcf_die("a record with fewer than 8 columns");
        auto ci = dict.contig_idx.find(std::string(col[0].first, col[0].second));
        if (ci == dict.contig_idx.end()) bcf_die("contig '" + std::string(col[0].first, col[0].second) + "' is not defined in the header");
        const int32_t pos0 = (int32_t)strtol(std::string(col[1].first, col[1].second).c_str(), nullptr, 10) - 1;
        std::string sh; // after the 24 fixed bytes
        if (col[2].second == 1 && col[2].first[0] == '.') enc_size(sh, 0, BT_CHAR);
        else enc_str(sh, col[2].first, col[2].second);
        enc_str(sh, col[3].first, col[3].second);
        int n_allele = 1;
        if (!(col[4].second == 1 && col[4].first[0] == '.'))
        {
            split(part, col[4].first, col[4].second, ',');
            for (const Tok& a : part) enc_str(sh, a.first, a.second);
            n_allele += (int)part.size();
        }
        std::vector<int32_t> iv;
        if (!(col[6].second == 1 && col[6].first[0] == '.'))
        {
            split(part, col[6].first, col[6].second, ';');
            for (const Tok& t : part) iv.push_back(id_of(t, "FILTER"));
        }
        enc_vint(sh, iv);
        int n_info = 0;
        int32_t rlen = (int32_t)col[3].second;
        if (!(col[7].second == 1 && col[7].first[0] == '.'))
        {
            split(part, col[7].first, col[7].second, ';');
            for (const Tok& kv : part)
            {
                if (!kv.second) continue;
                const char* eq = (const char*)memchr(kv.first, '=', kv.second);
                const Tok key(kv.first, eq ? (size_t)(eq - kv.first) : kv.second);
                const int idx = id_of(key, "INFO");
                auto it = dict.info.find(std::string(key.first, key.second));
                if (it == dict.info.end()) bcf_die("INFO '" + std::string(key.first, key.second) + "' is not defined in the header");
                iv.assign(1, idx);
                enc_vint(sh, iv);
                ++n_info;
                const char* vp = eq ? eq + 1 : nullptr;
                const size_t vn = eq ? kv.second - key.second - 1 : 0;
                const char ty = it->second.type;
                if (!eq || ty == 'L')
                    enc_size(sh, 0, BT_NULL);
                else if (ty == 'I')
                {
                    split(vals, vp, vn, ',');
                    iv.clear();
                    for (const Tok& v : vals) iv.push_back(int_of(v.first, v.second));
                    enc_vint(sh, iv);
                    if (key.second == 3 && !memcmp(key.first, "END", 3) && iv.size() == 1 && iv[0] > pos0) rlen = iv[0] - pos0; // vcf.c: rlen from INFO/END
                }
                else if (ty == 'F')
                {
                    split(vals, vp, vn, ',');
                    enc_size(sh, (int)vals.size(), BT_FLOAT);
                    for (const Tok& v : vals) put_u32(sh, float_bits(v.first, v.second));
                }
                else
                    enc_str(sh, vp, vn);
            }
        }
        std::string ind;
        int n_fmt = 0, n_sample = 0;
        if (col.size() > 9)
        {
            n_sample = (int)col.size() - 9;
            std::vector<Tok> keys;
            split(keys, col[8].first, col[8].second, ':');
            std::vector<std::vector<Tok>> S((size_t)n_sample);
            for (int s = 0; s < n_sample; ++s) split(S[(size_t)s], col[9 + (size_t)s].first, col[9 + (size_t)s].second, ':');
            static const char dot = '.';
            for (size_t k = 0; k < keys.size(); ++k)
            {
                const std::string kn(keys[k].first, keys[k].second);
                auto it = dict.fmt.find(kn);
                if (it == dict.fmt.end()) bcf_die("FORMAT '" + kn + "' is not defined in the header");
                iv.assign(1, id_of(keys[k], "FORMAT"));
                enc_vint(ind, iv);
                ++n_fmt;
                auto field = [&](int s) { return k < S[(size_t)s].size() ? S[(size_t)s][k] : Tok(&dot, 1); };
                const char ty = it->second.type;
                if (kn == "GT")
                { // vcf.c vcf_parse_format, the GT branch: (allele + 1) << 1 | phased
                    std::vector<std::vector<int32_t>> g((size_t)n_sample);
                    size_t w = 1;
                    int32_t mx = 0;
                    for (int s = 0; s < n_sample; ++s)
                    {
                        const Tok t = field(s);
                        int phased = 0;
                        size_t b = 0;
                        for (size_t q = 0; q <= t.second; ++q)
                            if (q == t.second || t.first[q] == '/' || t.first[q] == '|')
                            {
                                int32_t a = 0;
                                if (!(q - b == 1 && t.first[b] == '.') && q > b) a = (int32_t)strtol(std::string(t.first + b, q - b).c_str(), nullptr, 10) + 1;
                                g[(size_t)s].push_back(a << 1 | phased);
                                mx = std::max(mx, a << 1 | 1);
                                if (q < t.second) phased = t.first[q] == '|';
                                b = q + 1;
                            }
                        w = std::max(w, g[(size_t)s].size());
                    }
                    const int t = int_type(0, mx);
                    enc_size(ind, (int)w, t);
                    for (int s = 0; s < n_sample; ++s)
                        for (size_t a = 0; a < w; ++a) put_int(ind, a < g[(size_t)s].size() ? g[(size_t)s][a] : INT32_MIN + 1, t);
                }
                else if (ty == 'I' || ty == 'F')
                {
                    std::vector<std::vector<Tok>> V((size_t)n_sample);
                    size_t w = 1;
                    for (int s = 0; s < n_sample; ++s)
                    {
                        const Tok t = field(s);
                        split(V[(size_t)s], t.first, t.second, ',');
                        w = std::max(w, V[(size_t)s].size());
                    }
                    if (ty == 'I')
                    {
                        int32_t mn = 0, mx = 0;
                        bool any = false;
                        std::vector<int32_t> all;
                        for (int s = 0; s < n_sample; ++s)
                            for (size_t a = 0; a < w; ++a)
                            {
                                const int32_t x = a < V[(size_t)s].size() ? int_of(V[(size_t)s][a].first, V[(size_t)s][a].second) : INT32_MIN + 1;
                                all.push_back(x);
                                if (x != INT32_MIN && x != INT32_MIN + 1)
                                {
                                    mn = any ? std::min(mn, x) : x, mx = any ? std::max(mx, x) : x;
                                    any = true;
                                }
                            }
                        const int t = int_type(mn, mx);
                        enc_size(ind, (int)w, t);
                        for (int32_t x : all) put_int(ind, x, t);
                    }
                    else
                    {
                        enc_size(ind, (int)w, BT_FLOAT);
                        for (int s = 0; s < n_sample; ++s)
                            for (size_t a = 0; a < w; ++a)
                                put_u32(ind, a < V[(size_t)s].size() ? float_bits(V[(size_t)s][a].first, V[(size_t)s][a].second) : FLOAT_VEND);
                    }
                }
                else
                {
                    size_t w = 1;
                    for (int s = 0; s < n_sample; ++s) w = std::max(w, field(s).second);
                    enc_size(ind, (int)w, BT_CHAR);
                    for (int s = 0; s < n_sample; ++s)
                    {
                        const Tok t = field(s);
                        ind.append(t.first, t.second);
                        ind.append(w - t.second, '\0');
                    }
                }
            }
        }
        rec.clear();
        put_u32(rec, (uint32_t)(24 + sh.size()));
        put_u32(rec, (uint32_t)ind.size());
        put_u32(rec, (uint32_t)ci->second);
        put_u32(rec, (uint32_t)pos0);
        put_u32(rec, (uint32_t)rlen);
        put_u32(rec, float_bits(col[5].first, col[5].second));
        put_u32(rec, (uint32_t)n_allele << 16 | (uint32_t)n_info);
        put_u32(rec, (uint32_t)n_fmt << 24 | (uint32_t)n_sample);
        rec += sh;
        rec += ind;
        emit(rec.data(), rec.size());
    }
    bool write(const char* p, size_t n)
    {
        pend.append(p, n);
        size_t b = 0;
        for (;;)
        {
            const char* nl = (const char*)memchr(pend.data() + b, '\n', pend.size() - b);
            if (!nl) break;
            const size_t e = (size_t)(nl - pend.data());
            if (pend[b] == '#')
                header.append(pend.data() + b, e - b + 1);
            else
            {
                if (!header_done) write_header();
                if (e > b) encode(pend.data() + b, e - b);
            }
            b = e + 1;
        }
        pend.erase(0, b);
        return !failed;
    }
    bool close()
    {
        if (!pend.empty()) write("\n", 1);
        if (!header_done) write_header();
        if (bz) return bz->close() && !failed;
        return fflush(f) == 0 && !failed;
    }
};

} // namespace vtb2
