// Shared declarations of libccpack.so (host only): the inflated BAM, the packed batch, raw-record
// accessors and the key helpers of consensus_helper.py.  ccpack.cpp reads and packs, ccemit.cpp
// assembles and writes the output files.
#pragma once
#include <zlib.h>


#include <algorithm>
#include <atomic>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <string_view>
#include <thread>
#include <unordered_map>
#include <vector>

namespace ccp {

struct Fields {            // cc_fields of include/ccgpu.h
    uint32_t bc, cig;
    int32_t tid, pos, mtid, mpos, tlen;
    uint16_t flag;
    uint8_t mapq, rg;
};
static_assert(sizeof(Fields) == 32, "cc_fields is 32 bytes");

inline uint32_t rd32(const uint8_t *p) { uint32_t v; memcpy(&v, p, 4); return v; }
inline int32_t rdi32(const uint8_t *p) { int32_t v; memcpy(&v, p, 4); return v; }
inline uint16_t rd16(const uint8_t *p) { uint16_t v; memcpy(&v, p, 2); return v; }

inline void set_err(char *err, int len, const std::string &msg)
{
    if (err && len > 0) snprintf(err, len, "%s", msg.c_str());
}

}  // namespace ccp
using namespace ccp;

// byte buffer that is not zero-filled on allocation (the inflate threads write every byte of it)
struct RawBytes {
    uint8_t *p = nullptr;
    size_t n = 0;
    RawBytes() = default;
    RawBytes(const RawBytes &) = delete;
    RawBytes &operator=(const RawBytes &) = delete;
    ~RawBytes() { free(p); }
    bool resize(size_t bytes) { free(p); p = static_cast<uint8_t *>(malloc(bytes ? bytes : 1)); n = p ? bytes : 0; return p != nullptr; }
    uint8_t *data() { return p; }
    const uint8_t *data() const { return p; }
    size_t size() const { return n; }
};

struct ccp_bam {
    RawBytes data;                        // the whole uncompressed BAM stream
    std::string text;
    std::vector<std::string> ref_names;
    std::vector<int32_t> ref_lens;
    std::vector<uint64_t> offsets;        // offset of every record's block_size field
};

struct ccp_pack {
    std::vector<Fields> fields;
    std::vector<uint8_t> payload;
    std::vector<int64_t> src;
    std::vector<int32_t> lens;
    std::vector<int64_t> bad, leftover;
    std::vector<uint32_t> dup_bc;
    std::string barcodes, cigars, rgs;    // NUL-separated dictionaries in id order
    uint32_t n_bc = 0, n_cig = 0, n_rg = 0, lpad = 32, stride = 64;
    int64_t counters[5] = {0, 0, 0, 0, 0};   // total, unmapped, unmapped_mate, multiple_mapping, bad_spacer
};


// ---- record access and the key helpers of consensus_helper.py ------------------------------------
namespace ccp {

struct Rec {
    const uint8_t *p;       // points at block_size
    int32_t tid() const { return rdi32(p + 4); }
    int32_t pos() const { return rdi32(p + 8); }
    uint32_t l_name() const { return p[12]; }
    uint8_t mapq() const { return p[13]; }
    uint32_t n_cigar() const { return rd16(p + 16); }
    uint16_t flag() const { return rd16(p + 18); }
    uint32_t l_seq() const { return rd32(p + 20); }
    int32_t mtid() const { return rdi32(p + 24); }
    int32_t mpos() const { return rdi32(p + 28); }
    int32_t tlen() const { return rdi32(p + 32); }
    std::string_view qname() const { return {reinterpret_cast<const char *>(p + 36), l_name() ? l_name() - 1 : 0}; }
    const uint8_t *cigar() const { return p + 36 + l_name(); }
    const uint8_t *seq() const { return cigar() + 4 * n_cigar(); }
    const uint8_t *qual() const { return seq() + (l_seq() + 1) / 2; }
    const uint8_t *aux() const { return qual() + l_seq(); }
    const uint8_t *end() const { return p + 4 + rd32(p); }
};

inline std::string cigar_string(const Rec &r)
{
    static const char ops[] = "MIDNSHP=X";
    std::string s;
    for (uint32_t i = 0; i < r.n_cigar(); i++) {
        uint32_t w = rd32(r.cigar() + 4 * i);
        s += std::to_string(w >> 4);
        s += ops[w & 15];
    }
    return s;
}

// which_read (consensus_helper.py:57-81): 1 = R1, 2 = R2, 0 = none of the 16 paired flags
inline int read_number(uint16_t f)
{
    switch (f) {
    case 99: case 83: case 67: case 115: case 81: case 97: case 65: case 113: return 1;
    case 147: case 163: case 131: case 179: case 161: case 145: case 129: case 177: return 2;
    default: return 0;
    }
}

// which_strand (consensus_helper.py:84-155): 1 = pos, 0 = neg
inline int strand_of(const Rec &r)
{
    uint16_t f = r.flag();
    if (f == 99 || f == 147 || f == 67 || f == 131) return 1;
    if (f == 83 || f == 163 || f == 115 || f == 179) return 0;
    bool first = read_number(f) == 1;
    if (r.tid() != r.mtid()) return ((r.tid() < r.mtid()) == first) ? 1 : 0;
    if (r.pos() == r.mpos()) return 0;
    return ((r.pos() < r.mpos()) == first) ? 1 : 0;
}

inline bool is_mate_unmapped_flag(uint16_t f) { return f == 73 || f == 89 || f == 121 || f == 153 || f == 185 || f == 137; }

// value of the RG:Z tag, empty view with data() == nullptr when absent
inline std::string_view find_rg(const Rec &r)
{
    const uint8_t *p = r.aux(), *e = r.end();
    auto size_of = [](char t) { return t == 'A' || t == 'c' || t == 'C' ? 1 : t == 's' || t == 'S' ? 2 : 4; };
    while (p + 3 <= e) {
        char t = (char)p[2];
        const uint8_t *v = p + 3;
        if (t == 'Z' || t == 'H') {
            const uint8_t *z = v;
            while (z < e && *z) z++;
            if (p[0] == 'R' && p[1] == 'G' && t == 'Z') return {reinterpret_cast<const char *>(v), (size_t)(z - v)};
            p = z + 1;
        } else if (t == 'B') {
            uint32_t cnt = rd32(v + 1);
            p = v + 5 + (size_t)cnt * size_of((char)v[0]);
        } else {
            p = v + size_of(t);
        }
    }
    return {};
}

inline std::string duplex_barcode(const std::string &b)
{
    size_t dot = b.find('.');
    if (dot != std::string::npos) return b.substr(dot + 1) + "." + b.substr(0, dot);
    size_t h = b.size() / 2;
    return b.substr(h) + b.substr(0, h);
}

struct Dict {
    std::unordered_map<std::string, uint32_t> ids;
    std::vector<std::string> names;
    uint32_t get(const std::string &s)
    {
        auto it = ids.find(s);
        if (it != ids.end()) return it->second;
        uint32_t id = (uint32_t)names.size();
        ids.emplace(s, id);
        names.push_back(s);
        return id;
    }
    std::string joined() const
    {
        std::string out;
        for (auto &n : names) { out += n; out.push_back('\0'); }
        return out;
    }
};

}  // namespace ccp
