// libccpack.so - native host side of the packing step (no CUDA): BGZF inflate on all cores,
// BAM record index, and the filter / mate-pairing / dictionary-encoding half of read_bam
// (consensus_helper.py:376-441) straight into the SoA that libccgpu.so consumes.
// It produces byte-for-byte what consensuscruncher_b200/pack.py:pack_reads produces
// (tests/test_fastpack.py); pack.py remains the readable statement of the rules.
// Build: g++ -O2 -std=c++17 -shared -fPIC ccpack.cpp -lz -lpthread
#include "ccpack.hpp"

#include <chrono>
#include <cstdlib>

namespace {

struct Block { uint64_t coff, clen, uoff, ulen; };

bool inflate_block(const uint8_t *src, uint64_t clen, uint8_t *dst, uint64_t ulen)
{
    z_stream zs;
    memset(&zs, 0, sizeof(zs));
    if (inflateInit2(&zs, -15) != Z_OK) return false;
    zs.next_in = const_cast<uint8_t *>(src);
    zs.avail_in = (uInt)clen;
    zs.next_out = dst;
    zs.avail_out = (uInt)ulen;
    int rc = inflate(&zs, Z_FINISH);
    inflateEnd(&zs);
    return rc == Z_STREAM_END && zs.total_out == ulen;
}

}  // namespace

extern "C" {

ccp_bam *ccp_bam_open(const char *path, int threads, char *err, int errlen)
{
    const bool timing = getenv("CCP_TIMING") != nullptr;
    auto t_start = std::chrono::steady_clock::now();
    auto lap = [&](const char *what) {
        if (!timing) return;
        auto now = std::chrono::steady_clock::now();
        fprintf(stderr, "ccp_bam_open %-22s %.3f s\n", what, std::chrono::duration<double>(now - t_start).count());
        t_start = now;
    };
    FILE *fh = fopen(path, "rb");
    if (!fh) { set_err(err, errlen, std::string("cannot open ") + path); return nullptr; }
    fseeko(fh, 0, SEEK_END);
    const int64_t size = (int64_t)ftello(fh);
    fseeko(fh, 0, SEEK_SET);
    if (size < 0) { fclose(fh); set_err(err, errlen, std::string("cannot size ") + path); return nullptr; }
    std::vector<uint8_t> comp((size_t)size);
    if (size && fread(comp.data(), 1, (size_t)size, fh) != (size_t)size) {
        fclose(fh);
        set_err(err, errlen, "short read");
        return nullptr;
    }
    fclose(fh);
    // block table: the BC extra field gives the compressed size, the trailer the uncompressed one
    std::vector<Block> blocks;
    uint64_t off = 0, total = 0;
    while (off + 18 <= (uint64_t)size) {
        const uint8_t *p = comp.data() + off;
        if (!(p[0] == 0x1f && p[1] == 0x8b && p[2] == 8 && (p[3] & 4))) { set_err(err, errlen, "not a BGZF block"); return nullptr; }
        uint32_t xlen = rd16(p + 10), bsize = 0;
        bool found = false;
        if (off + 12 + xlen > (uint64_t)size) { set_err(err, errlen, "damaged BGZF block (extra field runs past the file)"); return nullptr; }
        for (uint32_t q = 0; q + 4 <= xlen;) {
            uint32_t slen = rd16(p + 12 + q + 2);
            if (p[12 + q] == 'B' && p[12 + q + 1] == 'C') { bsize = rd16(p + 12 + q + 4); found = true; }
            q += 4 + slen;
        }
        if (!found || off + bsize + 1 > (uint64_t)size || (uint64_t)bsize + 1 < 12ull + xlen + 8) {
            set_err(err, errlen, "damaged BGZF block");
            return nullptr;
        }
        uint64_t blen = (uint64_t)bsize + 1;
        uint32_t isize = rd32(p + blen - 4);
        blocks.push_back({off + 12 + xlen, blen - 12 - xlen - 8, total, isize});
        total += isize;
        off += blen;
    }
    lap("read + block table");
    ccp_bam *b = new ccp_bam();
    if (!b->data.resize(total)) { delete b; set_err(err, errlen, "out of memory"); return nullptr; }
    lap("allocate");
    std::atomic<size_t> next{0};
    std::atomic<bool> ok{true};
    int nt = std::max(1, std::min(threads, 64));
    std::vector<std::thread> pool;
    for (int t = 0; t < nt; t++)
        pool.emplace_back([&]() {
            for (size_t i; (i = next.fetch_add(1)) < blocks.size();) {
                const Block &k = blocks[i];
                if (k.ulen && !inflate_block(comp.data() + k.coff, k.clen, b->data.data() + k.uoff, k.ulen)) ok = false;
            }
        });
    for (auto &t : pool) t.join();
    lap("inflate");
    if (!ok) { delete b; set_err(err, errlen, "inflate failed"); return nullptr; }
    const uint8_t *d = b->data.data();
    const uint64_t n = b->data.size();
    if (n < 12 || memcmp(d, "BAM\1", 4) != 0) { delete b; set_err(err, errlen, "not a BAM file"); return nullptr; }
    // every length field is checked against the inflated size: a truncated or corrupt file gives
    // "damaged BAM", as bam.py raises, instead of reads past the buffer
    auto damaged = [&](const char *what) { delete b; set_err(err, errlen, std::string("damaged BAM: ") + what); return nullptr; };
    uint32_t l_text = rd32(d + 4);
    if (8 + (uint64_t)l_text + 4 > n) return damaged("header text runs past the data");
    b->text.assign(reinterpret_cast<const char *>(d + 8), strnlen(reinterpret_cast<const char *>(d + 8), l_text));
    uint64_t p = 8 + (uint64_t)l_text;
    uint32_t n_ref = rd32(d + p);
    p += 4;
    if ((uint64_t)n_ref * 8 > n - p) return damaged("reference count exceeds the data");
    for (uint32_t i = 0; i < n_ref; i++) {
        if (p + 4 > n) return damaged("reference table is truncated");
        uint32_t l_name = rd32(d + p);
        if (p + 8 + (uint64_t)l_name > n) return damaged("reference name runs past the data");
        b->ref_names.emplace_back(reinterpret_cast<const char *>(d + p + 4), l_name ? l_name - 1 : 0);
        b->ref_lens.push_back(rdi32(d + p + 4 + l_name));
        p += 8 + l_name;
    }
    while (p + 4 <= n) {
        const uint64_t block_size = rd32(d + p);
        if (block_size < 32 || p + 4 + block_size > n) return damaged("alignment record runs past the data");
        const Rec r{d + p};
        if (32ull + r.l_name() + 4ull * r.n_cigar() + ((uint64_t)r.l_seq() + 1) / 2 + r.l_seq() > block_size)
            return damaged("alignment record fields exceed its block_size");
        b->offsets.push_back(p);
        p += 4 + block_size;
    }
    if (p != n) return damaged("trailing bytes after the last alignment record");
    lap("record index");
    return b;
}

void ccp_bam_close(ccp_bam *b) { delete b; }
uint64_t ccp_bam_n_records(ccp_bam *b) { return b->offsets.size(); }
const uint8_t *ccp_bam_data(ccp_bam *b, uint64_t *len) { *len = b->data.size(); return b->data.data(); }
const uint64_t *ccp_bam_offsets(ccp_bam *b) { return b->offsets.data(); }
const char *ccp_bam_header_text(ccp_bam *b) { return b->text.c_str(); }
int ccp_bam_n_refs(ccp_bam *b) { return (int)b->ref_names.size(); }
const char *ccp_bam_ref_name(ccp_bam *b, int i) { return b->ref_names[i].c_str(); }
int32_t ccp_bam_ref_len(ccp_bam *b, int i) { return b->ref_lens[i]; }

}  // extern "C"

// ---- pack: the rules of pack.py:pair_reads / pack_reads -------------------------------------------
namespace {

// pack.py:visit_order - record indices in the order read_bam (consensus_helper.py:376-381) sees them
std::vector<int64_t> visit_order(const ccp_bam *b, const int32_t *regions, int n_regions)
{
    const uint8_t *d = b->data.data();
    const int64_t n = (int64_t)b->offsets.size();
    std::vector<int64_t> out;
    if (n_regions < 0) {
        out.resize(n);
        for (int64_t i = 0; i < n; i++) out[i] = i;
        return out;
    }
    struct Contig { std::vector<int64_t> idx; std::vector<int32_t> pos; bool sorted = true; };
    std::unordered_map<int32_t, Contig> by_tid;
    for (int64_t i = 0; i < n; i++) {
        Rec r{d + b->offsets[i]};
        Contig &c = by_tid[r.tid()];
        if (!c.pos.empty() && c.pos.back() > r.pos()) c.sorted = false;
        c.idx.push_back(i);
        c.pos.push_back(r.pos());
    }
    for (int g = 0; g < n_regions; g++) {
        const int32_t tid = regions[3 * g], start = regions[3 * g + 1], end = regions[3 * g + 2];
        auto it = by_tid.find(tid);
        if (it == by_tid.end()) continue;
        const Contig &c = it->second;
        if (c.sorted) {
            size_t lo = std::lower_bound(c.pos.begin(), c.pos.end(), start) - c.pos.begin();
            size_t hi = std::lower_bound(c.pos.begin(), c.pos.end(), end) - c.pos.begin();
            for (size_t k = lo; k < hi; k++) out.push_back(c.idx[k]);
        } else {
            for (size_t k = 0; k < c.idx.size(); k++)
                if (start <= c.pos[k] && c.pos[k] < end) out.push_back(c.idx[k]);
        }
    }
    return out;
}

// qname.split(delim)[1] (sscs stage) / qname.split('_')[0] (dcs stage), consensus_helper.py:423-431
std::string_view barcode_view(std::string_view q, bool sscs, std::string_view delim)
{
    if (!sscs) return q.substr(0, q.find('_'));
    size_t a = q.find(delim);
    if (a == std::string_view::npos) return std::string_view();
    a += delim.size();
    size_t e = q.find(delim, a);
    return q.substr(a, e == std::string_view::npos ? e : e - a);
}

}  // namespace

extern "C" {

// stage 0 = sscs (bad reads diverted, barcode after `delim`), 1 = dcs.  regions: n_regions x
// (tid, start, end), n_regions < 0 for file order.  nullptr + message on error; a message that
// starts with "domain:" is pack.DomainError on the Python side.
ccp_pack *ccp_pack_run(ccp_bam *b, int stage, const char *delim_c, const int32_t *regions, int n_regions,
                       int threads, char *err, int errlen)
{
    const bool sscs = stage == 0;
    const std::string_view delim(delim_c ? delim_c : "");
    const uint8_t *d = b->data.data();
    ccp_pack *out = new ccp_pack();
    int64_t *ct = out->counters;
    const bool timing = getenv("CCP_TIMING") != nullptr;
    auto t_start = std::chrono::steady_clock::now();
    auto lap = [&](const char *what) {
        if (!timing) return;
        auto now = std::chrono::steady_clock::now();
        fprintf(stderr, "ccp_pack_run %-22s %.3f s\n", what, std::chrono::duration<double>(now - t_start).count());
        t_start = now;
    };
    std::vector<int64_t> order = visit_order(b, regions, n_regions);
    lap("visit order");
    // Mates meet through an open-addressing table keyed by the query name (a slot: hash, record, visit
    // step; a record leaves the table when its mate arrives, the slot stays as a tombstone so that probe
    // chains of other names are not cut).
    struct Waiting { uint64_t hash; int64_t idx, step; };     // idx: -1 empty, -2 tombstone
    size_t cap = 64;
    while (cap < order.size() + order.size() / 2 + 16) cap <<= 1;
    std::vector<Waiting> table(cap, Waiting{0, -1, 0});
    std::vector<int64_t> &src = out->src;
    src.reserve(order.size());
    int64_t step = 0;
    for (int64_t i : order) {
        Rec r{d + b->offsets[i]};
        const uint16_t f = r.flag();
        std::string_view q = r.qname();
        ct[0]++;
        bool is_bad = true;
        if (sscs && q.find(delim) == std::string_view::npos) ct[4]++;
        else if (f & 0x4) { ct[1]++; ct[0]--; }
        else if (is_mate_unmapped_flag(f)) ct[2]++;
        else if (f & 0x900) ct[3]++;
        else is_bad = false;
        step++;
        if (is_bad && sscs) { out->bad.push_back(i); continue; }
        const uint64_t h = std::hash<std::string_view>{}(q) | 1;
        size_t s = (size_t)(h * 0x9E3779B97F4A7C15ull >> 17) & (cap - 1), grave = cap;
        for (;; s = (s + 1) & (cap - 1)) {
            Waiting &w = table[s];
            if (w.idx == -1) {                               // not waiting: this record waits (in the first free slot seen)
                Waiting &dst = grave != cap ? table[grave] : w;
                dst = Waiting{h, i, step};
                break;
            }
            if (w.idx == -2) { if (grave == cap) grave = s; continue; }
            if (w.hash == h && Rec{d + b->offsets[w.idx]}.qname() == q) {
                src.push_back(w.idx);                        // the pair is complete: first-seen mate, then this one
                src.push_back(i);
                w.idx = -2;
                break;
            }
        }
    }
    {
        std::vector<Waiting> left;
        for (auto &w : table)
            if (w.idx >= 0) left.push_back(w);
        std::sort(left.begin(), left.end(), [](const Waiting &x, const Waiting &y) { return x.step < y.step; });
        for (auto &w : left) out->leftover.push_back(w.idx);
    }
    lap("filter + pairing");
    const int64_t n = (int64_t)src.size();
    out->fields.resize(n);
    out->lens.resize(n);
    Dict bcs, cigs, rgs;
    uint32_t max_len = 0;
    std::string ckey, raw, last_raw;
    // The dictionaries are keyed by strings, but almost every pair repeats what an earlier one had: look
    // the raw bytes up first (views into the BAM / binary cigars) and build a string only for a new value.
    std::unordered_map<std::string_view, uint32_t> bc_seen, rg_seen;
    std::unordered_map<std::string, uint32_t> cig_seen;
    uint32_t last_cg = 0;
    std::string_view last_rg;
    uint32_t last_rg_id = 0;
    for (int64_t k = 0; k < n; k += 2) {
        Rec read{d + b->offsets[src[k]]}, mate{d + b->offsets[src[k + 1]]};
        for (const Rec *r : {&read, &mate})
            if (!read_number(r->flag())) {
                set_err(err, errlen, "domain: flag " + std::to_string(r->flag()) + " of " + std::string(r->qname()) +
                                         " is not one of the 16 paired flags the reference can key "
                                         "(consensus_helper.py:69-70)");
                delete out;
                return nullptr;
            }
        const std::string_view bview = barcode_view(read.qname(), sscs, delim);
        uint32_t bc;
        if (auto it = bc_seen.find(bview); it != bc_seen.end()) bc = it->second;
        else bc_seen.emplace(bview, bc = bcs.get(std::string(bview)));
        // cigar_order (consensus_helper.py:158-172): the read's cigar leads iff (pos and R1) or (neg and R2)
        const bool lead = (strand_of(read) == 1) == (read_number(read.flag()) == 1);
        const Rec &c1 = lead ? read : mate, &c2 = lead ? mate : read;
        raw.assign(reinterpret_cast<const char *>(c1.cigar()), 4 * c1.n_cigar());
        raw.push_back('_');
        raw.append(reinterpret_cast<const char *>(c2.cigar()), 4 * c2.n_cigar());
        raw.push_back((char)(c1.n_cigar() & 0xFF));          // the split point (n_cigar is 16 bits), so that two
        raw.push_back((char)(c1.n_cigar() >> 8));            // different pairs never give the same raw key
        uint32_t cg;
        if (k && raw == last_raw) cg = last_cg;
        else if (auto it = cig_seen.find(raw); it != cig_seen.end()) cg = it->second;
        else {
            ckey = cigar_string(c1) + "_" + cigar_string(c2);
            cig_seen.emplace(raw, cg = cigs.get(ckey));
        }
        last_raw = raw; last_cg = cg;
        for (int j = 0; j < 2; j++) {
            const Rec &r = j ? mate : read;
            Fields &f = out->fields[k + j];
            f.bc = bc; f.cig = cg;
            f.tid = r.tid(); f.pos = r.pos(); f.mtid = r.mtid(); f.mpos = r.mpos();
            f.tlen = r.tlen(); f.flag = r.flag(); f.mapq = r.mapq();
            std::string_view g = find_rg(r);
            if (g.data() == nullptr) f.rg = 0xff;
            else {
                uint32_t id;
                if (last_rg.data() != nullptr && g == last_rg) id = last_rg_id;
                else if (auto it = rg_seen.find(g); it != rg_seen.end()) id = it->second;
                else rg_seen.emplace(g, id = rgs.get(std::string(g)));
                last_rg = g; last_rg_id = id;
                if (id >= 0xff) { set_err(err, errlen, "domain: more than 255 distinct read groups"); delete out; return nullptr; }
                f.rg = (uint8_t)id;
            }
            out->lens[k + j] = (int32_t)r.l_seq();
            max_len = std::max(max_len, r.l_seq());
        }
    }
    lap("keys + fields");
    out->lpad = std::max(32u, (max_len + 31) / 32 * 32);
    out->stride = (out->lpad / 2 + out->lpad + 63) / 64 * 64;
    out->payload.assign((size_t)n * out->stride, 0);
    {
        // BAM's 4-bit codes with the nibbles of every byte swapped (payload: even position low)
        const uint32_t lpad = out->lpad, stride = out->stride;
        std::atomic<int64_t> next{0};
        const int64_t chunk = 4096;
        auto work = [&]() {
            for (int64_t lo; (lo = next.fetch_add(chunk)) < n;)
                for (int64_t x = lo; x < std::min(n, lo + chunk); x++) {
                    Rec r{d + b->offsets[src[x]]};
                    uint8_t *row = out->payload.data() + (size_t)x * stride;
                    const uint32_t L = r.l_seq();
                    const uint8_t *sq = r.seq();
                    for (uint32_t y = 0; y < L / 2; y++) row[y] = (uint8_t)((sq[y] << 4) | (sq[y] >> 4));
                    if (L & 1) row[L / 2] = (uint8_t)(sq[L / 2] >> 4);
                    memcpy(row + lpad / 2, r.qual(), L);
                }
        };
        int nt = std::max(1, std::min(threads, 64));
        std::vector<std::thread> pool;
        for (int t = 0; t < nt; t++) pool.emplace_back(work);
        for (auto &t : pool) t.join();
    }
    lap("payload");
    out->n_bc = (uint32_t)bcs.names.size();
    out->n_cig = (uint32_t)cigs.names.size();
    out->n_rg = (uint32_t)rgs.names.size();
    out->dup_bc.assign(out->n_bc, 0xffffffffu);
    for (uint32_t i = 0; i < out->n_bc; i++) {   // duplex_tag's barcode half swap (consensus_helper.py:629-639)
        auto it = bcs.ids.find(duplex_barcode(bcs.names[i]));
        if (it != bcs.ids.end()) out->dup_bc[i] = it->second;
    }
    out->barcodes = bcs.joined();
    out->cigars = cigs.joined();
    out->rgs = rgs.joined();
    return out;
}

void ccp_pack_free(ccp_pack *p) { delete p; }
uint64_t ccp_pack_n(ccp_pack *p) { return p->fields.size(); }
void ccp_pack_geometry(ccp_pack *p, uint32_t *lpad, uint32_t *stride) { *lpad = p->lpad; *stride = p->stride; }
const void *ccp_pack_fields(ccp_pack *p) { return p->fields.data(); }
const uint8_t *ccp_pack_payload(ccp_pack *p) { return p->payload.data(); }
const int64_t *ccp_pack_src(ccp_pack *p) { return p->src.data(); }
const int32_t *ccp_pack_lens(ccp_pack *p) { return p->lens.data(); }
const int64_t *ccp_pack_counters(ccp_pack *p) { return p->counters; }
const int64_t *ccp_pack_bad(ccp_pack *p, uint64_t *n) { *n = p->bad.size(); return p->bad.data(); }
const int64_t *ccp_pack_leftover(ccp_pack *p, uint64_t *n) { *n = p->leftover.size(); return p->leftover.data(); }
const uint32_t *ccp_pack_dup_bc(ccp_pack *p, uint32_t *n) { *n = p->n_bc; return p->dup_bc.data(); }
// which: 0 barcodes, 1 cigar pairs, 2 read groups; NUL-separated, in id order
const char *ccp_pack_dict(ccp_pack *p, int which, uint32_t *count, uint64_t *bytes)
{
    const std::string &s = which == 0 ? p->barcodes : which == 1 ? p->cigars : p->rgs;
    *count = which == 0 ? p->n_bc : which == 1 ? p->n_cig : p->n_rg;
    *bytes = s.size();
    return s.data();
}

}  // extern "C"
