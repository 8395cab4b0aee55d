// libccpack.so, output side (no CUDA): builds every record list `ConsensusCruncher.py consensus`
// leaves on disk from the family-level results of the GPU and writes them as BAM + BAI, with the
// BGZF deflate spread over the host cores.  It is the C++ statement of emit.py / bam.py's writer
// (which stay the readable version and the parity check, tests/test_fastemit.py):
//   record assembly     consensus_helper.py:538-585  create_aligned_segment
//   SSCS / singleton    SSCS_maker.py:274-298
//   DCS / passthrough   DCS_maker.py:226-259, dcs_consensus_tag :60-96
//   rescue outputs      singleton_correction.py:249-285
//   sort / merge        ConsensusCruncher.py:10-29, :247-251, :281-286
#include "ccpack.hpp"

#include <chrono>
#include <cstdlib>
#include <deque>
#include <map>
#include <tuple>
#include <unordered_set>

namespace {

// ---- one output record ------------------------------------------------------------------------
// Coordinates and cigar always come from a record of the input BAM (`tmpl`); bases and qualities
// either from that record too (passthrough: singletons, bad reads) or from a consensus row.
struct OutRec {
    int64_t fam = -1;
    std::string qname;
    int32_t tid = -1, pos = -1, mtid = -1, mpos = -1, tlen = 0;
    uint16_t flag = 0;
    uint8_t mapq = 0;
    const uint8_t *tmpl = nullptr;      // raw input record: cigar (+ bases, qualities, aux when row == nullptr)
    const uint8_t *row = nullptr;       // consensus row in the payload layout, or nullptr
    uint32_t l_seq = 0, lpad = 0;
    bool has_rg = false;
    std::string_view rg;                // RG:Z value (points into the input BAM or the rg dictionary)
};

inline int reg2bin(int64_t beg, int64_t end)
{
    end -= 1;
    if (beg >> 14 == end >> 14) return ((1 << 15) - 1) / 7 + (int)(beg >> 14);
    if (beg >> 17 == end >> 17) return ((1 << 12) - 1) / 7 + (int)(beg >> 17);
    if (beg >> 20 == end >> 20) return ((1 << 9) - 1) / 7 + (int)(beg >> 20);
    if (beg >> 23 == end >> 23) return ((1 << 6) - 1) / 7 + (int)(beg >> 23);
    if (beg >> 26 == end >> 26) return ((1 << 3) - 1) / 7 + (int)(beg >> 26);
    return 0;
}

// reference bases a cigar covers (M, D, N, =, X)
inline int64_t reference_length(const Rec &r)
{
    int64_t total = 0;
    for (uint32_t i = 0; i < r.n_cigar(); i++) {
        uint32_t w = rd32(r.cigar() + 4 * i), op = w & 15;
        if (op == 0 || op == 2 || op == 3 || op == 7 || op == 8) total += w >> 4;
    }
    return total;
}

inline void put32(std::vector<uint8_t> &out, uint32_t v) { uint8_t b[4]; memcpy(b, &v, 4); out.insert(out.end(), b, b + 4); }
inline void put16(std::vector<uint8_t> &out, uint16_t v) { uint8_t b[2]; memcpy(b, &v, 2); out.insert(out.end(), b, b + 2); }

// end of the interval the index files a record under (bam.py:encode_record / _BaiBuilder.add)
inline void index_interval(const OutRec &o, int64_t &beg, int64_t &end)
{
    int64_t ref_len = reference_length(Rec{o.tmpl});
    beg = std::max<int64_t>(o.pos, 0);
    end = beg + ((ref_len > 0 && !(o.flag & 4)) ? ref_len : 1);
}

// bam.py:encode_record - one alignment record appended to `out`
bool encode_record(const OutRec &o, std::vector<uint8_t> &out, std::string &err)
{
    const Rec t{o.tmpl};
    if (o.qname.size() > 254) { err = "query name longer than 254 characters: " + o.qname; return false; }
    const uint32_t l_name = (uint32_t)o.qname.size() + 1, n_cig = t.n_cigar(), l_seq = o.l_seq;
    // aux: RG first (as create_aligned_segment writes it), then - passthrough only - every other tag
    const uint8_t *aux = nullptr, *aux_end = nullptr;
    if (!o.row) { aux = t.aux(); aux_end = t.end(); }
    const size_t start = out.size();
    put32(out, 0);                                           // block_size, patched below
    put32(out, (uint32_t)o.tid);
    put32(out, (uint32_t)o.pos);
    int64_t ref_len = reference_length(t);
    int64_t end = (int64_t)o.pos + ((ref_len > 0 && !(o.flag & 4)) ? ref_len : 1);
    out.push_back((uint8_t)l_name);
    out.push_back(o.mapq);
    put16(out, (uint16_t)reg2bin(std::max<int64_t>(o.pos, 0), std::max<int64_t>(end, 1)));
    put16(out, (uint16_t)n_cig);
    put16(out, o.flag);
    put32(out, l_seq);
    put32(out, (uint32_t)o.mtid);
    put32(out, (uint32_t)o.mpos);
    put32(out, (uint32_t)o.tlen);
    out.insert(out.end(), o.qname.begin(), o.qname.end());
    out.push_back(0);
    out.insert(out.end(), t.cigar(), t.cigar() + 4 * n_cig);
    const size_t seq_at = out.size();
    out.resize(seq_at + (l_seq + 1) / 2 + l_seq);
    uint8_t *seq = out.data() + seq_at, *qual = seq + (l_seq + 1) / 2;
    if (o.row) {
        // payload rows keep the even position in the low nibble, BAM wants it in the high one
        for (uint32_t y = 0; y < (l_seq + 1) / 2; y++) seq[y] = (uint8_t)((o.row[y] << 4) | (o.row[y] >> 4));
        if (l_seq & 1) seq[l_seq / 2] &= 0xF0;
        memcpy(qual, o.row + o.lpad / 2, l_seq);
    } else {
        memcpy(seq, t.seq(), (l_seq + 1) / 2);
        if (l_seq & 1) seq[l_seq / 2] &= 0xF0;
        memcpy(qual, t.qual(), l_seq);
    }
    if (o.has_rg) {
        out.insert(out.end(), {'R', 'G', 'Z'});
        out.insert(out.end(), o.rg.begin(), o.rg.end());
        out.push_back(0);
    }
    if (aux) {                                               // every tag but RG:Z, in the order they came
        auto size_of = [](char c) { return c == 'A' || c == 'c' || c == 'C' ? 1 : c == 's' || c == 'S' ? 2 : 4; };
        const uint8_t *p = aux;
        while (p + 3 <= aux_end) {
            const uint8_t *tag = p;
            char ty = (char)p[2];
            const uint8_t *v = p + 3;
            if (ty == 'Z' || ty == 'H') { while (v < aux_end && *v) v++; v++; }
            else if (ty == 'B') v = v + 5 + (size_t)rd32(v + 1) * size_of((char)v[0]);
            else v += size_of(ty);
            if (v > aux_end) v = aux_end;
            if (!(tag[0] == 'R' && tag[1] == 'G' && ty == 'Z')) out.insert(out.end(), tag, v);
            p = v;
        }
    }
    uint32_t block = (uint32_t)(out.size() - start - 4);
    memcpy(out.data() + start, &block, 4);
    return true;
}

// ---- BAI (bam.py:_BaiBuilder; SAMv1 section 5.2) ------------------------------------------------
struct BaiBuilder {
    struct Meta { bool any = false; uint64_t first = 0, last = 0, mapped = 0, unmapped = 0; };
    std::vector<std::map<uint32_t, std::vector<std::pair<uint64_t, uint64_t>>>> bins;
    std::vector<std::map<int64_t, uint64_t>> linear;
    std::vector<Meta> meta;
    uint64_t no_coor = 0;
    explicit BaiBuilder(size_t n_ref) : bins(n_ref), linear(n_ref), meta(n_ref) {}

    void add(const OutRec &o, uint64_t beg_v, uint64_t end_v)
    {
        if (o.tid < 0 || (size_t)o.tid >= bins.size()) { no_coor++; return; }
        int64_t beg, end;
        index_interval(o, beg, end);
        auto &chunks = bins[o.tid][(uint32_t)reg2bin(beg, end)];
        if (!chunks.empty() && chunks.back().second == beg_v) chunks.back().second = end_v;   // adjacent records share a chunk
        else chunks.emplace_back(beg_v, end_v);
        auto &lin = linear[o.tid];
        for (int64_t win = beg >> 14; win <= ((end - 1) >> 14); win++) lin.emplace(win, beg_v);
        Meta &m = meta[o.tid];
        if (!m.any) { m.any = true; m.first = beg_v; }
        m.last = end_v;
        ((o.flag & 4) ? m.unmapped : m.mapped)++;
    }

    std::vector<uint8_t> dump() const
    {
        std::vector<uint8_t> out = {'B', 'A', 'I', 1};
        auto put64 = [&](uint64_t v) { uint8_t b[8]; memcpy(b, &v, 8); out.insert(out.end(), b, b + 8); };
        put32(out, (uint32_t)bins.size());
        for (size_t r = 0; r < bins.size(); r++) {
            const Meta &m = meta[r];
            put32(out, (uint32_t)(bins[r].size() + (m.any ? 1 : 0)));
            for (auto &kv : bins[r]) {
                put32(out, kv.first);
                put32(out, (uint32_t)kv.second.size());
                for (auto &c : kv.second) { put64(c.first); put64(c.second); }
            }
            if (m.any) {                                     // pseudo-bin with the mapped / unmapped counts
                put32(out, 37450); put32(out, 2);
                put64(m.first); put64(m.last); put64(m.mapped); put64(m.unmapped);
            }
            int64_t n_intv = linear[r].empty() ? 0 : linear[r].rbegin()->first + 1;
            put32(out, (uint32_t)n_intv);
            uint64_t last = 0;
            auto it = linear[r].begin();
            for (int64_t win = 0; win < n_intv; win++) {
                if (it != linear[r].end() && it->first == win) { last = it->second; ++it; }
                put64(last);
            }
        }
        put64(no_coor);
        return out;
    }
};

// ---- BGZF writer: blocks of 0xFF00 bytes deflated in parallel (bam.py:BgzfWriter) -------------------
constexpr size_t BGZF_BLOCK = 0xFF00;

bool deflate_block(const uint8_t *src, size_t len, int level, std::vector<uint8_t> &out)
{
    z_stream zs;
    memset(&zs, 0, sizeof(zs));
    if (deflateInit2(&zs, level, Z_DEFLATED, -15, 8, Z_DEFAULT_STRATEGY) != Z_OK) return false;
    out.resize(18 + deflateBound(&zs, (uLong)len) + 8);
    zs.next_in = const_cast<uint8_t *>(src);
    zs.avail_in = (uInt)len;
    zs.next_out = out.data() + 18;
    zs.avail_out = (uInt)(out.size() - 18 - 8);
    int rc = deflate(&zs, Z_FINISH);
    size_t clen = zs.total_out;
    deflateEnd(&zs);
    if (rc != Z_STREAM_END) return false;
    static const uint8_t head[16] = {0x1f, 0x8b, 8, 4, 0, 0, 0, 0, 0, 0xff, 6, 0, 'B', 'C', 2, 0};
    memcpy(out.data(), head, 16);
    uint16_t bsize = (uint16_t)(clen + 25);
    memcpy(out.data() + 16, &bsize, 2);
    uint32_t crc = (uint32_t)crc32(crc32(0L, Z_NULL, 0), src, (uInt)len), isize = (uint32_t)len;
    memcpy(out.data() + 18 + clen, &crc, 4);
    memcpy(out.data() + 18 + clen + 4, &isize, 4);
    out.resize(18 + clen + 8);
    return true;
}

// One output file on its way to disk: records -> uncompressed stream -> BGZF blocks -> file (+ index).
struct FileJob {
    std::string path;
    std::vector<const OutRec *> recs;
    bool coordinate = false, index = false;
    std::vector<uint8_t> stream;
    std::vector<uint64_t> rec_off;
    std::vector<std::vector<uint8_t>> comp;
    std::string err;
};

template <typename Fn>
void parallel_for(size_t n, int threads, Fn fn)
{
    std::atomic<size_t> next{0};
    auto work = [&]() { for (size_t i; (i = next.fetch_add(1)) < n;) fn(i); };
    int nt = (int)std::max<size_t>(1, std::min<size_t>({(size_t)std::max(threads, 1), n, 64}));
    std::vector<std::thread> pool;
    for (int t = 0; t < nt; t++) pool.emplace_back(work);
    for (auto &t : pool) t.join();
}

// bam.py:_header_bytes
std::vector<uint8_t> header_bytes(const ccp_bam *b, bool coordinate)
{
    std::string text = b->text;
    if (coordinate) {                                        // samtools sort rewrites / adds the @HD line
        std::string body;
        size_t p = 0;
        while (p <= text.size()) {
            size_t e = text.find('\n', p);
            std::string line = text.substr(p, e == std::string::npos ? std::string::npos : e - p);
            if (!line.empty() && line.compare(0, 3, "@HD") != 0) body += "\n" + line;
            if (e == std::string::npos) break;
            p = e + 1;
        }
        text = "@HD\tVN:1.6\tSO:coordinate" + body + "\n";
    }
    std::vector<uint8_t> out = {'B', 'A', 'M', 1};
    put32(out, (uint32_t)text.size());
    out.insert(out.end(), text.begin(), text.end());
    put32(out, (uint32_t)b->ref_names.size());
    for (size_t i = 0; i < b->ref_names.size(); i++) {
        put32(out, (uint32_t)b->ref_names[i].size() + 1);
        out.insert(out.end(), b->ref_names[i].begin(), b->ref_names[i].end());
        out.push_back(0);
        put32(out, (uint32_t)b->ref_lens[i]);
    }
    return out;
}

// bam.py:write_bam for a set of files at once: every file's records are encoded (one task per file),
// then the 0xFF00-byte blocks of all files are deflated by one pool, then files and indexes are written.
bool write_bams(const ccp_bam *b, std::vector<FileJob> &jobs, int threads, std::string &err)
{
    // records are encoded in pieces of 16 k by the whole pool (one task per file would leave the pool waiting for
    // the largest file), then every file's pieces are stitched behind its header
    const bool timing = getenv("CCP_TIMING") != nullptr;
    auto t_start = std::chrono::steady_clock::now();
    auto lap = [&](const char *what) {
        if (!timing) return;
        auto now = std::chrono::steady_clock::now();
        fprintf(stderr, "write_bams %-26s %.3f s\n", what, std::chrono::duration<double>(now - t_start).count());
        t_start = now;
    };
    constexpr size_t PIECE = 16384;
    struct Piece { uint32_t file; size_t first, last; std::vector<uint8_t> bytes; std::vector<uint64_t> off; std::string err; };
    std::vector<Piece> pieces;
    for (size_t j = 0; j < jobs.size(); j++)
        for (size_t a = 0; a < jobs[j].recs.size(); a += PIECE)
            pieces.push_back(Piece{(uint32_t)j, a, std::min(a + PIECE, jobs[j].recs.size()), {}, {}, {}});
    parallel_for(pieces.size(), threads, [&](size_t x) {
        Piece &pc = pieces[x];
        const FileJob &f = jobs[pc.file];
        pc.off.resize(pc.last - pc.first);
        for (size_t i = pc.first; i < pc.last && pc.err.empty(); i++) {
            pc.off[i - pc.first] = pc.bytes.size();
            encode_record(*f.recs[i], pc.bytes, pc.err);
        }
    });
    lap("encode");
    std::vector<size_t> first_piece(jobs.size() + 1, pieces.size());
    for (size_t x = pieces.size(); x-- > 0;) first_piece[pieces[x].file] = x;
    for (size_t j = jobs.size(); j-- > 0;) first_piece[j] = std::min(first_piece[j], first_piece[j + 1]);
    parallel_for(jobs.size(), threads, [&](size_t j) {
        FileJob &f = jobs[j];
        f.stream = header_bytes(b, f.coordinate);
        f.rec_off.resize(f.recs.size() + 1);
        size_t total = f.stream.size();
        for (size_t x = first_piece[j]; x < first_piece[j + 1]; x++) total += pieces[x].bytes.size();
        f.stream.reserve(total);
        for (size_t x = first_piece[j]; x < first_piece[j + 1] && f.err.empty(); x++) {
            Piece &pc = pieces[x];
            if (!pc.err.empty()) { f.err = pc.err; break; }
            const uint64_t at = f.stream.size();
            for (size_t i = pc.first; i < pc.last; i++) f.rec_off[i] = at + pc.off[i - pc.first];
            f.stream.insert(f.stream.end(), pc.bytes.begin(), pc.bytes.end());
            std::vector<uint8_t>().swap(pc.bytes);
        }
        f.rec_off[f.recs.size()] = f.stream.size();
        f.comp.resize((f.stream.size() + BGZF_BLOCK - 1) / BGZF_BLOCK);
    });
    lap("stitch");
    std::vector<std::pair<uint32_t, uint32_t>> blocks;       // (file, block) over all files
    for (size_t j = 0; j < jobs.size(); j++) {
        if (!jobs[j].err.empty()) { err = jobs[j].err; return false; }
        for (size_t k = 0; k < jobs[j].comp.size(); k++) blocks.emplace_back((uint32_t)j, (uint32_t)k);
    }
    std::atomic<bool> ok{true};
    // zlib level of the output BAMs: 6 like samtools / pysam unless CCP_DEFLATE_LEVEL says otherwise (the records
    // are the same; at level 6 the deflate is most of the command's wall time once the GPU has done its part)
    int level = 6;
    if (const char *e = getenv("CCP_DEFLATE_LEVEL")) { int v = atoi(e); if (v >= 0 && v <= 9) level = v; }
    parallel_for(blocks.size(), threads, [&](size_t x) {
        FileJob &f = jobs[blocks[x].first];
        size_t a = (size_t)blocks[x].second * BGZF_BLOCK, len = std::min(BGZF_BLOCK, f.stream.size() - a);
        if (!deflate_block(f.stream.data() + a, len, level, f.comp[blocks[x].second])) ok = false;
    });
    if (!ok) { err = "deflate failed"; return false; }
    lap("deflate");
    parallel_for(jobs.size(), threads, [&](size_t j) {
        FileJob &f = jobs[j];
        FILE *fh = fopen(f.path.c_str(), "wb");
        if (!fh) { f.err = "cannot create " + f.path; return; }
        const size_t n_blocks = f.comp.size();
        std::vector<uint64_t> coff(n_blocks + 1, 0);
        uint64_t off = 0;
        for (size_t i = 0; i < n_blocks; i++) {
            coff[i] = off;
            if (fwrite(f.comp[i].data(), 1, f.comp[i].size(), fh) != f.comp[i].size()) f.err = "short write to " + f.path;
            off += f.comp[i].size();
        }
        coff[n_blocks] = off;
        static const uint8_t eof[28] = {0x1f, 0x8b, 8, 4, 0, 0, 0, 0, 0, 0xff, 6, 0, 0x42, 0x43, 2, 0, 0x1b, 0, 3, 0, 0, 0, 0, 0, 0, 0, 0, 0};
        fwrite(eof, 1, 28, fh);
        fclose(fh);
        if (!f.index || !f.err.empty()) return;
        // BgzfWriter.tell() once the bytes before u are written: inside the block still being filled,
        // or - exactly on a block boundary - at the start of the next block (the EOF block at the end)
        auto voff = [&](uint64_t u) {
            uint64_t blk = u / BGZF_BLOCK, within = u % BGZF_BLOCK;
            return blk >= n_blocks ? (coff[n_blocks] << 16) : ((coff[blk] << 16) | within);
        };
        BaiBuilder bai(b->ref_names.size());
        for (size_t i = 0; i < f.recs.size(); i++) bai.add(*f.recs[i], voff(f.rec_off[i]), voff(f.rec_off[i + 1]));
        std::vector<uint8_t> dump = bai.dump();
        FILE *bh = fopen((f.path + ".bai").c_str(), "wb");
        if (!bh) { f.err = "cannot create " + f.path + ".bai"; return; }
        fwrite(dump.data(), 1, dump.size(), bh);
        fclose(bh);
    });
    for (auto &f : jobs)
        if (!f.err.empty()) { err = f.err; return false; }
    return true;
}

}  // namespace

// ---- emit: emit.py:emit_all -------------------------------------------------------------------------
extern "C" {

// Host copies of the result arrays of one GPU run (include/ccgpu.h: enum cc_array) plus what the
// packer kept about the batch.  Plain pointers, nothing is owned.
struct ccp_emit_in {
    uint64_t n_reads, n_fam;
    uint32_t lpad, stride;
    const int64_t *src;                  // packed read -> record of the input BAM
    const void *fields;                  // cc_fields[n_reads]
    const char *barcodes, *cigar_pairs, *rgs;   // NUL-separated dictionaries, id order
    uint32_t n_bc, n_cig, n_rg;
    const int64_t *bad;                  // records for badReads.bam
    uint64_t n_bad;
    const uint32_t *read_fam, *fam_size, *fam_tmpl;
    const uint8_t *fam_mapq; const uint16_t *fam_flag; const int32_t *fam_tlen; const uint8_t *fam_rg;
    const int32_t *partner, *sscs_row, *corr_row, *dcs_row, *dcssc_row;
    const uint8_t *sscs_rows, *corr_rows, *dcs_rows, *dcssc_rows;
    const uint8_t *dcs_mapq; const uint16_t *dcs_flag; const int32_t *dcs_tlen; const uint8_t *dcs_rg;
    const uint8_t *dcssc_mapq; const uint16_t *dcssc_flag; const int32_t *dcssc_tlen; const uint8_t *dcssc_rg;
};

}  // extern "C"

namespace {

enum { F_SSCS, F_SINGLETON, F_DCS, F_SSCS_SINGLETON, F_SSCS_COR, F_SING_COR, F_UNCORRECTED, F_SSCS_SC, F_DCS_SC,
       F_SSCS_SC_SINGLETON, F_ALL_UNIQUE, F_BAD, F_COUNT };

std::vector<std::string_view> split_dict(const char *joined, uint32_t count)
{
    std::vector<std::string_view> out;
    const char *p = joined;
    for (uint32_t i = 0; i < count; i++) {
        size_t len = strlen(p);
        out.emplace_back(p, len);
        p += len + 1;
    }
    return out;
}

struct DomainFault { std::string msg; };

struct Emitter {
    const ccp_bam *b;
    const ccp_emit_in &in;
    const Fields *fields;
    std::vector<std::string_view> barcodes, cigar_pairs, rgs;
    std::deque<OutRec> arena;
    std::vector<int64_t> mate_fam;

    Emitter(const ccp_bam *bam, const ccp_emit_in &i)
        : b(bam), in(i), fields(reinterpret_cast<const Fields *>(i.fields)),
          barcodes(split_dict(i.barcodes, i.n_bc)), cigar_pairs(split_dict(i.cigar_pairs, i.n_cig)),
          rgs(split_dict(i.rgs, i.n_rg)) {}

    Rec raw_record(int64_t record) const { return Rec{b->data.data() + b->offsets[record]}; }
    Rec raw(int64_t packed_read) const { return raw_record(in.src[packed_read]); }

    // keys.sscs_qname (consensus_helper.py:197-247) over (tid, pos, flag, mtid, mpos, tlen) of `read` and the
    // coordinates of `mate`
    static std::string sscs_qname(int32_t tid, int32_t pos, int32_t mate_tid, int32_t mate_pos, bool strand_pos,
                                  int32_t tlen, std::string_view barcode, std::string_view cigar)
    {
        int32_t a0 = tid, a1 = pos, b0 = mate_tid, b1 = mate_pos;
        if ((tid == mate_tid && pos > mate_pos) || tid > mate_tid) { std::swap(a0, b0); std::swap(a1, b1); }
        std::string s(barcode);
        s += "_" + std::to_string(a0) + "_" + std::to_string(a1) + "_" + std::to_string(b0) + "_" + std::to_string(b1) + "_";
        s += cigar;
        s += strand_pos ? "_pos_" : "_neg_";
        s += std::to_string(tlen < 0 ? -(int64_t)tlen : (int64_t)tlen);
        return s;
    }

    OutRec *passthrough(int64_t record, int64_t fam)
    {
        OutRec &o = arena.emplace_back();
        fill_passthrough(o, record, fam);
        return &o;
    }

    void fill_passthrough(OutRec &o, int64_t record, int64_t fam) const
    {
        Rec t = raw_record(record);
        o.fam = fam;
        o.qname = std::string(t.qname());
        o.tid = t.tid(); o.pos = t.pos(); o.mtid = t.mtid(); o.mpos = t.mpos(); o.tlen = t.tlen();
        o.flag = t.flag(); o.mapq = t.mapq();
        o.tmpl = t.p; o.row = nullptr; o.l_seq = t.l_seq(); o.lpad = in.lpad;
        std::string_view g = find_rg(t);
        o.has_rg = g.data() != nullptr;
        o.rg = g;
    }

    // consensus pair name of family g, computed from scratch (the SSCS stage runs on all host threads)
    std::string pair_name(int64_t g) const
    {
        const int64_t k = in.fam_tmpl[g] >> 1;
        Rec read = raw(2 * k), mate = raw(2 * k + 1);
        const Fields &f = fields[2 * k];
        return sscs_qname(read.tid(), read.pos(), mate.tid(), mate.pos(), strand_of(read) == 1, read.tlen(),
                          barcodes[f.bc], cigar_pairs[f.cig]);
    }

    // emit.py:consensus_record - coordinates and cigar of `tpl`, bases and qualities of a consensus row
    OutRec *consensus(int64_t fam, std::string qname, const uint8_t *rows, int64_t row, uint8_t mapq, int32_t tlen,
                      uint16_t flag, bool has_rg, std::string_view rg, const OutRec &tpl)
    {
        OutRec &o = arena.emplace_back();
        o.fam = fam;
        o.qname = std::move(qname);
        o.tid = tpl.tid; o.pos = tpl.pos; o.mtid = tpl.mtid; o.mpos = tpl.mpos;
        o.tlen = tlen; o.flag = flag; o.mapq = mapq;
        o.tmpl = tpl.tmpl; o.row = rows + (size_t)row * in.stride; o.l_seq = tpl.l_seq; o.lpad = in.lpad;
        o.has_rg = has_rg; o.rg = rg;
        return &o;
    }

    // records.sort_key / sort_records: (contig, position, strand), stable, no contig last
    static void sort_records(std::vector<OutRec *> &v)
    {
        auto key = [](const OutRec *o) {
            uint64_t t = o->tid < 0 ? (1ull << 31) : (uint64_t)o->tid;
            return std::make_tuple(t, (int64_t)o->pos, (o->flag & 0x10) ? 1 : 0);
        };
        std::stable_sort(v.begin(), v.end(), [&](const OutRec *x, const OutRec *y) { return key(x) < key(y); });
    }

    static std::vector<OutRec *> merged(std::initializer_list<const std::vector<OutRec *> *> lists)
    {
        std::vector<OutRec *> out;
        for (auto *l : lists) out.insert(out.end(), l->begin(), l->end());
        sort_records(out);
        return out;
    }

    // emit.py:_pair_walk - pairs in order of completion, first-seen mate first
    template <typename Fn>
    void pair_walk(const std::vector<OutRec *> &outs, Fn fn)
    {
        std::unordered_map<int64_t, OutRec *> waiting;
        waiting.reserve(outs.size());
        for (OutRec *o : outs) {
            int64_t k = std::min(o->fam, mate_fam[o->fam]);
            auto it = waiting.find(k);
            if (it == waiting.end()) { waiting.emplace(k, o); continue; }
            OutRec *first = it->second;
            waiting.erase(it);
            fn(first, o);
        }
    }
};

// keys.dcs_consensus_tag (DCS_maker.py:60-96)
std::string dcs_consensus_tag(const std::string &tag, const std::string &ds)
{
    auto before = [](const std::string &s, char c) { return s.substr(0, s.find(c)); };
    auto after_colon = [](const std::string &s) {           // s.split(":")[1]
        size_t a = s.find(':');
        if (a == std::string::npos) return std::string();
        size_t e = s.find(':', a + 1);
        return s.substr(a + 1, e == std::string::npos ? std::string::npos : e - a - 1);
    };
    std::string bc = before(tag, '_'), dbc = before(ds, '_');
    size_t first = tag.find('_'), last = tag.rfind('_');
    std::string coor = (first == std::string::npos) ? std::string() : tag.substr(first + 1, last > first ? last - first - 1 : std::string::npos);
    std::string n = after_colon(tag), dn = after_colon(ds);
    if (tag.find("pos") != std::string::npos) return bc + "_" + dbc + "_" + coor + ":" + n + "_" + dn;
    return dbc + "_" + bc + "_" + coor + ":" + dn + "_" + n;
}

}  // namespace

extern "C" {

// paths[F_*] (nullptr = do not write that file); counts[F_*] = records of each list.
// Returns 0, or -1 with a message; a message starting with "domain:" is pack.DomainError.
int ccp_emit(ccp_bam *b, const ccp_emit_in *in_p, int scorrect, const char *const *paths, int threads,
             int64_t *counts, char *err, int errlen)
{
    const ccp_emit_in &in = *in_p;
    Emitter E(b, in);
    const int64_t nf = (int64_t)in.n_fam;
    std::string msg;
    const bool timing = getenv("CCP_TIMING") != nullptr;
    auto t_start = std::chrono::steady_clock::now();
    auto lap = [&](const char *what) {
        if (!timing) return;
        auto now = std::chrono::steady_clock::now();
        fprintf(stderr, "ccp_emit %-28s %.3f s\n", what, std::chrono::duration<double>(now - t_start).count());
        t_start = now;
    };
    try {
        // the two families a consensus pair name owns (SSCS_maker.py:275 needs exactly two)
        E.mate_fam.resize(nf);
        for (int64_t g = 0; g < nf; g++) {
            const uint32_t mate_read = in.fam_tmpl[g] ^ 1u;
            const int64_t m = in.read_fam[mate_read];
            E.mate_fam[g] = m;
            if (in.fam_tmpl[m] != mate_read || in.fam_size[m] != in.fam_size[g])
                throw DomainFault{"a read pair does not found/join its two families together (orientation edge case, "
                                  "SURVEY.md App. B Q8/Q9): outside the GPU path's domain"};
        }
        auto rg_of = [&](uint8_t code, bool &has) { has = code != 0xFF; return has ? E.rgs[code] : std::string_view(); };

        // ---- SSCS stage (SSCS_maker.py:274-298) ---------------------------------------------------
        std::vector<OutRec *> L[F_COUNT];
        std::vector<OutRec> stage((size_t)nf);               // one record per family, built on all host threads
        const int64_t chunk = 2048;
        parallel_for((size_t)((nf + chunk - 1) / chunk), threads, [&](size_t c) {
            for (int64_t g = (int64_t)c * chunk; g < std::min(nf, (int64_t)(c + 1) * chunk); g++) {
                OutRec &o = stage[g];
                E.fill_passthrough(o, in.src[in.fam_tmpl[g]], g);      // coordinates / cigar of the first member
                o.qname = E.pair_name(g) + ":" + std::to_string(in.fam_size[g]);
                if (in.fam_size[g] != 1) {                             // consensus: row, mode fields, RG only
                    bool has;
                    o.rg = rg_of(in.fam_rg[g], has);
                    o.has_rg = has;
                    o.row = in.sscs_rows + (size_t)in.sscs_row[g] * in.stride;
                    o.mapq = in.fam_mapq[g]; o.tlen = in.fam_tlen[g]; o.flag = in.fam_flag[g];
                }
            }
        });
        for (int64_t g = 0; g < nf; g++) L[in.fam_size[g] == 1 ? F_SINGLETON : F_SSCS].push_back(&stage[g]);
        lap("sscs stage");
        Emitter::sort_records(L[F_SSCS]);
        Emitter::sort_records(L[F_SINGLETON]);
        lap("sort sscs, singleton");

        // ---- DCS stage (DCS_maker.py:226-259), replayed over a sorted record list ---------------------
        auto dcs_stage = [&](const std::vector<OutRec *> &inp, const std::unordered_map<int64_t, OutRec *> &rec_of,
                             bool second, std::vector<OutRec *> &dcs, std::vector<OutRec *> &single) {
            const int32_t *row_of = second ? in.dcssc_row : in.dcs_row;
            const uint8_t *rows = second ? in.dcssc_rows : in.dcs_rows;
            const uint8_t *mq = second ? in.dcssc_mapq : in.dcs_mapq;
            const uint16_t *fl = second ? in.dcssc_flag : in.dcs_flag;
            const int32_t *tl = second ? in.dcssc_tlen : in.dcs_tlen;
            const uint8_t *rgc = second ? in.dcssc_rg : in.dcs_rg;
            std::unordered_set<int64_t> made;
            E.pair_walk(inp, [&](OutRec *a, OutRec *bb) {
                for (OutRec *o : {a, bb}) {
                    const int64_t g = o->fam, p = in.partner[g];
                    if (p >= 0 && made.count(p)) continue;                 // the partner already led
                    const bool in_set = p >= 0 && (second ? rec_of.count(p) != 0 : in.fam_size[p] >= 2);
                    if (in_set) {
                        const int64_t row = row_of[g];
                        if (row < 0) throw DomainFault{"duplex strands met in an order the GPU path did not predict"};
                        const OutRec *other = rec_of.at(p);
                        bool has;
                        std::string_view rg = rg_of(rgc[row], has);
                        dcs.push_back(E.consensus(g, dcs_consensus_tag(o->qname, other->qname), rows, row, mq[row], tl[row],
                                                  fl[row], has, rg, *o));
                        made.insert(g);
                    } else {
                        single.push_back(o);
                    }
                }
            });
        };
        std::unordered_map<int64_t, OutRec *> rec_of;
        for (OutRec *o : L[F_SSCS]) rec_of[o->fam] = o;
        dcs_stage(L[F_SSCS], rec_of, false, L[F_DCS], L[F_SSCS_SINGLETON]);
        Emitter::sort_records(L[F_DCS]);
        Emitter::sort_records(L[F_SSCS_SINGLETON]);
        lap("dcs stage + sorts");

        // ---- singleton correction (singleton_correction.py:249-285) + second DCS pass --------------------
        if (scorrect) {
            E.pair_walk(L[F_SINGLETON], [&](OutRec *a, OutRec *bb) {
                Rec ra{a->tmpl}, rb{bb->tmpl};
                // keys.cigar_order(a, b): a's cigar leads iff (pos and R1) or (neg and R2)
                const bool pos_strand = strand_of(ra) == 1;
                const int num = read_number(a->flag);
                const bool lead = (pos_strand && num == 1) || (!pos_strand && num == 2);
                std::string cig = lead ? cigar_string(ra) + "_" + cigar_string(rb) : cigar_string(rb) + "_" + cigar_string(ra);
                std::string barcode = a->qname.substr(0, a->qname.find('_'));
                std::string qn = Emitter::sscs_qname(a->tid, a->pos, bb->tid, bb->pos, pos_strand, a->tlen, barcode, cig) + ":1";
                for (OutRec *o : {a, bb}) {
                    const int64_t g = o->fam, p = in.partner[g];
                    if (p < 0) { L[F_UNCORRECTED].push_back(o); continue; }
                    OutRec *rec = E.consensus(g, qn, in.corr_rows, in.corr_row[g], o->mapq, o->tlen, o->flag, o->has_rg,
                                              o->rg, *o);
                    L[in.fam_size[p] >= 2 ? F_SSCS_COR : F_SING_COR].push_back(rec);
                }
            });
            Emitter::sort_records(L[F_SSCS_COR]);
            Emitter::sort_records(L[F_SING_COR]);
            Emitter::sort_records(L[F_UNCORRECTED]);
            L[F_SSCS_SC] = Emitter::merged({&L[F_SSCS], &L[F_SSCS_COR], &L[F_SING_COR]});
            rec_of.clear();
            for (OutRec *o : L[F_SSCS_SC]) rec_of[o->fam] = o;
            dcs_stage(L[F_SSCS_SC], rec_of, true, L[F_DCS_SC], L[F_SSCS_SC_SINGLETON]);
            Emitter::sort_records(L[F_DCS_SC]);
            Emitter::sort_records(L[F_SSCS_SC_SINGLETON]);
            L[F_ALL_UNIQUE] = Emitter::merged({&L[F_DCS_SC], &L[F_SSCS_SC_SINGLETON], &L[F_UNCORRECTED]});
        }
        for (uint64_t i = 0; i < in.n_bad; i++) L[F_BAD].push_back(E.passthrough(in.bad[i], -1));
        lap("rescue + second dcs + merges");

        std::vector<FileJob> jobs;
        for (int f = 0; f < F_COUNT; f++) {
            counts[f] = (int64_t)L[f].size();
            if (!paths[f]) continue;
            FileJob &job = jobs.emplace_back();
            job.path = paths[f];
            job.recs.assign(L[f].begin(), L[f].end());
            job.coordinate = job.index = f != F_BAD;         // every list but badReads is sorted and indexed
        }
        if (!write_bams(b, jobs, threads, msg)) { set_err(err, errlen, msg); return -1; }
        lap("encode + deflate + write");
    } catch (const DomainFault &d) {
        set_err(err, errlen, "domain: " + d.msg);
        return -1;
    } catch (const std::exception &e) {
        set_err(err, errlen, e.what());
        return -1;
    }
    return 0;
}

}  // extern "C"
