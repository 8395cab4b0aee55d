"""Minimal BGZF / BAM / BAI codec (pysam and samtools are not available on the target boxes).

Covers exactly what the consensus path needs:
  * read a whole BAM into `Read` records (header text + reference list kept verbatim),
  * write records to BAM (BGZF blocks + EOF marker),
  * write the `.bai` index `samtools index` would leave next to a coordinate-sorted BAM.

Written from the SAMv1 specification.  Records keep their auxiliary fields as raw bytes, so a
passthrough record (singleton, uncorrected, SSCS singleton) is re-emitted with every tag intact;
the RG:Z tag is additionally decoded because the consensus path takes its mode
(consensus_helper.py:580-583).
"""
from __future__ import annotations

import struct
import zlib

import numpy as np

from .records import BASE_CODES, Read

_EOF = bytes.fromhex("1f8b08040000000000ff0600424302001b0003000000000000000000")
_CIGAR_OPS = "MIDNSHP=X"
_OP_CODE = {c: i for i, c in enumerate(_CIGAR_OPS)}
_SEQ_LUT = np.array([ord(c) for c in BASE_CODES], np.uint8)
_CODE_LUT = np.full(256, 15, np.uint8)
for _i, _c in enumerate(BASE_CODES):
    _CODE_LUT[ord(_c)] = _i
    _CODE_LUT[ord(_c.lower())] = _i


class BamHeader:
    def __init__(self, text="", references=(), lengths=()):
        self.text = text
        self.references = list(references)
        self.lengths = list(lengths)

    def tid(self, name):
        try:
            return self.references.index(name)
        except ValueError:
            return -1


# ---- BGZF --------------------------------------------------------------------------------
def bgzf_decompress(data: bytes) -> bytes:
    """Concatenated BGZF blocks -> the uncompressed stream."""
    out = []
    off, n = 0, len(data)
    while off < n:
        if data[off:off + 4] != b"\x1f\x8b\x08\x04":
            raise ValueError("not a BGZF block at offset %d" % off)
        xlen = struct.unpack_from("<H", data, off + 10)[0]
        extra = data[off + 12:off + 12 + xlen]
        bsize = None
        p = 0
        while p + 4 <= len(extra):
            si1, si2, slen = extra[p], extra[p + 1], struct.unpack_from("<H", extra, p + 2)[0]
            if si1 == 66 and si2 == 67:
                bsize = struct.unpack_from("<H", extra, p + 4)[0]
            p += 4 + slen
        if bsize is None:
            raise ValueError("BGZF block without BC subfield")
        cdata = data[off + 12 + xlen:off + bsize + 1 - 8]
        out.append(zlib.decompress(cdata, -15))
        off += bsize + 1
    return b"".join(out)


class BgzfWriter:
    """Block writer that also reports virtual file offsets (needed for the BAI)."""

    BLOCK = 0xFF00

    def __init__(self, fh, level=6):
        self.fh, self.level = fh, level
        self.buf = bytearray()
        self.block_start = 0          # compressed offset of the block being filled

    def tell(self):
        return (self.block_start << 16) | len(self.buf)

    def write(self, data: bytes):
        self.buf += data
        while len(self.buf) >= self.BLOCK:
            self._flush(self.BLOCK)

    def _flush(self, n):
        chunk = bytes(self.buf[:n])
        del self.buf[:n]
        comp = zlib.compressobj(self.level, zlib.DEFLATED, -15)
        cdata = comp.compress(chunk) + comp.flush()
        block = (b"\x1f\x8b\x08\x04\x00\x00\x00\x00\x00\xff\x06\x00BC\x02\x00" +
                 struct.pack("<H", len(cdata) + 25) + cdata +
                 struct.pack("<II", zlib.crc32(chunk) & 0xFFFFFFFF, len(chunk)))
        self.fh.write(block)
        self.block_start += len(block)

    def close(self):
        while self.buf:
            self._flush(min(len(self.buf), self.BLOCK))
        self.fh.write(_EOF)


# ---- records -----------------------------------------------------------------------------------
def _cigar_string(words):
    return "".join("%d%s" % (w >> 4, _CIGAR_OPS[w & 15]) for w in words)


def _cigar_words(cigar):
    words, num = [], 0
    for ch in cigar or "":
        if ch.isdigit():
            num = num * 10 + ord(ch) - 48
        else:
            words.append((num << 4) | _OP_CODE[ch])
            num = 0
    return words


def reference_length(cigar):
    num, total = 0, 0
    for ch in cigar or "":
        if ch.isdigit():
            num = num * 10 + ord(ch) - 48
        else:
            if ch in "MDN=X":
                total += num
            num = 0
    return total


def reg2bin(beg, end):
    end -= 1
    if beg >> 14 == end >> 14:
        return ((1 << 15) - 1) // 7 + (beg >> 14)
    if beg >> 17 == end >> 17:
        return ((1 << 12) - 1) // 7 + (beg >> 17)
    if beg >> 20 == end >> 20:
        return ((1 << 9) - 1) // 7 + (beg >> 20)
    if beg >> 23 == end >> 23:
        return ((1 << 6) - 1) // 7 + (beg >> 23)
    if beg >> 26 == end >> 26:
        return ((1 << 3) - 1) // 7 + (beg >> 26)
    return 0


_AUX_SIZE = {"A": 1, "c": 1, "C": 1, "s": 2, "S": 2, "i": 4, "I": 4, "f": 4}


def _find_rg(aux: bytes):
    """Value of the RG:Z tag in raw aux bytes, or None."""
    p, n = 0, len(aux)
    while p + 3 <= n:
        tag, typ = aux[p:p + 2], chr(aux[p + 2])
        p += 3
        if typ in _AUX_SIZE:
            p += _AUX_SIZE[typ]
        elif typ in "ZH":
            e = aux.index(b"\0", p)
            if tag == b"RG" and typ == "Z":
                return aux[p:e].decode()
            p = e + 1
        elif typ == "B":
            sub = chr(aux[p])
            cnt = struct.unpack_from("<i", aux, p + 1)[0]
            p += 5 + cnt * _AUX_SIZE[sub]
        else:
            raise ValueError("unknown aux type %r" % typ)
    return None


def _strip_rg(aux: bytes):
    """Raw aux bytes without the RG tag (kept separately on the record)."""
    p, n = 0, len(aux)
    out = bytearray()
    while p + 3 <= n:
        start = p
        tag, typ = aux[p:p + 2], chr(aux[p + 2])
        p += 3
        if typ in _AUX_SIZE:
            p += _AUX_SIZE[typ]
        elif typ in "ZH":
            p = aux.index(b"\0", p) + 1
        elif typ == "B":
            sub = chr(aux[p])
            cnt = struct.unpack_from("<i", aux, p + 1)[0]
            p += 5 + cnt * _AUX_SIZE[sub]
        if not (tag == b"RG" and typ == "Z"):
            out += aux[start:p]
    return bytes(out)


def _aux_bytes(read: Read) -> bytes:
    """RG first (as create_aligned_segment writes it), then whatever else the record carries."""
    out = b""
    if read.rg is not None:
        out += b"RGZ" + read.rg.encode() + b"\0"
    aux = read.aux
    if isinstance(aux, dict):
        for k, v in sorted(aux.items()):
            if isinstance(v, int):
                out += k.encode() + b"i" + struct.pack("<i", v)
            elif isinstance(v, float):
                out += k.encode() + b"f" + struct.pack("<f", v)
            else:
                out += k.encode() + b"Z" + str(v).encode() + b"\0"
    elif aux:
        out += bytes(aux)
    return out


def read_bam(path):
    """-> (BamHeader, [Read]) in file order."""
    with open(path, "rb") as fh:
        data = bgzf_decompress(fh.read())
    if data[:4] != b"BAM\1":
        raise ValueError("%s is not a BAM file" % path)
    l_text = struct.unpack_from("<i", data, 4)[0]
    text = data[8:8 + l_text].split(b"\0", 1)[0].decode()
    off = 8 + l_text
    n_ref = struct.unpack_from("<i", data, off)[0]
    off += 4
    names, lengths = [], []
    for _ in range(n_ref):
        l_name = struct.unpack_from("<i", data, off)[0]
        names.append(data[off + 4:off + 4 + l_name - 1].decode())
        lengths.append(struct.unpack_from("<i", data, off + 4 + l_name)[0])
        off += 8 + l_name
    header = BamHeader(text, names, lengths)
    reads = []
    n = len(data)
    view = np.frombuffer(data, np.uint8)
    while off + 4 <= n:
        r, off = decode_record(data, view, off)
        reads.append(r)
    return header, reads


def decode_record(data, view, off):
    """The alignment record whose block_size field sits at `off` -> (Read, offset of the next one)."""
    block_size = struct.unpack_from("<i", data, off)[0]
    (tid, pos, l_name, mapq, _bin, n_cig, flag, l_seq, mtid, mpos, tlen) = struct.unpack_from(
        "<iiBBHHHiiii", data, off + 4)
    p = off + 36
    qname = bytes(data[p:p + l_name - 1]).decode()
    p += l_name
    cigar = _cigar_string(struct.unpack_from("<%dI" % n_cig, data, p)) if n_cig else ""
    p += 4 * n_cig
    packed = view[p:p + (l_seq + 1) // 2]
    codes = np.empty(2 * len(packed), np.uint8)
    codes[0::2] = packed >> 4
    codes[1::2] = packed & 15
    seq = _SEQ_LUT[codes[:l_seq]].tobytes().decode()
    p += (l_seq + 1) // 2
    qual = view[p:p + l_seq].copy()
    p += l_seq
    aux = bytes(data[p:off + 4 + block_size])
    return (Read(qname, flag, tid, pos, mapq, cigar, mtid, mpos, tlen, seq, qual,
                 rg=_find_rg(aux), aux=_strip_rg(aux)), off + 4 + block_size)


def encode_record(r: Read) -> bytes:
    name = r.qname.encode() + b"\0"
    if len(name) > 255:
        raise ValueError("query name longer than 254 characters: %s" % r.qname)
    cig = _cigar_words(r.cigar)
    l_seq = len(r.seq)
    codes = _CODE_LUT[np.frombuffer(r.seq.encode(), np.uint8)]
    if l_seq & 1:
        codes = np.append(codes, np.uint8(0))
    packed = ((codes[0::2] << 4) | codes[1::2]).astype(np.uint8).tobytes()
    qual = np.asarray(r.qual, np.uint8)
    qual = qual.tobytes() if len(qual) == l_seq else b"\xff" * l_seq
    ref_len = reference_length(r.cigar)
    end = r.pos + (ref_len if ref_len > 0 and not (r.flag & 4) else 1)
    body = struct.pack("<iiBBHHHiiii", r.tid, r.pos, len(name), r.mapq, reg2bin(max(r.pos, 0), max(end, 1)),
                       len(cig), r.flag, l_seq, r.mtid, r.mpos, r.tlen)
    body += name + struct.pack("<%dI" % len(cig), *cig) + packed + qual + _aux_bytes(r)
    return struct.pack("<i", len(body)) + body


def _header_bytes(header: BamHeader, sort_order=None):
    text = header.text
    if sort_order:                      # samtools sort rewrites / adds the @HD line
        lines = [l for l in text.split("\n") if l and not l.startswith("@HD")]
        text = "\n".join(["@HD\tVN:1.6\tSO:%s" % sort_order] + lines) + "\n"
    t = text.encode()
    out = b"BAM\1" + struct.pack("<i", len(t)) + t + struct.pack("<i", len(header.references))
    for name, length in zip(header.references, header.lengths):
        nm = name.encode() + b"\0"
        out += struct.pack("<i", len(nm)) + nm + struct.pack("<i", length)
    return out


def write_bam(path, header: BamHeader, reads, sort_order=None, index=False, level=6):
    """Write `reads` to `path`; with index=True also `path + '.bai'` (input must be coordinate sorted)."""
    builder = _BaiBuilder(len(header.references)) if index else None
    with open(path, "wb") as fh:
        w = BgzfWriter(fh, level)
        w.write(_header_bytes(header, sort_order))
        for r in reads:
            rec = encode_record(r)
            beg_v = w.tell()
            w.write(rec)
            if builder is not None:
                builder.add(r, beg_v, w.tell())
        w.close()
    if builder is not None:
        with open(path + ".bai", "wb") as fh:
            fh.write(builder.dump())


class _BaiBuilder:
    """Binning + linear index of `samtools index` (SAMv1 section 5.2)."""

    def __init__(self, n_ref):
        self.bins = [dict() for _ in range(n_ref)]
        self.linear = [dict() for _ in range(n_ref)]
        self.meta = [[None, None, 0, 0] for _ in range(n_ref)]     # first offset, last offset, mapped, unmapped
        self.no_coor = 0

    def add(self, r, beg_v, end_v):
        if r.tid < 0:
            self.no_coor += 1
            return
        ref_len = reference_length(r.cigar)
        beg = max(r.pos, 0)
        end = beg + (ref_len if ref_len > 0 and not (r.flag & 4) else 1)
        chunks = self.bins[r.tid].setdefault(reg2bin(beg, end), [])
        if chunks and chunks[-1][1] == beg_v:
            chunks[-1][1] = end_v                          # adjacent records share a chunk
        else:
            chunks.append([beg_v, end_v])
        lin = self.linear[r.tid]
        for win in range(beg >> 14, ((end - 1) >> 14) + 1):
            if win not in lin:
                lin[win] = beg_v
        m = self.meta[r.tid]
        if m[0] is None:
            m[0] = beg_v
        m[1] = end_v
        m[3 if (r.flag & 4) else 2] += 1

    def dump(self):
        out = b"BAI\1" + struct.pack("<i", len(self.bins))
        for bins, lin, m in zip(self.bins, self.linear, self.meta):
            n_bin = len(bins) + (1 if m[0] is not None else 0)
            out += struct.pack("<i", n_bin)
            for b in sorted(bins):
                out += struct.pack("<Ii", b, len(bins[b]))
                for beg_v, end_v in bins[b]:
                    out += struct.pack("<QQ", beg_v, end_v)
            if m[0] is not None:                           # pseudo-bin with the mapped / unmapped counts
                out += struct.pack("<IiQQQQ", 37450, 2, m[0], m[1], m[2], m[3])
            n_intv = (max(lin) + 1) if lin else 0
            out += struct.pack("<i", n_intv)
            last = 0
            for win in range(n_intv):
                last = lin.get(win, last)
                out += struct.pack("<Q", last)
        out += struct.pack("<Q", self.no_coor)
        return out
