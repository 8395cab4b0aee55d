"""Host packing: records -> the SoA batch the CUDA kernels consume.

This is the part of `read_bam` (consensus_helper.py:305-480) that stays on host
cores: region visit order, the bad-read filter (`:390-410`), mate pairing by
query name (`:412-420,478`), barcode extraction (`:423-431`) and dictionary
encoding of the two string-valued key fields (barcode, ordered cigar pair).
Grouping reads into families - the dictionaries `read_dict` / `tag_dict` /
`csn_pair_dict` of `:447-475` - is done on the GPU from the packed fields.

Layout handed to the device (also described in DESIGN.md):

  fields[N]   32-byte records, reads in pair-completion order (2k = first-seen
              mate of pair k, 2k+1 = second-seen mate):
                u32 bc     barcode dictionary id
                u32 cig    ordered cigar-pair dictionary id
                i32 tid, pos, mtid, mpos
                i32 tlen ; u16 flag ; u8 mapq ; u8 rg (0xff = no RG tag)
              The family key of `unique_tag` (`:293-301`) is the first 24 bytes
              plus flag bits 0x10 (orientation) and 0x40/0x80 (read number).
  payload[N]  `stride` bytes per read: lpad/2 bytes of BAM 4-bit base codes, two
              per byte with the EVEN position in the LOW nibble (BAM itself
              has it in the high nibble; the swap makes nibble i of a 16-bit
              group position i, which the vote kernel uses as a PRMT selector),
              zero padded, followed by lpad quality bytes (zero padded),
              lpad = L rounded up to 32, stride a multiple of 64.
  dup_bc[B]   barcode id of the complementary strand's barcode (`duplex_tag`,
              `:629-639`), 0xffffffff when that barcode never occurs.
"""
from __future__ import annotations

import numpy as np

from . import keys
from .records import Read, seq_to_codes

FIELDS_DT = np.dtype([("bc", "<u4"), ("cig", "<u4"), ("tid", "<i4"), ("pos", "<i4"),
                      ("mtid", "<i4"), ("mpos", "<i4"), ("tlen", "<i4"),
                      ("flag", "<u2"), ("mapq", "u1"), ("rg", "u1")])
assert FIELDS_DT.itemsize == 32
NO_ID = 0xFFFFFFFF
NO_RG = 0xFF


class DomainError(ValueError):
    """Input outside the domain the reference itself handles."""


def row_geometry(max_len):
    """(lpad, stride) for reads of up to `max_len` bases."""
    lpad = max(32, (int(max_len) + 31) // 32 * 32)
    stride = (lpad // 2 + lpad + 63) // 64 * 64
    return lpad, stride


def pack_row(codes, qual, lpad, out):
    """Write one payload row (4-bit bases then qualities) into `out` (u8[stride])."""
    n = len(codes)
    c = np.zeros(lpad, np.uint8)
    c[:n] = codes
    out[:lpad // 2] = c[0::2] | (c[1::2] << 4)
    out[lpad // 2:lpad // 2 + n] = qual
    out[lpad // 2 + n:] = 0


def unpack_row(row, lpad, n):
    """Inverse of `pack_row`: (codes u8[n], qual u8[n])."""
    s = row[:lpad // 2]
    codes = np.empty(lpad, np.uint8)
    codes[0::2] = s & 15
    codes[1::2] = s >> 4
    return codes[:n].copy(), row[lpad // 2:lpad // 2 + n].copy()


class PackedReads:
    """SoA batch of the good, paired reads of one stage (+ host-side bookkeeping)."""

    def __init__(self, fields, payload, lpad, barcodes, dup_bc, cigar_pairs, rgs,
                 src=None, lens=None, counters=None, bad=None, leftover=None):
        self.fields = fields
        self.payload = payload
        self.lpad = int(lpad)
        self.stride = int(payload.shape[1]) if payload.ndim == 2 else 0
        self.barcodes = barcodes
        self.dup_bc = dup_bc
        self.cigar_pairs = cigar_pairs
        self.rgs = rgs
        self.src = src                    # index of each packed read in the input record list
        self.lens = lens                  # query length of each packed read
        self.counters = counters or {}
        self.bad = bad or []              # indices of records that go to badReads.bam
        self.leftover = leftover or []    # indices of records whose mate never showed up

    @property
    def n(self):
        return int(self.fields.shape[0])


def build_dup_table(barcodes):
    idx = {b: i for i, b in enumerate(barcodes)}
    dup = np.full(len(barcodes), NO_ID, np.uint32)
    for i, b in enumerate(barcodes):
        j = idx.get(keys.duplex_barcode(b))
        if j is not None:
            dup[i] = j
    return dup


def visit_order(reads, regions):
    """Indices of `reads` in the order `read_bam` visits them.

    regions None: file order (`fetch(until_eof=True)`).  Otherwise a list of
    (tid, start, end): for each, records of that contig overlapping the
    half-open window, kept iff start <= pos <= end (consensus_helper.py:381).
    """
    if regions is None:
        return range(len(reads))
    by_tid = {}
    for i, r in enumerate(reads):
        by_tid.setdefault(r.tid, []).append(i)
    pos_of = {t: np.array([reads[i].pos for i in ix]) for t, ix in by_tid.items()}
    out = []
    for tid, start, end in regions:
        ix = by_tid.get(tid)
        if not ix:
            continue
        p = pos_of[tid]
        if np.all(p[:-1] <= p[1:]):
            lo = int(np.searchsorted(p, start, "left"))
            hi = int(np.searchsorted(p, end, "left"))     # pos == end does not overlap [start, end)
            out.extend(ix[lo:hi])
        else:
            out.extend(i for i, q in zip(ix, p) if start <= q < end)
    return out


def pair_reads(reads, stage="sscs", bdelim="|", regions=None):
    """Filter + mate pairing.  Returns (pairs, barcodes_of_pairs, counters, bad, leftover).

    stage "sscs": barcode is `qname.split(bdelim)[1]`, bad reads are diverted;
    stage "dcs":  barcode is `qname.split('_')[0]`, nothing is diverted
                  (`badRead_bam=None`, consensus_helper.py:409-412).
    """
    sscs = stage == "sscs"
    counters = dict(total=0, unmapped=0, unmapped_mate=0, multiple_mapping=0, bad_spacer=0)
    waiting = {}
    pairs, pair_bc, bad = [], [], []
    for i in visit_order(reads, regions):
        r = reads[i]
        counters["total"] += 1
        is_bad = True
        if sscs and bdelim not in r.qname:
            counters["bad_spacer"] += 1
        elif r.flag & 0x4:
            counters["unmapped"] += 1
            counters["total"] -= 1
        elif r.flag in keys.MATE_UNMAPPED_FLAGS:
            counters["unmapped_mate"] += 1
        elif r.flag & 0x900:
            counters["multiple_mapping"] += 1
        else:
            is_bad = False
        if is_bad and sscs:
            bad.append(i)
            continue
        j = waiting.pop(r.qname, None)
        if j is None:
            waiting[r.qname] = i
            continue
        q = reads[j].qname
        pair_bc.append(q.split(bdelim)[1] if sscs else q.split("_")[0])
        pairs.append((j, i))
    return pairs, pair_bc, counters, bad, list(waiting.values())


def pack_reads(reads, stage="sscs", bdelim="|", regions=None, row_source=None):
    """Records -> PackedReads (see module docstring).

    `row_source(i) -> (codes, qual)` lets a caller supply bases/qualities that
    are already arrays; by default they come from `reads[i].seq/.qual`.
    """
    pairs, pair_bc, counters, bad, leftover = pair_reads(reads, stage, bdelim, regions)
    n = 2 * len(pairs)
    fields = np.zeros(n, FIELDS_DT)
    bc_ids, cig_ids, rg_ids = {}, {}, {}
    src = np.empty(n, np.int64)
    lens = np.empty(n, np.int32)
    max_len = 0
    for k, (a, b) in enumerate(pairs):
        read, mate = reads[a], reads[b]
        for r in (read, mate):
            if keys.which_read(r.flag) is None:
                raise DomainError("flag %d of %s is not one of the 16 paired flags the "
                                  "reference can key (consensus_helper.py:69-70)" % (r.flag, r.qname))
        bc = bc_ids.setdefault(pair_bc[k], len(bc_ids))
        cg = cig_ids.setdefault(keys.cigar_order(read, mate), len(cig_ids))
        for j, (i, r) in enumerate(((a, read), (b, mate))):
            f = fields[2 * k + j]
            f["bc"], f["cig"] = bc, cg
            f["tid"], f["pos"], f["mtid"], f["mpos"] = r.tid, r.pos, r.mtid, r.mpos
            f["tlen"], f["flag"], f["mapq"] = r.tlen, r.flag, r.mapq
            if r.rg is None:
                f["rg"] = NO_RG
            else:
                g = rg_ids.setdefault(r.rg, len(rg_ids))
                if g >= NO_RG:
                    raise DomainError("more than 255 distinct read groups")
                f["rg"] = g
            src[2 * k + j] = i
            lens[2 * k + j] = len(r.seq)
            max_len = max(max_len, len(r.seq))
    lpad, stride = row_geometry(max_len)
    payload = np.zeros((n, stride), np.uint8)
    for x in range(n):
        i = int(src[x])
        if row_source is None:
            codes, qual = seq_to_codes(reads[i].seq), reads[i].qual
        else:
            codes, qual = row_source(i)
        pack_row(codes, qual, lpad, payload[x])
    barcodes = list(bc_ids)
    return PackedReads(fields, payload, lpad, barcodes, build_dup_table(barcodes),
                       list(cig_ids), list(rg_ids), src=src, lens=lens,
                       counters=counters, bad=bad, leftover=leftover)


def to_q4(packed):
    """Compact rows for `cc_load_reads_q4`: (payload4 u8[n][stride4], qual_table u8[16]), or None
    when the batch has more than 15 distinct non-zero quality values.

    A row is lpad/2 base bytes (unchanged) + lpad/2 bytes of 4-bit quality codes, even position in
    the low nibble; code 0 is quality 0 (row padding)."""
    lpad, half = packed.lpad, packed.lpad // 2
    q = packed.payload[:, half:half + lpad]
    present = np.flatnonzero(np.bincount(q.reshape(-1), minlength=256)) if q.size else np.zeros(0, np.int64)
    values = [0] + [int(v) for v in present if v != 0]
    if len(values) > 16:
        return None
    table = np.zeros(16, np.uint8)
    table[:len(values)] = values
    code_of = np.zeros(256, np.uint8)
    code_of[values] = np.arange(len(values), dtype=np.uint8)
    codes = code_of[q]
    stride4 = (lpad + 15) // 16 * 16
    out = np.zeros((packed.n, stride4), np.uint8)
    out[:, :half] = packed.payload[:, :half]
    out[:, half:lpad] = codes[:, 0::2] | (codes[:, 1::2] << 4)
    return out, table


PAIR_DT = np.dtype([("bc", "<u4"), ("cig", "<u4"), ("tid", "<i4"), ("pos", "<i4"), ("mtid", "<i4"), ("mpos", "<i4"),
                    ("tlen1", "<i4"), ("tlen2", "<i4"), ("flag1", "<u2"), ("flag2", "<u2"),
                    ("mapq1", "u1"), ("mapq2", "u1"), ("rg1", "u1"), ("rg2", "u1")])       # include/ccgpu.h cc_pair
assert PAIR_DT.itemsize == 40


def to_pairs(fields):
    """cc_pair records of a packed batch (one per read pair), or None when some pair's mates do not share
    barcode / cigar pair or do not mirror each other's coordinates."""
    a, b = fields[0::2], fields[1::2]
    ok = ((a["bc"] == b["bc"]) & (a["cig"] == b["cig"]) & (a["tid"] == b["mtid"]) & (a["pos"] == b["mpos"]) &
          (a["mtid"] == b["tid"]) & (a["mpos"] == b["pos"]))
    if not ok.all():
        return None
    out = np.zeros(a.shape[0], PAIR_DT)
    for nm in ("bc", "cig", "tid", "pos", "mtid", "mpos"):
        out[nm] = a[nm]
    out["tlen1"], out["tlen2"] = a["tlen"], b["tlen"]
    out["flag1"], out["flag2"] = a["flag"], b["flag"]
    out["mapq1"], out["mapq2"] = a["mapq"], b["mapq"]
    out["rg1"], out["rg2"] = a["rg"], b["rg"]
    return out


def to_compact(packed):
    """Input of `cc_load_reads_compact`: (pairs, planes u8[n][stride_c], qbits, qual_table u8[16], base_table u8[16])
    or None when the batch does not fit (pairs that do not mirror, more than 16 codes).

    A position is a 2-bit base (A,C,G,T) plus a code that names its quality and - for anything that is not
    A/C/G/T: N, the padding after the read, IUPAC letters - its literal 4-bit base code."""
    pairs = to_pairs(packed.fields)
    if pairs is None:
        return None
    lpad, half = packed.lpad, packed.lpad // 2
    n = packed.n
    seq = packed.payload[:, :half]
    base = np.empty((n, lpad), np.uint8)
    base[:, 0::2] = seq & 15
    base[:, 1::2] = seq >> 4
    qual = packed.payload[:, half:half + lpad]
    two_bit = np.full(16, 255, np.uint8)
    two_bit[[1, 2, 4, 8]] = [0, 1, 2, 3]
    b2 = two_bit[base]
    literal = np.where(b2 == 255, base, 255).astype(np.uint16)          # 255: an A/C/G/T position
    combo = (qual.astype(np.uint16) << 8) | literal
    present = np.flatnonzero(np.bincount(combo.reshape(-1), minlength=65536)) if combo.size else np.zeros(0, np.int64)
    if len(present) > 16:
        return None
    qbits = max(1, int(np.ceil(np.log2(max(2, len(present))))))
    qual_table, base_table = np.zeros(16, np.uint8), np.full(16, 255, np.uint8)
    code_of = np.zeros(65536, np.uint8)
    for c, v in enumerate(present.tolist()):
        qual_table[c], base_table[c] = v >> 8, v & 255
        code_of[v] = c
    codes = code_of[combo]
    stride_c = (lpad // 4 + lpad * qbits // 8 + 15) // 16 * 16
    planes = np.zeros((n, stride_c), np.uint8)
    b2 = np.where(b2 == 255, 0, b2)
    planes[:, :lpad // 4] = b2[:, 0::4] | (b2[:, 1::4] << 2) | (b2[:, 2::4] << 4) | (b2[:, 3::4] << 6)
    bits = ((codes[:, :, None] >> np.arange(qbits, dtype=np.uint8)) & 1).reshape(n, lpad * qbits)
    planes[:, lpad // 4:lpad // 4 + lpad * qbits // 8] = np.packbits(bits, axis=1, bitorder="little")
    return pairs, planes, qbits, qual_table, base_table


PAIR16_DT = np.dtype([("bc", "<u4"), ("pos", "<i4"), ("mpos", "<i4"), ("tlen1", "<i2"), ("cls", "<u2")])   # cc_pair16
CLASS_DT = np.dtype([("cig", "<u4"), ("tid", "<i4"), ("mtid", "<i4"), ("flag1", "<u2"), ("flag2", "<u2"),
                     ("mapq1", "u1"), ("mapq2", "u1"), ("rg1", "u1"), ("rg2", "u1"), ("reserved", "<u4", (3,))])  # cc_pair_class
assert PAIR16_DT.itemsize == 16 and CLASS_DT.itemsize == 32
_CLASS_KEYS = ("cig", "tid", "mtid", "flag1", "flag2", "mapq1", "mapq2", "rg1", "rg2")


def to_classes(pairs):
    """Input of `cc_load_reads_classes` from cc_pair records: (pairs16, classes), or None when the batch does
    not fit - a pair whose TLENs are not each other's negatives or do not fit 16 bits, more than 65 536
    distinct (cigar pair, contigs, flags, MAPQs, read groups) combinations."""
    if not ((pairs["tlen2"] == -pairs["tlen1"]) & (np.abs(pairs["tlen1"]) <= 32767)).all():
        return None
    # classes by a 64-bit mix of their fields (np.unique on one integer column), then checked field by field
    h = np.zeros(len(pairs), np.uint64)
    for nm in _CLASS_KEYS:
        h = (h ^ pairs[nm].astype(np.int64).astype(np.uint64)) * np.uint64(0x9E3779B97F4A7C15)
        h ^= h >> np.uint64(29)
    _, first, inverse = np.unique(h, return_index=True, return_inverse=True)
    if len(first) > 65536:
        return None
    classes = np.zeros(len(first), CLASS_DT)
    for nm in _CLASS_KEYS:
        classes[nm] = pairs[nm][first]
        if not (classes[nm][inverse] == pairs[nm]).all():
            return None                                       # two classes under one hash value: not worth a second try
    out = np.zeros(len(pairs), PAIR16_DT)
    for nm in ("bc", "pos", "mpos", "tlen1"):
        out[nm] = pairs[nm]
    out["cls"] = inverse
    return out, classes


def from_classes(pairs16, classes):
    """cc_pair records of short pair records (the host statement of expand_pairs16_kernel)."""
    out = np.zeros(len(pairs16), PAIR_DT)
    c = classes[pairs16["cls"]]
    for nm in ("bc", "pos", "mpos", "tlen1"):
        out[nm] = pairs16[nm]
    out["tlen2"] = -pairs16["tlen1"].astype(np.int32)
    for nm in _CLASS_KEYS:
        out[nm] = c[nm]
    return out


def from_compact(pairs, planes, qbits, qual_table, base_table, lpad):
    """Inverse of `to_compact`: (fields, standard payload rows) - what the device expands to."""
    n = 2 * pairs.shape[0]
    f = np.zeros(n, FIELDS_DT)
    a, b = f[0::2], f[1::2]
    for nm in ("bc", "cig"):
        a[nm] = b[nm] = pairs[nm]
    a["tid"], a["pos"], a["mtid"], a["mpos"] = pairs["tid"], pairs["pos"], pairs["mtid"], pairs["mpos"]
    b["tid"], b["pos"], b["mtid"], b["mpos"] = pairs["mtid"], pairs["mpos"], pairs["tid"], pairs["pos"]
    a["tlen"], b["tlen"] = pairs["tlen1"], pairs["tlen2"]
    a["flag"], b["flag"] = pairs["flag1"], pairs["flag2"]
    a["mapq"], b["mapq"] = pairs["mapq1"], pairs["mapq2"]
    a["rg"], b["rg"] = pairs["rg1"], pairs["rg2"]
    return f, planes_to_rows(planes, qbits, qual_table, base_table, lpad)


def planes_to_rows(planes, qbits, qual_table, base_table, lpad):
    """Standard payload rows from rows in bit planes (the host statement of expand_planes_kernel; also how a host
    reads the CC_*_PLANES arrays of cc_pack_rows)."""
    n = planes.shape[0]
    half = lpad // 2
    _, stride = row_geometry(lpad)
    bp = planes[:, :lpad // 4]
    b2 = np.empty((n, lpad), np.uint8)
    for k in range(4):
        b2[:, k::4] = (bp >> (2 * k)) & 3
    bits = np.unpackbits(planes[:, lpad // 4:lpad // 4 + lpad * qbits // 8], axis=1, bitorder="little").reshape(n, lpad, qbits)
    codes = (bits << np.arange(qbits, dtype=np.uint8)).sum(axis=2).astype(np.uint8)
    lit = base_table[codes]
    base = np.where(lit != 255, lit, (1 << b2).astype(np.uint8)).astype(np.uint8)
    out = np.zeros((n, stride), np.uint8)
    out[:, :half] = base[:, 0::2] | (base[:, 1::2] << 4)
    out[:, half:half + lpad] = qual_table[codes]
    return out


def from_q4(payload4, table, lpad):
    """Inverse of `to_q4`: the standard payload rows (what the device expands to)."""
    half = lpad // 2
    n = payload4.shape[0]
    _, stride = row_geometry(lpad)
    out = np.zeros((n, stride), np.uint8)
    out[:, :half] = payload4[:, :half]
    c = payload4[:, half:lpad]
    out[:, half:half + lpad:2] = table[c & 15]
    out[:, half + 1:half + lpad:2] = table[c >> 4]
    return out
