"""Host-side alignment record used on both sides of the GPU path.

A `Read` carries exactly the attributes the reference touches on a
`pysam.AlignedSegment` (`consensus_helper.py:564-581` for the
setters, `:133-144,188-192,225-245,287-297,381-404` for the getters).  It is a
plain Python object: decode/encode to BAM lives in `bam.py`, packing into the
device SoA lives in `pack.py`.
"""
from __future__ import annotations

import numpy as np

# BAM 4-bit base codes ("=ACMGRSVTWYHKDBN"): the device payload keeps BAM's codes
# (A/C/G/T one-hot 1,2,4,8, N = 15) with the even position in the low nibble.
BASE_CODES = "=ACMGRSVTWYHKDBN"
_CODE_OF = np.full(256, 255, dtype=np.uint8)
for _i, _c in enumerate(BASE_CODES):
    _CODE_OF[ord(_c)] = _i
    _CODE_OF[ord(_c.lower())] = _i
_CHAR_OF = np.frombuffer(BASE_CODES.encode(), dtype=np.uint8)


def seq_to_codes(seq: str) -> np.ndarray:
    """ASCII bases -> BAM 4-bit codes (one byte per base)."""
    codes = _CODE_OF[np.frombuffer(seq.encode(), dtype=np.uint8)]
    if (codes == 255).any():
        raise ValueError("sequence has a character outside '=ACMGRSVTWYHKDBN': %r" % seq)
    return codes


def codes_to_seq(codes: np.ndarray) -> str:
    return _CHAR_OF[np.asarray(codes, dtype=np.uint8)].tobytes().decode()


class Read:
    """One alignment record (the subset of pysam.AlignedSegment the path uses)."""

    __slots__ = ("qname", "flag", "tid", "pos", "mapq", "cigar", "mtid", "mpos",
                 "tlen", "seq", "qual", "rg", "aux")

    def __init__(self, qname="", flag=0, tid=-1, pos=-1, mapq=0, cigar="", mtid=-1,
                 mpos=-1, tlen=0, seq="", qual=None, rg=None, aux=b""):
        self.qname = qname
        self.flag = int(flag)
        self.tid = int(tid)
        self.pos = int(pos)            # 0-based, as pysam reports reference_start
        self.mapq = int(mapq)
        self.cigar = cigar             # SAM string, e.g. "98M"
        self.mtid = int(mtid)
        self.mpos = int(mpos)
        self.tlen = int(tlen)
        self.seq = seq                 # ASCII
        self.qual = (np.zeros(0, np.uint8) if qual is None
                     else np.asarray(qual, dtype=np.uint8))
        self.rg = rg                   # RG:Z value or None
        # aux fields other than RG: raw BAM aux bytes (from the codec) or a {tag: value} dict
        self.aux = aux

    # -- flag helpers (consensus_helper.py:287,396-404) -------------------
    @property
    def is_reverse(self):
        return bool(self.flag & 0x10)

    @property
    def is_unmapped(self):
        return bool(self.flag & 0x4)

    @property
    def is_secondary(self):
        return bool(self.flag & 0x100)

    @property
    def is_supplementary(self):
        return bool(self.flag & 0x800)

    def copy(self):
        r = Read.__new__(Read)
        for s in Read.__slots__:
            setattr(r, s, getattr(self, s))
        return r

    def key(self):
        """Everything that ends up in a BAM record, for record-level diffs."""
        aux = self.aux
        if isinstance(aux, dict):
            aux = tuple(sorted(aux.items()))
        return (self.qname, self.flag, self.tid, self.pos, self.mapq, self.cigar,
                self.mtid, self.mpos, self.tlen, self.seq,
                bytes(np.asarray(self.qual, np.uint8)), self.rg, aux or ())

    def __repr__(self):
        return "Read(%s flag=%d %d:%d mapq=%d %s mate=%d:%d tlen=%d %s)" % (
            self.qname, self.flag, self.tid, self.pos, self.mapq, self.cigar,
            self.mtid, self.mpos, self.tlen, self.seq[:24])


def sort_key(r: Read):
    """Coordinate-sort key of `sort_index` (ConsensusCruncher.py:10-29).

    Classic samtools ordering: (tid, pos, is_reverse), stable otherwise;
    records without a reference go last.
    """
    tid = r.tid if r.tid >= 0 else (1 << 31)
    return (tid, r.pos, 1 if (r.flag & 0x10) else 0)


def sort_records(records, sorter=None):
    """Stable coordinate sort used in place of `samtools view -bu | samtools sort -`.

    `sorter(records) -> list` does the ordering elsewhere (`GpuConsensus.sort_records`: the same
    key, sorted on the GPU); the default is the host statement of the rule."""
    if sorter is not None:
        return sorter(records)
    return sorted(records, key=sort_key)


def merge_records(*files, sorter=None):
    """`samtools merge OUT IN1 IN2 IN3`: ties broken by input-file order
    (ConsensusCruncher.py:248,282)."""
    out = []
    for f in files:
        out.extend(f)
    return sort_records(out, sorter)
