"""Family-key helpers: the host mirror of `consensus_helper.py`'s naming functions.

Same names, arguments and results as the reference so parity tests read like
its docstrings, but table-driven instead of list scans:

  which_read        consensus_helper.py:57-81
  which_strand      consensus_helper.py:84-155
  cigar_order       consensus_helper.py:158-194
  sscs_qname        consensus_helper.py:197-247
  unique_tag        consensus_helper.py:250-302
  duplex_tag        consensus_helper.py:605-648
  dcs_consensus_tag DCS_maker.py:60-96

The GPU path never sees these strings: it works on the exact field tuple
(barcode id, cigar-pair id, tid, pos, mtid, mpos, orientation, read number)
that `unique_tag` would print.  The strings are only built when records are
written.
"""
from __future__ import annotations

# flag -> 'R1' / 'R2' (16 legal flags; anything else is outside the reference's domain)
_R1_FLAGS = (99, 83, 67, 115, 81, 97, 65, 113)
_R2_FLAGS = (147, 163, 131, 179, 161, 145, 129, 177)
READ_NUM = {f: "R1" for f in _R1_FLAGS}
READ_NUM.update({f: "R2" for f in _R2_FLAGS})

_POS_FLAGS = frozenset((99, 147, 67, 131))
_NEG_FLAGS = frozenset((83, 163, 115, 179))
_NO_ORI_FLAGS = frozenset((65, 129, 113, 177, 81, 161, 97, 145))

MATE_UNMAPPED_FLAGS = frozenset((73, 89, 121, 153, 185, 137))  # consensus_helper.py:390


def which_read(flag):
    """'R1' / 'R2' for the 16 legal paired flags, else None."""
    return READ_NUM.get(flag)


def which_strand(read):
    """'pos' / 'neg' strand of origin; coordinates decide for the no-orientation flags."""
    f = read.flag
    if f in _POS_FLAGS:
        return "pos"
    if f in _NEG_FLAGS:
        return "neg"
    if f in _NO_ORI_FLAGS:
        first = READ_NUM[f] == "R1"
        here = (read.tid, read.pos)
        there = (read.mtid, read.mpos)
        if read.tid != read.mtid:
            lower = read.tid < read.mtid
            return "pos" if lower == first else "neg"
        if here == there:
            return "neg"
        lower = read.pos < read.mpos
        return "pos" if lower == first else "neg"
    return None


def cigar_order(read, mate):
    """Ordered cigar pair: the read's cigar leads iff (pos and R1) or (neg and R2)."""
    strand = which_strand(read)
    num = which_read(read.flag)
    if (strand == "pos" and num == "R1") or (strand == "neg" and num == "R2"):
        return read.cigar + "_" + mate.cigar
    return mate.cigar + "_" + read.cigar


def pair_coords(read, mate):
    """(chrLo, posLo, chrHi, posHi) as `sscs_qname` orders them."""
    a = (read.tid, read.pos)
    b = (mate.tid, mate.pos)
    if (read.tid == mate.tid and read.pos > mate.pos) or read.tid > mate.tid:
        a, b = b, a
    return a + b


def sscs_qname(read, mate, barcode, cigar):
    """Consensus pair name `barcode_chrLo_posLo_chrHi_posHi_cigarpair_strand_|tlen|`."""
    c = pair_coords(read, mate)
    return "%s_%d_%d_%d_%d_%s_%s_%d" % (barcode, c[0], c[1], c[2], c[3], cigar,
                                        which_strand(read), abs(read.tlen))


def unique_tag(read, barcode, cigar):
    """Family key string `barcode_tid_pos_mtid_mpos_cigarpair_{fwd|rev}_{R1|R2}`."""
    return "%s_%d_%d_%d_%d_%s_%s_%s" % (
        barcode, read.tid, read.pos, read.mtid, read.mpos, cigar,
        "rev" if read.flag & 0x10 else "fwd", which_read(read.flag))


def duplex_barcode(barcode):
    """Barcode of the complementary strand: halves swapped at '.', else at len//2."""
    dot = barcode.find(".")
    if dot >= 0:
        return barcode[dot + 1:] + "." + barcode[:dot]
    h = len(barcode) // 2
    return barcode[h:] + barcode[:h]


def duplex_tag(tag):
    """Key of the complementary strand: barcode halves swapped, R1 <-> R2."""
    f = tag.split("_")
    f[0] = duplex_barcode(f[0])
    f[8] = "R2" if f[8] == "R1" else "R1"
    return "_".join(f)


def dcs_consensus_tag(tag, ds):
    """DCS query name from the two SSCS query names (pos-strand barcode first)."""
    bc = tag.split("_")[0]
    dbc = ds.split("_")[0]
    coor = tag.split("_", 1)[1].rsplit("_", 1)[0]
    n = tag.split(":")[1]
    dn = ds.split(":")[1]
    if "pos" in tag:
        return "%s_%s_%s:%s_%s" % (bc, dbc, coor, n, dn)
    return "%s_%s_%s:%s_%s" % (dbc, bc, coor, dn, n)


def bed_separator(bedfile):
    """Region list of a cytoband/bed file (consensus_helper.py:38-54).

    Returns an ordered dict `chr_band -> (start, end)`; duplicate keys keep the
    first position and the last value, like the reference's OrderedDict.
    """
    coor = {}
    with open(bedfile) as fh:
        for line in fh:
            f = line.split("\t")
            coor["%s_%s" % (f[0], f[3])] = (int(f[1]), int(f[2]))
    return coor
