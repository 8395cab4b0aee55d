"""TEST INFRASTRUCTURE - a stand-in for `pysam`, so the UNMODIFIED reference
scripts under /root/reference can be executed in this container (pysam is not
installable here).  Only the attributes the reference's
consensus path touches are provided.

"BAM files" are pickles of `{"header": {...}, "records": [tuple, ...]}` with one
tuple per record: (qname, flag, tid, pos, mapq, cigar, mtid, mpos, tlen, seq,
qual-bytes, tags-dict).  Nothing in the product imports this module.
"""
import array
import os
import pickle
import re

_CIG = re.compile(r"(\d+)([MIDNSHP=X])")
_OPS = "MIDNSHP=X"
_OPEN = []          # every AlignmentFile opened for writing (SSCS_maker never closes one of its)


def _cig_tuples(s):
    return [(_OPS.index(o), int(n)) for n, o in _CIG.findall(s or "")]


def _ref_len(s):
    return sum(int(n) for n, o in _CIG.findall(s or "") if o in "MDN=X")


def _query_len(s):
    return sum(int(n) for n, o in _CIG.findall(s or "") if o in "MIS=X")


class AlignedSegment(object):
    def __init__(self, header=None):
        self.query_name = None
        self.flag = 0
        self.reference_id = -1
        self.reference_start = -1
        self.mapping_quality = 0
        self.cigarstring = None
        self.next_reference_id = -1
        self.next_reference_start = -1
        self.template_length = 0
        self._seq = None
        self._qual = None
        self._tags = {}

    # --- names -----------------------------------------------------------
    @property
    def qname(self):
        return self.query_name

    @qname.setter
    def qname(self, v):
        self.query_name = v

    # --- sequence / qualities (assigning the sequence drops the qualities) -
    @property
    def query_sequence(self):
        return self._seq

    @query_sequence.setter
    def query_sequence(self, v):
        self._seq = v
        self._qual = None

    @property
    def query_qualities(self):
        return None if self._qual is None else array.array("B", self._qual)

    @query_qualities.setter
    def query_qualities(self, v):
        self._qual = None if v is None else bytes(bytearray(v))

    @property
    def query_length(self):
        return len(self._seq) if self._seq else 0

    def infer_query_length(self, always=False):
        n = _query_len(self.cigarstring)
        return n if n else None

    # --- cigar -----------------------------------------------------------
    @property
    def cigar(self):
        return _cig_tuples(self.cigarstring)

    @cigar.setter
    def cigar(self, v):
        self.cigarstring = "".join("%d%s" % (n, _OPS[o]) for o, n in v) if v else None

    cigartuples = cigar

    # --- flags -----------------------------------------------------------
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

    # --- aux tags --------------------------------------------------------
    def get_tag(self, tag):
        return self._tags[tag]          # KeyError when absent, like pysam

    def set_tag(self, tag, value, value_type=None, replace=True):
        self._tags[tag] = value

    def has_tag(self, tag):
        return tag in self._tags

    # --- plumbing --------------------------------------------------------
    def _tuple(self):
        return (self.query_name, self.flag, self.reference_id, self.reference_start,
                self.mapping_quality, self.cigarstring, self.next_reference_id,
                self.next_reference_start, self.template_length, self._seq, self._qual,
                dict(self._tags))

    @classmethod
    def _from_tuple(cls, t):
        s = cls()
        (s.query_name, s.flag, s.reference_id, s.reference_start, s.mapping_quality,
         s.cigarstring, s.next_reference_id, s.next_reference_start, s.template_length,
         s._seq, s._qual, tags) = t
        s._tags = dict(tags)
        return s

    def __eq__(self, other):
        return isinstance(other, AlignedSegment) and self._tuple() == other._tuple()

    def __ne__(self, other):
        return not self.__eq__(other)

    __hash__ = None

    def __str__(self):
        return "\t".join(str(x) for x in self._tuple()[:10])


class AlignmentFile(object):
    def __init__(self, path, mode="rb", template=None, header=None, **kw):
        self.filename = path
        self.mode = mode
        self._closed = False
        if "r" in mode:
            with open(path, "rb") as fh:
                blob = pickle.load(fh)
            self.header = blob["header"]
            self._records = [AlignedSegment._from_tuple(t) for t in blob["records"]]
        else:
            self.header = dict(template.header) if template is not None else dict(header or {})
            self._records = []
            _OPEN.append(self)

    # --- header ----------------------------------------------------------
    @property
    def references(self):
        return tuple(self.header.get("references", ()))

    def get_tid(self, name):
        refs = self.header.get("references", ())
        return refs.index(name) if name in refs else -1

    @property
    def mapped(self):
        return sum(1 for r in self._records if not r.is_unmapped)

    # --- reading ---------------------------------------------------------
    def fetch(self, contig=None, start=None, stop=None, until_eof=False, **kw):
        if contig is None:
            return iter(list(self._records))
        tid = self.get_tid(contig)
        if tid < 0:
            raise ValueError("invalid contig `%s`" % contig)
        out = []
        for r in self._records:
            if r.reference_id != tid or r.is_unmapped:
                continue
            end = r.reference_start + max(1, _ref_len(r.cigarstring))
            if r.reference_start < stop and end > start:      # half-open overlap
                out.append(r)
        return iter(out)

    def mate(self, seg):
        raise ValueError("mate not found")

    # --- writing ---------------------------------------------------------
    def write(self, seg):
        self._records.append(AlignedSegment._from_tuple(seg._tuple()))   # snapshot

    def close(self):
        if self._closed:
            return
        self._closed = True
        if "w" in self.mode:
            with open(self.filename, "wb") as fh:
                pickle.dump({"header": self.header,
                             "records": [r._tuple() for r in self._records]}, fh)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def flush_all():
    """Close every writer that the calling script left open."""
    for f in list(_OPEN):
        f.close()
    del _OPEN[:]
