"""TEST INFRASTRUCTURE - stand-in for Bio.SeqIO, enough for the unmodified reference
`extract_barcodes.py` (`SeqIO.parse(handle, "fastq")`, `record.seq[:n]`, `record[n:]`, `.id`,
`.description`, `SeqIO.write(record, handle, "fastq")`), restating Biopython's documented behaviour:

  * a FASTQ title line `@id rest of line`: `.id` is the first word, `.description` the whole title;
  * slicing a record slices the sequence and the per-letter qualities and keeps id / description;
  * writing: the title is the description when that starts with the id (as one word), else
    "id description"; the qualities go out as Phred+33.
"""


class _Record:
    def __init__(self, rid, description, seq, qual):
        self.id, self.description, self.seq, self._qual = rid, description, seq, qual
        self.name = rid

    def __getitem__(self, key):
        if not isinstance(key, slice):
            raise TypeError("only slices of a record are supported here")
        return _Record(self.id, self.description, self.seq[key], self._qual[key])

    def __len__(self):
        return len(self.seq)


def parse(handle, fmt):
    if fmt != "fastq":
        raise ValueError("only fastq")
    while True:
        title = handle.readline()
        if not title:
            return
        title = title.rstrip("\r\n")
        if not title:
            continue
        if not title.startswith("@"):
            raise ValueError("FASTQ record does not start with '@': %r" % title)
        seq = handle.readline().rstrip("\r\n")
        plus = handle.readline()
        if not plus.startswith("+"):
            raise ValueError("multi-line FASTQ is not supported by this stand-in")
        qual = handle.readline().rstrip("\r\n")
        if len(qual) != len(seq):
            raise ValueError("sequence and quality lengths differ in %r" % title)
        text = title[1:]
        yield _Record(text.split(None, 1)[0] if text.split() else "", text, seq, qual)


def write(record, handle, fmt):
    if fmt != "fastq":
        raise ValueError("only fastq")
    rid, desc = record.id, record.description
    if desc and desc.split(None, 1)[0] == rid:
        title = desc
    elif desc:
        title = "%s %s" % (rid, desc)
    else:
        title = rid
    handle.write("@%s\n%s\n+\n%s\n" % (title, record.seq, record._qual))
    return 1
