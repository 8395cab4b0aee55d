"""Barcode extraction: the step in front of the consensus path.

Same command line, same output files and the same text in them as the reference's
`ConsensusCruncher/extract_barcodes.py` (`:140-160` arguments, `:228-329` the two modes, `:340-396`
the stats file), without Biopython: FASTQ records are read as four lines and handled as byte
strings.  The plot of the barcode-list mode is not drawn (no matplotlib on the boxes).

  pattern mode  --bpattern NNT     N = barcode base, A/C/G/T = spacer base that has to match;
                                   the first len(pattern) bases are cut off both mates, the N
                                   positions of mate 1 and mate 2 go into the read name as
                                   `name|B1.B2/1` and `/2`
  list mode     --blist FILE       barcodes of any length ending in T; tried longest first, the
                                   last (shortest) length that matches wins, the T is dropped
Parity is pinned by tests/golden/extract_*.json.gz (oracle/make_golden_extract.py runs the
unmodified reference on the reference's own test FASTQ files).
"""
from __future__ import annotations

import argparse
import gzip
import re
import sys

import numpy as np

NUC = "ACGTN"
_NOT_ACGT = re.compile("[^ACGT]")
_NOT_ACGTN = re.compile("[^ACGTN]")


def read_fastq(path):
    """(title without '@', sequence, quality) of every 4-line record; gzip when 'gz' is in the name."""
    opener = gzip.open if "gz" in path else open
    with opener(path, "rt") as fh:
        while True:
            title = fh.readline()
            if not title:
                return
            title = title.rstrip("\r\n")
            if not title:
                continue
            seq = fh.readline().rstrip("\r\n")
            plus = fh.readline()
            qual = fh.readline().rstrip("\r\n")
            if not title.startswith("@") or not plus.startswith("+") or len(seq) != len(qual):
                raise ValueError("%s: malformed FASTQ record %r" % (path, title))
            yield title[1:], seq, qual


def _read_id(title):
    """Biopython's record.id: the first word of the title line."""
    words = title.split(None, 1)
    return words[0] if words else ""


def check_overlap(blist):
    """extract_barcodes.py:57-90 returns None whatever it finds (the flag it sets is never returned),
    so overlapping barcodes are never rejected; kept as a function for the same reason it is there."""
    return None


def _stats_head(outfile, total, bad_spacer, bad_barcode, good):
    return ("##########\n%s\n##########\nTotal sequences: %d\nMissing spacer: %d\nBad barcodes: %d\n"
            "Passing barcodes: %d\n" % (outfile.split("/")[-1], total, bad_spacer, bad_barcode, good))


def extract_pattern(read1, read2, outfile, bpattern, stats_path=None):
    """Pattern mode (extract_barcodes.py:233-267).  -> (total, bad_spacer, bad_barcode, good)."""
    import pandas as pd
    if _NOT_ACGTN.search(bpattern):
        raise ValueError("Invalid barcode pattern inputted. Please specify pattern with A|C|G|T = fixed, "
                         "N = variable (e.g. 'ATNNGCT').")
    plen = len(bpattern)
    b_index = [i for i, c in enumerate(bpattern) if c == "N"]
    s_index = [i for i in range(plen) if bpattern[i] != "N"]
    spacer = "".join(bpattern[i] for i in s_index)
    counts = [np.zeros((plen, 5), np.int64), np.zeros((plen, 5), np.int64)]      # [mate][position][A,C,G,T,N]
    col = {c: i for i, c in enumerate(NUC)}
    seen = [{}, {}]                              # prefix -> times seen; turned into per-position counts at the end
    total = bad_spacer = bad_barcode = good = 0
    with open("%s_barcode_R1.fastq" % outfile, "w") as o1, open("%s_barcode_R2.fastq" % outfile, "w") as o2:
        for (t1, s1, q1), (t2, s2, q2) in zip(read_fastq(read1), read_fastq(read2)):
            total += 1
            id1, id2 = _read_id(t1), _read_id(t2)
            if id1 != id2:
                raise AssertionError("read names differ: %s / %s" % (id1, id2))
            b1, b2 = s1[:plen], s2[:plen]
            if _NOT_ACGT.search(b1) or _NOT_ACGT.search(b2):
                bad_barcode += 1
                continue
            seen[0][b1] = seen[0].get(b1, 0) + 1
            seen[1][b2] = seen[1].get(b2, 0) + 1
            bc1 = "".join(b1[i] for i in b_index if i < len(b1))
            bc2 = "".join(b2[i] for i in b_index if i < len(b2))
            if "".join(b1[i] for i in s_index if i < len(b1)) == spacer and \
               "".join(b2[i] for i in s_index if i < len(b2)) == spacer:
                good += 1
                o1.write("@%s|%s.%s/1\n%s\n+\n%s\n" % (id1, bc1, bc2, s1[plen:], q1[plen:]))
                o2.write("@%s|%s.%s/2\n%s\n+\n%s\n" % (id2, bc1, bc2, s2[plen:], q2[plen:]))
            else:
                bad_spacer += 1
    for m in (0, 1):
        for b, k in seen[m].items():
            for i, c in enumerate(b):
                counts[m][i, col[c]] += k
    if stats_path:
        frames = [pd.DataFrame(c, index=np.arange(plen), columns=list(NUC)).apply(lambda x: x / x.sum(), axis=1)
                  for c in counts]
        with open(stats_path, "a") as fh:
            fh.write(_stats_head(outfile, total, bad_spacer, bad_barcode, good))
            fh.write("---BARCODE---\n%s\n-----------\n%s\n" % (frames[0], frames[1]))
    return total, bad_spacer, bad_barcode, good


def extract_list(read1, read2, outfile, blist, stats_path=None):
    """Barcode-list mode (extract_barcodes.py:270-329).  -> (total, 0, bad_barcode, good)."""
    import pandas as pd
    if _NOT_ACGTN.search("".join(blist)):
        raise ValueError("Invalid barcode list inputted. Please specify barcodes with A|C|G|T.")
    if [s for s in blist if not s.endswith("T")]:
        raise ValueError("There is one or more barcodes in the list that do not end with 'T'.")
    check_overlap(blist)
    known = set(blist)
    lengths = sorted(set(len(b) for b in blist), reverse=True)      # longest first: a shorter match found later wins
    tally = [dict.fromkeys(blist, 0), dict.fromkeys(blist, 0)]
    total = bad_barcode = good = 0
    with open("%s_barcode_R1.fastq" % outfile, "w") as o1, open("%s_barcode_R2.fastq" % outfile, "w") as o2, \
            open("%s_r1_bad_barcodes.txt" % outfile, "w") as bad1, open("%s_r2_bad_barcodes.txt" % outfile, "w") as bad2:
        for (t1, s1, q1), (t2, s2, q2) in zip(read_fastq(read1), read_fastq(read2)):
            total += 1
            id1, id2 = _read_id(t1), _read_id(t2)
            if id1 != id2:
                raise AssertionError("read names differ: %s / %s" % (id1, id2))
            hit1 = hit2 = None                  # (barcode without its T, bases cut off)
            b1 = b2 = ""
            for blen in lengths:
                b1, b2 = s1[:blen], s2[:blen]
                n1, n2 = _NOT_ACGT.search(b1), _NOT_ACGT.search(b2)
                if n1 or n2:                    # counted once per length tried, like the reference
                    bad_barcode += 1
                    if n1:
                        bad1.write(b1 + "\n")
                    if n2:
                        bad2.write(b2 + "\n")
                    continue
                if b1 in known:
                    hit1 = (b1[:blen - 1], blen)
                if b2 in known:
                    hit2 = (b2[:blen - 1], blen)
            if hit1 and hit2:
                good += 1
                tally[0][hit1[0] + "T"] += 1
                tally[1][hit2[0] + "T"] += 1
                o1.write("@%s|%s.%s/1\n%s\n+\n%s\n" % (id1, hit1[0], hit2[0], s1[hit1[1]:], q1[hit1[1]:]))
                o2.write("@%s|%s.%s/2\n%s\n+\n%s\n" % (id2, hit1[0], hit2[0], s2[hit2[1]:], q2[hit2[1]:]))
            else:
                bad_barcode += 1                # the barcodes written are those of the shortest length
                if not hit1:
                    bad1.write(b1 + "\n")
                if not hit2:
                    bad2.write(b2 + "\n")
    if stats_path:
        order = lambda d: sorted(d.items(), key=lambda kv: (len(kv[0]), kv[0]))
        r1 = pd.DataFrame(order(tally[0]), columns=["Barcode", "R1_Count"])
        r2 = pd.DataFrame(order(tally[1]), columns=["Barcode", "R2_Count"])
        both = pd.merge(r1, r2, on="Barcode")
        both["Total"] = both["R1_Count"] + both["R2_Count"]
        both = both.sort_values(by="Total", ascending=False)
        with open(stats_path, "a") as fh:
            fh.write(_stats_head(outfile, total, 0, bad_barcode, good))
            fh.write("---BARCODE---\n%s\n" % both)
    return total, 0, bad_barcode, good


def main(argv=None):
    p = argparse.ArgumentParser()
    p.add_argument("--read1", required=True, help="Input FASTQ file for Read 1 (gzipped if 'gz' is in the name)")
    p.add_argument("--read2", required=True, help="Input FASTQ file for Read 2")
    p.add_argument("--outfile", required=True, help="Prefix of the output files")
    p.add_argument("--bpattern", default=None,
                   help="Barcode pattern (N = random barcode bases, A|C|G|T = fixed spacer bases), e.g. NNT")
    p.add_argument("--blist", default=None, help="List of correct barcodes, one per line")
    a = p.parse_args(argv)
    if a.blist is None and a.bpattern is None:
        raise ValueError("No barcode specifications inputted. Please specify barcode list or pattern.")
    # the stats file sits next to the output directory, named after it (extract_barcodes.py:168)
    stats_path = "%s_barcode_stats.txt" % a.outfile.rsplit("/", 1)[0]
    if a.bpattern is not None:                   # the pattern wins when both are given (:233)
        counts = extract_pattern(a.read1, a.read2, a.outfile, a.bpattern, stats_path)
    else:
        with open(a.blist) as fh:
            blist = fh.read().splitlines()
        counts = extract_list(a.read1, a.read2, a.outfile, blist, stats_path)
    sys.stderr.write("Total sequences: %d\nMissing spacer: %d\nBad barcodes: %d\nPassing barcodes: %d\n" % counts)
    return 0


if __name__ == "__main__":
    sys.exit(main())
