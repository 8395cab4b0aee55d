"""Function-level mirrors of the reference's consensus primitives, running on the GPU.

Same names, argument meaning and results as the reference functions so that
tests read like the reference's docstring examples:

  consensus_maker(readList, cutoff)   SSCS_maker.py:75-157
  duplex_consensus(read1, read2)      DCS_maker.py:99-121
  sc_duplex_consensus(read1, read2)   singleton_correction.py:61-84

Reads are anything with `.seq` (str) and `.qual` (sequence of ints), e.g.
`records.Read`.  Errors follow the reference: a base other than A/C/G/T with
Q >= 30 raises (IndexError for N, ValueError otherwise in the reference;
`engine.BaseError` here).
"""
from __future__ import annotations

import numpy as np

from .engine import GpuConsensus
from .pack import pack_row, row_geometry, unpack_row
from .records import codes_to_seq, seq_to_codes

_default_engine = None


def default_engine():
    global _default_engine
    if _default_engine is None:
        _default_engine = GpuConsensus(0)
    return _default_engine


def rows_of(reads, lpad=None, stride=None):
    n = max(len(r.seq) for r in reads)
    if lpad is None:
        lpad, stride = row_geometry(n)
    rows = np.zeros((len(reads), stride), np.uint8)
    for i, r in enumerate(reads):
        pack_row(seq_to_codes(r.seq), np.asarray(r.qual, np.uint8), lpad, rows[i])
    return rows, lpad


def row_to_seq_qual(row, lpad, n):
    codes, qual = unpack_row(row, lpad, n)
    return codes_to_seq(codes), [int(q) for q in qual]


def consensus_maker(readList, cutoff, engine=None):
    """(list of reads, float) -> (consensus sequence, consensus qualities)."""
    eng = engine or default_engine()
    rows, lpad = rows_of(readList)
    out = eng.consensus_maker_rows(rows, lpad, cutoff)
    return row_to_seq_qual(out, lpad, len(readList[0].seq))


def duplex_consensus(read1, read2, engine=None):
    """DCS rule: equal bases keep the base with min(60, q1+q2), anything else is N with 0."""
    eng = engine or default_engine()
    rows, lpad = rows_of([read1, read2])
    out = eng.duplex_consensus_rows(rows[0:1], rows[1:2], lpad, sc=False)
    return row_to_seq_qual(out[0], lpad, len(read1.seq))


def sc_duplex_consensus(read1, read2, engine=None):
    """Singleton-correction rule: as duplex_consensus, and both qualities must exceed 29."""
    eng = engine or default_engine()
    rows, lpad = rows_of([read1, read2])
    out = eng.duplex_consensus_rows(rows[0:1], rows[1:2], lpad, sc=True)
    return row_to_seq_qual(out[0], lpad, len(read1.seq))
