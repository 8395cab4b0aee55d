"""Result of one pass of the consensus path over a packed batch.

Everything is indexed by family id in order of family creation (the order in
which `read_bam` first meets each `unique_tag`, consensus_helper.py:447-449),
which is also the reference's emission order within a stage
(SSCS_maker.py:274-276).  Row arrays are compact: `sscs_row[g]` etc. give the
row of family g or -1.
"""
from __future__ import annotations

import numpy as np

STAT_NAMES = ("sscs", "singletons", "dcs", "sscs_singletons", "sc_by_sscs", "sc_by_singleton",
              "uncorrected", "dcs_sc", "sscs_sc_singletons")

PER_FAMILY = ("fam_size", "fam_tmpl", "fam_mapq", "fam_flag", "fam_tlen", "fam_rg", "partner",
              "sscs_row", "corr_row", "dcs_row", "dcssc_row")
ROWS = ("sscs_rows", "corr_rows", "dcs_rows", "dcssc_rows")
PAIR_META = ("dcs_mapq", "dcs_flag", "dcs_tlen", "dcs_rg",
             "dcssc_mapq", "dcssc_flag", "dcssc_tlen", "dcssc_rg")
ALL_ARRAYS = ("read_fam",) + PER_FAMILY + ROWS + PAIR_META + ("stats",)


class ConsensusResult:
    __slots__ = ALL_ARRAYS + ("lpad", "stride", "scorrect")

    def __init__(self, lpad, stride, scorrect, **arrays):
        self.lpad, self.stride, self.scorrect = int(lpad), int(stride), bool(scorrect)
        for k in ALL_ARRAYS:
            setattr(self, k, arrays[k])

    @property
    def n_fam(self):
        return int(self.fam_size.shape[0])

    def stat(self, name):
        return int(self.stats[STAT_NAMES.index(name)])

    def family_size_table(self):
        """[(size, count)] in first-appearance order of sizes over family creation
        order - the body of read_families.txt (SSCS_maker.py:361-365)."""
        if self.n_fam == 0:
            return []
        sizes, first, counts = np.unique(self.fam_size, return_index=True, return_counts=True)
        order = np.argsort(first, kind="stable")
        return [(int(sizes[i]), int(counts[i])) for i in order]

    def diff(self, other):
        """Names of arrays that differ (bit-exact comparison)."""
        bad = []
        for k in ALL_ARRAYS:
            a, b = getattr(self, k), getattr(other, k)
            if a.shape != b.shape or a.dtype != b.dtype or not np.array_equal(a, b):
                bad.append(k)
        return bad
