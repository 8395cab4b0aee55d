"""CPU: the host logic of the three stage commands (region loops, order of emission, what a region loop forgets,
file names, stats text) against the reference's golden output tree, with the CPU oracle standing in for the device:
the stand-in answers exactly the calls `stages.py` makes on a GpuConsensus.  The GPU test of the same chain is
tests/test_stage_commands.py; this one keeps the host side covered where there is no GPU."""
import numpy as np
import pytest

from consensuscruncher_b200 import pack, records
from oracle import oracle

import test_stage_commands as chain


class OracleEngine:
    """Test double: every device call of stages.py answered by the CPU restatement (oracle/), or by a dictionary."""

    def __init__(self):
        self.cutoff, self.scorrect = 0.7, True
        self.packed = self.res = None

    def set_params(self, cutoff=None, scorrect=None):
        self.cutoff = self.cutoff if cutoff is None else cutoff
        self.scorrect = self.scorrect if scorrect is None else scorrect

    def load(self, packed):
        self.packed = packed

    def group(self):
        self.res = oracle.run(self.packed, self.cutoff, self.scorrect)

    def sscs_vote(self):
        pass

    def fetch(self, name):
        return getattr(self.res, name)

    def duplex_consensus_rows(self, rows1, rows2, lpad, sc):
        rows1, rows2 = np.atleast_2d(rows1), np.atleast_2d(rows2)        # (the oracle's call takes one pair of rows)
        return np.stack([oracle.duplex_consensus_rows(a, b, lpad, sc) for a, b in zip(rows1, rows2)])

    def sort_records(self, recs):
        return records.sort_records(recs)

    def join_records(self, fields, dup_bc):
        """What cc_join_records computes (include/ccgpu.h), stated with a dictionary."""
        key = lambda r, bc, first: (int(bc), int(r["cig"]), int(r["tid"]), int(r["pos"]), int(r["mtid"]), int(r["mpos"]),
                                    int(r["flag"]) >> 4 & 1, first)
        seen = {}
        for i, r in enumerate(fields):
            seen.setdefault(key(r, r["bc"], int(r["flag"]) >> 6 & 1), i)
        out = np.full(len(fields), -1, np.int32)
        for i, r in enumerate(fields):
            d = dup_bc[r["bc"]] if r["bc"] < len(dup_bc) else pack.NO_ID
            if d != pack.NO_ID:
                out[i] = seen.get(key(r, d, (int(r["flag"]) >> 6 & 1) ^ 1), -1)
        return out


@pytest.mark.parametrize("name", ["c1prime_real_fastq", "duplex_small", "mode_ties", "bed_three_chr", "crowded", "undotted_umi"])
def test_stage_chain_host_logic_reproduces_reference_tree(tmp_path, name):
    chain.test_stage_chain_reproduces_reference_tree(tmp_path, name, OracleEngine())
