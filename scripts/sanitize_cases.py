#!/usr/bin/env python
"""Small batches that reach every kernel of the path, for compute-sanitizer (memcheck / racecheck / synccheck /
initcheck): smoke-sized duplex batch, crowded keys, one huge family, shuffled input (all-through-the-table
grouping pass), a deep run spanning tiles, the compact upload, the sort and hash grouping paths.
Every result is still compared with the oracle."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from consensuscruncher_b200 import synth                      # noqa: E402
from consensuscruncher_b200.engine import GpuConsensus         # noqa: E402
from consensuscruncher_b200.pack import PackedReads             # noqa: E402
from oracle import oracle                                       # noqa: E402


def shuffled(p, seed=1):
    order = np.random.default_rng(seed).permutation(p.n // 2)
    idx = np.stack([2 * order, 2 * order + 1], 1).reshape(-1)
    return PackedReads(p.fields[idx].copy(), p.payload[idx].copy(), p.lpad, p.barcodes, p.dup_bc, p.cigar_pairs, p.rgs)


cases = {
    "smoke": synth.generate(seed=7, n_molecules=400, L=128).to_packed(),
    "crowded keys": synth.generate(seed=25, n_molecules=1500, L=36, umi_len=1, region=(0, 600), insert_sd=2.0).to_packed(),
    "one huge family": synth.generate(seed=95, n_molecules=2, L=64, dist="const", fam_mean=1500, p_duplex=1.0).to_packed(),
    "deep run across tiles": synth.generate(seed=72, n_molecules=900, L=40, umi_len=4, region=(0, 500), insert_sd=1.0, fam_mean=6.0).to_packed(),
    "mixed mode fields": synth.generate(seed=26, n_molecules=300, L=100, mapq_mixed=True).to_packed(),
}
cases["shuffled"] = shuffled(cases["smoke"])
mode = os.environ.get("CCGPU_GROUP", "tile")
with GpuConsensus(0, cutoff=0.7, scorrect=True) as eng:
    for name, packed in cases.items():
        for compact in (False, "auto"):
            got = eng.consensus(packed, compact=compact)
            bad = got.diff(oracle.run(packed, 0.7, True))
            assert not bad, (name, compact, bad)
        print("ok [%s] %s: %d reads, %d families" % (mode, name, packed.n, got.n_fam), flush=True)
