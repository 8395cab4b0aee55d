#!/usr/bin/env python
"""Soak: many random small and medium batches through the GPU path against the oracle, each batch several times
(members land in a family's segment in arrival order, so repeats exercise different interleavings of the atomics),
every third one also with its pairs shuffled (the all-through-the-table grouping pass), through the most compact
upload the batch fits and with the rows read back in bit planes.  compute-sanitizer is closed on this pool (its wrapper refuses to run), so this - bit-exact agreement with
the CPU oracle over hundreds of cases and repeats, in every grouping mode (CCGPU_GROUP) - stands in for racecheck.
python scripts/gpu_soak.py [first_seed] [count]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from consensuscruncher_b200 import pack, synth
from consensuscruncher_b200.engine import GpuConsensus, GpuError
from oracle import oracle
import test_gpu_parity as T

first = int(sys.argv[1]) if len(sys.argv) > 1 else 100
count = int(sys.argv[2]) if len(sys.argv) > 2 else 300
bad = 0
with GpuConsensus(0) as eng:
    for seed in range(first, first + count):
        packed, cutoff, scorrect = T._fuzz_case(seed)
        if seed % 10 == 0:                      # now and then something bigger
            packed = synth.generate(seed=seed, n_molecules=20000 + 1000 * (seed % 7), L=100 + seed % 60, mapq_mixed=True,
                                    region=(0, 50000)).to_packed()
        eng.set_params(cutoff=cutoff, scorrect=scorrect)
        try:
            exp, faults = oracle.run(packed, cutoff, scorrect, return_faults=True)
        except oracle.OracleError:
            continue
        if faults:
            continue
        for rep in range(3):
            got = eng.consensus(packed, compact="auto" if rep == 2 else False)
            d = got.diff(exp)
            if not d and rep == 1:                  # the same rows read back in bit planes (cc_pack_rows), when they fit
                d = eng.result(planes=True).diff(exp)
            if d:
                bad += 1
                print("seed", seed, "rep", rep, "differs in", d)
                break
        if seed % 3 == 0 and packed.n >= 4:
            order = np.random.default_rng(seed).permutation(packed.n // 2)
            idx = np.stack([2 * order, 2 * order + 1], 1).reshape(-1)
            mixed = pack.PackedReads(packed.fields[idx].copy(), packed.payload[idx].copy(), packed.lpad, packed.barcodes,
                                     packed.dup_bc, packed.cigar_pairs, packed.rgs)
            d = eng.consensus(mixed).diff(oracle.run(mixed, cutoff, scorrect))
            if d:
                bad += 1
                print("seed", seed, "shuffled differs in", d)
print("soak done:", count, "cases,", bad, "bad")
sys.exit(1 if bad else 0)
