#!/usr/bin/env python
"""Wall time of the `consensus` command on a synthetic BAM: python scripts/cli_timing.py [n_molecules]."""
import os, sys, tempfile, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from consensuscruncher_b200 import bam, cli, synth

n = int(sys.argv[1]) if len(sys.argv) > 1 else 150000
d = tempfile.mkdtemp()
t = time.time()
s = synth.generate(seed=11, n_molecules=n, L=128)
reads = s.to_reads()
path = os.path.join(d, "sample.bam")
bam.write_bam(path, bam.BamHeader("", ["chr1", "chr2"], [10 ** 8, 10 ** 8]), reads)
print("input: %d reads, %.0f MB BAM (generated in %.0f s)" % (len(reads), os.path.getsize(path) / 1e6, time.time() - t))
del reads, s
for k in range(2):
    t = time.time()
    out = cli.run_consensus(path, os.path.join(d, "out%d" % k), log=lambda *_: None)
    wall = time.time() - t
    print("run %d: %.2f s wall" % (k, wall))
    print(open(os.path.join(out, "sample.time_tracker.txt")).read().replace("\n", " | "), "(minutes)")

# where the GPU step's wall time goes
from consensuscruncher_b200 import fastpack
from consensuscruncher_b200.engine import GpuConsensus
header, f = fastpack.open_bam(path)
packed = fastpack.pack_reads(f)
t = time.time(); eng = GpuConsensus(0); t1 = time.time()
eng.run_host(packed); t2 = time.time()
res = eng.result(); t3 = time.time()
eng.run_host(packed); t4 = time.time()
print("create %.2f s, first run_host %.2f s, result() %.2f s, second run_host %.2f s" % (t1 - t, t2 - t1, t3 - t2, t4 - t3))
