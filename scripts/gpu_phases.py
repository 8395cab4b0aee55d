#!/usr/bin/env python
"""Phase times of one pass over a sharded workload (c3 / c4) on one GPU: python scripts/gpu_phases.py c3 200000000 [world]
(rank 0's share of `world` ranks; cc_profile events per phase, 3 passes after one warm-up)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from consensuscruncher_b200 import workloads
from consensuscruncher_b200.engine import GpuConsensus

wl, total = sys.argv[1], int(sys.argv[2])
world = int(sys.argv[3]) if len(sys.argv) > 3 else 1
pl = workloads.plan(wl, total // 2, seed=1234)
batch, info = workloads.generate_shard(pl, 0, world, "cuda:0")
with GpuConsensus(0, cutoff=0.7, scorrect=True) as eng:
    def go():
        eng.load_device(batch.fields.data_ptr(), batch.payload.data_ptr(), batch.n, batch.lpad, batch.stride,
                        batch.dup_bc.data_ptr(), batch.n_bc)
        eng.run()
    go()
    eng.profile(True)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    for _ in range(3):
        go()
    ph = eng.phase_ms()
    eng.profile(False)
    torch.cuda.synchronize(); ev0.record()
    for _ in range(3):
        go()
    ev1.record(); torch.cuda.synchronize()
    fs = eng.fetch("fam_size")
    print(wl, "reads", batch.n, "families", len(fs), "mean size %.2f" % (batch.n / max(1, len(fs))),
          ">15: %d  >127: %d  max %d" % ((fs > 15).sum(), (fs > 127).sum(), fs.max()))
    print("ms per pass (no profiling): %.3f" % (ev0.elapsed_time(ev1) / 3))
    print({k: round(v[0] / 3, 3) for k, v in ph.items() if v[1]})
