#!/usr/bin/env python
"""What the host gives N GPUs at once: N processes, one per GPU, each copying 3.2 GB pinned -> device with a
1.5 GB read-back running (the end-to-end leg's traffic pattern), all started together.

  python scripts/h2d_probe_ranks.py 1 2 4 8      -> one JSON line per N: per-rank and aggregate GB/s

Names the ceiling of the end-to-end leg at N GPUs: its aggregate cannot exceed what this prints."""
import json
import multiprocessing as mp
import sys
import time


def worker(rank, n_ranks, start, out):
    import torch
    torch.cuda.set_device(rank)
    n, m = 3_200_000_000, 1_500_000_000
    h = torch.empty(n, dtype=torch.uint8, pin_memory=True)
    d = torch.empty(n, dtype=torch.uint8, device="cuda")
    h2 = torch.empty(m, dtype=torch.uint8, pin_memory=True)
    d2 = torch.empty(m, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    with torch.cuda.stream(s1):
        d.copy_(h, non_blocking=True)
    torch.cuda.synchronize()
    start.wait()
    t = time.perf_counter()
    reps = 4
    for _ in range(reps):
        with torch.cuda.stream(s1):
            d.copy_(h, non_blocking=True)
        with torch.cuda.stream(s2):
            h2.copy_(d2, non_blocking=True)
    s1.synchronize()
    t_up = time.perf_counter() - t
    s2.synchronize()
    t_all = time.perf_counter() - t
    out.put((rank, reps * n / t_up / 1e9, reps * m / t_all / 1e9))


if __name__ == "__main__":
    mp.set_start_method("spawn")
    for n_ranks in [int(x) for x in sys.argv[1:]] or [1]:
        start, out = mp.Barrier(n_ranks), mp.Queue()
        procs = [mp.Process(target=worker, args=(r, n_ranks, start, out)) for r in range(n_ranks)]
        for p in procs:
            p.start()
        res = sorted(out.get() for _ in procs)
        for p in procs:
            p.join()
        print(json.dumps({"ranks": n_ranks, "h2d_gbs_per_rank": [round(r[1], 1) for r in res],
                          "d2h_gbs_per_rank": [round(r[2], 1) for r in res],
                          "h2d_gbs_aggregate": round(sum(r[1] for r in res), 1),
                          "d2h_gbs_aggregate": round(sum(r[2] for r in res), 1)}), flush=True)
