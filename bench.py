#!/usr/bin/env python
"""Benchmark of the consensus hot path (BASELINE.json metric: consensus reads/sec, SSCS+DCS+SC).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Workload (BASELINE.json configs[1]): synthetic duplex-UMI batch, 10 M read pairs
(2e7 reads) per GPU, one region (chr1:0-2.3 Mb), L = 128, family size ~ zero-truncated
Poisson(4), both strands present w.p. 0.8, cutoff 0.7, singleton correction on.
A step is one pass of the whole path (group -> SSCS vote -> DCS join -> rescue -> DCS after
rescue) over that batch.  With N > 1 every rank owns its own batch (genomic shards are
independent), so scaling is weak; the only collective is the all-reduce of
the stats counters and the family-size histogram, made once per run by libccgpu.so itself
(cc_stats_allreduce -> ncclAllReduce on one fused buffer).

sharded   extra legs on the same line: ONE dataset of BASELINE.json's stated total size - configs[2]
          (2e8-read panel) always, configs[3] (1e9-read WGS set) when a rank's share fits (8 GPUs) -
          defined band by band over the 863 hg19 cytobands, every rank generating and owning the
          bands shard.assign_regions deals it: fixed total, strong scaling.  Each leg checks itself
          inside the run (window oracle on every rank, NCCL-reduced stats and family-size histogram
          against the generator's own prediction) and the process exits 3 on a mismatch.
sweep_c5  configs[4] at N = 1: cutoff 0.5-1.0 x scorrect on/off over the panel set, window-checked.

value   reads/s with the batch resident in HBM when the timed region starts
e2e     reads/s through cc_run_host / cc_run_host_q4: pinned host buffers in, every result array
        read back to pinned host memory, copies inside the timed region.  The compact rows of
        cc_run_host_q4 (4-bit quality codes) are used when the batch has at most 15 distinct
        non-zero qualities - binned instruments; the 8-bit-row figure is reported beside it
Inputs (4.5 GB per GPU) are far larger than L2 (126 MB), so no L2 flush is needed
between iterations.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "consensus reads/sec (SSCS+DCS+SC)"
UNIT = "reads/s"
PAIRS_PER_GPU = 10_000_000
L = 128
CUTOFF = 0.7
CPU_SAMPLE_PAIRS = 8_000_000       # cpu_baseline leg: head of the GPU batch, ~10 s of one host thread
REF_SAMPLE_PAIRS = 2_000_000       # --impl reference: generated on the host, split over all host threads


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--pairs", type=int, default=PAIRS_PER_GPU, help="read pairs per GPU")
    ap.add_argument("--cpu-pairs", type=int, default=CPU_SAMPLE_PAIRS, help="read pairs of the CPU sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--e2e-contexts", type=int, default=2, help="contexts the end-to-end leg rotates over")
    ap.add_argument("--e2e-output", default="auto", choices=["auto", "rows", "planes"],
                    help="end-to-end leg: read the consensus rows back as they are, or in bit planes (cc_pack_rows) when they fit")
    ap.add_argument("--e2e-format", default="auto", choices=["auto", "q8", "q4", "planes", "classes"],
                    help="host row format of the end-to-end leg: 8-bit qualities (cc_run_host), 4-bit quality "
                         "codes (cc_run_host_q4), or both with the compact one reported when the batch allows it")
    ap.add_argument("--workload", default="c2", choices=["c2", "c3", "c4"],
                    help="c2 = BASELINE.json configs[1] (the bench line); c3 / c4: panel / WGS-like batches for profiling")
    ap.add_argument("--sharded", default="auto",
                    help="cytoband-sharded fixed-total legs added to the line: 'auto' (c3 always; c4 when a rank's share "
                         "fits, i.e. 8 GPUs), 'none', or a comma list of c3,c4")
    ap.add_argument("--c3-reads", type=int, default=200_000_000, help="total reads of the sharded panel set (configs[2])")
    ap.add_argument("--c4-reads", type=int, default=1_000_000_000, help="total reads of the sharded WGS set (configs[3])")
    ap.add_argument("--shard-steps", type=int, default=5)
    ap.add_argument("--sweep", default="auto", choices=["auto", "none"],
                    help="configs[4]: cutoff 0.5-1.0 x scorrect on/off over the panel set on one GPU (N = 1 only)")
    return ap.parse_args()


def workload_config(pairs):
    return {"workload": "synthetic duplex UMI batch, %d read pairs per GPU, single region chr1:0-2.3Mb, "
                        "L=128, family size ~ zero-truncated Poisson(4), p(duplex)=0.8, cutoff 0.7, "
                        "scorrect True (BASELINE.json configs[1])" % pairs,
            "reads_per_gpu": 2 * pairs, "read_length": L, "cutoff": CUTOFF, "scorrect": True,
            "l2_policy": "inputs (%.1f GB/GPU) exceed L2; no flush needed" % (2 * pairs * (32 + 192) / 1e9)}


# ---- clocks -------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()
        sm = sorted(int(r[1]) for r in self.rows if len(r) > 2 and r[1].isdigit())
        mx = [int(r[2]) for r in self.rows if len(r) > 2 and r[2].isdigit()]
        reasons = set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for r in self.rows:
            for nm, v in zip(names, r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def bind_to_gpu_numa_node(local):
    """Run this rank (and first-touch its pinned buffers) on the CPUs next to its GPU.

    With one rank per GPU every rank streams gigabytes per step between pinned host memory and
    its GPU; leaving the placement to chance sends half of that across the socket interconnect."""
    try:
        out = subprocess.run(["nvidia-smi", "-i", str(local), "--query-gpu=pci.bus_id", "--format=csv,noheader"],
                             capture_output=True, text=True, timeout=20).stdout.strip()
        bus = out.lower()
        if bus.startswith("00000000:"):
            bus = bus[4:]
        node = int(open("/sys/bus/pci/devices/%s/numa_node" % bus).read().strip())
        if node < 0:
            return None
        cpus = set()
        for part in open("/sys/devices/system/node/node%d/cpulist" % node).read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        allowed = os.sched_getaffinity(0) & cpus
        if allowed:
            os.sched_setaffinity(0, allowed)
            return {"numa_node": node, "cpus": len(allowed)}
    except Exception:
        pass
    return None


# ---- CPU legs (the only places bench.py may execute oracle/) ---------------------------------
def cpu_sample(pairs, seed):
    """A bounded sample of the same workload, generated on the host with the numpy generator."""
    from consensuscruncher_b200 import synth
    mean = 4.0 / (1 - 2.718281828 ** -4.0)
    mol = max(1, int(pairs / (mean * 1.8)))
    return synth.generate(seed=seed, n_molecules=mol, L=L).to_packed()


def run_oracle_threads(shards, threads):
    from concurrent.futures import ThreadPoolExecutor
    from oracle import oracle
    t0 = time.perf_counter()
    with ThreadPoolExecutor(max_workers=threads) as ex:
        list(ex.map(lambda p: oracle.run(p, CUTOFF, True), shards))
    return time.perf_counter() - t0


def reference_arm(args):
    """--impl reference: the CPU implementation of the path on the host cores.

    The reference is pure Python and cannot run on the GPU box (no pysam, /root/reference is
    absent there), so this times the oracle port (oracle/cc_oracle.c), one independent shard
    per host thread - the reference's own scale-out model is independent jobs
    (docs/source/FAQ.rst:4-7)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import oracle
    oracle.build()
    cores = os.cpu_count() or 1
    shards = None
    try:
        # the sample is a batch of the bench's own generator (made on the GPU when there is one: that is
        # input preparation, not the timed path), cut into one range of whole pairs per host thread
        import torch
        if torch.cuda.is_available():
            from consensuscruncher_b200.pack import PackedReads
            from consensuscruncher_b200.synth_device import generate_device
            whole = generate_device(seed=1234, n_pairs=min(args.pairs, args.cpu_pairs), L=L, device="cuda:0").to_packed()
            per = whole.n // 2 // cores
            shards = [PackedReads(whole.fields[2 * per * i:2 * per * (i + 1)], whole.payload[2 * per * i:2 * per * (i + 1)],
                                  whole.lpad, whole.barcodes, whole.dup_bc, whole.cigar_pairs, whole.rgs)
                      for i in range(cores)]
            torch.cuda.empty_cache()
    except Exception:
        shards = None
    if shards is None:
        per = max(1000, min(args.cpu_pairs, REF_SAMPLE_PAIRS) // cores)
        shards = [cpu_sample(per, 9000 + i) for i in range(cores)]
    reads = sum(p.n for p in shards)
    for _ in range(max(1, min(args.warmup, 1))):
        run_oracle_threads(shards[:cores], cores)
    t = 0.0
    for _ in range(args.steps):
        t += run_oracle_threads(shards, cores)
    value = reads * args.steps / t
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
            "data": "synthetic", "config": workload_config(args.pairs),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": "%d reads of the same distribution per step, %d shards of %d pairs, "
                                       "one oracle thread per shard" % (reads, cores, per)},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


# ---- our arm ---------------------------------------------------------------------------------
def vote_algorithmic_bytes(res_sizes, n_sscs, stride_l):
    """N*(1.5L+4) + F2*(1.5L+16), N = reads in families of size >= 2."""
    n_big = int(res_sizes[res_sizes >= 2].sum())
    return n_big * (1.5 * stride_l + 4) + n_sscs * (1.5 * stride_l + 16), n_big


def device_q4(batch):
    """(payload4 u8[n][stride4], table u8[16]) on the device for `cc_run_host_q4`, or None when the
    batch has more than 15 distinct non-zero quality values (pack.to_q4 is the host statement)."""
    import torch
    lpad, half = batch.lpad, batch.lpad // 2
    pay = batch.payload
    n = pay.shape[0]
    rows = 1 << 20
    hist = torch.zeros(256, dtype=torch.int64, device=pay.device)
    for a in range(0, n, rows):
        hist += torch.bincount(pay[a:a + rows, half:half + lpad].reshape(-1).to(torch.int32), minlength=256)
    values = [0] + [int(v) for v in torch.nonzero(hist).reshape(-1).tolist() if v != 0]
    if len(values) > 16:
        return None
    table = torch.zeros(16, dtype=torch.uint8, device=pay.device)
    table[:len(values)] = torch.tensor(values, dtype=torch.uint8, device=pay.device)
    code_of = torch.zeros(256, dtype=torch.uint8, device=pay.device)
    code_of[torch.tensor(values, dtype=torch.int64, device=pay.device)] = torch.arange(len(values), dtype=torch.uint8,
                                                                                      device=pay.device)
    stride4 = (lpad + 15) // 16 * 16
    out = torch.zeros((n, stride4), dtype=torch.uint8, device=pay.device)
    for a in range(0, n, rows):
        codes = code_of[pay[a:a + rows, half:half + lpad].to(torch.int64)]
        out[a:a + rows, :half] = pay[a:a + rows, :half]
        out[a:a + rows, half:lpad] = codes[:, 0::2] | (codes[:, 1::2] << 4)
    return out, table


# ---- sharded legs -----------------------------------------------------------------------------
def _all_ranks_agree(flag_bad, world, dev):
    """True if any rank reports a failure (so that every rank leaves a leg the same way)."""
    if world == 1:
        return bool(flag_bad)
    import torch
    import torch.distributed as dist
    t = torch.tensor([1 if flag_bad else 0], dtype=torch.int32, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return bool(int(t.item()))


def sharded_leg(args, wl, eng, rank, world, local, dev, stream, barrier):
    """One dataset of BASELINE.json's stated total size (configs[2] / configs[3]), defined band by band
    over the 863 hg19 cytobands; every rank generates and owns the bands shard.assign_regions gives it,
    runs the whole path over them as one batch, and the stats are summed by cc_stats_allreduce.
    Fixed total => strong scaling.  Checked inside the run, outside the timed region:
      * size-independent properties of every rank's result,
      * bit-exact comparison with the CPU oracle on a genomic window of every rank's shard,
      * the NCCL-reduced counters and family-size histogram equal what the generator's molecule
        model predicts for the whole dataset."""
    import numpy as np
    import torch
    import torch.distributed as dist
    from consensuscruncher_b200 import workloads
    total_reads = args.c3_reads if wl == "c3" else args.c4_reads
    pl = workloads.plan(wl, total_reads // 2, seed=1234)
    t_gen = time.perf_counter()
    batch, info = workloads.generate_shard(pl, rank, world, str(dev))
    torch.cuda.synchronize(dev)
    t_gen = time.perf_counter() - t_gen
    n = batch.n
    eng.set_params(cutoff=CUTOFF, scorrect=True)

    def step():
        eng.load_device(batch.fields.data_ptr(), batch.payload.data_ptr(), n, batch.lpad, batch.stride,
                        batch.dup_bc.data_ptr(), batch.n_bc)
        eng.run()

    out = {"workload": pl.describe, "total_reads": 2 * pl.total_pairs, "n_gpus": world, "scaling": "strong",
           "shard_map": "hg19 cytobands (863) -> ranks by read count (shard.assign_regions)", "steps": args.shard_steps}
    failed = None
    try:
        for _ in range(3):
            step()
        torch.cuda.synchronize(dev)
    except Exception as exc:                     # e.g. a shard that does not fit: no rank enters the collective then
        failed = "run: %r" % (exc,)
    if _all_ranks_agree(failed is not None, world, dev):
        # a shard that cannot be run here (memory) is reported, not counted as a failed check
        return dict(out, skipped=failed or "another rank could not run its shard")
    try:
        eng.stats_allreduce()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(args.shard_steps):
            step()
        eng.stats_allreduce()
        e1.record(stream)
        barrier()
        ms = e0.elapsed_time(e1)
        reduced_stats = eng.fetch("stats").astype(np.int64)
        reduced_hist = eng.fetch("fam_hist").astype(np.int64)
    except Exception as exc:                     # a rank that cannot run its shard fails the leg for everybody
        failed = "run: %r" % (exc,)
        ms = 0.0
    if _all_ranks_agree(failed is not None, world, dev):
        return dict(out, failed=failed or "another rank failed")
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    per = torch.tensor([n, info["n_bands"]], dtype=torch.int64, device=dev)
    gathered = [torch.zeros_like(per) for _ in range(world)]
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_gather(gathered, per)
    else:
        gathered = [per]
    ms_max = float(t.item())

    # ---- verification (not timed) ----------------------------------------------------------------
    parity = {}
    try:
        from oracle import window as owin            # the checker; never on the measured path
        owin.oracle.build()
        step()                                       # a fresh, un-reduced run of this rank's shard
        counts, _ = owin.check_shard_properties(eng, n)
        win = workloads.pick_window(batch)
        parity["window"] = owin.check_window(eng, batch, win, CUTOFF, True) if win is not None else None
        parity["largest_family"] = int(eng.fetch("fam_size").max()) if n else 0
        # the generator's prediction for the WHOLE dataset = sum of every rank's shard prediction
        pred = torch.from_numpy(np.concatenate([info["stats"], info["hist"]])).to(dev)
        if world > 1:
            dist.all_reduce(pred)
        pred = pred.cpu().numpy()
        if not (reduced_stats[:9] == pred[:9]).all():
            raise AssertionError("reduced stats %s != generator's %s" % (reduced_stats[:9].tolist(), pred[:9].tolist()))
        if not (reduced_hist == pred[16:]).all():
            raise AssertionError("reduced family-size histogram differs from the generator's")
        parity["reduced_stats_equal_generator"] = True
        parity["stats"] = {nm: int(reduced_stats[i]) for i, nm in
                           enumerate(("sscs", "singletons", "dcs", "sscs_singletons", "sc_by_sscs", "sc_by_singleton",
                                      "uncorrected", "dcs_sc", "sscs_sc_singletons"))}
    except Exception as exc:
        failed = "verification on rank %d: %s" % (rank, exc)
    if _all_ranks_agree(failed is not None, world, dev):
        out["failed"] = failed or "verification failed on another rank"
    out.update({"ms_per_step": ms_max / args.shard_steps,
                "value": 2 * pl.total_pairs * args.shard_steps / (ms_max * 1e-3), "unit": UNIT,
                "reads_per_rank": [int(g[0]) for g in gathered], "bands_per_rank": [int(g[1]) for g in gathered],
                "generate_s": round(t_gen, 2), "parity_rank0": parity})
    batch = None
    torch.cuda.empty_cache()
    return out


def sweep_leg(args, eng, dev, stream):
    """BASELINE.json configs[4]: cutoff 0.5 - 1.0 with and without singleton correction over the panel
    set, one GPU; every case is checked against the oracle on a window of the batch."""
    import torch
    from consensuscruncher_b200 import workloads
    pl = workloads.plan("c3", args.c3_reads // 2, seed=1234)
    batch, _ = workloads.generate_shard(pl, 0, 1, str(dev), predict=False)
    n = batch.n
    win = workloads.pick_window(batch, max_reads=400_000)
    from oracle import window as owin
    owin.oracle.build()
    cases = []
    for scorrect in (True, False):
        for cutoff in (0.5, 0.6, 0.7, 0.8, 0.9, 1.0):
            eng.set_params(cutoff=cutoff, scorrect=scorrect)
            case = {"cutoff": cutoff, "scorrect": scorrect}
            try:
                ms = []
                for _ in range(3):
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record(stream)
                    eng.load_device(batch.fields.data_ptr(), batch.payload.data_ptr(), n, batch.lpad, batch.stride,
                                    batch.dup_bc.data_ptr(), batch.n_bc)
                    eng.run()
                    e1.record(stream)
                    torch.cuda.synchronize(dev)
                    ms.append(e0.elapsed_time(e1))
                c = eng.counts()
                case.update({"ms_per_step": min(ms), "value": n / (min(ms) * 1e-3), "sscs": int(c["sscs"]),
                             "dcs": int(c["dcs"]), "dcs_sc": int(c["dcssc"])})
                w = owin.check_window(eng, batch, win, cutoff, scorrect)
                case["window"] = {"reads": w["reads"], "families": w["families"], "rows_compared": w["rows_compared"]}
            except Exception as exc:
                case["failed"] = "cutoff %.1f scorrect %s: %s" % (cutoff, scorrect, exc)
            cases.append(case)
    eng.set_params(cutoff=CUTOFF, scorrect=True)
    batch = None
    torch.cuda.empty_cache()
    return {"workload": pl.describe + "; cutoff 0.5-1.0 x scorrect on/off (BASELINE.json configs[4])",
            "reads": n, "cases": cases}


def device_compact(batch):
    """Input of `cc_run_host_compact` built on the device (pack.to_compact is the host statement):
    (pairs i32[n/2][10], planes u8[n][stride_c], qbits, qual_table u8[16], base_table u8[16]) or None."""
    import torch
    f = batch.fields
    a, b = f[0::2], f[1::2]
    mirror = (a[:, 0] == b[:, 0]) & (a[:, 1] == b[:, 1]) & (a[:, 2] == b[:, 4]) & (a[:, 3] == b[:, 5]) & \
             (a[:, 4] == b[:, 2]) & (a[:, 5] == b[:, 3])
    if not bool(mirror.all()):
        return None
    pairs = torch.empty((a.shape[0], 10), dtype=torch.int32, device=f.device)
    pairs[:, 0:6] = a[:, 0:6]
    pairs[:, 6], pairs[:, 7] = a[:, 6], b[:, 6]
    ma, mb = a[:, 7].to(torch.int64) & 0xFFFFFFFF, b[:, 7].to(torch.int64) & 0xFFFFFFFF
    flags = (ma & 0xFFFF) | ((mb & 0xFFFF) << 16)
    meta = ((ma >> 16) & 0xFF) | (((mb >> 16) & 0xFF) << 8) | ((ma >> 24) << 16) | ((mb >> 24) << 24)
    as_i32 = lambda x: torch.where(x >= 2 ** 31, x - 2 ** 32, x).to(torch.int32)
    pairs[:, 8], pairs[:, 9] = as_i32(flags), as_i32(meta)

    lpad, half = batch.lpad, batch.lpad // 2
    pay = batch.payload
    n = pay.shape[0]
    rows = 1 << 19
    two_bit = torch.full((16,), 255, dtype=torch.int64, device=pay.device)
    two_bit[torch.tensor([1, 2, 4, 8], device=pay.device)] = torch.arange(4, device=pay.device)

    def split(chunk):                          # -> 2-bit base (255: none), literal base (255: A/C/G/T), quality
        seq = chunk[:, :half].to(torch.int64)
        base = torch.empty((chunk.shape[0], lpad), dtype=torch.int64, device=pay.device)
        base[:, 0::2], base[:, 1::2] = seq & 15, seq >> 4
        b2 = two_bit[base]
        return b2, torch.where(b2 == 255, base, torch.full_like(base, 255)), chunk[:, half:half + lpad].to(torch.int64)

    hist = torch.zeros(65536, dtype=torch.int64, device=pay.device)
    for s0 in range(0, n, rows):
        b2, lit, q = split(pay[s0:s0 + rows])
        hist += torch.bincount(((q << 8) | lit).reshape(-1), minlength=65536)
    present = torch.nonzero(hist).reshape(-1)
    if present.numel() > 16:
        return None
    qbits = max(1, int(math.ceil(math.log2(max(2, present.numel())))))
    qual_table = torch.zeros(16, dtype=torch.uint8)
    base_table = torch.full((16,), 255, dtype=torch.uint8)
    qual_table[:present.numel()] = (present >> 8).to(torch.uint8).cpu()
    base_table[:present.numel()] = (present & 255).to(torch.uint8).cpu()
    code_of = torch.zeros(65536, dtype=torch.int64, device=pay.device)
    code_of[present] = torch.arange(present.numel(), device=pay.device)
    stride_c = (lpad // 4 + lpad * qbits // 8 + 15) // 16 * 16
    planes = torch.zeros((n, stride_c), dtype=torch.uint8, device=pay.device)
    w8 = (1 << torch.arange(8, device=pay.device)).to(torch.int64)
    for s0 in range(0, n, rows):
        b2, lit, q = split(pay[s0:s0 + rows])
        codes = code_of[(q << 8) | lit]
        b2 = torch.where(b2 == 255, torch.zeros_like(b2), b2)
        planes[s0:s0 + rows, :lpad // 4] = (b2[:, 0::4] | (b2[:, 1::4] << 2) | (b2[:, 2::4] << 4) | (b2[:, 3::4] << 6)).to(torch.uint8)
        bits = ((codes.unsqueeze(2) >> torch.arange(qbits, device=pay.device)) & 1).reshape(codes.shape[0], lpad * qbits // 8, 8)
        planes[s0:s0 + rows, lpad // 4:lpad // 4 + lpad * qbits // 8] = (bits * w8).sum(2).to(torch.uint8)
    return pairs, planes, qbits, qual_table, base_table


def device_classes(pairs):
    """Input of `cc_run_host_classes` from device_compact's pair records (pack.to_classes is the host statement):
    (pairs16 i32[n/2][4], classes i32[C][8]) or None."""
    import torch
    t1, t2 = pairs[:, 6].to(torch.int64), pairs[:, 7].to(torch.int64)
    if not bool(((t2 == -t1) & (t1.abs() <= 32767)).all()):
        return None
    keys = pairs[:, [1, 2, 4, 8, 9]].contiguous()           # cig, tid, mtid, flags, meta
    classes, inverse = torch.unique(keys, dim=0, return_inverse=True)
    if classes.shape[0] > 65536:
        return None
    out = torch.empty((pairs.shape[0], 4), dtype=torch.int32, device=pairs.device)
    out[:, 0], out[:, 1], out[:, 2] = pairs[:, 0], pairs[:, 3], pairs[:, 5]
    w = (t1 & 0xFFFF) | (inverse.to(torch.int64) << 16)
    out[:, 3] = torch.where(w >= 2 ** 31, w - 2 ** 32, w).to(torch.int32)
    table = torch.zeros((classes.shape[0], 8), dtype=torch.int32, device=pairs.device)
    table[:, 0:5] = classes
    return out, table


def main():
    args = parse()
    if args.impl == "reference":
        return reference_arm(args)

    import numpy as np
    import torch
    import torch.distributed as dist
    from consensuscruncher_b200 import build
    from consensuscruncher_b200.engine import GpuConsensus
    from consensuscruncher_b200.synth_device import generate_device

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d" % (args.gpus, world))
    if args.warmup < 3:
        args.warmup = 3
    numa = bind_to_gpu_numa_node(local)
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    build.build_lib()

    gen_kw = {"c2": {},
              "c3": dict(dist="lognormal", fam_mean=12.0, region=(0, 200_000), umi_len=4, p_duplex=0.9),
              "c4": dict(dist="wgs", p_duplex=0.3, region=(0, 240_000_000))}[args.workload]
    batch = generate_device(seed=1234 + rank, n_pairs=args.pairs, L=L, device=str(dev), **gen_kw)
    n_reads = batch.n
    eng = GpuConsensus(local, cutoff=CUTOFF, scorrect=True)
    stream = torch.cuda.current_stream(dev)
    eng.set_stream(stream.cuda_stream)

    def exchange(payload):
        """Rank 0's communicator id to every rank, over the launcher's process group (host plumbing)."""
        box = [payload]
        dist.broadcast_object_list(box, src=0)
        return box[0]

    # the path's one collective - sum of the stage counters and the family-size histogram - is made by
    # libccgpu.so itself (cc_stats_allreduce: one ncclAllReduce on the context's stream), once per run
    from consensuscruncher_b200 import shard
    shard.connect(eng, rank, world, exchange)

    def step():
        eng.load_device(batch.fields.data_ptr(), batch.payload.data_ptr(), n_reads, batch.lpad, batch.stride,
                        batch.dup_bc.data_ptr(), batch.n_bc)
        eng.run()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    for _ in range(args.warmup):
        step()
    counts = eng.counts()
    sizes = eng.fetch("fam_size")
    stats_once = eng.fetch("stats")
    eng.stats_allreduce()                      # warm the collective up too (first call builds NCCL's channels)

    # ---- device-resident timing -------------------------------------------------------
    clocks = ClockSampler(local)
    eng.profile(True)
    l0 = eng.kernel_launches()
    barrier()
    if rank == 0:
        clocks.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        step()
    eng.stats_allreduce()                      # once per run, as the reference writes its stats once per stage
    e1.record(stream)
    barrier()
    ms = e0.elapsed_time(e1)
    launches = eng.kernel_launches() - l0
    phases = eng.phase_ms()
    eng.profile(False)
    clk = clocks.stop() if rank == 0 else None
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t.item())
    value = n_reads * world * args.steps / (ms_max * 1e-3)

    # ---- end to end: pinned host in, results back to pinned host ------------------------
    e2e = None
    if not args.no_e2e:
        h_fields = torch.empty(batch.fields.shape, dtype=batch.fields.dtype, pin_memory=True)
        h_payload = torch.empty(batch.payload.shape, dtype=torch.uint8, pin_memory=True)
        h_dup = torch.empty(batch.dup_bc.shape, dtype=batch.dup_bc.dtype, pin_memory=True)
        h_fields.copy_(batch.fields); h_payload.copy_(batch.payload); h_dup.copy_(batch.dup_bc)
        torch.cuda.synchronize(dev)
        # every result array a caller needs to write the output files; the DCS rows / fields are the
        # first n_dcs entries of the DCS-after-rescue arrays, so they are not fetched twice
        out_names = ("fam_size", "fam_tmpl", "fam_mapq", "fam_flag", "fam_tlen", "fam_rg", "partner",
                     "sscs_row", "corr_row", "dcs_row", "dcssc_row", "sscs_rows", "corr_rows",
                     "dcssc_rows", "dcssc_mapq", "dcssc_flag", "dcssc_tlen", "dcssc_rg", "stats", "fam_hist")
        from consensuscruncher_b200.engine import ARRAYS
        from concurrent.futures import ThreadPoolExecutor
        # Two contexts alternate (double buffering): while one batch's results travel back to the
        # host, the next batch is already being uploaded, so both PCIe directions stay busy.
        # Each context works on its own stream; each step is still one cc_run_host call plus
        # the read-back of every result array.
        n_ctx = max(1, args.e2e_contexts)
        engines = [eng] + [GpuConsensus(local, cutoff=CUTOFF, scorrect=True) for _ in range(n_ctx - 1)]
        eng.set_stream(0)
        h_outs = [{} for _ in range(n_ctx)]

        # Compact rows (cc_run_host_q4): 4-bit quality codes when the batch has at most 15 distinct
        # non-zero qualities, as binned instruments produce; built here, outside every timed region.
        q4 = device_q4(batch) if args.e2e_format in ("auto", "q4") else None
        if args.e2e_format == "q4" and q4 is None:
            raise SystemExit("--e2e-format q4: the batch has more than 15 distinct non-zero qualities")
        h_q4 = h_table = None
        if q4 is not None:
            h_q4 = torch.empty(q4[0].shape, dtype=torch.uint8, pin_memory=True)
            h_q4.copy_(q4[0])
            h_table = q4[1].cpu().contiguous()
            torch.cuda.synchronize(dev)
            q4 = None
        # Most compact (cc_run_host_compact): 40-byte pair records + 2-bit bases + few-bit quality codes
        planes = device_compact(batch) if args.e2e_format in ("auto", "planes", "classes") else None
        if args.e2e_format in ("planes", "classes") and planes is None:
            raise SystemExit("--e2e-format %s: the batch does not fit the compact format" % args.e2e_format)
        h_pairs = h_planes = h_pairs16 = h_classes = None
        if planes is not None:
            h_pairs = torch.empty(planes[0].shape, dtype=torch.int32, pin_memory=True)
            h_planes = torch.empty(planes[1].shape, dtype=torch.uint8, pin_memory=True)
            h_pairs.copy_(planes[0]); h_planes.copy_(planes[1])
            p_qbits, p_qual, p_base = planes[2], planes[3].contiguous(), planes[4].contiguous()
            # Shorter still (cc_run_host_classes): 16-byte pair records + a table of what many pairs share
            short = None
            if args.e2e_format in ("auto", "classes"):
                try:
                    short = device_classes(planes[0])
                except RuntimeError:            # (no room for torch.unique's scratch next to the batch: the 40-byte records then)
                    short = None
                    torch.cuda.empty_cache()
            if args.e2e_format == "classes" and short is None:
                raise SystemExit("--e2e-format classes: the batch does not fit the short pair records")
            if short is not None:
                h_pairs16 = torch.empty(short[0].shape, dtype=torch.int32, pin_memory=True)
                h_classes = torch.empty(short[1].shape, dtype=torch.int32, pin_memory=True)
                h_pairs16.copy_(short[0]); h_classes.copy_(short[1])
            torch.cuda.synchronize(dev)
            planes = short = None
        fmt = {"name": "q8", "out": "rows"}
        import ctypes
        plane_of = {"sscs_rows": "sscs_planes", "corr_rows": "corr_planes", "dcssc_rows": "dcssc_planes"}
        pk_qbits, pk_stride = ctypes.c_uint32(0), ctypes.c_uint32(0)
        pk_qual, pk_base = (ctypes.c_uint8 * 16)(), (ctypes.c_uint8 * 16)()

        def e2e_step(which, step_no):
            e = engines[which]
            if fmt["name"] == "classes":
                e._check(e.lib.cc_run_host_classes(e.ctx, h_pairs16.data_ptr(), h_classes.data_ptr(), h_classes.shape[0],
                                                   h_planes.data_ptr(), n_reads, batch.lpad, h_planes.shape[1], p_qbits,
                                                   p_qual.data_ptr(), p_base.data_ptr(), h_dup.data_ptr(), batch.n_bc))
            elif fmt["name"] == "planes":
                e._check(e.lib.cc_run_host_compact(e.ctx, h_pairs.data_ptr(), h_planes.data_ptr(), n_reads, batch.lpad,
                                                   h_planes.shape[1], p_qbits, p_qual.data_ptr(), p_base.data_ptr(),
                                                   h_dup.data_ptr(), batch.n_bc))
            elif fmt["name"] == "q4":
                e._check(e.lib.cc_run_host_q4(e.ctx, h_fields.data_ptr(), h_q4.data_ptr(), n_reads, batch.lpad,
                                              h_q4.shape[1], h_table.data_ptr(), h_dup.data_ptr(), batch.n_bc))
            else:
                e._check(e.lib.cc_run_host(e.ctx, h_fields.data_ptr(), h_payload.data_ptr(), n_reads, batch.lpad,
                                           batch.stride, h_dup.data_ptr(), batch.n_bc))
            total = 0
            whats, dsts, sizes = [], [], []
            in_planes = False
            if fmt["out"] == "planes":          # rows once more in bit planes on the device: half the bytes to read back
                e._check(e.lib.cc_pack_rows(e.ctx, ctypes.byref(pk_qbits), ctypes.byref(pk_stride), pk_qual, pk_base))
                in_planes = pk_qbits.value > 0
            for nm in out_names:
                what = ARRAYS[plane_of[nm] if in_planes and nm in plane_of else nm][0]
                nbytes = e.lib.cc_array_bytes(e.ctx, what)
                if nbytes == 0:
                    continue
                buf = h_outs[which].get(nm)
                if buf is None or buf.numel() < nbytes:
                    buf = h_outs[which][nm] = torch.empty(int(nbytes * 1.05) + 64, dtype=torch.uint8, pin_memory=True)
                whats.append(what); dsts.append(buf.data_ptr()); sizes.append(nbytes)
                total += nbytes
            k = len(whats)                     # every copy is issued, then one wait (cc_fetch_many)
            e._check(e.lib.cc_fetch_many(e.ctx, k, (ctypes.c_int * k)(*whats), (ctypes.c_void_p * k)(*dsts),
                                         (ctypes.c_uint64 * k)(*sizes)))
            return total

        def worker(which, first, count):
            torch.cuda.set_device(local)
            out = 0
            for k in range(count):
                out = e2e_step(which, first + n_ctx * k)
            return out

        def run_steps(count):                  # `count` steps in total, alternating between the contexts
            with ThreadPoolExecutor(max_workers=n_ctx) as ex:
                futs = [ex.submit(worker, w, w, (count + n_ctx - 1 - w) // n_ctx) for w in range(n_ctx)]
                return max(f.result() for f in futs)

        e2e_steps = max(4, min(args.steps, 8)) // n_ctx * n_ctx or n_ctx

        def measure(name, out="rows"):
            fmt["name"], fmt["out"] = name, out
            run_steps(2 * n_ctx)               # warm-up: allocations, pinning of the result buffers
            barrier()
            t0 = time.perf_counter()
            d2h = run_steps(e2e_steps)
            eng.stats_allreduce()              # the run's one collective (context 0 holds the communicator)
            torch.cuda.synchronize(dev)
            wall = (time.perf_counter() - t0) * 1e3
            barrier()
            t = torch.tensor([wall], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            if name == "classes":
                h2d = (h_pairs16.numel() + h_classes.numel()) * 4 + h_planes.numel() + h_dup.numel() * h_dup.element_size()
            elif name == "planes":
                h2d = h_pairs.numel() * 4 + h_planes.numel() + h_dup.numel() * h_dup.element_size()
            else:
                payload_bytes = h_q4.numel() if name == "q4" else h_payload.numel()
                h2d = h_fields.numel() * h_fields.element_size() + payload_bytes + h_dup.numel() * h_dup.element_size()
            return {"value": n_reads * world * e2e_steps / (float(t.item()) * 1e-3), "unit": UNIT,
                    "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "steps": e2e_steps, "ms_per_step": float(t.item()) / e2e_steps}

        legs = {}
        if args.e2e_format in ("auto", "q8"):
            legs["q8"] = measure("q8")
        if h_q4 is not None:
            legs["q4"] = measure("q4")
        if h_planes is not None and args.e2e_format != "classes":
            legs["planes"] = measure("planes")
        if h_pairs16 is not None:
            legs["classes"] = measure("classes")
        best = "classes" if "classes" in legs else "planes" if "planes" in legs else "q4" if "q4" in legs else "q8"
        rows_back = None
        if best in ("classes", "planes") and args.e2e_output in ("auto", "planes"):
            packed_leg = measure(best, "planes")
            if pk_qbits.value > 0:              # (more than 16 quality / base combinations in the results: rows as they are)
                rows_back = {k: legs[best][k] for k in ("value", "ms_per_step", "d2h_bytes_per_step")}
                legs[best] = packed_leg
        e2e = dict(legs[best])
        e2e["output_format"] = ("cc_pack_rows: consensus rows read back as 2-bit bases + %d-bit (quality, literal base) codes, "
                                "%d bytes per row instead of %d" % (pk_qbits.value, pk_stride.value, batch.stride)
                                if rows_back else "rows as the kernels write them (%d bytes each)" % batch.stride)
        if rows_back:
            e2e["plain_rows_back"] = rows_back
        e2e["input_format"] = {
            "classes": "cc_run_host_classes: 16-byte pair records + %d classes + rows of 2-bit bases and %d-bit quality "
                       "codes (%d bytes per read)" % (h_classes.shape[0] if h_classes is not None else 0,
                                                      p_qbits if h_planes is not None else 0,
                                                      8 + (h_planes.shape[1] if h_planes is not None else 0)),
            "planes": "cc_run_host_compact: 40-byte pair records + rows of 2-bit bases and %d-bit quality codes "
                      "(%d bytes per read)" % (p_qbits if h_planes is not None else 0,
                                               20 + (h_planes.shape[1] if h_planes is not None else 0)),
            "q4": "cc_run_host_q4: rows with 4-bit quality codes",
            "q8": "cc_run_host: rows with 8-bit qualities"}[best]
        if best != "q8" and "q8" in legs:
            e2e["q8_rows"] = {k: legs["q8"][k] for k in ("value", "ms_per_step", "h2d_bytes_per_step")}
        if best in ("planes", "classes") and "q4" in legs:
            e2e["q4_rows"] = {k: legs["q4"][k] for k in ("value", "ms_per_step", "h2d_bytes_per_step")}
        if best == "classes" and "planes" in legs:
            e2e["pair40_rows"] = {k: legs["planes"][k] for k in ("value", "ms_per_step", "h2d_bytes_per_step")}
        e2e["pipeline"] = ("%d contexts take turns: read-back of batch k overlaps upload of batch k+1; "
                           "host wall clock" % n_ctx)
        e2e["numa"] = numa
        for extra in engines[1:]:
            extra.close()

    # ---- cytoband-sharded legs: ONE dataset of fixed total size, dealt to the ranks by hg19 band -------
    sharded, sweep, verify_fail = {}, None, []
    which = []
    if args.sharded == "auto":
        which = ["c3"] + (["c4"] if args.c4_reads // world <= 130_000_000 else [])
    elif args.sharded != "none":
        which = [w for w in args.sharded.split(",") if w]
    cpu_packed = None
    if rank == 0 and not args.no_cpu_baseline:
        # the CPU sample is the head of the batch the GPU just processed (whole pairs, so every kept pair still
        # founds / joins its two families together), downloaded - not generated again on the host
        from consensuscruncher_b200.synth_device import DeviceBatch
        n_cpu = min(batch.n, 2 * args.cpu_pairs)
        cpu_packed = DeviceBatch(batch.fields[:n_cpu], batch.payload[:n_cpu], batch.dup_bc, batch.lpad, batch.stride,
                                 batch.L, batch.umi_len).to_packed()
    if which or (args.sweep == "auto" and world == 1):
        # release the replica batch and the end-to-end leg's pinned buffers first
        batch = None
        h_fields = h_payload = h_dup = h_q4 = h_outs = h_pairs = h_planes = h_pairs16 = h_classes = None
        import gc
        gc.collect()
        torch.cuda.empty_cache()
        eng.set_stream(stream.cuda_stream)
    for wl in which:
        res = sharded_leg(args, wl, eng, rank, world, local, dev, stream, barrier)
        if rank == 0:
            sharded[wl] = res
        if res.get("failed"):
            verify_fail.append("%s: %s" % (wl, res["failed"]))
    if args.sweep == "auto" and world == 1:
        sweep = sweep_leg(args, eng, dev, stream)
        verify_fail += ["c5 %s" % x["failed"] for x in sweep["cases"] if x.get("failed")]

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        if verify_fail:
            sys.exit(3)
        return

    # ---- roofline of the dominant kernel (SSCS vote) ----------------------------------------
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    vote_ms, vote_runs = phases["vote"]
    alg_bytes, n_big = vote_algorithmic_bytes(sizes, counts["sscs"], L)
    achieved = alg_bytes / (vote_ms / max(1, vote_runs) * 1e-3) / 1e9 if vote_ms > 0 else 0.0
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "vote_traffic.json")
    if os.path.exists(tpath):
        try:
            traffic = json.load(open(tpath)).get("dram_bytes_per_launch")
        except Exception:
            traffic = None
    roofline = {"bound": "hbm", "kernel": "vote_small_kernel (SSCS vote: consensus_maker over every family of 2..127 reads)",
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "traffic_source": "profiles/vote_traffic.json (ncu --set full, dram__bytes_read+write per launch)",
                "peak_source": peak_src, "algorithmic_bytes_per_launch": alg_bytes,
                "ms_per_launch": vote_ms / max(1, vote_runs),
                "phase_ms_per_step": {k: v[0] / max(1, v[1]) for k, v in phases.items() if v[1]}}

    cpu = None
    if not args.no_cpu_baseline:
        from oracle import oracle
        oracle.build()
        sample = cpu_packed
        t0 = time.perf_counter()
        oracle.run(sample, CUTOFF, True)
        dt = time.perf_counter() - t0
        cpu = {"value": sample.n / dt, "unit": UNIT, "cores": 1, "kind": "port",
               "sample": "the first %d reads (%d pairs) of the GPU's batch, oracle/cc_oracle.c single thread, %.1f s"
                         % (sample.n, sample.n // 2, dt),
               "host_cores_available": os.cpu_count()}

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_max / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": dict(workload_config(args.pairs), **({} if args.workload == "c2" else {"workload": "profiling batch " + args.workload})), "clocks": clk, "e2e": e2e, "gpu_launches": int(launches),
            "roofline": roofline, "cpu_baseline": cpu,
            "counts": {k: int(v) for k, v in counts.items()},
            "stats": {"sscs": int(stats_once[0]), "singletons": int(stats_once[1]), "dcs": int(stats_once[2]),
                      "dcs_sc": int(stats_once[7])},
            "collective": {"call": "cc_stats_allreduce (libccgpu.so -> ncclAllReduce, int64 x %d, once per run)" % (16 + 65536),
                           "comm": eng.comm_info()},
            "sharded": sharded or None, "sweep_c5": sweep,
            "definitions": {
                "value": "device-resident: the batch is in HBM when the timed region starts; whole path, CUDA events, max over ranks",
                "e2e": "the same path through the C ABI from pinned host buffers, every result array read back, copies timed; "
                       "this is the number to set against the CPU arm (bench.py --impl reference)",
                "cpu_baseline": "oracle/cc_oracle.c, the builder's C restatement of the reference's algorithm (kind 'port'), pinned "
                                "against outputs of the unmodified reference; the reference's own Python runs ~227x slower per "
                                "core (BASELINE.md section 2), so ratios against this arm understate the ratio against the reference",
                "not_included": "BAM decode / pack and record assembly / BAM writing on the host (the consensus command; DESIGN.md section 8)"}}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    if verify_fail:
        sys.stderr.write("bench.py: verification FAILED: %s\n" % "; ".join(verify_fail))
        sys.exit(3)


if __name__ == "__main__":
    main()
