"""The synthetic configurations of BASELINE.json as cytoband-sharded datasets.

`north_star`: "Work is partitioned across the 8 GPUs of one box by the reference's own cytoband bed
regions, with each GPU owning independent genomic chunks."  The reference walks the 863 bands of
its bed file one after the other (SSCS_maker.py:229-257, DCS_maker.py:185-211,
singleton_correction.py:175-213); here a dataset is DEFINED band by band - band b's reads come from
a generator seeded with (seed, b) only - so every rank can make exactly the bands it owns, no rank
ever holds the whole set, and the union over any number of ranks is the same dataset:

  c2  BASELINE.json configs[1]: one region (chr1 p36.33), 10 M pairs, family size ~ ztPoisson(4)
  c3  configs[2]: targeted panel, 2e8 reads on ~100 targets of 2 kb (~200 kb), family size
      ~ lognormal(mean 12) with a tail to 10^3 - 10^4, 90 % duplex
  c4  configs[3]: whole genome, 1e9 reads spread over all 863 hg19 bands in proportion to band
      length, ~90 % singletons, duplex partner present for 30 % of the molecules

torch is used for data creation only (synth_device.py); nothing here is on the measured path.
"""
from __future__ import annotations

import os

import numpy as np

from . import keys, shard

HERE = os.path.dirname(os.path.abspath(__file__))
L = 128


def cytobands(genome="hg19"):
    """[(name, tid, start, end)] in bed order; tid = index of the contig's first appearance
    (the bed's own contig order stands in for a BAM header here)."""
    path = os.path.join(HERE, "data", "%s_cytoBand.txt" % genome)
    contigs, out = [], []
    for key, (a, b) in keys.bed_separator(path).items():
        chrom = key.split("_", 1)[0]
        if chrom not in contigs:
            contigs.append(chrom)
        out.append((key, contigs.index(chrom), a, b))
    return out


class Plan:
    """What every band holds: pairs[b] read pairs inside region[b] = (tid, lo, hi), made with gen_kw."""

    def __init__(self, name, bands, pairs, regions, gen_kw, seed, describe):
        self.name, self.bands, self.pairs, self.regions = name, bands, np.asarray(pairs, np.int64), regions
        self.gen_kw, self.seed, self.describe = gen_kw, int(seed), describe

    @property
    def total_pairs(self):
        return int(self.pairs.sum())

    def owners(self, world):
        """Rank of every band: the reference's region list dealt out by read count (shard.assign_regions)."""
        return np.asarray(shard.assign_regions(self.pairs.tolist(), world), np.int64)


def plan(workload, total_pairs, seed=1234, genome="hg19"):
    bands = cytobands(genome)
    nb = len(bands)
    margin = 2 * L + 400                                   # generate_device keeps both mates inside the region
    if workload == "c2":
        pairs = np.zeros(nb, np.int64)
        b = next(i for i, x in enumerate(bands) if x[0] == "chr1_p36.33")
        pairs[b] = total_pairs
        regions = [(t, a, e) for _, t, a, e in bands]
        return Plan("c2", bands, pairs, regions, {}, seed,
                    "synthetic duplex UMI batch, %d read pairs, single region chr1:0-2.3Mb (band p36.33), L=128, "
                    "family size ~ zero-truncated Poisson(4), p(duplex)=0.8 (BASELINE.json configs[1])" % total_pairs)
    if workload == "c3":
        # ~100 targets of 2 kb: every 8th band that is long enough carries one, 1 Mb into the band
        cand = [i for i, (_, _, a, e) in enumerate(bands) if e - a > 1_200_000][::8][:100]
        pairs = np.zeros(nb, np.int64)
        share, extra = divmod(total_pairs, len(cand))
        for k, i in enumerate(cand):
            pairs[i] = share + (1 if k < extra else 0)
        regions = [(t, a + 1_000_000, a + 1_000_000 + 2_000 + margin) if i in set(cand) else (t, a, e)
                   for i, (_, t, a, e) in enumerate(bands)]
        kw = dict(dist="lognormal", fam_mean=12.0, umi_len=4, p_duplex=0.9, tail=(2e-6, 400))
        return Plan("c3", bands, pairs, regions, kw, seed,
                    "synthetic targeted panel, %d reads on %d targets of 2 kb (one per cytoband, %d hg19 bands), L=128, "
                    "family size ~ lognormal(mean 12) with a 2e-6 tail of 400x families (10^3-10^4 reads), "
                    "p(duplex)=0.9, 8-nt UMIs (BASELINE.json configs[2])" % (2 * total_pairs, len(cand), nb))
    if workload == "c4":
        length = np.array([e - a for _, _, a, e in bands], np.float64)
        raw = length / length.sum() * total_pairs
        pairs = np.floor(raw).astype(np.int64)
        short = total_pairs - int(pairs.sum())
        pairs[np.argsort(-(raw - pairs), kind="stable")[:short]] += 1
        small = length < 4 * margin                        # a band too short to hold a pair population (none in hg19)
        pairs[np.argmax(length)] += int(pairs[small].sum())
        pairs[small] = 0
        regions = [(t, a, e) for _, t, a, e in bands]
        kw = dict(dist="wgs", p_duplex=0.3, umi_len=2)
        return Plan("c4", bands, pairs, regions, kw, seed,
                    "synthetic whole-genome low-duplication set, %d reads over the %d hg19 cytobands in proportion "
                    "to band length, L=128, ~90 %% singletons, duplex partner for 30 %% of the molecules "
                    "(BASELINE.json configs[3])" % (2 * total_pairs, nb))
    raise ValueError(workload)


def band_seed(seed, band):
    """Generator seed of one band: depends on (seed, band) only, never on the rank that makes it."""
    return (int(seed) * 1_000_003 + int(band) * 7919 + 17) % (2 ** 62)


def generate_shard(pl, rank=0, world=1, device="cuda:0", predict=True):
    """This rank's bands, in bed order, as one DeviceBatch (pair-completion order within a band, bands
    one after the other - the order in which the reference's region loop meets them).

    Returns (batch, info): info has the owned bands, reads per band and - with `predict` - the stage
    counters / size histogram the molecule model implies for THIS shard (`stats`, `hist`)."""
    import torch
    from .pack import row_geometry
    from .synth_device import DeviceBatch, generate_device
    owner = pl.owners(world)
    mine = [b for b in range(len(pl.bands)) if owner[b] == rank and pl.pairs[b] > 0]
    n_pairs = int(pl.pairs[mine].sum()) if mine else 0
    dev = torch.device(device)
    lpad, stride = row_geometry(L)
    fields = torch.empty((2 * n_pairs, 8), dtype=torch.int32, device=dev)
    payload = torch.empty((2 * n_pairs, stride), dtype=torch.uint8, device=dev)
    stats = torch.zeros(16, dtype=torch.int64)
    hist = torch.zeros(65536, dtype=torch.int64)
    dup_bc, at = None, 0
    umi_len = pl.gen_kw.get("umi_len", 2)
    for b in mine:
        k = int(pl.pairs[b])
        tid, lo, hi = pl.regions[b]
        part = generate_device(seed=band_seed(pl.seed, b), n_pairs=k, L=L, region=(lo, hi), tid=tid, device=str(dev),
                               fields_out=fields[2 * at:2 * (at + k)], payload_out=payload[2 * at:2 * (at + k)],
                               predict=predict, **pl.gen_kw)
        if predict:
            stats += part.predicted[0]
            hist += part.predicted[1]
        dup_bc = part.dup_bc if dup_bc is None else dup_bc
        at += k
    if dup_bc is None:
        nbc = 4 ** umi_len
        ids = torch.arange(nbc * nbc, device=dev)
        dup_bc = ((ids % nbc) * nbc + ids // nbc).to(torch.int32)
    batch = DeviceBatch(fields, payload, dup_bc, lpad, stride, L, umi_len, (stats, hist) if predict else None)
    info = {"bands": mine, "n_bands": len(mine), "pairs": n_pairs,
            "stats": stats.numpy().copy() if predict else None, "hist": hist.numpy().copy() if predict else None}
    return batch, info


def pick_window(batch, max_reads=1_500_000, min_reads=20_000):
    """Read indices (whole pairs, ascending) of a genomic window of the batch for the oracle: all pairs of
    the first contig present whose two mates start inside [lo, hi).  Families never straddle a window -
    every member, the mate family and the duplex partner share both coordinates."""
    import torch
    f = batch.fields
    if f.shape[0] == 0:
        return None
    tid_a, pos_a, pos_b = f[0::2, 2], f[0::2, 3], f[1::2, 3]
    t0 = int(tid_a[f.shape[0] // 4])                        # a contig from the middle of the shard
    on = tid_a == t0
    centre = int(torch.median(pos_a[on]))
    width = 64
    idx = None
    everything = 2 * int(on.sum())
    while width < (1 << 29):
        lo, hi = max(0, centre - width), centre + width
        m = on & (pos_a >= lo) & (pos_a < hi) & (pos_b >= lo) & (pos_b < hi)
        k = 2 * int(m.sum())
        if k > max_reads:
            break
        idx = (t0, lo, hi, m)
        if (k >= min_reads and k * 4 > max_reads) or k == everything:   # big enough (the next doubling might overshoot)
            break
        width *= 2
    if idx is None:
        return None
    t0, lo, hi, m = idx
    pairs = torch.nonzero(m).reshape(-1)
    reads = torch.stack([2 * pairs, 2 * pairs + 1], 1).reshape(-1)
    return {"tid": t0, "lo": lo, "hi": hi, "reads": reads}
