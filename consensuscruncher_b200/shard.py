"""Multi-GPU sharding of the consensus path (one process per GPU).

A family key contains the read's own (tid, pos) and `duplex_tag` leaves coordinates
untouched (consensus_helper.py:293-301, :629-648), so the SSCS families, the DCS partner
and the rescue partner of a read all start at that read's position: any partition of the
genome into disjoint start-position ranges gives independent work.  The
reference's own cytoband list (863 contiguous bands, `bed_separator`,
consensus_helper.py:38-54) is the default range set; it only ever processes them one
after the other (SSCS_maker.py:229-257).

Ranks never exchange reads.  The only collective is the sum of the stage counters and of
the family-size histogram at the end: on GPUs `engine.stats_allreduce()` - libccgpu.so calls
ncclAllReduce itself on one fused buffer (`cc_stats_allreduce`, include/ccgpu.h), no torch in
between; `connect` hands the communicator id from rank 0 to the other ranks.  `reduce_stats`
is the same sum for host tensors (gloo) and is what the CPU tests use.
"""
from __future__ import annotations

import numpy as np

from .result import STAT_NAMES


def assign_regions(weights, world):
    """Greedy balance: heaviest region first onto the lightest rank.

    weights: per-region read counts.  Returns rank of every region (list of int).
    Deterministic: ties go to the lower rank / earlier region.
    """
    order = sorted(range(len(weights)), key=lambda i: (-weights[i], i))
    load = [0] * world
    owner = [0] * len(weights)
    for i in order:
        r = min(range(world), key=lambda k: (load[k], k))
        owner[i] = r
        load[r] += weights[i]
    return owner


def region_of_reads(tid, pos, regions):
    """Index of the region (tid, start, end) holding each read start, -1 if none
    (reads on contigs absent from the bed are never visited).

    Disjoint regions (the cytoband tables: 863 contiguous bands) are resolved by one binary
    search over the (contig, start) keys; overlapping ones keep the last match, region by region."""
    tid = np.asarray(tid)
    pos = np.asarray(pos)
    if len(regions) == 0:
        return np.full(tid.shape, -1, np.int64)
    reg = np.asarray(regions, np.int64).reshape(-1, 3)
    key = (reg[:, 0] << 32) | reg[:, 1]
    by_start = np.argsort(key, kind="stable")
    skey, sreg = key[by_start], reg[by_start]
    disjoint = bool(np.all((sreg[1:, 0] != sreg[:-1, 0]) | (sreg[1:, 1] >= sreg[:-1, 2]))) and bool(np.all(reg[:, 0] >= 0))
    if disjoint:
        rkey = (tid.astype(np.int64) << 32) | np.maximum(pos.astype(np.int64), 0)
        j = np.searchsorted(skey, rkey, side="right") - 1
        jj = np.maximum(j, 0)
        hit = (j >= 0) & (tid >= 0) & (pos >= 0) & (sreg[jj, 0] == tid) & (pos < sreg[jj, 2])
        return np.where(hit, by_start[jj], -1).astype(np.int64)
    out = np.full(tid.shape, -1, np.int64)
    for i, (t, a, b) in enumerate(regions):
        m = (tid == t) & (pos >= a) & (pos < b)
        out[m] = i
    return out


def shard_pairs(packed_fields, regions, world):
    """Which rank owns each read PAIR of a packed batch.

    A pair goes where its first-seen mate starts; both families the pair feeds are keyed
    by that mate's and its partner's coordinates, and every pair with the same key lands
    on the same rank because the key contains both coordinates.  To keep the two mates'
    families (and their duplex partners, which share both coordinates) together, the pair
    is placed by the LOWER of its two (tid, pos) coordinates.
    """
    f = packed_fields
    tid_a, pos_a = f["tid"][0::2], f["pos"][0::2]
    tid_b, pos_b = f["tid"][1::2], f["pos"][1::2]
    lower_is_a = (tid_a < tid_b) | ((tid_a == tid_b) & (pos_a <= pos_b))
    tid = np.where(lower_is_a, tid_a, tid_b)
    pos = np.where(lower_is_a, pos_a, pos_b)
    reg = region_of_reads(tid, pos, regions)
    weights = np.bincount(reg[reg >= 0], minlength=len(regions)).tolist()
    owner = np.asarray(assign_regions(weights, world), np.int64)
    return np.where(reg >= 0, owner[np.maximum(reg, 0)], -1), reg


def take_pairs(packed, pair_mask):
    """Sub-batch holding the selected pairs (completion order is preserved)."""
    from .pack import PackedReads
    m = np.repeat(np.asarray(pair_mask, bool), 2)
    return PackedReads(packed.fields[m].copy(), packed.payload[m].copy(), packed.lpad, packed.barcodes,
                       packed.dup_bc, packed.cigar_pairs, packed.rgs,
                       src=None if packed.src is None else packed.src[m],
                       lens=None if packed.lens is None else packed.lens[m],
                       counters=dict(packed.counters))


def connect(engine, rank, world, exchange):
    """Join `engine`'s context to the communicator of `world` ranks.

    `exchange(payload_or_None) -> bytes`: the host's own way of handing rank 0's 128-byte id to every
    rank (rank 0 passes the id, the others None; all get the id back) - a torch.distributed /
    MPI broadcast, a shared file, a launcher's key-value store.  Nothing to do for one rank."""
    if world <= 1:
        return
    from .engine import GpuConsensus
    uid = exchange(GpuConsensus.comm_unique_id() if rank == 0 else None)
    engine.comm_init(uid, rank, world)


def file_exchange(path, rank, timeout=120.0):
    """An `exchange` for `connect` over a file every rank can see (one node: any temporary directory)."""
    import os
    import time

    def exchange(payload):
        if rank == 0:
            tmp = "%s.%d" % (path, os.getpid())
            with open(tmp, "wb") as fh:
                fh.write(payload)
            os.replace(tmp, path)
            return payload
        t0 = time.time()
        while not os.path.exists(path):
            if time.time() - t0 > timeout:
                raise TimeoutError("no communicator id at %s" % path)
            time.sleep(0.01)
        with open(path, "rb") as fh:
            return fh.read()
    return exchange


def reduce_stats(stats, hist, group=None):
    """Sum `stats` (i64[16]) and `hist` (family-size histogram) over all ranks, in place.

    Accepts torch tensors on the device of the process group's backend (cuda for nccl,
    cpu for gloo).  Returns (stats, hist)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(stats, group=group)
        dist.all_reduce(hist, group=group)
    return stats, hist


def stats_dict(stats):
    return {nm: int(stats[i]) for i, nm in enumerate(STAT_NAMES)}


class _DevicePointer:
    """Minimal __cuda_array_interface__ carrier so torch can wrap a raw device pointer."""

    def __init__(self, ptr, count, typestr):
        self.__cuda_array_interface__ = {"shape": (count,), "typestr": typestr, "data": (int(ptr), False),
                                         "version": 2, "strides": None}


def device_tensor(ptr, count, dtype, device_index):
    """torch view (no copy) of `count` elements of `dtype` at device address `ptr`."""
    import torch
    typestr = {torch.int64: "<i8", torch.int32: "<i4", torch.uint8: "|u1"}[dtype]
    if not ptr:
        raise ValueError("NULL device pointer")
    return torch.as_tensor(_DevicePointer(ptr, count, typestr), device=torch.device("cuda", device_index))
