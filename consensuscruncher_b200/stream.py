"""One packed batch through one or several GPUs in region chunks.

The reference splits a sample by the bands of its bed file "to offload the memory burden"
(`ConsensusCruncher.py:374-379`) and walks them one after the other (SSCS_maker.py:229-257).  Here the
packed reads - already in the order in which that walk completes the pairs - are cut into chunks of
consecutive bands (a family, its mate family and its duplex partner complete inside one band, so no
family straddles a cut), the chunks are dealt to the contexts (one per GPU, one host thread each; the
library releases the GIL) and the chunk results are joined back, offsets added, into exactly the arrays
one pass over the whole batch would have produced.  Device memory is bounded by the chunk size, and the
output files are byte-identical however many chunks or GPUs took part.

Pairs that leave the path's domain - a pair whose two mates do not found / join their two families
together (consensus_helper.py:456-468 prints "Consensus tag NOT UNIQUE" for them; downstream the reference
loses such families between stages) - no longer abort the sample: they are set
aside, reported, and written to badReads.bam; everything else is processed.
"""
from __future__ import annotations

from concurrent.futures import ThreadPoolExecutor

import numpy as np

from .pack import DomainError, PackedReads
from .result import ALL_ARRAYS, PAIR_META, ConsensusResult


def chunk_bounds(packed, regions=None, max_reads=20_000_000):
    """Cut points (read indices, even) of `packed` into chunks of at most ~max_reads reads, only where the
    completing mates (odd reads) pass from one region to the next - or, without regions, from one
    (contig, position) to a larger one.  Input whose completion coordinates are not in order is one chunk."""
    n = packed.n
    if n <= max_reads:
        return [0, n]
    tid, pos = packed.fields["tid"][1::2], packed.fields["pos"][1::2]
    if regions is not None:
        from .shard import region_of_reads
        key = region_of_reads(tid, pos, regions)
        if (key < 0).any() or (np.diff(key) < 0).any():      # pairs completed outside the bed / bands revisited
            return [0, n]
        change = np.flatnonzero(np.diff(key) != 0) + 1
    else:
        key = (tid.astype(np.int64) << 32) | pos.astype(np.int64)
        if (np.diff(key) < 0).any():
            return [0, n]
        change = np.flatnonzero(np.diff(key) != 0) + 1
    cuts, last = [0], 0
    budget = max_reads // 2                                   # in pairs
    for c in change.tolist():
        if c - last >= budget:
            cuts.append(2 * c)
            last = c
    cuts.append(n)
    return cuts


def take(packed, a, b):
    """Reads a..b of a packed batch (views, nothing is copied)."""
    return PackedReads(packed.fields[a:b], packed.payload[a:b], packed.lpad, packed.barcodes, packed.dup_bc,
                       packed.cigar_pairs, packed.rgs)


def concat(parts, sizes, lpad, stride, scorrect):
    """Join the results of consecutive chunks (`sizes`: reads per chunk) into the result of the whole batch."""
    if len(parts) == 1:
        return parts[0]
    out = {k: [] for k in ALL_ARRAYS}
    n_dcs_total = sum(int((p.dcs_row >= 0).sum()) for p in parts)
    read_off = fam_off = sscs_off = corr_off = dcs_off = extra_off = 0
    extras = {k: [] for k in ("dcssc_rows",) + tuple(x for x in PAIR_META if x.startswith("dcssc_"))}
    shared = {k: [] for k in extras}
    for p, n in zip(parts, sizes):
        nd = p.dcs_rows.shape[0]
        out["read_fam"].append(p.read_fam + np.uint32(fam_off))
        out["fam_tmpl"].append(p.fam_tmpl + np.uint32(read_off))
        for k in ("fam_size", "fam_mapq", "fam_flag", "fam_tlen", "fam_rg"):
            out[k].append(getattr(p, k))
        out["partner"].append(np.where(p.partner >= 0, p.partner + fam_off, -1).astype(np.int32))
        out["sscs_row"].append(np.where(p.sscs_row >= 0, p.sscs_row + sscs_off, -1).astype(np.int32))
        out["corr_row"].append(np.where(p.corr_row >= 0, p.corr_row + corr_off, -1).astype(np.int32))
        out["dcs_row"].append(np.where(p.dcs_row >= 0, p.dcs_row + dcs_off, -1).astype(np.int32))
        x = p.dcssc_row
        out["dcssc_row"].append(np.where(x < 0, -1, np.where(x < nd, x + dcs_off, x - nd + n_dcs_total + extra_off)).astype(np.int32))
        for k in ("sscs_rows", "corr_rows", "dcs_rows", "dcs_mapq", "dcs_flag", "dcs_tlen", "dcs_rg"):
            out[k].append(getattr(p, k))
        for k in extras:                                      # the pool: every chunk's shared DCS rows, then every chunk's extras
            a = getattr(p, k)
            if a.shape[0]:
                shared[k].append(a[:nd])
                extras[k].append(a[nd:])
        out["stats"].append(p.stats)
        read_off += n
        fam_off += p.n_fam
        sscs_off += p.sscs_rows.shape[0]
        corr_off += p.corr_rows.shape[0]
        dcs_off += nd
        extra_off += max(0, p.dcssc_rows.shape[0] - nd)
    arrays = {}
    for k in ALL_ARRAYS:
        if k == "stats":
            arrays[k] = np.sum(out[k], axis=0).astype(parts[0].stats.dtype)
        elif k in extras:
            pieces = shared[k] + extras[k]
            arrays[k] = np.concatenate(pieces) if pieces else getattr(parts[0], k)
        else:
            arrays[k] = np.concatenate(out[k])
    return ConsensusResult(lpad, stride, scorrect, **arrays)


def _faulty_pairs(engine, n_reads):
    """Pairs of the loaded batch with a read in a family that breaks the pairing rule: the mate of a family's
    template read must be the template of a family of the same size (then the two families were founded and
    grown together, SSCS_maker.py:275).  Decided family by family through the mate, not by position, so one
    orphan family does not drag its neighbours along."""
    tmpl, size = engine.fetch("fam_tmpl").astype(np.int64), engine.fetch("fam_size")
    read_fam = engine.fetch("read_fam").astype(np.int64)
    mate_fam = read_fam[tmpl ^ 1]
    ok = (tmpl[mate_fam] == (tmpl ^ 1)) & (size[mate_fam] == size)
    bad_read = ~ok[read_fam]
    return np.unique(np.flatnonzero(bad_read) // 2)


def run_chunk(engine, chunk, compact=False, max_rounds=8):
    """One chunk on one context -> (result, kept pair mask or None).  Pairs outside the domain are set aside."""
    keep = None
    cur = chunk
    for _ in range(max_rounds):
        try:
            return engine.consensus(cur, compact=compact), keep
        except DomainError as exc:
            if "found/join" not in str(exc):
                raise
            bad = _faulty_pairs(engine, cur.n)
            if bad.size == 0:
                raise
            alive = np.flatnonzero(keep) if keep is not None else np.arange(chunk.n // 2)
            if keep is None:
                keep = np.ones(chunk.n // 2, bool)
            keep[alive[bad]] = False
            m = np.repeat(keep, 2)
            cur = PackedReads(chunk.fields[m].copy(), chunk.payload[m].copy(), chunk.lpad, chunk.barcodes, chunk.dup_bc,
                              chunk.cigar_pairs, chunk.rgs)
    raise DomainError("pairs outside the domain could not be separated from the rest")


def consensus_chunked(packed, engines, regions=None, max_reads=20_000_000, compact=False, log=None):
    """The whole batch through `engines` (one GpuConsensus per GPU).  `compact`: upload format (False = plain rows; the
    compact formats pay off when the host side packs them natively - numpy's packing costs more than the link saves).
    Returns (result, packed') where packed'
    is `packed` without the pairs that were set aside (their records are added to packed'.bad)."""
    cuts = chunk_bounds(packed, regions, max_reads)
    chunks = [take(packed, a, b) for a, b in zip(cuts, cuts[1:])]
    results = [None] * len(chunks)

    def work(slot):
        eng = engines[slot]
        for k in range(slot, len(chunks), len(engines)):
            results[k] = run_chunk(eng, chunks[k], compact)
        return slot

    if len(engines) == 1 or len(chunks) == 1:
        for k in range(len(chunks)):
            results[k] = run_chunk(engines[0], chunks[k], compact)
    else:
        with ThreadPoolExecutor(max_workers=len(engines)) as ex:
            list(ex.map(work, range(len(engines))))
    keeps = [r[1] for r in results]
    if any(k is not None for k in keeps):
        mask = np.concatenate([np.repeat(k, 2) if k is not None else np.ones(c.n, bool) for k, c in zip(keeps, chunks)])
        dropped = np.flatnonzero(~mask)
        if log:
            log("%d read pairs do not found/join their two families together (consensus tag not unique, "
                "consensus_helper.py:456-468): set aside into badReads.bam" % (dropped.size // 2))
        bad = list(packed.bad) + ([] if packed.src is None else packed.src[dropped].tolist())
        counters = dict(packed.counters)
        packed = PackedReads(packed.fields[mask], packed.payload[mask], packed.lpad, packed.barcodes, packed.dup_bc,
                             packed.cigar_pairs, packed.rgs, src=None if packed.src is None else packed.src[mask],
                             lens=None if packed.lens is None else packed.lens[mask], counters=counters,
                             bad=sorted(bad), leftover=packed.leftover)
    sizes = [int(k.sum()) * 2 if k is not None else c.n for k, c in zip(keeps, chunks)]
    res = concat([r[0] for r in results], sizes, packed.lpad, packed.stride, results[0][0].scorrect)
    return res, packed, len(chunks)
