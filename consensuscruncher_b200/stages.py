"""The reference's three stage scripts as stand-alone commands on the GPU path.

`ConsensusCruncher.py consensus` (cli.py) runs the whole path in one pass on the GPU.  The
reference also lets a user run its stages one at a time; these entry points keep that
interface - same arguments, same files next to `--outfile` / `--singleton`:

  sscs_main      SSCS_maker.py  --cutoff --infile --outfile [--bdelim] [--bedfile]   (SSCS_maker.py:172-382)
  dcs_main       DCS_maker.py   --infile --outfile [--bedfile]                       (DCS_maker.py:128-284)
  sc_main        singleton_correction.py --singleton [--bedfile]                     (singleton_correction.py:111-311)

The SSCS stage is the GPU grouping + vote (cc_group, cc_sscs_vote).  The DCS and rescue
stages read consensus BAMs in which every family is a single record: which tag pairs with
which is the device's hash join over those records (cc_join_records, once per region, in place
of `duplex_tag` strings and dictionary probes), the per-base arithmetic of all pairs goes to
the GPU in one batch (cc_duplex_consensus), and the host keeps what is left of the reference's
bookkeeping - the order of emission and what a region loop forgets between regions.  Mode ties
take the first value, as everywhere in this package.
"""
from __future__ import annotations

import argparse
import os
import time
from collections import OrderedDict

import numpy as np

from . import bam, fastpack, keys, pack
from .api import row_to_seq_qual, rows_of
from .pack import NO_RG, row_geometry
from .records import Read, codes_to_seq


# ---- shared helpers -------------------------------------------------------------------------
def _regions(bedfile, header):
    """[(name, tid, start, end)] in bed order, or [None] for one pass over the whole file."""
    if bedfile is None:
        return [None]
    out = []
    for key, (start, end) in keys.bed_separator(bedfile).items():
        chrom = key.split("_", 1)[0]
        tid = header.tid(chrom)
        if tid < 0:
            raise ValueError("bed contig %r is not in the BAM header" % chrom)
        out.append((key, chrom, tid, start, end))
    return out


def _visit(reads, region, index):
    if region is None:
        return range(len(reads))
    _, _, tid, start, end = region
    pos, idx = index.get(tid, (None, None))
    if pos is None:
        return []
    lo = int(np.searchsorted(pos, start, "left"))
    hi = int(np.searchsorted(pos, end, "left"))
    return idx[lo:hi]


def _position_index(reads):
    by_tid = {}
    for i, r in enumerate(reads):
        by_tid.setdefault(r.tid, []).append(i)
    out = {}
    for t, ix in by_tid.items():
        p = np.array([reads[i].pos for i in ix])
        order = np.argsort(p, kind="stable")
        out[t] = (p[order], [ix[k] for k in order])
    return out


class _Families:
    """State that `read_bam` threads through its calls (consensus_helper.py:305-480): reads
    waiting for their mate, families by tag, family sizes, and the tags owned by each pair name."""

    def __init__(self):
        self.waiting = {}
        self.members = OrderedDict()      # tag -> [reads]            (read_dict)
        self.size = {}                    # tag -> count              (tag_dict)
        self.names = OrderedDict()        # pair name -> [tags]       (csn_pair_dict)
        self.key = {}                     # tag -> (barcode, cigar pair, first read): the tag as a tuple, for the device join
        self.total = self.unmapped_mate = self.multiple = 0

    def feed(self, reads, indices):
        for i in indices:
            r = reads[i]
            self.total += 1
            if r.flag & 0x4:
                self.total -= 1
            elif r.flag in keys.MATE_UNMAPPED_FLAGS:
                self.unmapped_mate += 1
            elif r.flag & 0x900:
                self.multiple += 1
            first = self.waiting.pop(r.qname, None)
            if first is None:
                self.waiting[r.qname] = r
                continue
            barcode = first.qname.split("_")[0]
            cigar = keys.cigar_order(first, r)
            name = keys.sscs_qname(first, r, barcode, cigar)
            for x in (first, r):
                tag = keys.unique_tag(x, barcode, cigar)
                if tag not in self.members and tag not in self.size:
                    self.members[tag] = [x]
                    self.size[tag] = 1
                    self.key[tag] = (barcode, cigar, x)
                    owned = self.names.setdefault(name, [])
                    if len(owned) < 2:
                        owned.append(tag)
                elif tag in self.size and tag in self.members:
                    self.members[tag].append(x)
                    self.size[tag] += 1


def _join(engine, groups):
    """[(families, tags)] -> partner index of every tag within the concatenation of the tag lists, or -1:
    the record whose key is the tag's with the barcode halves swapped and R1 <-> R2 (duplex_tag,
    consensus_helper.py:605-648), found by the device (cc_join_records); of several, the first."""
    n = sum(len(tags) for _, tags in groups)
    fields = np.zeros(n, pack.FIELDS_DT)
    bc_ids, cig_ids = {}, {}
    k = 0
    for fam, tags in groups:
        for tag in tags:
            barcode, cigar, r = fam.key[tag]
            f = fields[k]
            f["bc"] = bc_ids.setdefault(barcode, len(bc_ids))
            f["cig"] = cig_ids.setdefault(cigar, len(cig_ids))
            f["tid"], f["pos"], f["mtid"], f["mpos"], f["flag"] = r.tid, r.pos, r.mtid, r.mpos, r.flag
            k += 1
    return engine.join_records(fields, pack.build_dup_table(list(bc_ids)))


def _pair_mode(a, b, is_flag=False):
    if a == b:
        return a
    if is_flag:
        rank = {99: 0, 83: 1, 147: 2, 163: 3}
        if rank.get(b, 4) < rank.get(a, 4):
            return b
    return a


def _agree_batch(engine, pairs, sc):
    """[(read1, read2)] -> [(seq, qual)] with the GPU doing every base."""
    if not pairs:
        return []
    lpad, stride = row_geometry(max(len(a.seq) for a, _ in pairs))
    r1, _ = rows_of([a for a, _ in pairs], lpad, stride)
    r2, _ = rows_of([b for _, b in pairs], lpad, stride)
    out = engine.duplex_consensus_rows(r1, r2, lpad, sc)
    return [row_to_seq_qual(out[i], lpad, len(pairs[i][0].seq)) for i in range(len(pairs))]


def _base(outfile, marker):
    return outfile.split(marker)[0]


def _get_engine(engine, device=0):
    if engine is not None:
        return engine, False
    from .engine import GpuConsensus
    return GpuConsensus(device), True


# ---- SSCS_maker ------------------------------------------------------------------------------
def run_sscs(infile, outfile, cutoff, bdelim="|", bedfile=None, engine=None):
    t0 = time.time()
    from .emit import sscs_stage_records
    header, reads = fastpack.open_bam(infile)
    regions = None
    if bedfile is not None:
        regions = [(tid, a, b) for _, _, tid, a, b in _regions(bedfile, header)]
    packed = fastpack.pack_reads(reads, "sscs", bdelim, regions)
    eng, own = _get_engine(engine)
    try:
        eng.set_params(cutoff=cutoff, scorrect=False)
        eng.load(packed)
        eng.group()
        eng.sscs_vote()
        arrays = {k: eng.fetch(k) for k in ("fam_size", "fam_tmpl", "fam_mapq", "fam_flag", "fam_tlen", "fam_rg",
                                            "sscs_row", "sscs_rows", "read_fam")}
    finally:
        if own:
            eng.close()
    sscs, singleton = sscs_stage_records(reads, packed, arrays, packed.lpad)
    base = _base(outfile, ".sscs")
    bam.write_bam(outfile, header, sscs)
    bam.write_bam(base + ".singleton.bam", header, singleton)
    bam.write_bam(base + ".badReads.bam", header, [reads[i] for i in packed.bad])
    c = packed.counters
    with open(base + ".stats.txt", "w") as fh:
        fh.write("# === SSCS ===\nUncollapsed - Total reads: %d\nUncollapsed - Unmapped reads: %d\n"
                 "Uncollapsed - Secondary/Supplementary reads: %d\nSSCS reads: %d\nSingletons: %d\nBad spacers: %d\n"
                 % (c["total"], c["unmapped_mate"], c["multiple_mapping"], len(sscs), len(singleton), c["bad_spacer"]))
    sizes, first, counts = (np.unique(arrays["fam_size"], return_index=True, return_counts=True)
                            if len(arrays["fam_size"]) else ([], [], []))
    order = np.argsort(first, kind="stable") if len(first) else []
    with open(base + ".read_families.txt", "w") as fh:
        fh.write("family_size\tfrequency\n" + "\n".join("%d\t%d" % (sizes[i], counts[i]) for i in order))
    with open(base + ".time_tracker.txt", "w") as fh:
        fh.write("SSCS: %s\n" % ((time.time() - t0) / 60))
    return len(sscs), len(singleton)


# ---- DCS_maker -------------------------------------------------------------------------------
def run_dcs(infile, outfile, bedfile=None, engine=None):
    t0 = time.time()
    header, reads = bam.read_bam(infile)
    after_sc = "dcs.sc" in outfile
    base = _base(outfile, ".dcs.sc" if after_sc else ".dcs")
    single_path = base + (".sscs.sc.singleton.bam" if after_sc else ".sscs.singleton.bam")
    fam = _Families()
    index = _position_index(reads) if bedfile is not None else None
    made, made_before = set(), set()  # tags that led a DCS (in this region / in earlier regions)
    jobs, plan = [], []              # plan: ("dcs", job index, lead, partner) | ("single", read)
    eng, own = _get_engine(engine)
    try:
        for region in _regions(bedfile, header):
            fam.feed(reads, _visit(reads, region, index))
            tags = list(fam.members)
            where = {t: i for i, t in enumerate(tags)}
            partner = _join(eng, [(fam, tags)])
            for name in list(fam.names):
                for tag in fam.names[name]:
                    if tag not in fam.members:
                        continue
                    j = int(partner[where[tag]])
                    if j >= 0:
                        ds = tags[j]
                    elif made_before:
                        ds = keys.duplex_tag(tag)          # (not among this region's tags: only an earlier region can have made it)
                    else:
                        ds = None
                    if ds is not None and (ds in made or ds in made_before):
                        continue
                    if j >= 0 and ds in fam.members:       # (a member has a size: DCS_maker.py:236)
                        lead, other = fam.members[tag][0], fam.members[ds][0]
                        plan.append(("dcs", len(jobs), lead, other))
                        jobs.append((lead, other))
                        made.add(tag)
                    else:
                        plan.append(("single", fam.members[tag][0]))
                    del fam.members[tag]
                del fam.names[name]
            made_before |= made
            made = set()
        results = _agree_batch(eng, jobs, sc=False)
    finally:
        if own:
            eng.close()
    dcs, single = [], []
    for item in plan:
        if item[0] == "single":
            single.append(item[1])
            continue
        _, j, lead, other = item
        seq, qual = results[j]
        rg = None if (lead.rg is None or other.rg is None) else _pair_mode(lead.rg, other.rg)
        dcs.append(Read(keys.dcs_consensus_tag(lead.qname, other.qname), _pair_mode(lead.flag, other.flag, True),
                        lead.tid, lead.pos, _pair_mode(lead.mapq, other.mapq), lead.cigar, lead.mtid, lead.mpos,
                        _pair_mode(lead.tlen, other.tlen), seq, qual, rg=rg, aux=b""))
    bam.write_bam(outfile, header, dcs)
    bam.write_bam(single_path, header, single)
    tag = " SC" if after_sc else ""
    with open(base + ".stats.txt", "a") as fh:
        fh.write("# === %s ===\nSSCS%s - Total reads: %d\nSSCS%s - Unmapped reads: %d\n"
                 "SSCS%s - Secondary/Supplementary reads: %d\nDCS%s reads: %d\nSSCS%s singletons: %d \n"
                 % ("DCS - Singleton Correction" if after_sc else "DCS", tag, fam.total, tag, fam.unmapped_mate,
                    tag, 0, tag, len(dcs), tag, len(single)))
    with open(base + ".time_tracker.txt", "a") as fh:
        fh.write("DCS: %s\n" % ((time.time() - t0) / 60))
    return len(dcs), len(single)


# ---- singleton_correction -------------------------------------------------------------------------
def run_sc(singleton_path, bedfile=None, engine=None):
    head, tail = singleton_path.split(".singleton")[0], singleton_path.split(".singleton")[1]
    header, singles = bam.read_bam(singleton_path)
    _, sscs = bam.read_bam("%s.sscs%s" % (head, tail))
    s_fam, c_fam = _Families(), _Families()
    s_index = _position_index(singles) if bedfile is not None else None
    c_index = _position_index(sscs) if bedfile is not None else None
    last_chr = "chrM"
    jobs, plan = [], []              # ("sscs" | "singleton", job, read, qname) | ("uncorrected", read)
    pending = {}                     # singleton-by-singleton corrections waiting for their partner's turn
    counter = 0
    eng, own = _get_engine(engine)
    try:
        for region in _regions(bedfile, header):
            if region is not None and region[1] != last_chr:
                # a new chromosome forgets the SSCS seen so far and the singleton sizes (singleton_correction.py:192-200)
                s_fam.size = {}
                kept_total = c_fam.total
                c_fam = _Families()
                c_fam.total = kept_total
                last_chr = region[1]
            s_fam.feed(singles, _visit(singles, region, s_index))
            c_fam.feed(sscs, _visit(sscs, region, c_index))
            # the device's join: a singleton's partner among the SSCS first (they come first in the concatenation),
            # else among the singletons; the second call answers for singletons whose SSCS partner is already taken
            c_tags, s_tags = list(c_fam.members), list(s_fam.members)
            nc = len(c_tags)
            where = {t: i for i, t in enumerate(s_tags)}
            both = _join(eng, [(c_fam, c_tags), (s_fam, s_tags)])
            among_singles = _join(eng, [(s_fam, s_tags)])
            for name in list(s_fam.names):
                for tag in s_fam.names[name]:
                    counter += 1
                    qname = name + ":1"
                    read = s_fam.members[tag][0]
                    i = where[tag]
                    jc, js = int(both[nc + i]), int(among_singles[i])
                    by_sscs = c_tags[jc] if 0 <= jc < nc else None
                    by_single = s_tags[js] if js >= 0 else None
                    if by_sscs is not None and by_sscs in c_fam.members:
                        plan.append(("sscs", len(jobs), read, qname))
                        jobs.append((read, c_fam.members[by_sscs][0]))
                        del c_fam.members[by_sscs]
                        del s_fam.members[tag]
                    elif by_single is not None and by_single in s_fam.members:
                        ds = by_single
                        plan.append(("singleton", len(jobs), read, qname))
                        jobs.append((read, s_fam.members[ds][0]))
                        pending[tag] = ds
                        if ds in pending:
                            del s_fam.members[tag], s_fam.members[ds], pending[tag], pending[ds]
                    else:
                        plan.append(("uncorrected", read))
                        del s_fam.members[tag]
                del s_fam.names[name]
        results = _agree_batch(eng, jobs, sc=True)
    finally:
        if own:
            eng.close()
    out = {"sscs": [], "singleton": [], "uncorrected": []}
    for item in plan:
        if item[0] == "uncorrected":
            out["uncorrected"].append(item[1])
            continue
        kind, j, read, qname = item
        seq, qual = results[j]
        out[kind].append(Read(qname, read.flag, read.tid, read.pos, read.mapq, read.cigar, read.mtid, read.mpos,
                              read.tlen, seq, qual, rg=read.rg, aux=b""))
    bam.write_bam(head + ".sscs.correction.bam", header, out["sscs"])
    bam.write_bam(head + ".singleton.correction.bam", header, out["singleton"])
    bam.write_bam(head + ".uncorrected.bam", header, out["uncorrected"])
    if s_fam.total:                  # the reference stops here with a ZeroDivisionError otherwise (App. B Q20)
        with open(head + ".stats.txt", "a") as fh:
            fh.write("# === Singleton Correction ===\nTotal singletons: %d\nSingleton Correction by SSCS: %d\n"
                     "%% Singleton Correction by SSCS: %s\nSingleton Correction by Singletons: %d\n"
                     "%% Singleton Correction by Singletons : %s\nUncorrected Singletons: %d \n"
                     % (counter, len(out["sscs"]), repr(len(out["sscs"]) / s_fam.total * 100), len(out["singleton"]),
                        repr(len(out["singleton"]) / s_fam.total * 100), len(out["uncorrected"])))
    return len(out["sscs"]), len(out["singleton"]), len(out["uncorrected"])


# ---- command lines ---------------------------------------------------------------------------------
def sscs_main(argv=None):
    p = argparse.ArgumentParser(prog="SSCS_maker.py")
    p.add_argument("--cutoff", type=float, required=True, help="proportion of identical bases needed for a consensus")
    p.add_argument("--infile", required=True, help="input BAM")
    p.add_argument("--outfile", required=True, help="output SSCS BAM")
    p.add_argument("--bdelim", default="|", help="delimiter before the barcode in the read name")
    p.add_argument("--bedfile", required=False, help="bed file that divides the genome")
    a = p.parse_args(argv)
    run_sscs(a.infile, a.outfile, float(a.cutoff), a.bdelim, a.bedfile)
    return 0


def dcs_main(argv=None):
    p = argparse.ArgumentParser(prog="DCS_maker.py")
    p.add_argument("--infile", required=True, help="sorted SSCS BAM")
    p.add_argument("--outfile", required=True, help="output DCS BAM")
    p.add_argument("--bedfile", required=False)
    a = p.parse_args(argv)
    run_dcs(str(a.infile), str(a.outfile), a.bedfile)
    return 0


def sc_main(argv=None):
    p = argparse.ArgumentParser(prog="singleton_correction.py")
    p.add_argument("--singleton", required=True, type=str, help="sorted singleton BAM")
    p.add_argument("--bedfile", required=False)
    a = p.parse_args(argv)
    run_sc(a.singleton, a.bedfile)
    return 0


# ---- what the driver does between the stages (ConsensusCruncher.py:10-29, :247-251) ---------------
def sort_index(path, engine=None):
    """`samtools view -bu | samtools sort -`, remove the unsorted file, `samtools index`.
    With an engine the ordering runs on the GPU (`cc_sort_coordinates`)."""
    from .records import sort_records
    header, reads = bam.read_bam(path)
    out = path.split(".bam", 1)[0] + ".sorted.bam"
    bam.write_bam(out, header, sort_records(reads, engine.sort_records if engine else None),
                  sort_order="coordinate", index=True)
    os.remove(path)
    return out


def merge_bams(out, *inputs, engine=None):
    """`samtools merge OUT IN1 IN2 ...` (ties keep the order of the inputs)."""
    from .records import merge_records
    header = None
    lists = []
    for p in inputs:
        h, reads = bam.read_bam(p)
        header = header or h
        lists.append(reads)
    bam.write_bam(out, header, merge_records(*lists, sorter=engine.sort_records if engine else None))
    return out
