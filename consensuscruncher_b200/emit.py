"""Host emit: family-level results from the GPU -> the reference's output records.

Builds, from a `ConsensusResult`, every record list that
`ConsensusCruncher.py consensus` leaves on disk plus the
text of `stats.txt` and `read_families.txt` (App. A.7):

  record assembly     consensus_helper.py:538-585  create_aligned_segment
  SSCS / singleton    SSCS_maker.py:274-298
  DCS / passthrough   DCS_maker.py:226-259, dcs_consensus_tag :60-96
  rescue outputs      singleton_correction.py:249-285
  sort / merge        ConsensusCruncher.py:10-29, :247-251, :281-286

The consensus arithmetic, the family grouping, the duplex join and the
mode/flag fields all come from `res` (computed on the GPU).  What happens here
is string formatting and the bookkeeping of which strand of a duplex a later
stage meets first: stages are replayed over the sorted record lists exactly as
the reference re-reads its sorted BAMs, because that order decides the DCS
template and query name.
"""
from __future__ import annotations

import numpy as np

from . import keys
from .pack import DomainError, NO_RG, unpack_row
from .records import Read, codes_to_seq, merge_records, sort_records


class _Out:
    """A record together with the family it represents."""
    __slots__ = ("rec", "fam")

    def __init__(self, rec, fam):
        self.rec, self.fam = rec, fam


def _sorted(outs, sorter=None):
    recs = sort_records([o.rec for o in outs], sorter)
    by_id = {id(o.rec): o for o in outs}
    return [by_id[id(r)] for r in recs]


def _merged(*lists, sorter=None):
    by_id = {}
    for l in lists:
        for o in l:
            by_id[id(o.rec)] = o
    return [by_id[id(r)] for r in merge_records(*[[o.rec for o in l] for l in lists], sorter=sorter)]


def _pair_walk(outs, mate_fam):
    """Families in the order a stage's `read_bam` + emit loop meets them
    (consensus_helper.py:412-478; DCS_maker.py:226-227): pairs in order of
    completion, first-seen mate's family before the second's."""
    waiting = {}
    for o in outs:
        g = o.fam
        k = min(g, int(mate_fam[g]))
        first = waiting.pop(k, None)
        if first is None:
            waiting[k] = o
        else:
            yield first, o


def _row_seq_qual(rows, idx, lpad, n):
    codes, qual = unpack_row(rows[idx], lpad, n)
    return codes_to_seq(codes), qual


def sscs_stage_records(reads, packed, arrays, lpad, with_family=False):
    """Records of the SSCS stage in the reference's emission order (SSCS_maker.py:274-298):
    (consensus records, singleton records).  `arrays` holds fam_size, fam_tmpl, fam_mapq,
    fam_flag, fam_tlen, fam_rg, sscs_row, sscs_rows and read_fam of one GPU run."""
    f, src = packed.fields, packed.src
    size = arrays["fam_size"]
    tmpl = arrays["fam_tmpl"].astype(np.int64)
    nf = len(size)
    if nf:
        mate_fam = arrays["read_fam"][tmpl ^ 1].astype(np.int64)
        if np.any(arrays["fam_tmpl"][mate_fam] != (tmpl ^ 1)) or np.any(size[mate_fam] != size):
            raise DomainError("a read pair does not found/join its two families together "
                              "(orientation edge case, SURVEY.md App. B Q8/Q9): outside the GPU path's domain")
    names = {}
    sscs, singleton = [], []
    for g in range(nf):
        k = int(tmpl[g]) >> 1
        if k not in names:
            names[k] = keys.sscs_qname(reads[src[2 * k]], reads[src[2 * k + 1]], packed.barcodes[f["bc"][2 * k]],
                                       packed.cigar_pairs[f["cig"][2 * k]])
        t = reads[src[tmpl[g]]]
        qn = "%s:%d" % (names[k], size[g])
        if size[g] == 1:
            r = t.copy()
            r.qname = qn
            singleton.append((r, g) if with_family else r)
        else:
            seq, qual = _row_seq_qual(arrays["sscs_rows"], arrays["sscs_row"][g], lpad, len(t.seq))
            code = arrays["fam_rg"][g]
            r = Read(qn, int(arrays["fam_flag"][g]), t.tid, t.pos, int(arrays["fam_mapq"][g]), t.cigar, t.mtid,
                     t.mpos, int(arrays["fam_tlen"][g]), seq, qual, rg=None if code == NO_RG else packed.rgs[code],
                     aux=b"")
            sscs.append((r, g) if with_family else r)
    return sscs, singleton


def stats_text(c, n, scorrect=True):
    """stats.txt from the packer's counters `c` and the record counts `n` of the
    output lists (keys as in `emit_all`'s files)."""
    stats = ("# === SSCS ===\nUncollapsed - Total reads: %d\nUncollapsed - Unmapped reads: %d\n"
             "Uncollapsed - Secondary/Supplementary reads: %d\nSSCS reads: %d\nSingletons: %d\n"
             "Bad spacers: %d\n" % (c.get("total", 0), c.get("unmapped_mate", 0),
                                    c.get("multiple_mapping", 0), n["sscs.sorted"], n["singleton.sorted"],
                                    c.get("bad_spacer", 0)))
    dcs_block = ("# === %s ===\nSSCS%s - Total reads: %d\nSSCS%s - Unmapped reads: %d\n"
                 "SSCS%s - Secondary/Supplementary reads: %d\nDCS%s reads: %d\nSSCS%s singletons: %d \n")
    stats += dcs_block % ("DCS", "", n["sscs.sorted"], "", 0, "", 0, "", n["dcs.sorted"], "", n["sscs.singleton.sorted"])
    if scorrect:
        n_single = n["singleton.sorted"]
        n_cor, n_sing, n_unc = n["sscs.correction.sorted"], n["singleton.correction.sorted"], n["uncorrected.sorted"]
        if n_single:
            stats += ("# === Singleton Correction ===\nTotal singletons: %d\nSingleton Correction by SSCS: %d\n"
                      "%% Singleton Correction by SSCS: %s\nSingleton Correction by Singletons: %d\n"
                      "%% Singleton Correction by Singletons : %s\nUncorrected Singletons: %d \n"
                      % (n_cor + n_sing + n_unc, n_cor, repr((n_cor / n_single) * 100), n_sing,
                         repr((n_sing / n_single) * 100), n_unc))
        # else: the reference dies on a ZeroDivisionError before writing this block (App. B Q20)
        stats += dcs_block % ("DCS - Singleton Correction", " SC", n["sscs.sc.sorted"], " SC", 0, " SC", 0,
                              " SC", n["dcs.sc.sorted"], " SC", n["sscs.sc.singleton.sorted"])
    return stats


def families_text(res):
    """read_families.txt (SSCS_maker.py:360-365)."""
    return "family_size\tfrequency\n" + "\n".join("%d\t%d" % x for x in res.family_size_table())


def emit_all(reads, packed, res, scorrect=True, sorter=None):
    """-> (files: {name: [Read]}, stats_text, read_families_text).

    `reads` is the record list `packed` was built from.  File names follow
    App. A.6 without the sample prefix, e.g. "sscs.sorted", "dcs.sc.sorted".
    """
    f = packed.fields
    src, lens = packed.src, packed.lens
    nf = res.n_fam
    tmpl = res.fam_tmpl.astype(np.int64)
    size = res.fam_size
    partner = res.partner
    lpad = res.lpad
    # the two families a consensus pair name owns (SSCS_maker.py:275 needs exactly two)
    mate_fam = res.read_fam[tmpl ^ 1].astype(np.int64) if nf else np.zeros(0, np.int64)
    if nf and (np.any(res.fam_tmpl[mate_fam] != (tmpl ^ 1)) or np.any(size[mate_fam] != size)):
        raise DomainError("a read pair does not found/join its two families together "
                          "(orientation edge case, SURVEY.md App. B Q8/Q9): outside the GPU path's domain")

    def rg_of(code):
        return None if code == NO_RG else packed.rgs[code]

    def ctag_of(g):
        k = int(tmpl[g]) >> 1
        read, mate = reads[src[2 * k]], reads[src[2 * k + 1]]
        return keys.sscs_qname(read, mate, packed.barcodes[f["bc"][2 * k]],
                               packed.cigar_pairs[f["cig"][2 * k]])

    ctag = {}

    def family_ctag(g):
        k = int(tmpl[g]) >> 1
        if k not in ctag:
            ctag[k] = ctag_of(g)
        return ctag[k]

    def consensus_record(g, qname, rows, row, mapq, tlen, flag, rg, tpl):
        n = len(tpl.seq)
        seq, qual = _row_seq_qual(rows, row, lpad, n)
        return Read(qname, flag, tpl.tid, tpl.pos, mapq, tpl.cigar, tpl.mtid, tpl.mpos, tlen,
                    seq, qual, rg=rg, aux=b"")

    # ---- SSCS stage ---------------------------------------------------------
    arrays = {k: getattr(res, k) for k in ("fam_size", "fam_tmpl", "fam_mapq", "fam_flag", "fam_tlen", "fam_rg",
                                           "sscs_row", "sscs_rows", "read_fam")}
    sscs_pairs, singleton_pairs = sscs_stage_records(reads, packed, arrays, lpad, with_family=True)
    sscs = [_Out(r, g) for r, g in sscs_pairs]
    singleton = [_Out(r, g) for r, g in singleton_pairs]
    sscs_sorted, singleton_sorted = _sorted(sscs, sorter), _sorted(singleton, sorter)

    def dcs_stage(inp, in_set, row_of, rows, mq, fl, tl, rgc):
        dcs, single = [], []
        made = set()
        for a, b in _pair_walk(inp, mate_fam):
            for o in (a, b):
                g = o.fam
                p = int(partner[g])
                if p in made:                       # `ds in duplex_dict`: the partner already led
                    continue
                if p >= 0 and in_set(p):
                    row = int(row_of[g])
                    if row < 0:
                        raise DomainError("duplex strands met in an order the GPU path did not predict")
                    other = rec_of[p]
                    qn = keys.dcs_consensus_tag(o.rec.qname, other.qname)
                    dcs.append(_Out(consensus_record(g, qn, rows, row, int(mq[row]), int(tl[row]),
                                                     int(fl[row]), rg_of(rgc[row]), o.rec), g))
                    made.add(g)
                else:
                    single.append(o)
        return dcs, single

    rec_of = {o.fam: o.rec for o in sscs_sorted}
    dcs, sscs_single = dcs_stage(sscs_sorted, lambda p: size[p] >= 2, res.dcs_row, res.dcs_rows,
                                 res.dcs_mapq, res.dcs_flag, res.dcs_tlen, res.dcs_rg)
    files = {"sscs.sorted": sscs_sorted, "singleton.sorted": singleton_sorted,
             "dcs.sorted": _sorted(dcs, sorter), "sscs.singleton.sorted": _sorted(sscs_single, sorter)}

    # ---- singleton correction + second DCS pass ---------------------------------
    if scorrect:
        sscs_cor, sing_cor, uncorrected = [], [], []
        for a, b in _pair_walk(singleton_sorted, mate_fam):
            qn = keys.sscs_qname(a.rec, b.rec, a.rec.qname.split("_")[0],
                                 keys.cigar_order(a.rec, b.rec)) + ":1"
            for o in (a, b):
                g = o.fam
                p = int(partner[g])
                if p < 0:
                    uncorrected.append(o)
                    continue
                s = o.rec
                rec = consensus_record(g, qn, res.corr_rows, res.corr_row[g], s.mapq, s.tlen, s.flag, s.rg, s)
                (sscs_cor if size[p] >= 2 else sing_cor).append(_Out(rec, g))
        files["sscs.correction.sorted"] = _sorted(sscs_cor, sorter)
        files["singleton.correction.sorted"] = _sorted(sing_cor, sorter)
        files["uncorrected.sorted"] = _sorted(uncorrected, sorter)
        sscs_sc = _merged(sscs_sorted, files["sscs.correction.sorted"], files["singleton.correction.sorted"],
                          sorter=sorter)
        files["sscs.sc.sorted"] = sscs_sc
        rec_of = {o.fam: o.rec for o in sscs_sc}
        dcs_sc, sscs_sc_single = dcs_stage(sscs_sc, lambda p: p in rec_of, res.dcssc_row, res.dcssc_rows,
                                           res.dcssc_mapq, res.dcssc_flag, res.dcssc_tlen, res.dcssc_rg)
        files["dcs.sc.sorted"] = _sorted(dcs_sc, sorter)
        files["sscs.sc.singleton.sorted"] = _sorted(sscs_sc_single, sorter)
        files["all.unique.dcs.sorted"] = _merged(files["dcs.sc.sorted"], files["sscs.sc.singleton.sorted"],
                                                 files["uncorrected.sorted"], sorter=sorter)
    files["badReads"] = [_Out(reads[i], -1) for i in packed.bad]
    stats = stats_text(packed.counters, {k: len(v) for k, v in files.items()}, scorrect)
    return {k: [o.rec for o in v] for k, v in files.items()}, stats, families_text(res)
