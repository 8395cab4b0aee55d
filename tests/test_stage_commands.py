"""GPU: the three stage commands (SSCS_maker / DCS_maker / singleton_correction) chained the way
`ConsensusCruncher.py:148-303` chains them, against the reference's golden output tree."""
import os

import pytest

from consensuscruncher_b200 import bam, stages
from oracle import golden

pytestmark = pytest.mark.gpu


def _write_bed(path, bed):
    with open(path, "w") as fh:
        for c, a, b, nm in bed:
            fh.write("%s\t%d\t%d\t%s\tgneg\n" % (c, a, b, nm))


@pytest.mark.parametrize("name", ["c1prime_real_fastq", "duplex_small", "mode_ties", "bed_three_chr", "crowded", "undotted_umi"])
def test_stage_chain_reproduces_reference_tree(tmp_path, name, engine):
    g = golden.load(name)
    reads = [golden.tuple_to_read(t) for t in g["input"]]
    header = bam.BamHeader("", g["header"]["references"], g["header"]["lengths"])
    d = str(tmp_path)
    inp = os.path.join(d, "s.bam")
    bam.write_bam(inp, header, reads)
    bed = None
    if g["params"].get("bed"):
        bed = os.path.join(d, "r.bed")
        _write_bed(bed, g["params"]["bed"])
    sscs = os.path.join(d, "s.sscs.bam")
    stages.run_sscs(inp, sscs, g["params"]["cutoff"], "|", bed, engine=engine)
    sscs = stages.sort_index(sscs, engine=engine)
    sing = stages.sort_index(os.path.join(d, "s.singleton.bam"))
    dcs = os.path.join(d, "s.dcs.bam")
    stages.run_dcs(sscs, dcs, bed, engine=engine)
    out = {"sscs.sorted": sscs, "singleton.sorted": sing, "dcs.sorted": stages.sort_index(dcs),
           "sscs.singleton.sorted": stages.sort_index(os.path.join(d, "s.sscs.singleton.bam"))}
    stages.run_sc(sing, bed, engine=engine)
    for nm in ("sscs.correction", "singleton.correction", "uncorrected"):
        out[nm + ".sorted"] = stages.sort_index(os.path.join(d, "s.%s.bam" % nm))
    sc = stages.merge_bams(os.path.join(d, "s.sscs.sc.bam"), sscs, out["sscs.correction.sorted"],
                           out["singleton.correction.sorted"], engine=engine)
    out["sscs.sc.sorted"] = stages.sort_index(sc)
    dcs_sc = os.path.join(d, "s.dcs.sc.bam")
    stages.run_dcs(out["sscs.sc.sorted"], dcs_sc, bed, engine=engine)
    out["dcs.sc.sorted"] = stages.sort_index(dcs_sc)
    out["sscs.sc.singleton.sorted"] = stages.sort_index(os.path.join(d, "s.sscs.sc.singleton.bam"))
    allu = stages.merge_bams(os.path.join(d, "s.all.unique.dcs.bam"), out["dcs.sc.sorted"],
                             out["sscs.sc.singleton.sorted"], out["uncorrected.sorted"])
    out["all.unique.dcs.sorted"] = stages.sort_index(allu)
    for fname, recs in g["files"].items():
        if fname == "badReads":
            continue
        _, got = bam.read_bam(out[fname])
        exp = [golden.tuple_to_read(t).key()[:12] for t in recs]
        assert [r.key()[:12] for r in got] == exp, fname
    assert open(os.path.join(d, "s.stats.txt")).read() == g["stats"]
    assert open(os.path.join(d, "s.read_families.txt")).read() == g["read_families"]
