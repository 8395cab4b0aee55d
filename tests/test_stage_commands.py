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


def test_record_join_equals_the_dictionary_probe(engine):
    """cc_join_records against what the stage scripts do with strings: partner = the first record whose tuple is this
    record's with the barcode swapped (through the table) and R1 <-> R2; records without one, records whose barcode
    has no counterpart, several records under one key."""
    import numpy as np
    from consensuscruncher_b200 import pack
    rng = np.random.default_rng(12)
    n, n_bc = 20000, 37
    f = np.zeros(n, pack.FIELDS_DT)
    f["bc"] = rng.integers(0, n_bc, n)
    f["cig"] = rng.integers(0, 3, n)
    f["tid"] = rng.integers(0, 2, n)
    f["pos"] = rng.integers(0, 40, n)
    f["mtid"] = f["tid"]
    f["mpos"] = f["pos"] + rng.integers(0, 3, n)
    f["flag"] = rng.choice(np.array([99, 147, 83, 163], np.uint16), n)
    f["tlen"], f["mapq"] = rng.integers(-9, 9, n), 60             # not part of the key
    dup = rng.permutation(n_bc).astype(np.uint32)
    dup[rng.random(n_bc) < 0.2] = pack.NO_ID
    got = engine.join_records(f, dup)
    key = lambda r, bc, bits: (int(bc), int(r["cig"]), int(r["tid"]), int(r["pos"]), int(r["mtid"]), int(r["mpos"]), bits)
    bits = lambda r: (int(r["flag"]) >> 4 & 1, int(r["flag"]) >> 6 & 1)
    first = {}
    for i in range(n):
        first.setdefault(key(f[i], f[i]["bc"], bits(f[i])), i)
    exp = np.full(n, -1, np.int32)
    for i in range(n):
        d = dup[f[i]["bc"]]
        if d != pack.NO_ID:
            o, m = bits(f[i])
            exp[i] = first.get(key(f[i], d, (o, m ^ 1)), -1)
    assert (exp >= 0).sum() > 1000 and (exp < 0).sum() > 1000
    assert np.array_equal(got, exp)
    assert len(engine.join_records(f[:0], dup)) == 0
