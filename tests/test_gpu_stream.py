"""GPU: the chunked / multi-GPU walk of one sample (stream.py, `consensus --gpus N --chunk-reads R`)."""
import numpy as np
import pytest

from consensuscruncher_b200 import bam, cli, stream, synth
from consensuscruncher_b200.engine import GpuConsensus
from consensuscruncher_b200.pack import DomainError, PackedReads
from oracle import golden, oracle

import helpers

pytestmark = pytest.mark.gpu


def test_chunked_walk_equals_one_pass(engine):
    torch = pytest.importorskip("torch")
    s = synth.generate(seed=27, n_molecules=20000, L=64, n_tid=4, region=(0, 400000), fam_mean=3.0, p_duplex=0.7)
    packed = s.to_packed()
    engine.set_params(cutoff=0.7, scorrect=True)
    whole = engine.consensus(packed)
    helpers.assert_same_result(whole, oracle.run(packed, 0.7, True))
    regions = [(t, a, a + 20000) for t in range(4) for a in range(0, 400000, 20000)]
    engines = [engine]
    extra = None
    if torch.cuda.device_count() > 1:                       # a second GPU takes every other chunk
        extra = GpuConsensus(1, cutoff=0.7, scorrect=True)
        engines.append(extra)
    try:
        for regs, budget in ((regions, 12000), (None, 9000)):
            res, kept, n_chunks = stream.consensus_chunked(packed, engines, regs, budget)
            assert n_chunks > 5 and kept is packed
            helpers.assert_same_result(res, whole)
    finally:
        if extra:
            extra.close()


def test_pairs_outside_the_domain_are_set_aside_on_the_gpu(engine):
    packed = synth.generate(seed=19, n_molecules=400, L=32, p_duplex=1.0).to_packed()
    f = packed.fields.copy()
    k = int(np.flatnonzero((f["flag"][0::2] == 99))[5])
    twin_f = f[2 * k:2 * k + 2].copy()
    twin_f["flag"] = [67, 131]
    fields = np.concatenate([f[:2 * k + 2], twin_f, f[2 * k + 2:]])
    payload = np.concatenate([packed.payload[:2 * k + 2], packed.payload[2 * k:2 * k + 2], packed.payload[2 * k + 2:]])
    odd = PackedReads(fields, payload, packed.lpad, packed.barcodes, packed.dup_bc, packed.cigar_pairs, packed.rgs,
                      src=np.arange(fields.shape[0], dtype=np.int64), lens=np.full(fields.shape[0], 32, np.int32),
                      counters=dict(packed.counters))
    engine.set_params(cutoff=0.7, scorrect=True)
    with pytest.raises(DomainError):
        engine.consensus(odd)
    notes = []
    res, kept, _ = stream.consensus_chunked(odd, [engine], None, 10 ** 9, log=notes.append)
    # every pair with a read in one of the three families that no longer pair up goes (the twin's family mates too)
    assert notes and odd.n - 40 <= kept.n < odd.n and len(kept.bad) == odd.n - kept.n
    helpers.assert_same_result(res, oracle.run(kept, 0.7, True))


@pytest.mark.parametrize("name", ["bed_three_chr", "duplex_small"])
def test_consensus_command_in_chunks_reproduces_the_reference_tree(tmp_path, name, engine):
    g = golden.load(name)
    reads = [golden.tuple_to_read(t) for t in g["input"]]
    header = bam.BamHeader("@HD\tVN:1.0\tSO:coordinate\n", g["header"]["references"], g["header"]["lengths"])
    inp = str(tmp_path / "sample.bam")
    bam.write_bam(inp, header, reads)
    bed = "False"
    if g["params"].get("bed"):
        bed = str(tmp_path / "regions.bed")
        with open(bed, "w") as fh:
            for c, a, b, nm in g["params"]["bed"]:
                fh.write("%s\t%d\t%d\t%s\tgneg\n" % (c, a, b, nm))
    trees = {}
    for label, chunk in (("one", 10 ** 9), ("many", 40)):
        out = tmp_path / label
        out.mkdir()
        cli.run_consensus(inp, str(out), cutoff=g["params"]["cutoff"], scorrect=g["params"]["scorrect"], bedfile=bed,
                          engine=engine, log=lambda *_: None, chunk_reads=chunk)
        trees[label] = {p.relative_to(out): p.read_bytes() for p in sorted(out.rglob("*")) if p.is_file()
                        and not p.name.endswith("time_tracker.txt")}
    assert trees["one"].keys() == trees["many"].keys() and len(trees["one"]) > 10
    for k in trees["one"]:
        assert trees["one"][k] == trees["many"][k], k          # byte-identical files however the sample was cut
    assert (tmp_path / "many" / "sample" / "sample.stats.txt").read_text() == g["stats"]
    assert (tmp_path / "many" / "sample" / "sample.read_families.txt").read_text() == g["read_families"]
