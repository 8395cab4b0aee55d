"""CPU: BGZF/BAM/BAI codec round trips and the command-line front end's argument handling.
GPU: the whole `consensus` command against the reference's golden output tree."""
import gzip
import os
import struct

import numpy as np
import pytest

from consensuscruncher_b200 import bam, cli, synth
from consensuscruncher_b200.records import Read, sort_records
from oracle import golden


def _reads():
    s = synth.generate(seed=3, n_molecules=40, L=37, fam_mean=2.0, region=(0, 40000), n_tid=2)
    reads = s.to_reads()
    reads[0].aux = b"NMi\x02\x00\x00\x00XSZhello\x00"
    reads[1].rg = None
    reads.append(Read("unmapped|AA.CC", 77, -1, -1, 0, "", -1, -1, 0, "ACGTN", [2, 3, 4, 5, 6]))
    return bam.BamHeader("@HD\tVN:1.0\tSO:coordinate\n@SQ\tSN:chr1\tLN:50000\n@SQ\tSN:chr2\tLN:50000\n@RG\tID:1\n",
                         ["chr1", "chr2"], [50000, 50000]), reads


def test_bam_round_trip(tmp_path):
    header, reads = _reads()
    p = str(tmp_path / "a.bam")
    bam.write_bam(p, header, reads)
    h2, r2 = bam.read_bam(p)
    assert h2.text == header.text and h2.references == header.references and h2.lengths == header.lengths
    assert [r.key() for r in r2] == [r.key() for r in reads]
    raw = open(p, "rb").read()
    assert raw.endswith(bytes.fromhex("1f8b08040000000000ff0600424302001b0003000000000000000000"))
    assert gzip.decompress(raw)[:4] == b"BAM\x01"                  # BGZF is a valid multi-member gzip


def test_large_bam_spans_many_blocks(tmp_path):
    s = synth.generate(seed=4, n_molecules=600, L=100, region=(0, 200000))
    reads = s.to_reads()
    header = bam.BamHeader("", s.ref_names, s.ref_lengths)
    p = str(tmp_path / "b.bam")
    bam.write_bam(p, header, reads, sort_order="coordinate", index=True)
    assert os.path.getsize(p) > 3 * 65536 // 4
    h2, r2 = bam.read_bam(p)
    assert h2.text.startswith("@HD\tVN:1.6\tSO:coordinate")
    assert [r.key() for r in r2] == [r.key() for r in reads]


def test_bai_structure_and_counts(tmp_path):
    s = synth.generate(seed=5, n_molecules=300, L=50, region=(0, 400000), n_tid=2)
    reads = sort_records(s.to_reads())
    header = bam.BamHeader("", s.ref_names, s.ref_lengths)
    p = str(tmp_path / "c.bam")
    bam.write_bam(p, header, reads, sort_order="coordinate", index=True)
    idx = open(p + ".bai", "rb").read()
    assert idx[:4] == b"BAI\x01"
    n_ref = struct.unpack_from("<i", idx, 4)[0]
    assert n_ref == 2
    off, mapped = 8, []
    for _ in range(n_ref):
        n_bin = struct.unpack_from("<i", idx, off)[0]
        off += 4
        for _ in range(n_bin):
            b, n_chunk = struct.unpack_from("<Ii", idx, off)
            off += 8
            if b == 37450:
                mapped.append(struct.unpack_from("<QQQQ", idx, off)[2])
            else:
                assert b < 37450
                for k in range(n_chunk):
                    beg, end = struct.unpack_from("<QQ", idx, off + 16 * k)
                    assert beg < end
            off += 16 * n_chunk
        n_intv = struct.unpack_from("<i", idx, off)[0]
        ivs = struct.unpack_from("<%dQ" % n_intv, idx, off + 4)
        assert list(ivs) == sorted(ivs)
        off += 4 + 8 * n_intv
    assert struct.unpack_from("<Q", idx, off)[0] == 0 and off + 8 == len(idx)
    assert mapped == [sum(1 for r in reads if r.tid == t) for t in range(2)]
    assert bam.reg2bin(0, 1) == 4681 and bam.reg2bin(1 << 14, (1 << 14) + 1) == 4682 and bam.reg2bin(0, 1 << 29) == 0


def test_cli_argument_handling(tmp_path, capsys):
    assert cli.main([]) == 0                                       # prints help like the reference
    assert cli.main(["consensus"]) == 0                            # required arguments missing -> help
    with pytest.raises(SystemExit):                                # missing input: an argparse error, no traceback
        cli.main(["consensus", "-i", "x.bam", "-o", str(tmp_path)])
    assert cli.default_bedfile("hg19") and cli.default_bedfile("hg38")   # what `-b` / `-g hg38` default to
    cfg = tmp_path / "c.ini"
    cfg.write_text("[consensus]\nbam = from_config.bam\nc_output = %s\ncutoff = 0.9\nbedfile = False\n" % tmp_path)
    main_p, _ = cli.build_parser(str(cfg))
    a = main_p.parse_args(["-c", str(cfg), "consensus", "--cutoff", "0.5"])
    assert a.bam == "from_config.bam" and float(a.cutoff) == 0.5 and a.bedfile == "False" and a.scorrect == "True"


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["c1prime_real_fastq", "duplex_small", "bed_three_chr", "no_scorrect", "no_rg_aux"])
def test_consensus_command_reproduces_reference_tree(tmp_path, name, engine):
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
    out = tmp_path / "out"
    out.mkdir()
    cli.run_consensus(inp, str(out), cutoff=g["params"]["cutoff"], scorrect=g["params"]["scorrect"], bedfile=bed,
                      engine=engine, log=lambda *_: None)
    sub = {"sscs.sorted": "sscs", "singleton.sorted": "sscs", "badReads": "sscs", "dcs.sorted": "dcs",
           "sscs.singleton.sorted": "dcs", "sscs.correction.sorted": "sscs_sc", "singleton.correction.sorted": "sscs_sc",
           "uncorrected.sorted": "sscs_sc", "sscs.sc.sorted": "sscs_sc", "dcs.sc.sorted": "dcs_sc",
           "sscs.sc.singleton.sorted": "dcs_sc", "all.unique.dcs.sorted": "dcs_sc"}
    for fname, recs in g["files"].items():
        path = out / "sample" / sub[fname] / ("sample.%s.bam" % fname)
        assert path.exists(), fname
        _, got = bam.read_bam(str(path))
        exp = [golden.tuple_to_read(t) for t in recs]
        assert len(got) == len(exp), fname
        for a, b in zip(got, exp):
            ka, kb = a.key(), b.key()
            assert ka[:12] == kb[:12], (fname, ka, kb)            # aux compared below (raw bytes vs dict)
        if fname.endswith(".sorted"):
            assert (out / "sample" / sub[fname] / ("sample.%s.bam.bai" % fname)).exists()
    assert (out / "sample" / "sample.stats.txt").read_text() == g["stats"]
    assert (out / "sample" / "sample.read_families.txt").read_text() == g["read_families"]
    with pytest.raises(FileExistsError):                           # like the reference, a re-run needs a clean directory
        cli.run_consensus(inp, str(out), bedfile="False", engine=engine, log=lambda *_: None)
