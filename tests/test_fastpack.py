"""CPU: the C++ packer (csrc/ccpack.cpp via fastpack.py) against pack.py, the readable statement of
read_bam's filter / pairing / key rules (consensus_helper.py:376-441), byte for byte."""
import ctypes
import os
import random

import numpy as np
import pytest

from consensuscruncher_b200 import bam, build, fastpack, pack, synth
from consensuscruncher_b200.records import Read
from oracle import golden

CASES = ["duplex_small", "cutoff_050", "mode_ties", "undotted_umi", "crowded", "bed_three_chr", "big_families",
         "no_rg_aux"]


def _same(a, b):
    assert a.n == b.n and a.lpad == b.lpad and a.stride == b.stride
    assert a.fields.tobytes() == b.fields.tobytes()
    assert np.array_equal(a.payload, b.payload)
    assert a.barcodes == b.barcodes and a.cigar_pairs == b.cigar_pairs and a.rgs == b.rgs
    assert np.array_equal(a.dup_bc, b.dup_bc)
    assert np.array_equal(a.src, b.src) and np.array_equal(a.lens, b.lens)
    assert a.counters == b.counters
    assert list(a.bad) == list(b.bad) and list(a.leftover) == list(b.leftover)


def _both(path, stage="sscs", bdelim="|", regions=None):
    header, reads = bam.read_bam(path)
    exp = pack.pack_reads(reads, stage, bdelim, regions)
    with fastpack.BamFile(path, threads=3) as f:
        assert f.header.references == header.references and f.header.lengths == header.lengths
        assert f.header.text == header.text
        assert len(f) == len(reads)
        got = fastpack.pack_reads(f, stage, bdelim, regions)
        for i in list(exp.bad)[:5] + list(exp.src[:6]):
            assert f[i].key() == reads[i].key()
    _same(got, exp)
    return got


def test_library_exports_every_symbol():
    lib = ctypes.CDLL(build.build_pack_lib())
    for s in fastpack.SYMBOLS:
        assert hasattr(lib, s), s


@pytest.mark.parametrize("name", CASES)
def test_golden_inputs_pack_identically(tmp_path, name):
    g = golden.load(name)
    reads = [golden.tuple_to_read(t) for t in g["input"]]
    header = bam.BamHeader("", g["header"]["references"], g["header"]["lengths"])
    path = os.path.join(str(tmp_path), "in.bam")
    bam.write_bam(path, header, reads)
    regions = None
    if g["params"].get("bed"):
        refs = g["header"]["references"]
        regions = [(refs.index(c), a, b) for c, a, b, _ in g["params"]["bed"]]
    _both(path, "sscs", "|", regions)


def _messy_reads(seed, n_pairs=300):
    """Pairs with every filter outcome, odd lengths, multi-op cigars, several read groups, lone mates,
    a third read reusing a name, and no coordinate order."""
    rng = random.Random(seed)
    out = []
    flags = [(99, 147), (83, 163), (163, 83), (147, 99), (97, 145), (65, 129), (81, 161), (113, 177)]
    for k in range(n_pairs):
        umi = "".join(rng.choice("ACGT") for _ in range(2)) + "." + "".join(rng.choice("ACGT") for _ in range(2))
        name = "r%d|%s|x" % (k, umi) if rng.random() < 0.9 else "r%d" % k
        fa, fb = rng.choice(flags)
        tid, mtid = rng.randrange(3), rng.randrange(3)
        if rng.random() < 0.8:
            mtid = tid
        pos, mpos = rng.randrange(50, 400), rng.randrange(50, 400)
        roll = rng.random()
        if roll < 0.05:
            fa, fb = 77, 141
        elif roll < 0.10:
            fa, fb = 73, 133
        elif roll < 0.15:
            fa |= 0x100
        elif roll < 0.18:
            fb |= 0x800
        pair = []
        for flag, t, p, mt, mp in ((fa, tid, pos, mtid, mpos), (fb, mtid, mpos, tid, pos)):
            L = rng.choice([31, 32, 33, 64, 75])
            cig = rng.choice(["%dM" % L, "5S%dM" % (L - 5), "10M2I%dM" % (L - 12), "10M3D%dM" % (L - 10)])
            seq = "".join(rng.choice("ACGTN") for _ in range(L))
            qual = np.array([rng.randrange(2, 41) for _ in range(L)], np.uint8)
            rg = rng.choice([None, "lane1", "lane2", "lane3"])
            pair.append(Read(name, flag, t, p, rng.randrange(61), cig, mt, mp, rng.randrange(-300, 300), seq, qual,
                             rg=rg, aux={"NM": rng.randrange(5), "XS": "text"} if rng.random() < 0.5 else b""))
        if rng.random() < 0.05:
            pair = pair[:1]
        out.extend(pair)
    out.append(Read(out[0].qname, 99, 0, 77, 60, "32M", 0, 99, 54, "A" * 32, np.full(32, 30, np.uint8)))
    rng.shuffle(out)
    return out


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_messy_input_packs_identically(tmp_path, seed):
    reads = _messy_reads(seed)
    header = bam.BamHeader("@HD\tVN:1.6\n", ["chrA", "chrB", "chrC"], [1000, 2000, 3000])
    path = os.path.join(str(tmp_path), "messy.bam")
    bam.write_bam(path, header, reads)
    got = _both(path)
    assert got.bad and got.leftover and got.counters["bad_spacer"]
    _both(path, regions=[(1, 0, 200), (0, 100, 300), (1, 150, 420), (7, 0, 10)])     # overlapping + absent contig
    _both(path, regions=[])
    reads.sort(key=lambda r: (r.tid, r.pos))
    bam.write_bam(path, header, reads)
    _both(path, regions=[(0, 60, 250), (2, 0, 1000)])                                 # sorted: bisection branch


def test_dcs_stage_names_and_no_diversion(tmp_path):
    s = synth.generate(seed=5, n_molecules=60, L=40)
    reads = s.to_reads()
    for r in reads:
        r.qname = r.qname.split("|")[1].replace(".", "") + "_1_100_1_300_40M_40M_pos_200"
    reads[3].flag = 77
    reads[8].flag |= 0x100
    header = bam.BamHeader("", ["chr1", "chr2"], [10 ** 6, 10 ** 6])
    path = os.path.join(str(tmp_path), "d.bam")
    bam.write_bam(path, header, reads)
    header2, reads2 = bam.read_bam(path)
    with fastpack.BamFile(path) as f:
        try:
            exp = pack.pack_reads(reads2, "dcs", "|", None)
        except pack.DomainError:
            with pytest.raises(pack.DomainError):
                fastpack.pack_reads(f, "dcs", "|", None)
            return
        _same(fastpack.pack_reads(f, "dcs", "|", None), exp)


def test_unkeyable_flag_is_a_domain_error(tmp_path):
    a = Read("q|AA.CC", 99, 0, 10, 60, "32M", 0, 50, 72, "A" * 32, np.full(32, 35, np.uint8))
    b = Read("q|AA.CC", 1171, 0, 50, 60, "32M", 0, 10, -72, "A" * 32, np.full(32, 35, np.uint8))   # dup flag set
    path = os.path.join(str(tmp_path), "f.bam")
    bam.write_bam(path, bam.BamHeader("", ["c"], [100]), [a, b])
    with fastpack.BamFile(path) as f:
        with pytest.raises(pack.DomainError):
            fastpack.pack_reads(f)
    with pytest.raises(pack.DomainError):
        pack.pack_reads(bam.read_bam(path)[1])


def test_not_a_bam_is_rejected(tmp_path):
    path = os.path.join(str(tmp_path), "x.bam")
    open(path, "wb").write(b"hello")
    with pytest.raises(ValueError):
        fastpack.BamFile(path)


@pytest.mark.parametrize("name", ["duplex_small", "bed_three_chr", "no_rg_aux"])
def test_lazy_records_feed_the_emit_step(tmp_path, name):
    """oracle result + emit over the lazily decoded BamFile reproduces the reference's golden tree."""
    from oracle import oracle
    from consensuscruncher_b200 import emit
    g = golden.load(name)
    header = bam.BamHeader("", g["header"]["references"], g["header"]["lengths"])
    path = os.path.join(str(tmp_path), "in.bam")
    bam.write_bam(path, header, [golden.tuple_to_read(t) for t in g["input"]])
    regions = None
    if g["params"].get("bed"):
        refs = g["header"]["references"]
        regions = [(refs.index(c), a, b) for c, a, b, _ in g["params"]["bed"]]
    with fastpack.BamFile(path) as f:
        packed = fastpack.pack_reads(f, "sscs", "|", regions)
        res = oracle.run(packed, g["params"]["cutoff"], g["params"]["scorrect"])
        files, stats, fam = emit.emit_all(f, packed, res, g["params"]["scorrect"])
    for fname, recs in g["files"].items():       # aux travels as raw bytes here, so compare the 12 SAM fields + RG
        exp = [golden.tuple_to_read(t).key()[:12] for t in recs]
        assert [r.key()[:12] for r in files.get(fname, [])] == exp, fname
    assert stats == g["stats"] and fam == g["read_families"]
