"""CPU: the C++ emit + BAM/BAI writer (csrc/ccemit.cpp via fastpack.emit_files) against emit.py +
bam.write_bam, the readable statement of the reference's stage outputs."""
import os
import struct
import zlib

import numpy as np
import pytest

from consensuscruncher_b200 import bam, emit, fastpack, synth
from oracle import golden, oracle


def _bai_from_bam(path, n_ref):
    """The index bam.py's builder makes for the records of `path` at the virtual offsets they have there."""
    raw = open(path, "rb").read()
    blocks, off, uoff = [], 0, 0               # (compressed offset, uncompressed offset, uncompressed size)
    stream = bytearray()
    while off < len(raw):
        bsize = struct.unpack_from("<H", raw, off + 16)[0] + 1
        chunk = zlib.decompress(raw[off + 18:off + bsize - 8], -15)
        blocks.append((off, uoff, len(chunk)))
        stream += chunk
        uoff += len(chunk)
        off += bsize
    data = bytes(stream)

    def voff(u):                               # BgzfWriter.tell() once the bytes before u are written
        for coff, start, size in blocks:
            if start <= u < start + size:
                return (coff << 16) | (u - start)
        coff, start, size = [b for b in blocks if b[2]][-1]
        if size < bam.BgzfWriter.BLOCK:         # end of the data inside a block that is still open
            return (coff << 16) | size
        return blocks[-1][0] << 16              # ... or exactly at a block boundary: the EOF block

    header, reads = bam.read_bam(path)
    p = 12 + struct.unpack_from("<i", data, 4)[0]
    for _ in range(n_ref):
        p += 8 + struct.unpack_from("<i", data, p)[0]
    builder = bam._BaiBuilder(n_ref)
    for r in reads:
        size = 4 + struct.unpack_from("<i", data, p)[0]
        builder.add(r, voff(p), voff(p + size))
        p += size
    return builder.dump()


def _compare(tmp_path, header, reads, regions, cutoff, scorrect):
    d = str(tmp_path)
    inp = os.path.join(d, "in.bam")
    bam.write_bam(inp, header, reads)
    with fastpack.BamFile(inp, threads=3) as f:
        packed = fastpack.pack_reads(f, "sscs", "|", regions)
        res = oracle.run(packed, cutoff, scorrect)
        files, stats, fam = emit.emit_all(f, packed, res, scorrect)
        names = [k for k in fastpack.FILES if k in files]
        paths = {k: os.path.join(d, "native.%s.bam" % k) for k in names}
        n, stats2, fam2 = fastpack.emit_files(f, packed, res, scorrect, paths)
        assert stats2 == stats and fam2 == fam
        for k in names:
            assert n[k] == len(files[k]), k
            ref = os.path.join(d, "python.%s.bam" % k)
            is_sorted = k.endswith(".sorted")
            bam.write_bam(ref, f.header, files[k], sort_order="coordinate" if is_sorted else None, index=is_sorted)
            h1, r1 = bam.read_bam(ref)
            h2, r2 = bam.read_bam(paths[k])
            assert (h1.text, h1.references, h1.lengths) == (h2.text, h2.references, h2.lengths), k
            assert [r.key() for r in r1] == [r.key() for r in r2], k
            assert bam.bgzf_decompress(open(ref, "rb").read()) == bam.bgzf_decompress(open(paths[k], "rb").read()), k
            if is_sorted:
                assert open(paths[k] + ".bai", "rb").read() == _bai_from_bam(paths[k], len(h2.references)), k
                if open(ref, "rb").read() == open(paths[k], "rb").read():      # same zlib behind both writers
                    assert open(ref + ".bai", "rb").read() == open(paths[k] + ".bai", "rb").read(), k
            else:
                assert not os.path.exists(paths[k] + ".bai")
        for k in fastpack.FILES:
            if k not in files:
                assert n[k] == 0
    return n


@pytest.mark.parametrize("name", ["c1prime_real_fastq", "duplex_small", "mode_ties", "undotted_umi", "no_scorrect", "crowded", "bed_three_chr",
                                  "big_families", "no_rg_aux"])
def test_native_emit_equals_python_emit_on_golden_inputs(tmp_path, name):
    g = golden.load(name)
    reads = [golden.tuple_to_read(t) for t in g["input"]]
    header = bam.BamHeader("@HD\tVN:1.0\tSO:unsorted\n@RG\tID:1\n", g["header"]["references"], g["header"]["lengths"])
    regions = None
    if g["params"].get("bed"):
        refs = g["header"]["references"]
        regions = [(refs.index(c), a, b) for c, a, b, _ in g["params"]["bed"]]
    _compare(tmp_path, header, reads, regions, g["params"]["cutoff"], g["params"]["scorrect"])


@pytest.mark.parametrize("kw", [dict(seed=3, n_molecules=1500, L=101), dict(seed=4, n_molecules=400, L=150, n_tid=3),
                                dict(seed=5, n_molecules=3000, L=64, region=(0, 300000), mapq_mixed=True)])
def test_native_emit_equals_python_emit_on_synthetic(tmp_path, kw):
    s = synth.generate(**kw)
    reads = s.to_reads()
    header = bam.BamHeader("", ["chr%d" % (i + 1) for i in range(3)], [10 ** 7] * 3)
    n = _compare(tmp_path, header, reads, None, 0.7, True)
    assert n["dcs.sorted"] and n["sscs.correction.sorted"] and n["all.unique.dcs.sorted"]


def test_bad_reads_and_aux_tags_pass_through(tmp_path):
    """Diverted records (badReads.bam) and singletons keep every field and tag; RG moves to the front."""
    from consensuscruncher_b200.records import Read
    s = synth.generate(seed=6, n_molecules=300, L=50, fam_mean=1.6)
    reads = s.to_reads()
    for i, r in enumerate(reads):
        r.aux = {"NM": i % 4, "XA": "alt%d" % (i % 3)} if i % 2 else b""
        if i % 7 == 0:
            r.rg = None
    q = np.full(50, 30, np.uint8)
    extra = [Read("nodelim", 99, 0, 500, 60, "50M", 0, 700, 250, "A" * 50, q, rg="1"),
             Read("nodelim", 147, 0, 700, 60, "50M", 0, 500, -250, "C" * 50, q, rg="1"),
             Read("u|AA.CC", 77, -1, -1, 0, "", -1, -1, 0, "G" * 50, q, aux={"XS": "unmapped"}),
             Read("u|AA.CC", 141, -1, -1, 0, "", -1, -1, 0, "T" * 50, q),
             Read("m|AC.GT", 73, 0, 900, 60, "50M", 0, 900, 0, "A" * 50, q, rg="1"),
             Read("s|AC.GT", 99 | 0x100, 0, 950, 3, "20S30M", 0, 990, 90, "A" * 50, q, rg="1")]
    reads = sorted(reads + extra, key=lambda r: (r.tid if r.tid >= 0 else 1 << 30, r.pos))
    header = bam.BamHeader("@HD\tVN:1.6\tSO:coordinate\n", ["chr1"], [10 ** 7])
    n = _compare(tmp_path, header, reads, None, 0.7, True)
    assert n["badReads"] == 6 and n["singleton.sorted"]
