"""CPU: host-side mirror of the reference's helper functions, packing, synthetic data."""
import numpy as np
import pytest

from consensuscruncher_b200 import keys, pack, synth
from consensuscruncher_b200.records import Read, merge_records, sort_records


def _r(flag, tid, pos, mtid, mpos, cigar="98M", tlen=100, qname="x|AAAA"):
    return Read(qname, flag, tid, pos, 60, cigar, mtid, mpos, tlen, "A" * 5, [30] * 5)


def test_which_read_docstring():                      # consensus_helper.py:62-67
    assert keys.which_read(83) == "R1"
    assert keys.which_read(131) == "R2"
    assert keys.which_read(177) == "R2"
    assert keys.which_read(1024 + 99) is None


def test_which_strand_vectors():
    for f in (147, 67, 131, 99):
        assert keys.which_strand(_r(f, 1, 10, 1, 50)) == "pos"
    for f in (83, 163, 115, 179):
        assert keys.which_strand(_r(f, 1, 10, 1, 50)) == "neg"
    for f in (81, 97, 65, 113):                       # R1: pos iff (tid,pos) < (mtid,mpos)
        assert keys.which_strand(_r(f, 12, 10, 16, 5)) == "pos"
        assert keys.which_strand(_r(f, 16, 10, 12, 5)) == "neg"
        assert keys.which_strand(_r(f, 3, 10, 3, 50)) == "pos"
        assert keys.which_strand(_r(f, 3, 50, 3, 10)) == "neg"
    for f in (161, 145, 129, 177):                    # R2: the mirror image
        assert keys.which_strand(_r(f, 12, 10, 16, 5)) == "neg"
        assert keys.which_strand(_r(f, 16, 10, 12, 5)) == "pos"
    assert keys.which_strand(_r(97, 3, 10, 3, 10)) == "neg"   # equal coordinates
    assert keys.which_strand(_r(1024, 3, 10, 3, 10)) is None


def test_key_string_narratives():                     # consensus_helper.py:177-183, :210-220, :276-284
    r1 = _r(99, 24, 58847416, 24, 58847448, "137M10S", 148, "q|TTTG")
    r2 = _r(147, 24, 58847448, 24, 58847416, "147M", -148, "q|TTTG")
    for a, b in ((r1, r2), (r2, r1)):
        cig = keys.cigar_order(a, b)
        assert cig == "137M10S_147M"
        assert keys.sscs_qname(a, b, "TTTG", cig) == "TTTG_24_58847416_24_58847448_137M10S_147M_pos_148"
    assert keys.unique_tag(r1, "TTTG", "137M10S_147M") == "TTTG_24_58847416_24_58847448_137M10S_147M_fwd_R1"
    assert keys.unique_tag(r2, "TTTG", "137M10S_147M") == "TTTG_24_58847448_24_58847416_137M10S_147M_rev_R2"
    n1 = _r(83, 24, 58847448, 24, 58847416, "147M", -148)
    n2 = _r(163, 24, 58847416, 24, 58847448, "137M10S", 148)
    assert keys.cigar_order(n1, n2) == "137M10S_147M" == keys.cigar_order(n2, n1)


def test_duplex_tag_actual_outputs():                 # SURVEY.md App. D 7 (docstring is stale)
    assert keys.duplex_tag("GTCT_1_1507809_7_55224319_98M_98M_fwd_R1") == "CTGT_1_1507809_7_55224319_98M_98M_fwd_R2"
    assert keys.duplex_tag("AC.GT_1_2_3_4_5M_5M_fwd_R1") == "GT.AC_1_2_3_4_5M_5M_fwd_R2"
    assert keys.duplex_tag("ACG.T_1_2_3_4_5M_5M_rev_R2") == "T.ACG_1_2_3_4_5M_5M_rev_R1"
    assert keys.duplex_barcode("ACGTA") == "GTAAC"      # odd, undotted: not an involution (Q11)


def test_dcs_consensus_tag_docstring():               # DCS_maker.py:67-74
    f = keys.dcs_consensus_tag
    assert f("TTCA_7_55259315_7_55259454_98M_98M_neg:3", "CATT_7_55259315_7_55259454_98M_98M_pos:6") == \
        "CATT_TTCA_7_55259315_7_55259454_98M_98M:6_3"
    assert f("CTTC_23_74804535_23_74804611_98M_98M_pos:2", "TCCT_23_74804535_23_74804611_98M_98M_neg:3") == \
        "CTTC_TCCT_23_74804535_23_74804611_98M_98M:2_3"
    assert f("TTTC_7_140477735_7_140477790_98M_98M_neg:3", "TCTT_7_140477735_7_140477790_98M_98M_pos:2") == \
        "TCTT_TTTC_7_140477735_7_140477790_98M_98M:2_3"


def test_bed_separator_reads_cytoband(tmp_path):
    p = tmp_path / "b.txt"
    p.write_text("chr1\t0\t2300000\tp36.33\tgneg\nchr1\t2300000\t5400000\tp36.32\tgpos25\n")
    assert keys.bed_separator(str(p)) == {"chr1_p36.33": (0, 2300000), "chr1_p36.32": (2300000, 5400000)}


def test_row_roundtrip_and_geometry():
    rng = np.random.default_rng(0)
    for n in (1, 31, 32, 33, 100, 128, 150, 151):
        lpad, stride = pack.row_geometry(n)
        assert lpad % 32 == 0 and stride % 64 == 0 and stride >= lpad * 3 // 2 and lpad >= n
        codes = rng.integers(0, 16, n).astype(np.uint8)
        qual = rng.integers(0, 94, n).astype(np.uint8)
        row = np.full(stride, 0xAB, np.uint8)
        pack.pack_row(codes, qual, lpad, row)
        c2, q2 = pack.unpack_row(row, lpad, n)
        assert (c2 == codes).all() and (q2 == qual).all()
        assert (row[lpad // 2 + n:] == 0).all()


def test_filter_and_pairing_counters():
    good1 = _r(99, 0, 10, 0, 60, qname="a|AA.CC")
    good2 = _r(147, 0, 60, 0, 10, qname="a|AA.CC")
    nospacer = _r(99, 0, 11, 0, 61, qname="nodelim")
    unmapped = Read("u|AA.CC", 77, -1, -1, 0, "", -1, -1, 0, "A", [2])
    mate_unmapped = _r(73, 0, 12, 0, 12, qname="m|AA.CC")
    secondary = _r(99 | 256, 0, 13, 0, 63, qname="s|AA.CC")
    lonely = _r(99, 0, 14, 0, 64, qname="l|AA.CC")
    reads = [good1, nospacer, unmapped, mate_unmapped, secondary, lonely, good2]
    p = pack.pack_reads(reads)
    assert p.n == 2 and list(p.src) == [0, 6]
    assert p.counters == dict(total=6, unmapped=1, unmapped_mate=1, multiple_mapping=1, bad_spacer=1)
    assert p.bad == [1, 2, 3, 4] and p.leftover == [5]
    assert p.barcodes == ["AA.CC"] and p.dup_bc[0] == pack.NO_ID
    with pytest.raises(pack.DomainError):
        pack.pack_reads([_r(99 | 1024, 0, 1, 0, 2, qname="d|AA.CC"), _r(147, 0, 2, 0, 1, qname="d|AA.CC")])


def test_region_visit_order_matches_fetch_filter():
    reads = [_r(99, 0, p, 0, p + 50, qname="q%d|AA.CC" % p) for p in (0, 5, 10, 15, 20)]
    assert list(pack.visit_order(reads, [(0, 10, 20), (0, 0, 10), (1, 0, 100)])) == [2, 3, 0, 1]


def test_synth_packed_equals_packed_records():
    s = synth.generate(seed=5, n_molecules=120, L=40, fam_mean=2.5, region=(0, 3000))
    direct = s.to_packed()
    via = pack.pack_reads(s.to_reads())
    assert direct.n == via.n == s.n
    assert (direct.fields == via.fields).all()
    assert (direct.payload == via.payload).all()
    assert direct.barcodes == via.barcodes and (direct.dup_bc == via.dup_bc).all()
    assert direct.cigar_pairs == via.cigar_pairs and (direct.src == via.src).all()


def test_sort_and_merge_model():
    a = _r(163, 0, 10, 0, 60, qname="a")
    b = _r(99, 0, 10, 0, 60, qname="b")
    c = _r(83, 0, 10, 0, 0, qname="c")
    d = _r(99, 0, 5, 0, 60, qname="d")
    u = Read("u", 77, -1, -1)
    assert [r.qname for r in sort_records([u, a, c, b, d])] == ["d", "a", "b", "c", "u"]
    assert [r.qname for r in merge_records([b], [a], [d])] == ["d", "b", "a"]


def test_q4_rows_round_trip():
    s = synth.generate(seed=9, n_molecules=120, L=75)
    packed = s.to_packed()
    q4 = pack.to_q4(packed)
    assert q4 is not None
    payload4, table = q4
    assert table[0] == 0 and payload4.shape[1] % 16 == 0 and payload4.shape[1] >= packed.lpad
    assert np.array_equal(pack.from_q4(payload4, table, packed.lpad), packed.payload)
    # 16 distinct non-zero qualities do not fit (code 0 is reserved for the padding)
    s.qual[:, :] = (np.arange(s.qual.size).reshape(s.qual.shape) % 16 + 20).astype(np.uint8)
    assert pack.to_q4(s.to_packed()) is None
    s.qual[:, :] = (np.arange(s.qual.size).reshape(s.qual.shape) % 15 + 20).astype(np.uint8)
    p2 = s.to_packed()
    payload4, table = pack.to_q4(p2)
    assert np.array_equal(pack.from_q4(payload4, table, p2.lpad), p2.payload)


# ---- the shard map: the cytoband tables `-b` defaults to (ConsensusCruncher.py:382-384, :143-145) ----
@pytest.mark.parametrize("genome", ["hg19", "hg38"])
def test_cytoband_tables_are_863_contiguous_bands(genome):
    from consensuscruncher_b200 import cli
    path = cli.default_bedfile(genome)
    assert path is not None and path.endswith("%s_cytoBand.txt" % genome)
    bands = keys.bed_separator(path)                   # consensus_helper.py:38-54
    assert len(bands) == 863
    per_chrom = {}
    for key, (a, b) in bands.items():
        per_chrom.setdefault(key.split("_", 1)[0], []).append((a, b))
    assert len(per_chrom) == 25
    for chrom, spans in per_chrom.items():             # no gaps, no overlaps, bed order ascending
        assert spans[0][0] == 0, chrom
        for (a0, b0), (a1, b1) in zip(spans, spans[1:]):
            assert b0 == a1 and a1 < b1, chrom


def test_region_of_reads_binary_search_equals_the_region_loop():
    from consensuscruncher_b200 import cli, shard
    bands = keys.bed_separator(cli.default_bedfile("hg19"))
    chroms = []
    for key in bands:
        c = key.split("_", 1)[0]
        if c not in chroms:
            chroms.append(c)
    regions = [(chroms.index(k.split("_", 1)[0]), a, b) for k, (a, b) in bands.items()]
    rng = np.random.default_rng(5)
    tid = rng.integers(-1, len(chroms) + 2, 20000)
    pos = rng.integers(-1, 260_000_000, 20000)
    pick = rng.integers(0, len(regions), 2000)         # plus reads exactly on band edges
    tid = np.concatenate([tid, [regions[i][0] for i in pick], [regions[i][0] for i in pick]])
    pos = np.concatenate([pos, [regions[i][1] for i in pick], [regions[i][2] for i in pick]])
    got = shard.region_of_reads(tid, pos, regions)
    exp = np.full(tid.shape, -1, np.int64)
    for i, (t, a, b) in enumerate(regions):
        exp[(tid == t) & (pos >= a) & (pos < b)] = i
    assert np.array_equal(got, exp)
    # overlapping regions keep the loop's answer (last match)
    odd = [(0, 0, 100), (0, 50, 150), (1, 0, 10)]
    t2, p2 = np.array([0, 0, 0, 1, 1]), np.array([10, 60, 149, 5, 10])
    assert shard.region_of_reads(t2, p2, odd).tolist() == [0, 1, 1, 2, -1]


def test_compact_upload_format_round_trips():
    """pack.to_compact (input of cc_load_reads_compact): pair records + bit planes decode to the very same
    field records and payload rows; batches that do not fit are refused."""
    for kw in (dict(seed=5, n_molecules=300, L=100), dict(seed=6, n_molecules=200, L=33, n_tid=3),
               dict(seed=7, n_molecules=150, L=150, mapq_mixed=True)):
        packed = synth.generate(**kw).to_packed()
        c = pack.to_compact(packed)
        assert c is not None
        pairs, planes, qbits, qt, bt = c
        assert pairs.dtype.itemsize == 40 and pairs.shape[0] == packed.n // 2 and 1 <= qbits <= 4
        assert planes.shape[1] % 16 == 0 and planes.shape[1] >= packed.lpad // 4 + packed.lpad * qbits // 8
        f, rows = pack.from_compact(pairs, planes, qbits, qt, bt, packed.lpad)
        assert f.tobytes() == np.ascontiguousarray(packed.fields).tobytes()
        assert np.array_equal(rows, packed.payload)
    # too many (quality, literal base) combinations for 4 bits
    packed = synth.generate(seed=8, n_molecules=200, L=64).to_packed()
    half = packed.lpad // 2
    q = packed.payload[:, half:half + 64]
    q[...] = np.where(q > 2, np.random.default_rng(1).integers(3, 42, q.shape).astype(np.uint8), q)
    assert pack.to_compact(packed) is None
    # mates that do not mirror each other
    packed = synth.generate(seed=9, n_molecules=50, L=64).to_packed()
    packed.fields["mpos"][0] += 1
    assert pack.to_pairs(packed.fields) is None and pack.to_compact(packed) is None


def test_short_pair_records_round_trip():
    """pack.to_classes (input of cc_load_reads_classes): 16-byte pair records + a class table give back the very
    same cc_pair records; batches that do not fit are refused."""
    for kw in (dict(seed=5, n_molecules=300, L=100), dict(seed=6, n_molecules=200, L=33, n_tid=3),
               dict(seed=7, n_molecules=150, L=150, mapq_mixed=True)):
        pairs = pack.to_pairs(synth.generate(**kw).to_packed().fields)
        short = pack.to_classes(pairs)
        assert short is not None
        p16, classes = short
        assert p16.dtype.itemsize == 16 and classes.dtype.itemsize == 32 and len(p16) == len(pairs)
        assert 0 < len(classes) <= len(pairs) and int(p16["cls"].max()) == len(classes) - 1
        assert pack.from_classes(p16, classes).tobytes() == pairs.tobytes()
    odd = pairs.copy()
    odd["tlen2"][3] += 1                                   # TLENs that are not each other's negatives
    assert pack.to_classes(odd) is None
    far = pairs.copy()
    far["tlen1"][5], far["tlen2"][5] = 40000, -40000       # does not fit 16 bits
    assert pack.to_classes(far) is None
    many = np.zeros(70000, pack.PAIR_DT)
    many["cig"] = np.arange(70000)                          # more classes than a u16 names
    assert pack.to_classes(many) is None
    assert pack.to_classes(many[:65536]) is not None


@pytest.mark.parametrize("wl,total", [("c3", 60_000), ("c4", 120_000)])
def test_cytoband_sharded_dataset_is_the_same_whoever_makes_it(wl, total):
    """workloads.py: band b's reads depend on (seed, b) only, ranks own disjoint bands, and the generator's own
    prediction of the stage counters equals what the oracle computes (CPU tensors; the GPU tests repeat it at size)."""
    pytest.importorskip("torch")
    from consensuscruncher_b200 import workloads
    from oracle import oracle
    pl = workloads.plan(wl, total // 2, seed=5)
    assert pl.total_pairs == total // 2 and len(pl.bands) == 863
    whole, iw = workloads.generate_shard(pl, 0, 1, "cpu")
    res = oracle.run(whole.to_packed(), 0.7, True)
    assert (res.stats[:9] == iw["stats"][:9]).all()
    assert (np.bincount(res.fam_size, minlength=65536) == iw["hist"]).all()
    owner = pl.owners(3)
    parts = [workloads.generate_shard(pl, r, 3, "cpu") for r in range(3)]
    assert sorted(b for _, ip in parts for b in ip["bands"]) == iw["bands"]
    assert all(owner[b] == r for r, (_, ip) in enumerate(parts) for b in ip["bands"])
    assert (sum(ip["stats"] for _, ip in parts) == iw["stats"]).all() and (sum(ip["hist"] for _, ip in parts) == iw["hist"]).all()
    f = whole.fields.numpy()
    for part, ip in parts:                                   # a rank's reads are exactly its bands' reads of the whole set
        fp = part.fields.numpy()
        keep = np.zeros(f.shape[0], bool)
        for b in ip["bands"]:
            tid, lo, hi = pl.regions[b]
            keep |= (f[:, 2] == tid) & (f[:, 3] >= lo) & (f[:, 3] < hi) & (f[:, 5] >= lo) & (f[:, 5] < hi)
        assert np.array_equal(f[keep], fp)


def test_bench_builds_the_compact_forms_like_pack_does():
    """bench.py makes the inputs of cc_run_host_compact / cc_run_host_classes with torch (on the device, when there is
    one); pack.to_compact / pack.to_classes are the host statements: same bytes."""
    torch = pytest.importorskip("torch")
    import types
    import bench
    p = synth.generate(seed=5, n_molecules=1500, L=100, n_tid=3, mapq_mixed=True).to_packed()
    f = torch.from_numpy(np.ascontiguousarray(p.fields).view(np.int32).reshape(-1, 8).copy())
    b = types.SimpleNamespace(fields=f, payload=torch.from_numpy(p.payload.copy()), lpad=p.lpad)
    pairs_t, planes_t, qbits, qt, bt = bench.device_compact(b)
    pairs, planes, qbits_h, qt_h, bt_h = pack.to_compact(p)
    assert pairs_t.numpy().tobytes() == pairs.tobytes() and np.array_equal(planes_t.numpy(), planes)
    assert qbits == qbits_h and np.array_equal(qt.numpy(), qt_h) and np.array_equal(bt.numpy(), bt_h)
    p16_t, classes_t = bench.device_classes(pairs_t)
    p16 = p16_t.numpy().view(pack.PAIR16_DT).reshape(-1)
    classes = classes_t.numpy().view(pack.CLASS_DT).reshape(-1)
    assert len(classes) == len(pack.to_classes(pairs)[1])
    assert pack.from_classes(p16, classes).tobytes() == pairs.tobytes()
