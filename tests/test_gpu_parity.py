"""GPU: the CUDA path (through the C ABI) against the oracle and the reference's goldens.
Bit-exact: every result array must be identical, byte for byte."""
import numpy as np
import pytest

from consensuscruncher_b200 import api, pack, synth
from consensuscruncher_b200.engine import BaseError, GpuConsensus, GpuError
from consensuscruncher_b200.records import Read
from oracle import golden, oracle

import helpers

pytestmark = pytest.mark.gpu


def _both(engine, packed, cutoff=0.7, scorrect=True):
    engine.set_params(cutoff=cutoff, scorrect=scorrect)
    got = engine.consensus(packed)
    exp = oracle.run(packed, cutoff, scorrect)
    helpers.assert_same_result(got, exp)
    return got


@pytest.mark.parametrize("name", golden.names())
def test_gpu_reproduces_reference_outputs(engine, name):
    g, reads, packed = helpers.load_case(name)
    engine.set_params(cutoff=g["params"]["cutoff"], scorrect=g["params"]["scorrect"])
    res = engine.consensus(packed)
    helpers.assert_matches_golden(g, reads, packed, res)
    helpers.assert_same_result(res, oracle.run(packed, g["params"]["cutoff"], g["params"]["scorrect"]))


@pytest.mark.parametrize("kw", [
    dict(seed=21, n_molecules=3000, L=128),
    dict(seed=22, n_molecules=3000, L=150, fam_mean=2.0, p_duplex=0.5),
    dict(seed=23, n_molecules=2000, L=100, dist="wgs", p_duplex=0.3, n_tid=4),
    dict(seed=24, n_molecules=300, L=64, dist="lognormal", fam_mean=12.0, err=0.02),
    dict(seed=25, n_molecules=4000, L=36, umi_len=1, region=(0, 600), insert_sd=2.0),   # crowded keys
    dict(seed=26, n_molecules=1500, L=250, mapq_mixed=True),
    dict(seed=27, n_molecules=2, L=32, dist="const", fam_mean=700, p_duplex=1.0, err=0.05),   # > 255 reads / family
    dict(seed=28, n_molecules=40, L=40, dist="const", fam_mean=40, p_duplex=1.0, mapq_mixed=True),  # mode over > 32 reads
    dict(seed=29, n_molecules=60, L=600, fam_mean=3.0),                # lpad > 512: every family on the warp-per-family kernel
    dict(seed=30, n_molecules=300, L=100, dist="const", fam_mean=14),  # just below / above the counter-flush boundary
    dict(seed=31, n_molecules=300, L=100, dist="const", fam_mean=17),
    dict(seed=32, n_molecules=200, L=160, dist="lognormal", fam_mean=30.0, err=0.02),   # 10 of 16 lanes active, mid sizes
    dict(seed=33, n_molecules=30, L=128, dist="const", fam_mean=127, p_duplex=1.0),     # largest small-kernel family
    dict(seed=34, n_molecules=30, L=128, dist="const", fam_mean=128, p_duplex=1.0),     # smallest warp-per-family one
])
def test_gpu_equals_oracle_on_synthetic(engine, kw):
    packed = synth.generate(**kw).to_packed()
    res = _both(engine, packed)
    assert res.n_fam > 0


@pytest.mark.parametrize("cutoff", [0.5, 0.6, 0.7, 0.8, 0.9, 1.0])
@pytest.mark.parametrize("scorrect", [True, False])
def test_cutoff_sweep(engine, cutoff, scorrect):
    packed = synth.generate(seed=31, n_molecules=1500, L=64, err=0.05, fam_mean=3.0).to_packed()
    _both(engine, packed, cutoff, scorrect)


def test_medium_batch_and_properties(engine):
    s = synth.generate(seed=41, n_molecules=40000, L=128)
    packed = s.to_packed()
    res = _both(engine, packed)
    # size-independent properties
    assert int(res.fam_size.sum()) == packed.n
    assert (res.fam_tmpl[1:] > res.fam_tmpl[:-1]).all()                 # creation order
    p = res.partner
    has = p >= 0
    assert (p[p[has]] == np.nonzero(has)[0]).all()                       # the join is an involution
    assert res.stat("sscs") + res.stat("singletons") == res.n_fam
    assert res.stat("dcs") * 2 + res.stat("sscs_singletons") == res.stat("sscs")


def test_stagewise_calls_equal_one_shot(engine):
    packed = synth.generate(seed=42, n_molecules=2500, L=96).to_packed()
    engine.set_params(cutoff=0.7, scorrect=True)
    one = engine.consensus(packed)
    engine.load(packed)
    engine.group()
    assert engine.counts()["fam"] == one.n_fam
    engine.sscs_vote()
    engine.dcs_join()
    engine.sc_rescue()
    helpers.assert_same_result(engine.result(), one)


@pytest.mark.parametrize("mode", ["pair", "tile"])
def test_voting_a_batch_again_with_another_cutoff(mode, monkeypatch):
    """cc_set_params + cc_sscs_vote on a batch that was voted already: the lists of 16+ and 128+ read families and
    the rows queued for the exact kernel (quality bytes >= 128) are made afresh, not appended to."""
    from consensuscruncher_b200.engine import GpuConsensus
    monkeypatch.setenv("CCGPU_GROUP", mode)
    s = synth.generate(seed=43, n_molecules=160, L=64, dist="lognormal", fam_mean=40.0, max_family=400, p_n=0.0)
    s.qual[::7, 3] = 200
    packed = s.to_packed()
    with GpuConsensus(0, cutoff=0.7, scorrect=True) as eng:
        eng.load(packed)
        eng.group()
        assert (eng.fetch("fam_size") > 127).any() and ((eng.fetch("fam_size") > 15) & (eng.fetch("fam_size") < 128)).any()
        expected = {c: oracle.run(packed, c, True) for c in (0.7, 0.9, 0.51)}
        for cutoff in (0.7, 0.9, 0.7, 0.51) * 6:              # (more joins than the pool of scan tickets holds)
            eng.set_params(cutoff=cutoff, scorrect=True)
            eng.sscs_vote()
            eng.dcs_join()
            eng.sc_rescue()
            helpers.assert_same_result(eng.result(), expected[cutoff])


def test_empty_and_tiny_batches(engine):
    engine.set_params(cutoff=0.7, scorrect=True)
    empty = pack.pack_reads([])
    res = engine.consensus(empty)
    assert res.n_fam == 0 and res.sscs_rows.shape == (0, 64)
    one_pair = synth.generate(seed=3, n_molecules=1, dist="const", fam_mean=1, p_duplex=0.0, L=33).to_packed()
    _both(engine, one_pair)
    two = synth.generate(seed=4, n_molecules=1, dist="const", fam_mean=2, p_duplex=1.0, L=31).to_packed()
    _both(engine, two)


def test_repeated_runs_reuse_the_context(engine):
    a = synth.generate(seed=51, n_molecules=1200, L=128).to_packed()
    b = synth.generate(seed=52, n_molecules=300, L=64).to_packed()
    for p in (a, b, a):
        _both(engine, p)


def test_hash_collisions_are_regrouped_exactly(engine):
    # a tiny position range and 1-nt UMIs with many molecules force equal sort keys for
    # different tags (the sort only uses as many hash bits as the batch needs)
    packed = synth.generate(seed=61, n_molecules=30000, L=32, umi_len=2, region=(0, 1200),
                            fam_mean=1.5, insert_sd=8.0).to_packed()
    _both(engine, packed)


def test_n_base_with_high_quality_is_an_error(engine):
    s = synth.generate(seed=71, n_molecules=20, L=32, dist="const", fam_mean=3, p_n=0.0)
    s.codes[5, 7] = 15
    s.qual[5, 7] = 37
    packed = s.to_packed()
    engine.set_params(cutoff=0.7, scorrect=True)
    with pytest.raises(BaseError):
        engine.consensus(packed)
    with pytest.raises(oracle.OracleError):
        oracle.run(packed)


@pytest.mark.parametrize("fam_mean,dist", [(2, "const"), (4.0, "ztpoisson"), (24, "const")])
def test_quality_bytes_above_127_take_the_exact_path(engine, fam_mean, dist):
    # SAM stops at Q93, BAM bytes do not: the reference counts 128..255 as passing.  The byte-lane
    # pass test of the bulk kernel assumes q < 128 and hands such families to the exact kernel.
    s = synth.generate(seed=75, n_molecules=300, L=128, dist=dist, fam_mean=fam_mean, p_n=0.0)
    rng = np.random.default_rng(5)
    hit = rng.random(s.qual.shape) < 0.004
    s.qual[hit] = rng.choice(np.array([128, 140, 157, 158, 200, 255], np.uint8), int(hit.sum()))
    # a failing N right after a byte that overflows: a carry would make it look like a passing N
    rows = rng.choice(len(s.qual), 40, replace=False)
    s.qual[rows, 40] = 255
    s.qual[rows, 41] = 29
    s.codes[rows, 41] = 15
    packed = s.to_packed()
    _both(engine, packed)


def test_domain_fault_is_reported(engine):
    # pair 99/147 and pair 65/129 with the same UMI and coordinates: R1 tags merge, R2 tags do not (F11)
    def pair(q, f1, f2):
        return [Read(q, f1, 0, 100, 60, "30M", 0, 200, 130, "A" * 30, [35] * 30, rg="1"),
                Read(q, f2, 0, 200, 60, "30M", 0, 100, -130, "C" * 30, [35] * 30, rg="1")]
    reads = pair("a|AC.GT", 99, 147) + pair("b|AC.GT", 65, 129)
    reads = [reads[0], reads[2], reads[1], reads[3]]
    packed = pack.pack_reads(reads)
    engine.set_params(cutoff=0.7, scorrect=True)
    with pytest.raises(pack.DomainError):
        engine.consensus(packed)


def test_geometry_is_validated(engine):
    p = synth.generate(seed=81, n_molecules=5, L=32).to_packed()
    bad = pack.PackedReads(p.fields, p.payload[:, :48].copy(), p.lpad, p.barcodes, p.dup_bc, p.cigar_pairs, p.rgs)
    bad.stride = 48
    with pytest.raises(GpuError):
        engine.load(bad)


# ---- function-level mirrors (the reference's docstring examples) ----------------------------
def _reads(seqs, quals):
    return [Read("r%d" % i, seq=s, qual=q) for i, (s, q) in enumerate(zip(seqs, quals))]


def test_consensus_maker_known_answers(engine):
    seqs = ["ACTGATACTT", "ACTGAAACCT", "ACTGATACCT", "ACTGATACTT"]           # SSCS_maker.py:23-29
    assert api.consensus_maker(_reads(seqs, [[40] * 10] * 4), 0.7, engine) == ("ACTGATACNT", [60] * 10)
    r = _reads(["AC", "CC", "AA", "CA"], [[40, 40], [40, 10], [40, 40], [40, 40]])
    assert api.consensus_maker(r, 0.5, engine) == ("AA", [60, 60])
    assert api.consensus_maker(r, 0.7, engine) == ("NN", [60, 60])
    assert api.consensus_maker(_reads(["A", "A"], [[10], [20]]), 0.7, engine) == ("N", [0])
    assert api.consensus_maker(_reads(["A", "C"], [[35], [33]]), 0.7, engine) == ("N", [35])
    with pytest.raises(BaseError):
        api.consensus_maker(_reads(["N", "A"], [[35], [35]]), 0.7, engine)


def test_duplex_consensus_known_answers(engine):
    a, b = _reads(["ACGTN", "ACGAN"], [[40, 20, 35, 40, 2], [25, 20, 30, 40, 2]])
    assert api.duplex_consensus(a, b, engine) == ("ACGNN", [60, 40, 60, 0, 4])
    assert api.sc_duplex_consensus(a, b, engine) == ("NNGNN", [0, 0, 60, 0, 0])


def test_primitives_equal_oracle_on_random_rows(engine):
    rng = np.random.default_rng(7)
    for lpad in (32, 128, 160, 288):
        stride = pack.row_geometry(lpad)[1]
        n = 37
        rows = np.zeros((n, stride), np.uint8)
        codes = np.array([1, 2, 4, 8], np.uint8)[rng.integers(0, 4, (n, lpad))]
        codes[:, : lpad // 2] = codes[0, : lpad // 2]                  # half the columns mostly agree
        qual = rng.integers(0, 60, (n, lpad)).astype(np.uint8)
        nn = rng.random((n, lpad)) < 0.02
        codes[nn] = 15
        qual[nn] = 2
        for i in range(n):
            pack.pack_row(codes[i], qual[i], lpad, rows[i])
        for cutoff in (0.5, 0.7, 1.0):
            for k in (2, 3, 37):
                got = engine.consensus_maker_rows(rows[:k], lpad, cutoff)
                assert (got == oracle.consensus_maker_rows(rows[:k], lpad, cutoff)).all()
        for sc in (False, True):
            got = engine.duplex_consensus_rows(rows[:18], rows[18:36], lpad, sc)
            for i in range(18):
                assert (got[i] == oracle.duplex_consensus_rows(rows[i], rows[18 + i], lpad, sc)).all()


def test_region_shards_on_the_gpu_equal_the_whole(engine):
    """Multi-GPU semantics on the device path: split a batch by genomic regions into 8 shards the
    way `bench.py --gpus 8` / shard.py would, run every shard through the CUDA path, and compare
    the union with the unsharded run (families never straddle shards)."""
    from consensuscruncher_b200 import shard
    s = synth.generate(seed=95, n_molecules=6000, L=64, n_tid=4, region=(0, 200000), dist="wgs", p_duplex=0.4)
    packed = s.to_packed()
    engine.set_params(cutoff=0.7, scorrect=True)
    whole = engine.consensus(packed)
    regions = [(t, a, a + 25000) for t in range(4) for a in range(0, 200000, 25000)]
    owner, _ = shard.shard_pairs(packed.fields, regions, 8)
    assert (owner >= 0).all() and len(set(owner.tolist())) == 8
    stats = np.zeros(16, np.int64)
    hist = np.zeros(65536, np.uint64)
    seen = {}
    for r in range(8):
        mine = shard.take_pairs(packed, owner == r)
        res = engine.consensus(mine)
        stats += res.stats
        hist += engine.fetch("fam_hist")
        keys = mine.fields[res.fam_tmpl][["bc", "tid", "pos", "mtid", "mpos", "flag"]]
        for g, k in enumerate(keys.tolist()):
            row = bytes(res.sscs_rows[res.sscs_row[g]]) if res.sscs_row[g] >= 0 else None
            corr = bytes(res.corr_rows[res.corr_row[g]]) if res.corr_row[g] >= 0 else None
            seen[tuple(k)] = (int(res.fam_size[g]), row, corr, int(res.partner[g]) >= 0)
    assert (stats == whole.stats).all()
    assert (hist[:64] == np.bincount(whole.fam_size, minlength=64)[:64].astype(np.uint64)).all()
    keys = packed.fields[whole.fam_tmpl][["bc", "tid", "pos", "mtid", "mpos", "flag"]]
    assert len(seen) == whole.n_fam
    for g, k in enumerate(keys.tolist()):
        size, row, corr, has_partner = seen[tuple(k)]
        assert size == whole.fam_size[g] and has_partner == (whole.partner[g] >= 0)
        assert row == (bytes(whole.sscs_rows[whole.sscs_row[g]]) if whole.sscs_row[g] >= 0 else None)
        assert corr == (bytes(whole.corr_rows[whole.corr_row[g]]) if whole.corr_row[g] >= 0 else None)


@pytest.mark.parametrize("n,n_tid,max_pos", [(1, 1, 10), (2, 1, 0), (5000, 1, 200), (100000, 3, 5000),
                                             (300000, 700, 2 ** 31 - 2), (70000, 1, 70000)])
def test_coordinate_sort_equals_the_host_rule(engine, n, n_tid, max_pos):
    # sort_index / samtools merge (ConsensusCruncher.py:10-29,248): (contig, position, strand), stable,
    # records without a contig last
    rng = np.random.default_rng(n)
    tid = rng.integers(-1, n_tid, n).astype(np.int32)
    pos = rng.integers(-1, max_pos + 1, n, dtype=np.int64).astype(np.int32)
    flag = rng.choice(np.array([99, 147, 83, 163, 77, 141, 1123], np.uint16), n)
    got = engine.sort_coordinates(tid, pos, flag)
    t = np.where(tid < 0, np.int64(1) << 31, tid.astype(np.int64))
    exp = np.lexsort(((flag >> 4) & 1, pos.astype(np.int64), t))          # lexsort is stable, last key first
    assert np.array_equal(got.astype(np.int64), exp)
    recs = [Read("q%d" % i, int(flag[i]), int(tid[i]), int(pos[i]), 0, "", -1, -1, 0, "", []) for i in range(min(n, 3000))]
    from consensuscruncher_b200.records import sort_records
    assert [r.qname for r in engine.sort_records(recs)] == [r.qname for r in sort_records(recs)]


def test_coordinate_sort_rejects_positions_it_cannot_key(engine):
    with pytest.raises(Exception):
        engine.sort_coordinates(np.array([0], np.int32), np.array([-2], np.int32), np.array([0], np.uint16))
    assert len(engine.sort_coordinates(np.zeros(0, np.int32), np.zeros(0, np.int32), np.zeros(0, np.uint16))) == 0


@pytest.mark.parametrize("L", [32, 75, 128, 150])
def test_compact_quality_rows_give_the_same_result(engine, L):
    packed = synth.generate(seed=91, n_molecules=900, L=L).to_packed()
    engine.set_params(cutoff=0.7, scorrect=True)
    got = engine.consensus(packed, compact=True)
    helpers.assert_same_result(got, oracle.run(packed, 0.7, True))
    helpers.assert_same_result(got, engine.consensus(packed))


def test_compact_rows_need_a_zero_code(engine):
    packed = synth.generate(seed=92, n_molecules=10, L=32).to_packed()
    payload4, table = pack.to_q4(packed)
    table = table.copy()
    table[0] = 7
    rc = engine.lib.cc_load_reads_q4(engine.ctx, packed.fields.ctypes.data, payload4.ctypes.data, packed.n, packed.lpad,
                                     payload4.shape[1], table.ctypes.data, packed.dup_bc.ctypes.data, len(packed.dup_bc))
    assert rc == -2


def test_one_sided_duplex_partnerships_are_outside_the_domain(engine):
    # duplex_tag swaps an undotted barcode at len // 2: for an odd length the swap is not its own inverse,
    # a family can then have a partner that has another partner (or none).  What the reference writes
    # in that case depends on file order and it can die on read_dict[ds] (DCS_maker.py:231-247), so both
    # the GPU path and the oracle report the input instead of inventing an answer.
    s = synth.generate(seed=93, n_molecules=6000, L=32, umi_len=2, umi_dot=False, region=(0, 1500), insert_sd=3.0,
                       fam_mean=2.0)
    reads = s.to_reads()
    for r in reads:
        head, bc = r.qname.rsplit("|", 1)
        r.qname = head + "|" + bc[:3]
    packed = pack.pack_reads(reads)
    d = packed.dup_bc
    has = d != 0xFFFFFFFF
    assert np.any(d[d[has]] != np.flatnonzero(has))          # the swap really is not an involution here
    engine.set_params(cutoff=0.7, scorrect=True)
    with pytest.raises(pack.DomainError):
        engine.consensus(packed)
    exp, faults = oracle.run(packed, 0.7, True, return_faults=True)
    p = exp.partner
    linked = np.flatnonzero(p >= 0)
    assert faults == int(np.sum(p[p[linked]] != linked)) > 0
    # even-length undotted barcodes stay in the domain: the join takes its R1 / R2 shortcut
    even = synth.generate(seed=94, n_molecules=3000, L=32, umi_len=2, umi_dot=False, region=(0, 1500)).to_packed()
    got = _both(engine, even)
    q = got.partner
    linked = np.flatnonzero(q >= 0)
    assert len(linked) and np.array_equal(q[q[linked]], linked)


def _reordered(packed, order):
    """The same pairs in another order (pair k = reads 2k, 2k+1 stay together)."""
    from consensuscruncher_b200.pack import PackedReads
    idx = np.stack([2 * order, 2 * order + 1], 1).reshape(-1)
    return PackedReads(packed.fields[idx].copy(), packed.payload[idx].copy(), packed.lpad, packed.barcodes, packed.dup_bc,
                       packed.cigar_pairs, packed.rgs)


def test_inputs_out_of_order_and_pairs_across_contigs(engine):
    """The default grouping finishes families inside a tile when completion coordinates come in order, and
    checks that order itself.  Out-of-order inputs (shuffled pairs, a contig in two stretches, descending
    positions) must take the all-through-the-table pass; pairs whose mates lie on two contigs always go
    through the table.  The oracle does not care about order, so every variant is compared with it."""
    base = synth.generate(seed=71, n_molecules=9000, L=48, n_tid=3, region=(0, 60000), fam_mean=3.0).to_packed()
    n_pairs = base.n // 2
    rng = np.random.default_rng(71)
    ident = np.arange(n_pairs)
    variants = {"sorted": ident, "shuffled": rng.permutation(n_pairs), "reversed": ident[::-1].copy(),
                "two stretches": np.concatenate([ident[n_pairs // 2:], ident[:n_pairs // 2]]),
                "one pair moved to the front": np.concatenate([[n_pairs - 1], ident[:-1]])}
    for name, order in variants.items():
        packed = _reordered(base, order)
        engine.set_params(cutoff=0.7, scorrect=True)
        helpers.assert_same_result(engine.consensus(packed), oracle.run(packed, 0.7, True))
    # pairs across contigs: every pair starting in a window gets its second mate moved to another contig
    # (consistently for whole families and their duplex partners: the choice depends on coordinates only)
    cross = _reordered(base, ident)
    f = cross.fields
    sel = (f["pos"][0::2] % 7 == 3)
    odd, even = np.nonzero(np.repeat(sel, 2) & (np.arange(cross.n) % 2 == 1))[0], np.nonzero(np.repeat(sel, 2) & (np.arange(cross.n) % 2 == 0))[0]
    f["tid"][odd] = 9
    f["mtid"][even] = 9
    res = engine.consensus(cross)
    helpers.assert_same_result(res, oracle.run(cross, 0.7, True))
    assert res.n_fam > 1000 and (res.partner >= 0).sum() > 100
    # long runs: thousands of pairs complete at one coordinate, so whole tiles are edge runs
    deep = synth.generate(seed=72, n_molecules=3000, L=40, umi_len=4, region=(0, 700), insert_sd=1.0, fam_mean=6.0).to_packed()
    helpers.assert_same_result(engine.consensus(deep), oracle.run(deep, 0.7, True))


def test_mates_that_do_not_mirror_each_other_are_grouped_read_by_read(engine):
    """The default grouping works on pairs and needs mates that carry each other's coordinates; a batch where they
    do not (here: the second mates of some pairs name a shifted mate position, consistently per family) is
    detected on the device and grouped read by read - same result as the oracle either way."""
    packed = synth.generate(seed=73, n_molecules=4000, L=40, n_tid=2, region=(0, 30000), fam_mean=3.0).to_packed()
    f = packed.fields
    odd = np.arange(packed.n) % 2 == 1
    sel = odd & (f["mpos"] % 5 == 2)
    assert sel.sum() > 100
    f["mpos"][sel] += 7
    engine.set_params(cutoff=0.7, scorrect=True)
    res = engine.consensus(packed)
    helpers.assert_same_result(res, oracle.run(packed, 0.7, True))


@pytest.mark.parametrize("mode", ["tile", "hash", "sort"])
def test_other_grouping_paths_give_the_same_result(monkeypatch, mode):
    # CCGPU_GROUP=sort selects the radix-sort grouping (hash keys, stable onesweep sort, exact boundaries),
    # CCGPU_GROUP=hash the table with one trip per read (group_hash.cuh), CCGPU_GROUP=tile the read-by-read
    # tile grouping (group_tile.cuh); the default groups pair by pair (group_pair.cuh).  Same families, same
    # numbering, same everything.
    monkeypatch.setenv("CCGPU_GROUP", mode)
    cases = [synth.generate(seed=41, n_molecules=2500, L=128).to_packed(),
             synth.generate(seed=42, n_molecules=800, L=64, mapq_mixed=True, max_family=40, dist="const", fam_mean=20).to_packed(),
             synth.generate(seed=61, n_molecules=30000, L=32, umi_len=2, region=(0, 1200), fam_mean=1.5, insert_sd=8.0).to_packed(),
             helpers.load_case("mode_ties")[2]]
    cases.append(_reordered(cases[0], np.random.default_rng(3).permutation(cases[0].n // 2)))
    with GpuConsensus(0, cutoff=0.7, scorrect=True) as other:
        for packed in cases:
            helpers.assert_same_result(other.consensus(packed), oracle.run(packed, 0.7, True))


def test_one_huge_family_and_all_singletons(engine):
    # the two extremes of the family table: every read of a strand hammers one slot / every read founds its own
    huge = synth.generate(seed=95, n_molecules=2, L=64, dist="const", fam_mean=20000, p_duplex=1.0)
    _both(engine, huge.to_packed())
    lone = synth.generate(seed=96, n_molecules=40000, L=32, dist="const", fam_mean=1, p_duplex=0.3, umi_len=4)
    got = _both(engine, lone.to_packed())
    assert got.n_fam == got.fam_size.sum()


def _fuzz_case(seed):
    """A small random batch: random geometry and family structure, then tlen / mapq / RG / flag bit 0x20
    scrambled per read - none of them is part of the family key, all of them feed the mode fields,
    so most families get ties that only the first-appearance rule decides."""
    rng = np.random.default_rng(1000 + seed)
    L = int(rng.integers(20, 200))
    kw = dict(seed=seed, n_molecules=int(rng.integers(1, 400)), L=L,
              dist=str(rng.choice(["ztpoisson", "const", "lognormal", "wgs"])), fam_mean=float(rng.choice([1, 2, 3, 6, 20])),
              p_duplex=float(rng.choice([0.0, 0.5, 1.0])), umi_len=int(rng.integers(1, 4)), umi_dot=bool(rng.integers(0, 2)),
              n_tid=int(rng.integers(1, 4)), region=(0, int(rng.choice([600 + 2 * L, 5000, 200000]))),
              insert_sd=float(rng.choice([2.0, 20.0])), p_n=float(rng.choice([0.0, 0.01])), max_family=300)
    packed = synth.generate(**kw).to_packed()
    f = packed.fields
    n = packed.n
    if n:
        f["tlen"] = rng.integers(-3, 4, n).astype(np.int32)
        f["mapq"] = rng.choice(np.array([0, 20, 60], np.uint8), n)
        f["rg"] = rng.choice(np.array([0, 1, 2, 0xFF], np.uint8), n, p=[0.4, 0.3, 0.25, 0.05])
        f["flag"] = f["flag"] ^ (rng.integers(0, 2, n).astype(np.uint16) << 5)
    packed.rgs = ["rgA", "rgB", "rgC"]
    return packed, float(rng.choice([0.5, 0.6, 0.7, 0.9, 1.0])), bool(rng.integers(0, 2))


@pytest.mark.parametrize("block", range(4))
def test_fuzz_small_batches(engine, block):
    for seed in range(block * 15, block * 15 + 15):
        packed, cutoff, scorrect = _fuzz_case(seed)
        engine.set_params(cutoff=cutoff, scorrect=scorrect)
        try:
            exp, faults = oracle.run(packed, cutoff, scorrect, return_faults=True)
        except oracle.OracleError:
            with pytest.raises(GpuError):
                engine.consensus(packed)
            continue
        if faults:
            with pytest.raises(pack.DomainError):
                engine.consensus(packed)
            continue
        got = engine.consensus(packed)
        bad = got.diff(exp)
        assert not bad, "seed %d: %s" % (seed, bad)
