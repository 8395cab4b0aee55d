"""GPU: the configurations of BASELINE.json in their stated shape - deep-panel families of
10^3 - 10^4 reads at >= 10^6 reads, cytoband-sharded datasets whose shards together equal the
whole, and the path's one collective (cc_stats_allreduce: NCCL called by libccgpu.so itself)."""
import os
import sys

import numpy as np
import pytest

from consensuscruncher_b200 import shard, workloads
from consensuscruncher_b200.engine import GpuConsensus
from oracle import oracle
from oracle import window as owin

import helpers

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(engine, batch, cutoff=0.7, scorrect=True):
    engine.set_params(cutoff=cutoff, scorrect=scorrect)
    engine.load_device(batch.fields.data_ptr(), batch.payload.data_ptr(), batch.n, batch.lpad, batch.stride,
                       batch.dup_bc.data_ptr(), batch.n_bc)
    engine.run()


def test_deep_panel_families_up_to_1e4_reads_equal_the_oracle(engine):
    """configs[2] in small: 2.4e6 reads on one 2 kb target, lognormal sizes with a tail of families of
    thousands of reads (vote_big_kernel, one cursor / one table slot shared by 10^4 reads), every
    array compared with the oracle."""
    pytest.importorskip("torch")
    from consensuscruncher_b200.synth_device import generate_device
    batch = generate_device(seed=77, n_pairs=1_200_000, L=128, dist="lognormal", fam_mean=12.0, umi_len=4, p_duplex=0.9,
                            tail=(1e-4, 700), region=(5_000_000, 5_000_000 + 2_000 + 656), tid=3, device="cuda:0",
                            predict=True)
    _run(engine, batch)
    got = engine.result()
    assert got.fam_size.max() >= 2000 and (got.fam_size >= 1000).sum() >= 2, got.fam_size.max()
    exp = oracle.run(batch.to_packed(), 0.7, True)
    helpers.assert_same_result(got, exp)
    st, hist = batch.predicted
    assert (got.stats[:9] == st.numpy()[:9]).all()
    assert (engine.fetch("fam_hist").astype(np.int64) == hist.numpy()).all()


@pytest.mark.parametrize("wl,total", [("c3", 3_000_000), ("c4", 2_000_000)])
def test_cytoband_shards_together_equal_the_whole_dataset(engine, wl, total):
    """One dataset, dealt to 1 or 4 ranks by the 863 hg19 bands (here the 4 ranks take turns on one GPU):
    the shards' stats and histograms add up to the single-rank run and to the generator's prediction,
    every shard passes the window oracle, and a band's reads are the same whoever makes them."""
    torch = pytest.importorskip("torch")
    pl = workloads.plan(wl, total // 2, seed=99)
    whole, iw = workloads.generate_shard(pl, 0, 1, "cuda:0")
    _run(engine, whole)
    owin.check_shard_properties(engine, whole.n)
    w_stats, w_hist = engine.fetch("stats").copy(), engine.fetch("fam_hist").astype(np.int64)
    assert (w_stats[:9] == iw["stats"][:9]).all() and (w_hist == iw["hist"]).all()
    win = workloads.pick_window(whole, max_reads=300_000, min_reads=2_000)
    summary = owin.check_window(engine, whole, win, 0.7, True)
    assert summary["rows_compared"] > 0
    stats, hist, reads = np.zeros(16, np.int64), np.zeros(65536, np.int64), 0
    digest_whole = {}
    f = whole.fields.cpu().numpy()
    for b in iw["bands"]:
        tid, lo, hi = pl.regions[b]
        m = (f[:, 2] == tid) & (f[:, 3] >= lo) & (f[:, 3] < hi)
        digest_whole[b] = int(f[m].astype(np.int64).sum())
    for r in range(4):
        part, ip = workloads.generate_shard(pl, r, 4, "cuda:0")
        assert ip["n_bands"] > 0
        _run(engine, part)
        owin.check_shard_properties(engine, part.n)
        win = workloads.pick_window(part, max_reads=300_000, min_reads=2_000)
        owin.check_window(engine, part, win, 0.7, True)
        stats += engine.fetch("stats")
        hist += engine.fetch("fam_hist").astype(np.int64)
        reads += part.n
        fp = part.fields.cpu().numpy()
        for b in ip["bands"]:
            tid, lo, hi = pl.regions[b]
            m = (fp[:, 2] == tid) & (fp[:, 3] >= lo) & (fp[:, 3] < hi)
            assert int(fp[m].astype(np.int64).sum()) == digest_whole[b], "band %d differs between 1 and 4 ranks" % b
    assert reads == whole.n == total
    assert (stats == w_stats).all() and (hist == w_hist).all()


def test_stats_allreduce_is_a_no_op_on_one_rank_and_works_on_a_one_rank_communicator(engine):
    torch = pytest.importorskip("torch")
    pl = workloads.plan("c2", 20_000, seed=3)
    batch, info = workloads.generate_shard(pl, 0, 1, "cuda:0")
    _run(engine, batch)
    before = engine.fetch("stats").copy()
    engine.stats_allreduce()                                  # no communicator: nothing happens
    assert (engine.fetch("stats") == before).all()
    with GpuConsensus(0) as solo:
        solo.comm_init(GpuConsensus.comm_unique_id(), 0, 1)   # NCCL is loaded here (dlopen), one rank
        info_c = solo.comm_info()
        assert info_c["nranks"] == 1 and info_c["nccl_version"] >= 21800
        _run(solo, batch)
        solo.stats_allreduce()
        assert (solo.fetch("stats") == before).all()
        assert (solo.fetch("fam_hist").astype(np.int64) == info["hist"]).all()
        solo.comm_destroy()


def _two_gpu_worker(rank, world, uid_path, out_dir, wl, total):
    sys.path.insert(0, ROOT)
    import torch
    torch.cuda.set_device(rank)
    pl = workloads.plan(wl, total // 2, seed=41)
    batch, info = workloads.generate_shard(pl, rank, world, "cuda:%d" % rank)
    with GpuConsensus(rank, cutoff=0.7, scorrect=True) as eng:
        shard.connect(eng, rank, world, shard.file_exchange(uid_path, rank))
        _run(eng, batch)
        local = eng.fetch("stats").copy()
        owin.check_window(eng, batch, workloads.pick_window(batch, max_reads=200_000, min_reads=2_000), 0.7, True)
        _run(eng, batch)
        eng.stats_allreduce()
        np.savez(os.path.join(out_dir, "rank%d.npz" % rank), stats=eng.fetch("stats"), hist=eng.fetch("fam_hist"),
                 local=local, reads=batch.n, pred_stats=info["stats"], pred_hist=info["hist"])
        eng.comm_destroy()


@pytest.mark.parametrize("wl,total", [("c3", 2_000_000), ("c4", 1_500_000)])
def test_two_gpus_sharded_by_cytoband_equal_one_gpu(tmp_path, wl, total):
    """The CUDA path on two GPUs (one process per GPU, NCCL through cc_stats_allreduce) against the
    unsharded run on one GPU."""
    torch = pytest.importorskip("torch")
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp
    mp.spawn(_two_gpu_worker, args=(2, str(tmp_path / "uid"), str(tmp_path), wl, total), nprocs=2, join=True)
    pl = workloads.plan(wl, total // 2, seed=41)
    whole, iw = workloads.generate_shard(pl, 0, 1, "cuda:0")
    with GpuConsensus(0, cutoff=0.7, scorrect=True) as eng:
        _run(eng, whole)
        stats, hist = eng.fetch("stats"), eng.fetch("fam_hist")
    r0, r1 = np.load(tmp_path / "rank0.npz"), np.load(tmp_path / "rank1.npz")
    assert int(r0["reads"]) + int(r1["reads"]) == whole.n and int(r0["reads"]) > 0 and int(r1["reads"]) > 0
    for r in (r0, r1):                                        # every rank holds the same, reduced, numbers
        assert (r["stats"] == stats).all() and (r["hist"] == hist).all()
    assert (r0["local"] + r1["local"] == stats).all()
    assert (r0["pred_stats"] + r1["pred_stats"])[:9].tolist() == stats[:9].tolist()
    assert (r0["pred_hist"] + r1["pred_hist"] == hist.astype(np.int64)).all()


def test_batches_with_many_distinct_qualities_take_the_8_bit_rows(engine):
    """consensus(compact="auto"): more than 15 distinct non-zero qualities cannot use the 4-bit codes."""
    from consensuscruncher_b200 import pack, synth
    s = synth.generate(seed=61, n_molecules=800, L=100)
    packed = s.to_packed()
    rng = np.random.default_rng(8)
    half, lpad = packed.lpad // 2, packed.lpad
    q = packed.payload[:, half:half + 100]
    q[...] = np.where(q > 2, rng.integers(3, 42, q.shape).astype(np.uint8), q)     # ~39 distinct values
    assert pack.to_q4(packed) is None
    engine.set_params(cutoff=0.7, scorrect=True)
    got = engine.consensus(packed, compact="auto")
    helpers.assert_same_result(got, oracle.run(packed, 0.7, True))
    few = synth.generate(seed=62, n_molecules=800, L=100).to_packed()
    assert pack.to_q4(few) is not None
    helpers.assert_same_result(engine.consensus(few, compact="auto"), oracle.run(few, 0.7, True))


@pytest.mark.parametrize("kw", [dict(seed=81, n_molecules=1500, L=128), dict(seed=82, n_molecules=900, L=33, n_tid=3),
                                dict(seed=83, n_molecules=600, L=250, mapq_mixed=True), dict(seed=84, n_molecules=3, L=40)])
def test_compact_upload_gives_the_same_result(engine, kw):
    """cc_load_reads_compact: pair records + 2-bit bases + few-bit quality codes, expanded on the device."""
    from consensuscruncher_b200 import pack, synth
    packed = synth.generate(**kw).to_packed()
    assert pack.to_compact(packed) is not None
    engine.set_params(cutoff=0.7, scorrect=True)
    got = engine.consensus(packed, compact="planes")
    helpers.assert_same_result(got, oracle.run(packed, 0.7, True))
    helpers.assert_same_result(got, engine.consensus(packed))
    # cc_load_reads_classes: 16-byte pair records + class table
    assert pack.to_classes(pack.to_compact(packed)[0]) is not None
    helpers.assert_same_result(engine.consensus(packed, compact="classes"), got)


def test_short_pair_records_with_a_class_out_of_range_are_refused(engine):
    from consensuscruncher_b200 import pack, synth
    from consensuscruncher_b200.engine import GpuError
    packed = synth.generate(seed=85, n_molecules=400, L=64).to_packed()
    compact = pack.to_compact(packed)
    pairs, classes = pack.to_classes(compact[0])
    pairs["cls"][7] = len(classes)
    engine.set_params(cutoff=0.7, scorrect=True)
    dup = np.ascontiguousarray(packed.dup_bc, dtype=np.uint32)
    _, planes, qbits, qt, bt = compact
    rc = engine.lib.cc_run_host_classes(engine.ctx, pairs.ctypes.data, classes.ctypes.data, len(classes), planes.ctypes.data,
                                        packed.n, packed.lpad, planes.shape[1], int(qbits), qt.ctypes.data, bt.ctypes.data,
                                        dup.ctypes.data, len(dup))
    assert rc == -2 and b"class" in engine.lib.cc_last_error(engine.ctx)          # CC_ERR_ARG
    helpers.assert_same_result(engine.consensus(packed, compact="classes"), oracle.run(packed, 0.7, True))


@pytest.mark.parametrize("kw,scorrect", [(dict(seed=91, n_molecules=1500, L=128), True), (dict(seed=92, n_molecules=900, L=33, n_tid=3), True),
                                         (dict(seed=93, n_molecules=600, L=250, mapq_mixed=True), False),
                                         (dict(seed=94, n_molecules=40, L=100, dist="const", fam_mean=200), True)])
def test_rows_read_back_in_bit_planes_are_the_same_rows(engine, kw, scorrect):
    """cc_pack_rows: 2-bit bases + few-bit (quality, literal base) codes for every produced row; expanded on the
    host they are the rows cc_fetch returns."""
    from consensuscruncher_b200 import synth
    packed = synth.generate(**kw).to_packed()
    engine.set_params(cutoff=0.7, scorrect=scorrect)
    engine.run_host(packed)
    plain = engine.result()
    helpers.assert_same_result(plain, oracle.run(packed, 0.7, scorrect))
    via = engine.result(planes=True)
    assert engine.packing is not None and 1 <= engine.packing[0] <= 4
    helpers.assert_same_result(via, plain)
    assert engine.fetch("sscs_planes").shape == (plain.sscs_rows.shape[0], engine.packing[1])
    assert engine.packing[1] * 2 <= engine.stride + 32
    engine.set_params(cutoff=0.9, scorrect=scorrect)          # another cutoff: the planes of the old rows are gone
    assert engine.lib.cc_array_bytes(engine.ctx, 26) == 0


def test_rows_with_many_distinct_qualities_stay_as_they_are(engine):
    from consensuscruncher_b200 import synth
    s = synth.generate(seed=95, n_molecules=800, L=100)
    packed = s.to_packed()
    half = packed.lpad // 2
    q = packed.payload[:, half:half + 100]
    q[...] = np.where(q > 2, np.random.default_rng(8).integers(3, 42, q.shape).astype(np.uint8), q)
    engine.set_params(cutoff=0.7, scorrect=True)
    engine.run_host(packed)
    assert engine.pack_rows() is None and engine.lib.cc_array_bytes(engine.ctx, 26) == 0
    helpers.assert_same_result(engine.result(planes=True), oracle.run(packed, 0.7, True))
