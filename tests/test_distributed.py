"""CPU, world_size 2, gloo: the multi-GPU logic of the path - region sharding and the one
collective (sum of stage counters and family-size histogram).  Each rank runs the oracle on
its shard; the union must equal the unsharded result."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from consensuscruncher_b200 import shard, synth
    from oracle import oracle
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    packed = synth.generate(seed=91, n_molecules=1500, L=48, n_tid=3, region=(0, 30000)).to_packed()
    regions = [(t, a, a + 5000) for t in range(3) for a in range(0, 30000, 5000)]
    owner, _ = shard.shard_pairs(packed.fields, regions, world)
    mine = shard.take_pairs(packed, owner == rank)
    res = oracle.run(mine, 0.7, True)
    stats = torch.from_numpy(res.stats.copy())
    hist = torch.from_numpy(np.bincount(res.fam_size, minlength=256).astype(np.int64))
    shard.reduce_stats(stats, hist)
    np.savez(os.path.join(out_dir, "rank%d.npz" % rank), stats=stats.numpy(), hist=hist.numpy(),
             n_fam=res.n_fam, n_reads=mine.n,
             keys=mine.fields[res.fam_tmpl][["bc", "tid", "pos", "mtid", "mpos", "flag"]].copy(),
             sizes=res.fam_size, sscs=res.sscs_rows, order=np.argsort(res.sscs_row[res.sscs_row >= 0]))
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_equal_one(tmp_path):
    import torch.multiprocessing as mp
    from consensuscruncher_b200 import synth
    from oracle import oracle
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    full_in = synth.generate(seed=91, n_molecules=1500, L=48, n_tid=3, region=(0, 30000)).to_packed()
    full = oracle.run(full_in, 0.7, True)
    r0 = np.load(tmp_path / "rank0.npz")
    r1 = np.load(tmp_path / "rank1.npz")
    # the reduced counters / histogram are identical on both ranks and equal the unsharded run
    assert (r0["stats"] == r1["stats"]).all() and (r0["stats"] == full.stats).all()
    assert (r0["hist"] == np.bincount(full.fam_size, minlength=256)).all()
    assert int(r0["n_reads"]) + int(r1["n_reads"]) == full_in.n
    assert int(r0["n_fam"]) + int(r1["n_fam"]) == full.n_fam
    # same families with the same sizes
    def fam_set(keys, sizes):
        return sorted((tuple(int(x) for x in k), int(s)) for k, s in zip(keys.tolist(), sizes.tolist()))
    full_keys = full_in.fields[full.fam_tmpl][["bc", "tid", "pos", "mtid", "mpos", "flag"]]
    both = fam_set(np.concatenate([r0["keys"], r1["keys"]]), np.concatenate([r0["sizes"], r1["sizes"]]))
    assert both == fam_set(full_keys, full.fam_size)
    # and the same multiset of consensus rows
    rows = np.concatenate([r0["sscs"], r1["sscs"]])
    assert sorted(map(bytes, rows)) == sorted(map(bytes, full.sscs_rows))


def test_assign_regions_is_balanced_and_deterministic():
    from consensuscruncher_b200 import shard
    w = [100, 1, 1, 50, 50, 1, 97]
    a = shard.assign_regions(w, 2)
    assert a == shard.assign_regions(w, 2)
    load = [sum(x for x, r in zip(w, a) if r == k) for k in range(2)]
    assert abs(load[0] - load[1]) <= 3
    assert shard.assign_regions([5, 5, 5], 8) == [0, 1, 2]


def test_pairs_follow_their_lower_coordinate():
    from consensuscruncher_b200 import shard, synth
    p = synth.generate(seed=5, n_molecules=200, L=40, region=(0, 4000)).to_packed()
    regions = [(0, 0, 2000), (0, 2000, 5000)]
    owner, reg = shard.shard_pairs(p.fields, regions, 2)
    lo = np.minimum(p.fields["pos"][0::2], p.fields["pos"][1::2])
    assert ((lo < 2000) == (reg == 0)).all()
    assert set(owner.tolist()) <= {0, 1}
