"""TEST INFRASTRUCTURE (checker only): bit-exact comparison of a GPU run over a large batch with
the CPU oracle on a genomic window of that batch.

The oracle cannot run 10^8 reads in test time; families never straddle a genomic window (every
member of a family, its mate family and its duplex partner share both coordinates), so the
oracle's result on the window's reads must reappear, family by family and byte by byte, inside
the full run.  Used by tests/ and by bench.py's verification leg (outside every timed region).
"""
from __future__ import annotations

import numpy as np

from consensuscruncher_b200.pack import FIELDS_DT, PackedReads
from oracle import oracle


def check_window(engine, batch, window, cutoff, scorrect):
    """`engine` has just run `batch` (a synth_device.DeviceBatch); `window` comes from
    workloads.pick_window.  Raises AssertionError on the first difference; returns a summary dict."""
    import torch
    idx_t = window["reads"]
    idx = idx_t.cpu().numpy()
    n = batch.n
    f_win = batch.fields[idx_t].cpu().numpy().view(np.uint8).reshape(-1).view(FIELDS_DT)
    sub = PackedReads(f_win.copy(), batch.payload[idx_t].cpu().numpy(), batch.lpad, None,
                      batch.dup_bc.cpu().numpy().astype(np.uint32), None, None)
    exp = oracle.run(sub, cutoff, scorrect)

    size, tmpl, partner = engine.fetch("fam_size"), engine.fetch("fam_tmpl"), engine.fetch("partner")
    nf = size.shape[0]
    pos_in_sub = np.full(n, -1, np.int64)
    pos_in_sub[idx] = np.arange(idx.size)
    g_ids = np.nonzero(pos_in_sub[tmpl] >= 0)[0]            # families of the window inside the full run
    assert g_ids.size == exp.n_fam, "window: %d families on the GPU, %d in the oracle" % (g_ids.size, exp.n_fam)
    assert (pos_in_sub[tmpl[g_ids]] == exp.fam_tmpl).all(), "window: family order / templates differ"
    assert (size[g_ids] == exp.fam_size).all(), "window: family sizes differ"
    full_to_sub = np.full(nf, -1, np.int64)
    full_to_sub[g_ids] = np.arange(g_ids.size)
    p_sub = np.where(partner[g_ids] >= 0, full_to_sub[np.maximum(partner[g_ids], 0)], -1)
    assert (p_sub == exp.partner).all(), "window: duplex partners differ"
    rows_checked = 0
    names = [("sscs_rows", "sscs_row")]
    if scorrect:
        names += [("corr_rows", "corr_row"), ("dcssc_rows", "dcssc_row")]
    else:
        names += [("dcs_rows", "dcs_row")]
    for rows_name, map_name in names:
        exp_map = getattr(exp, map_name)
        owners = np.nonzero(exp_map >= 0)[0]
        if owners.size == 0:
            continue
        rowmap = engine.fetch(map_name)
        want = getattr(exp, rows_name)[exp_map[owners]]
        got = fetch_rows(engine, rows_name, rowmap[g_ids[owners]], batch.stride)
        assert got.shape == want.shape and (got == want).all(), "window: %s differ" % rows_name
        rows_checked += owners.size
    for name in ("fam_mapq", "fam_flag", "fam_tlen", "fam_rg"):
        assert (engine.fetch(name)[g_ids] == getattr(exp, name)).all(), "window: %s differ" % name
    return {"tid": int(window["tid"]), "lo": int(window["lo"]), "hi": int(window["hi"]), "reads": int(idx.size),
            "families": int(exp.n_fam), "rows_compared": int(rows_checked),
            "largest_family": int(exp.fam_size.max()) if exp.n_fam else 0}


def fetch_rows(engine, rows_name, rows, stride):
    """Selected rows of a device-resident row array (a gather on the device, not a download of the pool)."""
    import torch
    from consensuscruncher_b200.engine import ARRAYS
    from consensuscruncher_b200.shard import device_tensor
    ptr = engine.device_ptr(rows_name)
    total = engine.lib.cc_array_bytes(engine.ctx, ARRAYS[rows_name][0])
    if not ptr or total == 0:
        return np.zeros((0, stride), np.uint8)
    pool = device_tensor(ptr, total, torch.uint8, engine.device).reshape(-1, stride)
    sel = torch.from_numpy(np.asarray(rows, np.int64)).to(pool.device)
    torch.cuda.current_stream(pool.device).synchronize()
    return pool[sel].cpu().numpy()


def check_shard_properties(engine, n_reads):
    """Size-independent properties of a finished run (any size): families partition the reads, creation
    order, the duplex join is an involution, counters agree with the per-family arrays."""
    counts = engine.counts()
    size, tmpl, partner = engine.fetch("fam_size"), engine.fetch("fam_tmpl"), engine.fetch("partner")
    stats = engine.fetch("stats")
    nf = counts["fam"]
    assert int(size.sum(dtype=np.int64)) == n_reads and (size >= 1).all()
    assert (tmpl[1:] > tmpl[:-1]).all()
    has = partner >= 0
    assert (partner[partner[has]] == np.nonzero(has)[0]).all()
    big = size >= 2
    assert counts["sscs"] == int(big.sum()) == stats[0] and stats[1] == nf - counts["sscs"]
    pair_big = has & big & (size[np.maximum(partner, 0)] >= 2)
    assert stats[2] == counts["dcs"] == int(pair_big.sum()) // 2
    return counts, stats
