"""GPU: BASELINE.json config 2 at full size (10 M read pairs = 2e7 reads on one B200).

The oracle cannot run 2e7 reads in test time, so the full batch is checked through
size-independent properties, run-to-run determinism, and a bit-exact comparison with the
oracle on a genomic window of the same batch (families never straddle a window: every member
of a family, its mate family and its duplex partner share both coordinates)."""
import hashlib

import numpy as np
import pytest

from consensuscruncher_b200.pack import FIELDS_DT, PackedReads
from oracle import oracle

pytestmark = pytest.mark.gpu

PAIRS = 10_000_000


def _digest(a):
    return hashlib.sha1(np.ascontiguousarray(a).view(np.uint8)).hexdigest()


def test_config2_full_size(engine):
    torch = pytest.importorskip("torch")
    from consensuscruncher_b200.synth_device import generate_device
    batch = generate_device(seed=1234, n_pairs=PAIRS, L=128, device="cuda:0")
    n = batch.n
    engine.set_params(cutoff=0.7, scorrect=True)
    engine.load_device(batch.fields.data_ptr(), batch.payload.data_ptr(), n, batch.lpad, batch.stride,
                       batch.dup_bc.data_ptr(), batch.n_bc)
    engine.run()
    counts = engine.counts()
    size, tmpl, partner = engine.fetch("fam_size"), engine.fetch("fam_tmpl"), engine.fetch("partner")
    stats = engine.fetch("stats")
    sscs_row, dcs_row = engine.fetch("sscs_row"), engine.fetch("dcs_row")
    corr_row, dcssc_row = engine.fetch("corr_row"), engine.fetch("dcssc_row")
    nf = counts["fam"]

    # ---- size-independent properties ---------------------------------------------------
    assert int(size.sum()) == n and (size >= 1).all()
    assert (tmpl[1:] > tmpl[:-1]).all()                               # families in creation order
    has = partner >= 0
    assert (partner[partner[has]] == np.nonzero(has)[0]).all()        # the duplex join is an involution
    big = size >= 2
    assert counts["sscs"] == int(big.sum()) == stats[0] and stats[1] == nf - counts["sscs"]
    assert (np.sort(sscs_row[big]) == np.arange(counts["sscs"])).all() and (sscs_row[~big] == -1).all()
    pair_big = has & big & (size[np.maximum(partner, 0)] >= 2)
    assert stats[2] == counts["dcs"] == int(pair_big.sum()) // 2
    assert (dcs_row >= 0).sum() == counts["dcs"] and (dcs_row[(dcs_row >= 0)] < counts["dcs"]).all()
    assert stats[4] + stats[5] == counts["corr"] == int((~big & has).sum()) and stats[6] == int((~big & ~has).sum())
    assert stats[7] == counts["dcssc"] == int(has.sum()) // 2
    assert (np.sort(dcssc_row[dcssc_row >= 0]) == np.arange(counts["dcssc"])).all()
    assert (dcssc_row[dcs_row >= 0] == dcs_row[dcs_row >= 0]).all()     # shared rows of SSCS-SSCS pairs
    hist = engine.fetch("fam_hist")
    assert (hist[:256] == np.bincount(size, minlength=256)[:256]).all()

    # ---- determinism: a second pass over the same batch gives the same bytes ----------------
    first = {k: _digest(engine.fetch(k)) for k in ("sscs_rows", "dcssc_rows", "corr_rows", "fam_tmpl", "partner")}
    engine.load_device(batch.fields.data_ptr(), batch.payload.data_ptr(), n, batch.lpad, batch.stride,
                       batch.dup_bc.data_ptr(), batch.n_bc)
    engine.run()
    for k, d in first.items():
        assert _digest(engine.fetch(k)) == d, k

    # ---- oracle on a window of the same batch -------------------------------------------------
    f = batch.fields.cpu().numpy().view(np.uint8).reshape(-1).view(FIELDS_DT)
    lo, hi = 1_000_000, 1_030_000
    in_win = (f["pos"][0::2] >= lo) & (f["pos"][0::2] < hi) & (f["pos"][1::2] >= lo) & (f["pos"][1::2] < hi)
    rows = np.repeat(in_win, 2)
    idx = np.nonzero(rows)[0]
    assert 50_000 < idx.size < 2_000_000
    sub = PackedReads(f[idx].copy(), batch.payload[torch.from_numpy(idx).cuda()].cpu().numpy(), batch.lpad,
                      None, batch.dup_bc.cpu().numpy().astype(np.uint32), None, None)
    exp = oracle.run(sub, 0.7, True)
    # families of the window in the full run, matched by template read (global index -> window index)
    pos_in_sub = np.full(n, -1, np.int64)
    pos_in_sub[idx] = np.arange(idx.size)
    fam_in = pos_in_sub[tmpl] >= 0
    g_ids = np.nonzero(fam_in)[0]
    assert g_ids.size == exp.n_fam
    assert (pos_in_sub[tmpl[g_ids]] == exp.fam_tmpl).all()            # same families, same order, same templates
    assert (size[g_ids] == exp.fam_size).all()
    full_to_sub = np.full(nf, -1, np.int64)
    full_to_sub[g_ids] = np.arange(g_ids.size)
    p_sub = np.where(partner[g_ids] >= 0, full_to_sub[np.maximum(partner[g_ids], 0)], -1)
    assert (p_sub == exp.partner).all()
    for name, rowmap in (("sscs_rows", sscs_row), ("corr_rows", corr_row), ("dcssc_rows", dcssc_row)):
        exp_map = getattr(exp, name[:-1])
        owners = np.nonzero(exp_map >= 0)[0]
        got = engine.fetch(name)[rowmap[g_ids[owners]]]
        want = getattr(exp, name)[exp_map[owners]]
        assert got.shape == want.shape and (got == want).all(), name
    for name in ("fam_mapq", "fam_flag", "fam_tlen", "fam_rg"):
        assert (engine.fetch(name)[g_ids] == getattr(exp, name)).all(), name
