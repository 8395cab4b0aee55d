"""Synthetic duplex-UMI batches generated directly in device memory (torch, CUDA).

Same molecule model as `synth.py` (BASELINE.json configs 2-4) but
vectorised on the GPU so that the 10 M-pair configuration is ready in seconds.
torch is used here for data creation only - it is not on the measured path.

The result is already in packed form (pack.py): reads in pair-completion order,
32-byte field records, 4-bit bases + qualities per row.  For FR pairs the left
mate is always seen first, so pair k completes when its right mate is met:
completion order = pairs sorted by the right mate's position (stable).
"""
from __future__ import annotations

import math

import numpy as np
import torch

from .pack import FIELDS_DT, PackedReads, row_geometry
from .synth import QUAL_PROBS, QUAL_VALUES, _umi_str


class DeviceBatch:
    """Packed reads in device memory (torch tensors own the storage)."""

    def __init__(self, fields, payload, dup_bc, lpad, stride, L, umi_len, predicted=None):
        self.fields, self.payload, self.dup_bc = fields, payload, dup_bc
        self.lpad, self.stride, self.L, self.umi_len = lpad, stride, L, umi_len
        self.predicted = predicted          # what the molecule model says the stats must be (predict_stats)

    @property
    def n(self):
        return int(self.fields.shape[0])

    @property
    def n_bc(self):
        return int(self.dup_bc.shape[0])

    def nbytes(self):
        return self.fields.numel() * 4 + self.payload.numel() + self.dup_bc.numel() * 4

    def to_packed(self) -> PackedReads:
        """Download (for the oracle / host-side checks)."""
        f = self.fields.cpu().numpy().view(np.uint8).reshape(-1).view(FIELDS_DT)
        nb = 4 ** self.umi_len
        barcodes = [_umi_str(v // nb, self.umi_len) + "." + _umi_str(v % nb, self.umi_len)
                    for v in range(nb * nb)]
        cig = "%dM_%dM" % (self.L, self.L)
        return PackedReads(f, self.payload.cpu().numpy(), self.lpad, barcodes,
                           self.dup_bc.cpu().numpy().astype(np.uint32), [cig], ["1"],
                           src=None, lens=np.full(self.n, self.L, np.int32),
                           counters=dict(total=self.n, unmapped=0, unmapped_mate=0,
                                         multiple_mapping=0, bad_spacer=0))


def _family_sizes(gen, n, dist, mean, device):
    if dist == "ztpoisson":
        lam = torch.full((n,), float(mean), device=device)
        out = torch.poisson(lam, generator=gen)
        for _ in range(64):
            bad = out == 0
            k = int(bad.sum())
            if k == 0:
                break
            out[bad] = torch.poisson(lam[:k], generator=gen)
        return out.clamp_(min=1).to(torch.int64)
    if dist == "lognormal":
        sigma = 1.0
        mu = math.log(mean) - sigma * sigma / 2
        z = torch.randn(n, device=device, generator=gen)
        return torch.clamp(torch.round(torch.exp(mu + sigma * z)), min=1).to(torch.int64)
    if dist == "wgs":
        out = torch.ones(n, device=device, dtype=torch.int64)
        dup = torch.rand(n, device=device, generator=gen) < 0.10
        extra = torch.poisson(torch.full((n,), 0.5, device=device), generator=gen).to(torch.int64)
        return torch.where(dup, 2 + extra, out)
    raise ValueError(dist)


def predict_stats(start, insert, umi_a, umi_b, n_plus, n_minus, nb, hist_bins=65536):
    """Stage counters and family-size histogram the molecule model implies (int64[16], int64[hist_bins]).

    Molecules that agree in start, insert and both UMIs are one molecule to the family key
    (their strands' pairs share every key field), so they are merged first.  A strand with p pairs
    gives two tag families of p reads (its R1 and its R2 reads); the duplex partner of a strand's
    R1 / R2 family is the other strand's R2 / R1 family (duplex_tag swaps the barcode halves and
    the read number).  Counter indices: include/ccgpu.h `cc_stat`."""
    key = ((start.to(torch.int64) * 4096 + insert.to(torch.int64)) * nb + umi_a.to(torch.int64)) * nb + umi_b.to(torch.int64)
    uniq, inv = torch.unique(key, return_inverse=True)
    P = torch.zeros(uniq.numel(), dtype=torch.int64, device=key.device).index_add_(0, inv, n_plus)
    M = torch.zeros(uniq.numel(), dtype=torch.int64, device=key.device).index_add_(0, inv, n_minus)
    big_p, big_m, one_p, one_m = P >= 2, M >= 2, P == 1, M == 1
    cnt = lambda m: 2 * int(m.sum())
    st = torch.zeros(16, dtype=torch.int64)
    st[0] = cnt(big_p) + cnt(big_m)                                   # SSCS
    st[1] = cnt(one_p) + cnt(one_m)                                   # singletons
    st[2] = cnt(big_p & big_m)                                        # DCS
    st[3] = cnt(big_p & ~big_m) + cnt(big_m & ~big_p)                 # SSCS without a partner SSCS
    st[4] = cnt(one_p & big_m) + cnt(one_m & big_p)                   # singletons corrected by an SSCS
    st[5] = 2 * cnt(one_p & one_m)                                    # ... by the partner singleton
    st[6] = cnt(one_p & (M == 0)) + cnt(one_m & (P == 0))             # uncorrected
    st[7] = cnt((P > 0) & (M > 0))                                    # DCS after the rescue
    st[8] = cnt(big_p & (M == 0)) + cnt(big_m & (P == 0))             # SSCS still alone after the rescue
    sizes = torch.cat([P[P > 0], M[M > 0]]).clamp_(max=hist_bins - 1)
    hist = 2 * torch.bincount(sizes, minlength=hist_bins).to(torch.int64).cpu()
    return st, hist


def generate_device(seed=1234, n_pairs=10_000_000, L=128, umi_len=2, dist="ztpoisson", fam_mean=4.0,
                    p_duplex=0.8, region=(0, 2_300_000), tid=0, insert_mean=170.0, insert_sd=20.0,
                    err=0.005, p_n=0.001, mapq=60, device="cuda:0", chunk=1 << 20, tail=None,
                    fields_out=None, payload_out=None, predict=False):
    """Exactly `n_pairs` read pairs (2*n_pairs reads) of duplex-UMI data on `device`.

    tail = (probability, factor): that share of the strands has `factor` times its drawn size (the
    10^3 - 10^4-read families of a deep panel).  fields_out / payload_out: write the batch into these
    slices of a larger batch instead of allocating.  predict: also derive the expected stage counters
    and size histogram from the molecule list (`DeviceBatch.predicted`)."""
    dev = torch.device(device)
    gen = torch.Generator(device=dev)
    gen.manual_seed(int(seed))
    mean_size = {"ztpoisson": fam_mean / (1 - math.exp(-fam_mean)), "lognormal": fam_mean, "wgs": 1.15}[dist]
    m = int(n_pairs / (mean_size * (1 + p_duplex)) * 1.02) + 16
    if n_pairs < 2_000_000:                  # small populations (one cytoband of a sharded set) vary more
        m += int(8 * math.sqrt(n_pairs))
    n_plus = _family_sizes(gen, m, dist, fam_mean, dev)
    has_minus = torch.rand(m, device=dev, generator=gen) < p_duplex
    n_minus = torch.where(has_minus, _family_sizes(gen, m, dist, fam_mean, dev), torch.zeros_like(n_plus))
    if tail is not None:
        for sizes in (n_plus, n_minus):
            jumbo = torch.rand(m, device=dev, generator=gen) < float(tail[0])
            sizes[jumbo] *= int(tail[1])
    per_mol = n_plus + n_minus
    cum = torch.cumsum(per_mol, 0)
    keep = int((cum <= n_pairs).sum())
    if keep == m and int(cum[-1]) < n_pairs:
        raise RuntimeError("molecule budget too small; raise the 1.02 head-room")
    short = n_pairs - (int(cum[keep - 1]) if keep else 0)      # top up with single-pair molecules
    n_plus = torch.cat([n_plus[:keep], torch.ones(short, device=dev, dtype=torch.int64)])
    n_minus = torch.cat([n_minus[:keep], torch.zeros(short, device=dev, dtype=torch.int64)])
    m = keep + short

    insert = torch.clamp(torch.round(insert_mean + insert_sd * torch.randn(m, device=dev, generator=gen)),
                         min=L).to(torch.int32)
    span = max(1, int(region[1]) - int(region[0]) - 2 * L - 400)
    start = (int(region[0]) + torch.randint(0, span, (m,), device=dev, generator=gen)).to(torch.int32)
    nb = 4 ** umi_len
    umi_a = torch.randint(0, nb, (m,), device=dev, generator=gen)
    umi_b = torch.randint(0, nb, (m,), device=dev, generator=gen)
    acgt = torch.tensor([1, 2, 4, 8], dtype=torch.uint8, device=dev)
    tmpl = acgt[torch.randint(0, 4, (m, 2, L), device=dev, generator=gen)]

    reps = torch.stack([n_plus, n_minus], 1).reshape(-1)
    mol_of_pair = torch.repeat_interleave(torch.arange(m, device=dev).repeat_interleave(2), reps)
    strand_of_pair = torch.repeat_interleave(torch.arange(2, device=dev).repeat(m), reps)
    assert mol_of_pair.numel() == n_pairs
    lo = start[mol_of_pair]
    hi = lo + insert[mol_of_pair] - L
    order = torch.sort(hi, stable=True).indices                 # pair completion order
    mol_of_pair, strand_of_pair, lo, hi = mol_of_pair[order], strand_of_pair[order], lo[order], hi[order]
    ins = insert[mol_of_pair]
    del order

    n = 2 * n_pairs
    predicted = predict_stats(start, insert, umi_a, umi_b, n_plus, n_minus, nb) if predict else None
    fields = torch.empty((n, 8), dtype=torch.int32, device=dev) if fields_out is None else fields_out
    assert tuple(fields.shape) == (n, 8) and fields.dtype == torch.int32
    ua = torch.where(strand_of_pair == 0, umi_a[mol_of_pair], umi_b[mol_of_pair])
    ub = torch.where(strand_of_pair == 0, umi_b[mol_of_pair], umi_a[mol_of_pair])
    bc = (ua * nb + ub).to(torch.int32)
    for end in (0, 1):                                          # 0 = left (first seen), 1 = right mate
        v = fields[end::2]
        v[:, 0] = bc
        v[:, 1] = 0                                             # single cigar pair "LM_LM"
        v[:, 2] = tid
        v[:, 3] = lo if end == 0 else hi
        v[:, 4] = tid
        v[:, 5] = hi if end == 0 else lo
        v[:, 6] = ins if end == 0 else -ins
        # + strand: 99 / 147, - strand: 163 / 83 ; meta word = flag | mapq << 16 | rg(0) << 24
        flag = torch.where(strand_of_pair == 0, torch.full_like(bc, 99 if end == 0 else 147),
                           torch.full_like(bc, 163 if end == 0 else 83))
        v[:, 7] = flag | (int(mapq) << 16)
    ids = torch.arange(nb * nb, device=dev)
    dup_bc = ((ids % nb) * nb + ids // nb).to(torch.int32)

    lpad, stride = row_geometry(L)
    if payload_out is None:
        payload = torch.zeros((n, stride), dtype=torch.uint8, device=dev)
    else:
        payload = payload_out
        assert tuple(payload.shape) == (n, stride) and payload.dtype == torch.uint8
        payload.zero_()
    qv = torch.tensor(QUAL_VALUES, dtype=torch.uint8, device=dev)
    qcum = torch.tensor(np.cumsum(QUAL_PROBS)[:-1], dtype=torch.float32, device=dev)
    pairs_per_chunk = max(1, chunk // 2)
    for p0 in range(0, n_pairs, pairs_per_chunk):
        p1 = min(n_pairs, p0 + pairs_per_chunk)
        mol = mol_of_pair[p0:p1].repeat_interleave(2)
        end = torch.arange(2, device=dev).repeat(p1 - p0)
        codes = tmpl[mol, end]                                   # [reads, L]
        k = codes.shape[0]
        sub = torch.rand((k, L), device=dev, generator=gen) < err
        shift = torch.randint(1, 4, (k, L), device=dev, generator=gen)
        cur = (codes >> 1).to(torch.int64)
        cur = torch.where(cur == 4, torch.full_like(cur, 3), cur)     # 1,2,4,8 -> 0,1,2,3
        codes = torch.where(sub, acgt[(cur + shift) % 4], codes)
        qual = qv[torch.bucketize(torch.rand((k, L), device=dev, generator=gen), qcum)]
        nn = torch.rand((k, L), device=dev, generator=gen) < p_n
        codes = torch.where(nn, torch.full_like(codes, 15), codes)
        qual = torch.where(nn, torch.full_like(qual, 2), qual)
        c = torch.zeros((k, lpad), dtype=torch.uint8, device=dev)
        c[:, :L] = codes
        rows = payload[2 * p0:2 * p1]
        rows[:, :lpad // 2] = c[:, 0::2] | (c[:, 1::2] << 4)
        rows[:, lpad // 2:lpad // 2 + L] = qual
    if dev.type == "cuda":
        torch.cuda.synchronize(dev)
    return DeviceBatch(fields, payload, dup_bc, lpad, stride, L, umi_len, predicted)
