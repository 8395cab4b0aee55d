"""Seeded synthetic duplex-UMI data (BASELINE.json configs 2-4).

A molecule is (tid, start, insert, umiA, umiB) with two read templates (left
end, right end, both in reference orientation).  Its + strand gives pairs
R1 fwd @start (flag 99) / R2 rev @start+insert-L (flag 147) named `...|A.B`;
its - strand gives R1 rev @start+insert-L (flag 83) / R2 fwd @start (flag 163)
named `...|B.A`.  Per-strand family sizes follow the chosen distribution; each
copy is the template with substitution errors and drawn base qualities.

`SynthReads.to_reads()` gives coordinate-sorted `Read` records (small cases, the
reference / oracle side); `SynthReads.to_packed()` gives the same data packed
for the GPU without going through Python objects (large cases).  Both describe
the same reads in the same order, which a test checks.
"""
from __future__ import annotations

import numpy as np

from .pack import FIELDS_DT, PackedReads, row_geometry, NO_RG
from .records import Read, codes_to_seq

_ACGT = np.array([1, 2, 4, 8], np.uint8)          # BAM codes of A, C, G, T
_UMI_CH = "ACGT"
QUAL_VALUES = np.array([12, 22, 27, 32, 37, 41], np.uint8)
QUAL_PROBS = np.array([0.02, 0.03, 0.03, 0.12, 0.30, 0.50])   # P(Q < 30) = 0.08


def _family_sizes(rng, n, dist, mean):
    if dist == "ztpoisson":                     # zero-truncated Poisson(mean)
        out = rng.poisson(mean, n)
        bad = out == 0
        while bad.any():
            out[bad] = rng.poisson(mean, int(bad.sum()))
            bad = out == 0
        return out
    if dist == "lognormal":                     # heavy tail, mean ~ `mean`
        sigma = 1.0
        mu = np.log(mean) - sigma * sigma / 2
        return np.maximum(1, np.rint(rng.lognormal(mu, sigma, n))).astype(np.int64)
    if dist == "wgs":                           # ~90 % singletons
        out = np.ones(n, np.int64)
        dup = rng.random(n) < 0.10
        out[dup] = 2 + rng.poisson(0.5, int(dup.sum()))
        return out
    if dist == "const":
        return np.full(n, int(mean), np.int64)
    raise ValueError(dist)


class SynthReads:
    """Per-read arrays in coordinate-sorted ("BAM") order."""

    def __init__(self, **kw):
        self.__dict__.update(kw)

    @property
    def n(self):
        return int(self.flag.shape[0])

    # -- record view (small inputs) ----------------------------------------
    def qname(self, i):
        a, b = self.umi_a[self.mol[i]], self.umi_b[self.mol[i]]
        if self.strand[i]:
            a, b = b, a
        sep = "." if self.umi_dot else ""
        return "M%d:%d:%d|%s%s%s" % (self.mol[i], self.strand[i], self.copy[i],
                                      _umi_str(a, self.umi_len), sep, _umi_str(b, self.umi_len))

    def to_reads(self):
        out = []
        cig = "%dM" % self.L
        for i in range(self.n):
            out.append(Read(self.qname(i), self.flag[i], self.tid[i], self.pos[i], self.mapq[i],
                            cig, self.mtid[i], self.mpos[i], self.tlen[i],
                            codes_to_seq(self.codes[i]), self.qual[i].copy(),
                            rg=self.rg_name, aux=b""))
        return out

    def header(self):
        return {"references": list(self.ref_names), "lengths": list(self.ref_lengths)}

    # -- packed view (any size, vectorised) ---------------------------------
    def to_packed(self):
        n = self.n
        idx = np.arange(n)
        # pair k of read i; the mate is the other read with the same pair id
        order = np.argsort(self.pair, kind="stable")
        first = np.empty(n // 2, np.int64)
        second = np.empty(n // 2, np.int64)
        first[:] = order[0::2]
        second[:] = order[1::2]                    # stable => first < second in BAM order
        done = np.argsort(second, kind="stable")   # pairs complete when the later mate is seen
        seq_idx = np.empty(n, np.int64)
        seq_idx[0::2] = first[done]
        seq_idx[1::2] = second[done]
        fields = np.zeros(n, FIELDS_DT)
        a = self.umi_a[self.mol]
        b = self.umi_b[self.mol]
        sw = self.strand.astype(bool)
        ua = np.where(sw, b, a).astype(np.uint64)
        ub = np.where(sw, a, b).astype(np.uint64)
        nb = 4 ** self.umi_len
        bc = (ua * nb + ub)
        # dictionary ids in first-appearance order over completed pairs, like pack_reads
        bc_seq = bc[seq_idx[0::2]]
        uniq, first_at = np.unique(bc_seq, return_index=True)
        rank = np.empty(len(uniq), np.int64)
        rank[np.argsort(first_at, kind="stable")] = np.arange(len(uniq))
        ids = rank[np.searchsorted(uniq, bc)]
        by_first = uniq[np.argsort(first_at, kind="stable")]
        sep = "." if self.umi_dot else ""
        barcodes = [_umi_str(int(v) // nb, self.umi_len) + sep + _umi_str(int(v) % nb, self.umi_len)
                    for v in by_first]
        swapped = (by_first % nb) * nb + by_first // nb
        pos_in = np.searchsorted(uniq, swapped)
        pos_in = np.minimum(pos_in, len(uniq) - 1)
        found = uniq[pos_in] == swapped
        dup_bc = np.where(found, rank[pos_in], 0xFFFFFFFF).astype(np.uint32)
        s = seq_idx
        fields["bc"] = ids[s]
        fields["cig"] = 0
        fields["tid"] = self.tid[s]
        fields["pos"] = self.pos[s]
        fields["mtid"] = self.mtid[s]
        fields["mpos"] = self.mpos[s]
        fields["tlen"] = self.tlen[s]
        fields["flag"] = self.flag[s]
        fields["mapq"] = self.mapq[s]
        fields["rg"] = 0 if self.rg_name is not None else NO_RG
        lpad, stride = row_geometry(self.L)
        payload = np.zeros((n, stride), np.uint8)
        c = np.zeros((n, lpad), np.uint8)
        c[:, :self.L] = self.codes[s]
        payload[:, :lpad // 2] = c[:, 0::2] | (c[:, 1::2] << 4)
        payload[:, lpad // 2:lpad // 2 + self.L] = self.qual[s]
        cig = "%dM_%dM" % (self.L, self.L)
        return PackedReads(fields, payload, lpad, barcodes, dup_bc, [cig],
                           [self.rg_name] if self.rg_name is not None else [],
                           src=s.copy(), lens=np.full(n, self.L, np.int32),
                           counters=dict(total=n, unmapped=0, unmapped_mate=0,
                                         multiple_mapping=0, bad_spacer=0))


def _umi_str(v, k):
    return "".join(_UMI_CH[(int(v) >> (2 * (k - 1 - i))) & 3] for i in range(k))


def generate(seed=1234, n_molecules=1000, L=128, umi_len=2, umi_dot=True, dist="ztpoisson",
             fam_mean=4.0, p_duplex=0.8, n_tid=1, region=(0, 2_300_000), insert_mean=170.0,
             insert_sd=20.0, err=0.005, p_n=0.001, mapq=60, mapq_mixed=False,
             rg_name="1", ref_names=None, ref_lengths=None, max_family=None):
    """Draw molecules and expand them into coordinate-sorted reads (see module doc)."""
    rng = np.random.default_rng(seed)
    m = int(n_molecules)
    tid = rng.integers(0, n_tid, m).astype(np.int32)
    insert = np.maximum(L, np.rint(rng.normal(insert_mean, insert_sd, m))).astype(np.int32)
    start = rng.integers(region[0], max(region[0] + 1, region[1] - 2 * L - 400), m).astype(np.int32)
    umi_a = rng.integers(0, 4 ** umi_len, m)
    umi_b = rng.integers(0, 4 ** umi_len, m)
    n_plus = _family_sizes(rng, m, dist, fam_mean)
    has_minus = rng.random(m) < p_duplex
    n_minus = np.where(has_minus, _family_sizes(rng, m, dist, fam_mean), 0)
    if max_family:
        n_plus = np.minimum(n_plus, max_family)
        n_minus = np.minimum(n_minus, max_family)
    tmpl = _ACGT[rng.integers(0, 4, (m, 2, L))]              # [molecule][lo/hi end][L]

    # one row per read pair: (molecule, strand, copy)
    reps = np.stack([n_plus, n_minus], 1).reshape(-1)        # m*2 (strand-major within molecule)
    pm = np.repeat(np.repeat(np.arange(m), 2), reps)
    ps = np.repeat(np.tile(np.array([0, 1]), m), reps)
    starts = np.cumsum(reps) - reps
    pc = np.arange(len(pm)) - np.repeat(starts, reps)
    npairs = len(pm)
    # two reads per pair: end 0 = left (lo) read, end 1 = right (hi) read
    mol = np.repeat(pm, 2)
    strand = np.repeat(ps, 2).astype(np.uint8)
    copy = np.repeat(pc, 2)
    end = np.tile(np.array([0, 1]), npairs)
    pair = np.repeat(np.arange(npairs), 2)
    lo = start[mol]
    hi = lo + insert[mol] - L
    pos = np.where(end == 0, lo, hi).astype(np.int32)
    mpos = np.where(end == 0, hi, lo).astype(np.int32)
    tlen = np.where(end == 0, insert[mol], -insert[mol]).astype(np.int32)
    # + strand: lo = R1 fwd (99), hi = R2 rev (147); - strand: lo = R2 fwd (163), hi = R1 rev (83)
    flag = np.where(strand == 0, np.where(end == 0, 99, 147),
                    np.where(end == 0, 163, 83)).astype(np.uint16)
    n = 2 * npairs
    codes = tmpl[mol, end].copy()
    sub = rng.random((n, L)) < err
    shift = rng.integers(1, 4, int(sub.sum()))
    cur = np.log2(codes[sub]).astype(np.int64)
    codes[sub] = _ACGT[(cur + shift) % 4]
    qual = QUAL_VALUES[rng.choice(len(QUAL_VALUES), size=(n, L), p=QUAL_PROBS)]
    nn = rng.random((n, L)) < p_n                            # N calls only ever carry Q2
    codes[nn] = 15
    qual[nn] = 2
    if mapq_mixed:
        mq = rng.choice(np.array([20, 40, 60], np.uint8), n)
    else:
        mq = np.full(n, mapq, np.uint8)
    # coordinate sort (stable in generation order), as `samtools sort` leaves the input
    order = np.lexsort((pos, tid[mol]))
    take = lambda x: x[order]
    if ref_names is None:
        ref_names = ["chr%d" % (i + 1) for i in range(n_tid)]
    if ref_lengths is None:
        ref_lengths = [int(region[1]) + 1000] * len(ref_names)
    return SynthReads(L=L, umi_len=umi_len, umi_dot=umi_dot, umi_a=umi_a, umi_b=umi_b,
                      mol=take(mol), strand=take(strand), copy=take(copy), pair=take(pair),
                      flag=take(flag), tid=take(tid[mol]), pos=take(pos), mtid=take(tid[mol]),
                      mpos=take(mpos), tlen=take(tlen), mapq=take(mq), codes=take(codes),
                      qual=take(qual), rg_name=rg_name, ref_names=ref_names,
                      ref_lengths=ref_lengths, n_molecules=m)
