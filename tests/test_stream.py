"""CPU: region chunking of a packed batch and the joining of chunk results (stream.py), with the oracle
standing in for the GPU context - the chunked result must equal one pass over the whole batch."""
import numpy as np
import pytest

from consensuscruncher_b200 import stream, synth
from consensuscruncher_b200.pack import DomainError, PackedReads
from oracle import oracle

import helpers


class OracleEngine:
    """The slice of GpuConsensus that stream.py uses, answered by the CPU oracle."""

    def __init__(self, scorrect=True):
        self.scorrect, self.last = scorrect, None

    def consensus(self, packed, compact=False):
        res, faults = oracle.run(packed, 0.7, self.scorrect, return_faults=True)
        self.last = res
        if faults:
            raise DomainError("a read pair does not found/join its two families together")
        return res

    def fetch(self, name):
        return getattr(self.last, name)


@pytest.mark.parametrize("scorrect", [True, False])
def test_chunks_join_into_the_result_of_one_pass(scorrect):
    s = synth.generate(seed=17, n_molecules=5000, L=40, n_tid=3, region=(0, 90000), fam_mean=2.5, p_duplex=0.6)
    packed = s.to_packed()
    whole = oracle.run(packed, 0.7, scorrect)
    regions = [(t, a, a + 10000) for t in range(3) for a in range(0, 90000, 10000)]
    for regs, budget in ((regions, 4000), (None, 3000), (regions, 10 ** 9)):
        cuts = stream.chunk_bounds(packed, regs, budget)
        assert cuts[0] == 0 and cuts[-1] == packed.n and all(c % 2 == 0 for c in cuts)
        if budget < packed.n:
            assert len(cuts) > 4
        res, kept, n_chunks = stream.consensus_chunked(packed, [OracleEngine(scorrect), OracleEngine(scorrect)], regs, budget)
        assert n_chunks == len(cuts) - 1 and kept is packed
        helpers.assert_same_result(res, whole)


def test_input_out_of_order_is_one_chunk():
    packed = synth.generate(seed=18, n_molecules=800, L=32).to_packed()
    order = np.random.default_rng(1).permutation(packed.n // 2)
    idx = np.stack([2 * order, 2 * order + 1], 1).reshape(-1)
    shuffled = PackedReads(packed.fields[idx], packed.payload[idx], packed.lpad, packed.barcodes, packed.dup_bc,
                           packed.cigar_pairs, packed.rgs)
    assert stream.chunk_bounds(shuffled, None, 100) == [0, shuffled.n]


def test_pairs_outside_the_domain_are_set_aside_not_fatal():
    """A 99/147 pair and a 67/131 pair at the same coordinates, barcode and cigars: the R1 tags coincide, the
    R2 tags differ (consensus_helper.py:456-468 prints 'Consensus tag NOT UNIQUE').  The sample must go through."""
    packed = synth.generate(seed=19, n_molecules=400, L=32, p_duplex=1.0).to_packed()
    f = packed.fields.copy()
    k = int(np.flatnonzero((f["flag"][0::2] == 99))[5])                # one + strand pair ...
    twin_f = f[2 * k:2 * k + 2].copy()
    twin_f["flag"] = [67, 131]                                         # ... and its twin with the other orientation flags
    fields = np.concatenate([f[:2 * k + 2], twin_f, f[2 * k + 2:]])
    payload = np.concatenate([packed.payload[:2 * k + 2], packed.payload[2 * k:2 * k + 2], packed.payload[2 * k + 2:]])
    src = np.arange(fields.shape[0], dtype=np.int64)
    odd = PackedReads(fields, payload, packed.lpad, packed.barcodes, packed.dup_bc, packed.cigar_pairs, packed.rgs, src=src,
                      lens=np.full(fields.shape[0], 32, np.int32), counters=dict(packed.counters))
    with pytest.raises(DomainError):
        OracleEngine().consensus(odd)
    messages = []
    res, kept, _ = stream.consensus_chunked(odd, [OracleEngine()], None, 10 ** 9, log=messages.append)
    assert messages and "set aside" in messages[0]
    assert kept.n < odd.n and kept.n >= odd.n - 40 and len(kept.bad) == odd.n - kept.n
    helpers.assert_same_result(res, oracle.run(kept, 0.7, True))
