"""CPU: the oracle (oracle/cc_oracle.c) against the reference's own outputs and
known-answer vectors.  Goldens were produced by the unmodified reference
(oracle/make_golden.py); the KATs are those of SURVEY.md App. D."""
import numpy as np
import pytest

from consensuscruncher_b200 import pack
from consensuscruncher_b200.records import Read, codes_to_seq
from oracle import golden, oracle

import helpers


@pytest.mark.parametrize("name", golden.names())
def test_oracle_reproduces_reference_outputs(name):
    g, reads, packed = helpers.load_case(name)
    res, faults = oracle.run(packed, g["params"]["cutoff"], g["params"]["scorrect"], return_faults=True)
    assert faults == 0
    helpers.assert_matches_golden(g, reads, packed, res)


def _family(seqs, quals):
    lpad, stride = pack.row_geometry(len(seqs[0]))
    rows = np.zeros((len(seqs), stride), np.uint8)
    for i, (s, q) in enumerate(zip(seqs, quals)):
        pack.pack_row(pack.seq_to_codes(s), np.asarray(q, np.uint8), lpad, rows[i])
    return rows, lpad


def _vote(seqs, quals, cutoff):
    rows, lpad = _family(seqs, quals)
    out = oracle.consensus_maker_rows(rows, lpad, cutoff)
    codes, qual = pack.unpack_row(out, lpad, len(seqs[0]))
    return codes_to_seq(codes), qual.tolist()


def test_kat_sscs_maker_docstring():
    # SSCS_maker.py:23-29 / --cutoff help :176-184
    seqs = ["ACTGATACTT", "ACTGAAACCT", "ACTGATACCT", "ACTGATACTT"]
    assert _vote(seqs, [[40] * 10] * 4, 0.7) == ("ACTGATACNT", [60] * 10)


def test_kat_ties_and_cutoff():
    seqs = ["AC", "CC", "AA", "CA"]
    quals = [[40, 40], [40, 10], [40, 40], [40, 40]]
    assert _vote(seqs, quals, 0.5) == ("AA", [60, 60])
    assert _vote(seqs, quals, 0.7) == ("NN", [60, 60])


def test_kat_all_fail_and_n_keeps_quality():
    assert _vote(["A", "A"], [[10], [20]], 0.7) == ("N", [0])
    # one passing read out of two: winner has count 1, N keeps its quality when the cutoff fails
    assert _vote(["A", "C"], [[35], [33]], 0.7) == ("N", [35])
    assert _vote(["A", "C"], [[35], [33]], 0.5) == ("A", [35])


def test_kat_n_with_high_quality_is_an_error():
    with pytest.raises(oracle.OracleError):
        _vote(["N", "A"], [[35], [35]], 0.7)


def test_kat_duplex_rules():
    rows, lpad = _family(["ACGTN", "ACGAN"], [[40, 20, 35, 40, 2], [25, 20, 30, 40, 2]])
    out = oracle.duplex_consensus_rows(rows[0], rows[1], lpad, sc=False)
    codes, qual = pack.unpack_row(out, lpad, 5)
    assert (codes_to_seq(codes), qual.tolist()) == ("ACGNN", [60, 40, 60, 0, 4])
    out = oracle.duplex_consensus_rows(rows[0], rows[1], lpad, sc=True)
    codes, qual = pack.unpack_row(out, lpad, 5)
    assert (codes_to_seq(codes), qual.tolist()) == ("NNGNN", [0, 0, 60, 0, 0])


def test_oracle_empty_and_single_pair():
    from consensuscruncher_b200 import synth
    s = synth.generate(seed=3, n_molecules=1, dist="const", fam_mean=1, p_duplex=0.0, L=33)
    p = s.to_packed()
    res = oracle.run(p)
    assert res.n_fam == 2 and res.stat("singletons") == 2 and res.stat("uncorrected") == 2
    empty = pack.pack_reads([])
    res = oracle.run(empty)
    assert res.n_fam == 0 and res.sscs_rows.shape[0] == 0
