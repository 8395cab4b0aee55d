"""Shared by the CPU and GPU tests: golden cases -> (reads, packed, params)."""
import numpy as np

from consensuscruncher_b200 import emit, pack
from oracle import golden


def load_case(name):
    g = golden.load(name)
    reads = [golden.tuple_to_read(t) for t in g["input"]]
    regions = None
    if g["params"].get("bed"):
        refs = g["header"]["references"]
        regions = [(refs.index(c), a, b) for c, a, b, _ in g["params"]["bed"]]
    packed = pack.pack_reads(reads, "sscs", "|", regions)
    return g, reads, packed


def assert_matches_golden(g, reads, packed, res):
    files, stats, fam = emit.emit_all(reads, packed, res, g["params"]["scorrect"])
    for name, recs in g["files"].items():
        exp = [golden.tuple_to_read(t).key() for t in recs]
        got = [r.key() for r in files.get(name, [])]
        assert len(exp) == len(got), "%s: %d records expected, %d produced" % (name, len(exp), len(got))
        for i, (a, b) in enumerate(zip(exp, got)):
            assert a == b, "%s record %d differs:\n  reference %r\n  produced  %r" % (name, i, a, b)
    assert stats == g["stats"]
    assert fam == g["read_families"]


def assert_same_result(a, b):
    bad = a.diff(b)
    if bad:
        k = bad[0]
        x, y = getattr(a, k), getattr(b, k)
        detail = "shape %s vs %s" % (x.shape, y.shape)
        if x.shape == y.shape:
            idx = np.argwhere(x != y)[:5].tolist()
            detail = "first differences at %s" % idx
        raise AssertionError("arrays differ: %s (%s: %s)" % (bad, k, detail))
