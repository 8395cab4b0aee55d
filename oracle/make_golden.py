"""TEST INFRASTRUCTURE - writes tests/golden/*.json.gz by running the UNMODIFIED
reference (/root/reference, through oracle/run_reference.py) on seeded
synthetic inputs.  Run in the build container only:

    python -m oracle.make_golden

Each golden holds the input records, every output file of the reference's
consensus mode as record tuples, stats.txt and read_families.txt.
"""
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from consensuscruncher_b200 import synth            # noqa: E402
from oracle import golden, run_reference           # noqa: E402

BED3 = [("chr1", 0, 3000, "p1"), ("chr1", 3000, 12000, "p2"), ("chr2", 0, 6000, "p1"),
        ("chr2", 6000, 12000, "q1"), ("chr3", 0, 12000, "q")]

CASES = {
    # name: (generator kwargs, run kwargs)
    "duplex_small": (dict(seed=11, n_molecules=70, L=50, fam_mean=2.2, p_duplex=0.7, region=(0, 4000)),
                     dict(cutoff=0.7, scorrect=True)),
    "cutoff_050": (dict(seed=12, n_molecules=50, L=40, fam_mean=2.5, p_duplex=0.7, region=(0, 3000), err=0.08),
                   dict(cutoff=0.5, scorrect=True)),
    "cutoff_100": (dict(seed=12, n_molecules=50, L=40, fam_mean=2.5, p_duplex=0.7, region=(0, 3000), err=0.08),
                   dict(cutoff=1.0, scorrect=True)),
    "mode_ties": (dict(seed=13, n_molecules=60, L=36, fam_mean=3.0, p_duplex=0.8, region=(0, 3000),
                       mapq_mixed=True),
                  dict(cutoff=0.7, scorrect=True)),
    "undotted_umi": (dict(seed=14, n_molecules=50, L=33, fam_mean=2.0, p_duplex=0.8, region=(0, 2500),
                          umi_dot=False, umi_len=3),
                     dict(cutoff=0.7, scorrect=True)),
    "no_scorrect": (dict(seed=15, n_molecules=50, L=40, fam_mean=2.0, p_duplex=0.7, region=(0, 2500)),
                    dict(cutoff=0.7, scorrect=False)),
    "crowded": (dict(seed=16, n_molecules=150, L=32, fam_mean=3.0, p_duplex=0.9, region=(0, 700),
                     umi_len=1, insert_sd=3.0),
                dict(cutoff=0.7, scorrect=True)),
    "bed_three_chr": (dict(seed=17, n_molecules=90, L=40, fam_mean=2.0, p_duplex=0.7, region=(0, 10000),
                           n_tid=3),
                      dict(cutoff=0.7, scorrect=True, bed=BED3)),
    "big_families": (dict(seed=18, n_molecules=6, L=32, fam_mean=45.0, p_duplex=1.0, region=(0, 1500),
                          err=0.03),
                     dict(cutoff=0.7, scorrect=True)),
    "c1prime_real_fastq": (None, dict(cutoff=0.7, scorrect=True)),
    "no_rg_aux": (dict(seed=19, n_molecules=40, L=40, fam_mean=2.0, p_duplex=0.7, region=(0, 2500),
                       rg_name=None),
                  dict(cutoff=0.7, scorrect=True, aux=True)),
}


def c1prime_reads(seed=56, pairs=1500, molecules=140, fam_mean=2.5, p_duplex=0.75, err=0.01):
    """Config 1 of BASELINE.json cannot be produced here (no bwa / hg19 / picard), so this is its
    stand-in (SURVEY.md 8d "c1'"): the reference's own test FASTQ pairs, barcodes extracted with pattern
    NNT (extract.py, pinned against the reference's extract_barcodes.py), real 123-base sequences as
    molecule templates, real quality strings (the full HiSeq alphabet, Q2 'N' calls included) on every
    copy, synthetic positions and cigar 123M.  -> (reads in coordinate order, header)"""
    import numpy as np
    from consensuscruncher_b200 import extract
    from consensuscruncher_b200.records import Read
    rng = np.random.default_rng(seed)
    d = tempfile.mkdtemp()
    for r in ("R1", "R2"):
        with open(os.path.join(run_reference.REF_ROOT, "test", "fastq", "LargeMid_56_L005_%s.fastq" % r)) as fh, \
                open(os.path.join(d, r + ".fastq"), "w") as out:
            out.write("".join(fh.readline() for _ in range(4 * pairs)))
    extract.extract_pattern(os.path.join(d, "R1.fastq"), os.path.join(d, "R2.fastq"), os.path.join(d, "x"), "NNT")
    mates = [list(extract.read_fastq(os.path.join(d, "x_barcode_%s.fastq" % r))) for r in ("R1", "R2")]
    L = 123
    real = [(t1.split("|")[0], t1.split("|")[1].split("/")[0], s1, q1, s2, q2)
            for (t1, s1, q1), (_, s2, q2) in zip(*mates) if len(s1) == L and len(s2) == L]
    assert len(real) > 2 * molecules
    reads, donor = [], molecules                 # records [0, molecules) are templates, the rest lend names + qualities
    for m in range(molecules):
        _, umi, s1, _, s2, _ = real[m]
        a, b = umi.split(".")
        start = int(rng.integers(100, 60000))
        insert = int(max(L, rng.normal(170, 20)))
        for strand in (0, 1):
            if strand == 1 and rng.random() >= p_duplex:
                continue
            for _ in range(max(1, int(rng.poisson(fam_mean)))):
                name, _, _, q1, _, q2 = real[donor % len(real)]
                donor += 1
                seqs = []
                for s in (s1, s2):
                    s = list(s)
                    for i in np.flatnonzero(rng.random(L) < err):
                        if s[i] != "N":
                            s[i] = "ACGT"[("ACGT".index(s[i]) + int(rng.integers(1, 4))) % 4]
                    seqs.append("".join(s))
                quals = [np.frombuffer(q.encode(), np.uint8) - 33 for q in (q1, q2)]
                for i in (0, 1):                 # an N call carries Q2 in real data; a confident N kills the reference (App. B)
                    quals[i] = quals[i].copy()
                    quals[i][np.frombuffer(seqs[i].encode(), np.uint8) == ord("N")] = 2
                qname = "%s|%s.%s" % (name, a, b) if strand == 0 else "%s|%s.%s" % (name, b, a)
                lo, hi = start, start + insert - L
                # + strand: lo = R1 fwd (99), hi = R2 rev (147); - strand: lo = R2 fwd (163), hi = R1 rev (83)
                first, second = (0, 1) if strand == 0 else (1, 0)
                flags = (99, 147) if strand == 0 else (163, 83)
                reads.append(Read(qname, flags[0], 0, lo, 60, "%dM" % L, 0, hi, insert, seqs[first], quals[first], rg="1"))
                reads.append(Read(qname, flags[1], 0, hi, 60, "%dM" % L, 0, lo, -insert, seqs[second], quals[second], rg="1"))
    order = sorted(range(len(reads)), key=lambda i: reads[i].pos)
    return [reads[i] for i in order], {"references": ["chr1"], "lengths": [100000]}


def build_case(name):
    gen_kw, run_kw = CASES[name]
    run_kw = dict(run_kw)
    if gen_kw is None:                           # the stand-in for config 1
        reads, header = c1prime_reads()
        tuples = [golden.read_to_tuple(r) for r in reads]
        out = run_reference.run_reference(tuples, header, bedfile=None, **run_kw)
        return golden.save(name, dict(gen="c1prime_reads()", cutoff=run_kw["cutoff"], scorrect=run_kw["scorrect"], bed=None),
                           header, tuples, out)
    s = synth.generate(**gen_kw)
    reads = s.to_reads()
    if run_kw.pop("aux", False):
        for i, r in enumerate(reads):
            r.aux = {"NM": i % 3, "XS": "t%d" % (i % 5)}
    bed = run_kw.pop("bed", None)
    bedfile = None
    if bed:
        fd, bedfile = tempfile.mkstemp(suffix=".bed")
        with os.fdopen(fd, "w") as fh:
            for c, a, b, nm in bed:
                fh.write("%s\t%d\t%d\t%s\tgneg\n" % (c, a, b, nm))
    tuples = [golden.read_to_tuple(r) for r in reads]
    out = run_reference.run_reference(tuples, s.header(), bedfile=bedfile, **run_kw)
    if bedfile:
        os.remove(bedfile)
    params = dict(gen=gen_kw, cutoff=run_kw["cutoff"], scorrect=run_kw["scorrect"], bed=bed)
    return golden.save(name, params, s.header(), tuples, out)


def main(argv):
    for name in (argv or list(CASES)):
        path = build_case(name)
        g = golden.load(name)
        print("%-16s %6d input records  %s  %d bytes" % (
            name, len(g["input"]), {k: len(v) for k, v in g["files"].items()}, os.path.getsize(path)))


if __name__ == "__main__":
    main(sys.argv[1:])
