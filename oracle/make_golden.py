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
    "no_rg_aux": (dict(seed=19, n_molecules=40, L=40, fam_mean=2.0, p_duplex=0.7, region=(0, 2500),
                       rg_name=None),
                  dict(cutoff=0.7, scorrect=True, aux=True)),
}


def build_case(name):
    gen_kw, run_kw = CASES[name]
    run_kw = dict(run_kw)
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
