"""TEST INFRASTRUCTURE - wall-clock of the UNMODIFIED reference scripts (through oracle/run_reference.py's
pysam / samtools stand-ins) next to the C port of oracle/cc_oracle.c, on this machine's host cores.

  python oracle/time_reference.py [--reads 200000] [--procs N]

Cases: c1' (the stand-in for config 1: the reference's own test FASTQ reads at synthetic positions) and a
sample of config 2's distribution (L=128, ztPoisson(4) families, 80 % duplex).  "All cores" follows the
reference's own scale-out model - independent jobs (docs/source/FAQ.rst:4-7) - one process per core, each on
its own shard of the same distribution.  /root/reference exists only in the build container, so this
script runs there; its output is recorded in BASELINE.md section 2.
"""
import argparse
import json
import multiprocessing as mp
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _c2_sample(n_reads, seed):
    from consensuscruncher_b200 import synth
    mean = 4.0 / (1 - 2.718281828 ** -4.0)
    mol = max(1, int(n_reads / 2 / (mean * 1.8)))
    return synth.generate(seed=seed, n_molecules=mol, L=128)


def _time_python(args):
    kind, n_reads, seed = args
    from oracle import golden, make_golden, run_reference
    if kind == "c1prime":
        reads, header = make_golden.c1prime_reads()
    else:
        s = _c2_sample(n_reads, seed)
        reads, header = s.to_reads(), s.header()
    tuples = [golden.read_to_tuple(r) for r in reads]
    t0 = time.perf_counter()
    run_reference.run_reference(tuples, header, cutoff=0.7, scorrect=True, bedfile=None)
    return len(tuples), time.perf_counter() - t0


def _time_port(args):
    kind, n_reads, seed = args
    from oracle import oracle
    packed = _c2_sample(n_reads, seed).to_packed()
    t0 = time.perf_counter()
    oracle.run(packed, 0.7, True)
    return packed.n, time.perf_counter() - t0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reads", type=int, default=200_000)
    ap.add_argument("--procs", type=int, default=os.cpu_count() or 1)
    a = ap.parse_args()
    from oracle import oracle
    oracle.build()
    out = {"host_cores": os.cpu_count(), "python": sys.version.split()[0]}
    n, dt = _time_python(("c1prime", 0, 0))
    out["c1prime_python_1core"] = {"reads": n, "seconds": round(dt, 2), "reads_per_s": round(n / dt)}
    n, dt = _time_python(("c2", a.reads, 1))
    out["c2_python_1core"] = {"reads": n, "seconds": round(dt, 2), "reads_per_s": round(n / dt)}
    with mp.Pool(a.procs) as pool:
        t0 = time.perf_counter()
        res = pool.map(_time_python, [("c2", a.reads // a.procs, 100 + k) for k in range(a.procs)])
        wall = time.perf_counter() - t0
    n = sum(r[0] for r in res)
    out["c2_python_all_cores"] = {"processes": a.procs, "reads": n, "seconds": round(wall, 2), "reads_per_s": round(n / wall)}
    n, dt = _time_port(("c2", 10 * a.reads, 1))
    out["c2_port_1core"] = {"reads": n, "seconds": round(dt, 2), "reads_per_s": round(n / dt)}
    out["port_over_python_1core"] = round(out["c2_port_1core"]["reads_per_s"] / out["c2_python_1core"]["reads_per_s"], 1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
