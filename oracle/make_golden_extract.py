"""TEST INFRASTRUCTURE - golden vectors for the barcode extraction step.

Runs the UNMODIFIED reference `ConsensusCruncher/extract_barcodes.py` (its `main()`, through runpy) on
the head of the reference's own test FASTQ pairs, with the `Bio.SeqIO` / `matplotlib.pyplot` stand-ins
of oracle/refshim, and stores inputs and every output file in tests/golden/extract_*.json.gz.
Only usable where /root/reference exists (the build container); the tests read the stored vectors.

    python -m oracle.make_golden_extract
"""
import contextlib
import gzip
import io
import json
import os
import runpy
import shutil
import sys
import tempfile

from .run_reference import REF_PKG, REF_ROOT, SHIM_DIR

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")

CASES = {
    "extract_pattern_nnt": dict(sample="LargeMid_56_L005", records=400, bpattern="NNT"),
    "extract_pattern_long": dict(sample="LargeMid_58_L005", records=300, bpattern="NNNNT"),
    "extract_pattern_inner_spacer": dict(sample="LargeMid_61_L006", records=300, bpattern="NTN"),
    "extract_list": dict(sample="LargeMid_57_L005", records=400,
                         blist=["ACT", "CTT", "GAT", "TGT", "AGT", "CAT", "GGT", "TCT", "ACGT", "TTAT"]),
}


def _head(path, records):
    with open(path) as fh:
        return "".join(fh.readline() for _ in range(4 * records))


def build_case(name):
    spec = CASES[name]
    d = tempfile.mkdtemp()
    try:
        inputs = {}
        for r in ("R1", "R2"):
            text = _head(os.path.join(REF_ROOT, "test", "fastq", "%s_%s.fastq" % (spec["sample"], r)), spec["records"])
            inputs[r] = text
            with gzip.open(os.path.join(d, "%s.fastq.gz" % r), "wt") as fh:      # 'gz' in the name: the gzip branch
                fh.write(text)
        os.makedirs(os.path.join(d, "out"))
        outfile = os.path.join(d, "out", "sample")
        argv = ["extract_barcodes.py", "--read1", os.path.join(d, "R1.fastq.gz"), "--read2", os.path.join(d, "R2.fastq.gz"),
                "--outfile", outfile]
        if "bpattern" in spec:
            argv += ["--bpattern", spec["bpattern"]]
        if "blist" in spec:
            with open(os.path.join(d, "blist.txt"), "w") as fh:
                fh.write("\n".join(spec["blist"]) + "\n")
            argv += ["--blist", os.path.join(d, "blist.txt")]
        old_argv, old_path = sys.argv, list(sys.path)
        sys.argv = argv
        sys.path.insert(0, SHIM_DIR)
        for m in [m for m in sys.modules if m == "Bio" or m.startswith("Bio.") or m.startswith("matplotlib")]:
            del sys.modules[m]
        err = io.StringIO()
        try:
            with contextlib.redirect_stdout(io.StringIO()), contextlib.redirect_stderr(err):
                runpy.run_path(os.path.join(REF_PKG, "extract_barcodes.py"), run_name="__main__")
        finally:
            sys.argv, sys.path[:] = old_argv, old_path
        files = {}
        for root, _, names in os.walk(d):
            for nm in names:
                p = os.path.join(root, nm)
                rel = os.path.relpath(p, d)
                if rel in ("R1.fastq.gz", "R2.fastq.gz", "blist.txt"):
                    continue
                files[rel] = open(p).read()
        return {"spec": spec, "inputs": inputs, "files": files, "stderr": err.getvalue()}
    finally:
        shutil.rmtree(d, ignore_errors=True)


def main(argv):
    for name in (argv or list(CASES)):
        g = build_case(name)
        with gzip.open(os.path.join(OUT, name + ".json.gz"), "wt") as fh:
            json.dump(g, fh)
        print(name, {k: len(v) for k, v in g["files"].items()}, g["stderr"].strip().replace("\n", " | "))


if __name__ == "__main__":
    main(sys.argv[1:])
