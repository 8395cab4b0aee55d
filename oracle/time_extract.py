"""TEST INFRASTRUCTURE - throughput of the barcode extraction step: the repo's extract.py next to the UNMODIFIED
reference extract_barcodes.py (through oracle/refshim's Bio.SeqIO stand-in), same FASTQ pair, pattern NNT.
Build container only (/root/reference).   python -m oracle.time_extract [sample]"""
import contextlib
import io
import os
import runpy
import shutil
import sys
import tempfile
import time

from .run_reference import REF_PKG, REF_ROOT, SHIM_DIR

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main(sample="LargeMid_56_L005"):
    sys.path.insert(0, ROOT)
    from consensuscruncher_b200 import extract
    r1 = os.path.join(REF_ROOT, "test", "fastq", "%s_R1.fastq" % sample)
    r2 = os.path.join(REF_ROOT, "test", "fastq", "%s_R2.fastq" % sample)
    n = sum(1 for _ in open(r1)) // 4
    d = tempfile.mkdtemp()
    import gzip
    copies = 20                                             # the test pair is 10 000 records: repeat it
    n *= copies
    for k, src in (("R1", r1), ("R2", r2)):                 # the reference's plain-text branch opens with mode "rU",
        data = open(src, "rb").read()                       # gone in Python 3.11+: both run on gzipped copies
        with gzip.open(os.path.join(d, k + ".fastq.gz"), "wb", 1) as fo:
            for _ in range(copies):
                fo.write(data)
    r1, r2 = os.path.join(d, "R1.fastq.gz"), os.path.join(d, "R2.fastq.gz")
    try:
        for name, run in (("repo extract.py", "ours"), ("reference extract_barcodes.py (Bio.SeqIO stand-in)", "ref")):
            os.makedirs(os.path.join(d, run))
            argv = ["--read1", r1, "--read2", r2, "--outfile", os.path.join(d, run, "s"), "--bpattern", "NNT"]
            t0 = time.perf_counter()
            with contextlib.redirect_stdout(io.StringIO()), contextlib.redirect_stderr(io.StringIO()):
                if run == "ours":
                    extract.main(argv)
                else:
                    old_argv, old_path = sys.argv, list(sys.path)
                    sys.argv = ["extract_barcodes.py"] + argv
                    sys.path.insert(0, SHIM_DIR)
                    try:
                        runpy.run_path(os.path.join(REF_PKG, "extract_barcodes.py"), run_name="__main__")
                    finally:
                        sys.argv, sys.path[:] = old_argv, old_path
            dt = time.perf_counter() - t0
            print("%-52s %d read pairs in %.2f s = %.0f k pairs/s" % (name, n, dt, n / dt / 1e3))
    finally:
        shutil.rmtree(d, ignore_errors=True)


if __name__ == "__main__":
    main(*sys.argv[1:])
