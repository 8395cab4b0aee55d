"""TEST INFRASTRUCTURE - run the UNMODIFIED reference stage scripts in-process.

`run_reference()` drives /root/reference/ConsensusCruncher/{SSCS_maker,DCS_maker,
singleton_correction}.py in the stage order of `ConsensusCruncher.py:148-303`
(sort_index after every stage, the two `samtools merge` calls), with the
`pysam` / `matplotlib.pyplot` stand-ins of `oracle/refshim` on `sys.path`.  The
consensus arithmetic that runs is therefore the reference's own code.

Two things are pinned from outside, and only these:
  * `samtools sort` / `merge` are replaced by a stable (tid, pos, is_reverse)
    sort and by concatenation in argument order + the same sort (SURVEY.md
    App. A.8) - the binary is not available here;
  * `random.randint` is replaced by `lambda a, b: a`, so the reference's
    unseeded tie-break in `read_mode` / `consensus_flag`
    (consensus_helper.py:495,531) always takes the first-appearing of the tied
    values.  That is one of the outcomes the reference can produce, and the
    policy the GPU path implements.

/root/reference exists only in the build container: this module is used by
`oracle/make_golden.py` to write `tests/golden/*`, never at GPU-test time.
"""
import contextlib
import io
import os
import pickle
import random
import runpy
import shutil
import sys
import tempfile

REF_ROOT = "/root/reference"
REF_PKG = os.path.join(REF_ROOT, "ConsensusCruncher")
SHIM_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "refshim")


def reference_available():
    return os.path.isfile(os.path.join(REF_PKG, "SSCS_maker.py"))


def _load(path):
    with open(path, "rb") as fh:
        return pickle.load(fh)


def _dump(path, header, records):
    with open(path, "wb") as fh:
        pickle.dump({"header": header, "records": list(records)}, fh)


def _sort_key(t):
    tid = t[2] if t[2] >= 0 else (1 << 31)
    return (tid, t[3], 1 if (t[1] & 0x10) else 0)


def sort_index(path):
    """ConsensusCruncher.py:10-29 with the sort model of SURVEY.md App. A.8."""
    blob = _load(path)
    out = path.split(".bam", 1)[0] + ".sorted.bam"
    _dump(out, blob["header"], sorted(blob["records"], key=_sort_key))
    os.remove(path)
    return out


def merge(out, *ins):
    recs, header = [], None
    for p in ins:
        blob = _load(p)
        header = header or blob["header"]
        recs.extend(blob["records"])
    _dump(out, header, recs)


@contextlib.contextmanager
def _reference_env():
    saved_path = list(sys.path)
    saved_mods = {k: sys.modules.get(k) for k in
                  ("pysam", "matplotlib", "matplotlib.pyplot", "consensus_helper")}
    saved_randint = random.randint
    saved_argv = list(sys.argv)
    for k in saved_mods:
        sys.modules.pop(k, None)
    sys.path[:0] = [SHIM_DIR, REF_PKG]
    random.randint = lambda a, b: a
    try:
        yield
    finally:
        random.randint = saved_randint
        sys.argv[:] = saved_argv
        sys.path[:] = saved_path
        for k, v in saved_mods.items():
            sys.modules.pop(k, None)
            if v is not None:
                sys.modules[k] = v


def _run_script(name, argv, quiet=True):
    import warnings
    sys.argv[:] = [name] + [str(a) for a in argv]
    sink = io.StringIO()
    status = "ok"
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        with contextlib.redirect_stdout(sink if quiet else sys.stdout):
            try:
                runpy.run_path(os.path.join(REF_PKG, name), run_name="__main__")
            except ZeroDivisionError:
                status = "ZeroDivisionError"     # singleton_correction.py:291 with no singletons (Q20)
            finally:
                import pysam                     # the shim
                pysam.flush_all()
    return status, sink.getvalue()


def run_reference(records, header, cutoff=0.7, scorrect=True, bedfile=None, bdelim="|",
                  workdir=None, keep=False):
    """Run the reference's consensus mode on `records` (tuples, file order).

    Returns {"files": {name: [record tuples]}, "stats": str, "read_families": str,
             "status": {stage: "ok"|...}}; file names are those of SURVEY.md App. A.6
    without the sample prefix (e.g. "sscs.sorted", "all.unique.dcs.sorted").
    """
    if not reference_available():
        raise RuntimeError("/root/reference is not present on this machine")
    tmp = workdir or tempfile.mkdtemp(prefix="ccref_")
    sid = "s"
    d = os.path.join(tmp, sid)
    for sub in ("sscs", "dcs", "sscs_sc", "dcs_sc"):
        os.makedirs(os.path.join(d, sub), exist_ok=True)
    inp = os.path.join(tmp, sid + ".bam")
    _dump(inp, header, records)
    bed = [] if bedfile in (None, "False", False) else ["--bedfile", bedfile]
    status = {}
    with _reference_env():
        sscs = "%s/sscs/%s.sscs.bam" % (d, sid)
        sing = "%s/sscs/%s.singleton.bam" % (d, sid)
        a = ["--infile", inp, "--outfile", sscs, "--cutoff", cutoff] + bed
        if bdelim != "|":
            a += ["--bdelim", bdelim]
        status["sscs"], _ = _run_script("SSCS_maker.py", a)
        sscs = sort_index(sscs)
        sing = sort_index(sing)
        stats = "%s/sscs/%s.stats.txt" % (d, sid)
        tt = "%s/sscs/%s.time_tracker.txt" % (d, sid)
        os.rename(stats, "%s/dcs/%s.stats.txt" % (d, sid))
        os.rename(tt, "%s/dcs/%s.time_tracker.txt" % (d, sid))
        dcs = "%s/dcs/%s.dcs.bam" % (d, sid)
        status["dcs"], _ = _run_script("DCS_maker.py", ["--infile", sscs, "--outfile", dcs] + bed)
        dcs = sort_index(dcs)
        sort_index("%s/dcs/%s.sscs.singleton.bam" % (d, sid))
        stats_final = "%s/dcs/%s.stats.txt" % (d, sid)
        if scorrect:
            os.rename("%s/dcs/%s.stats.txt" % (d, sid), stats)
            os.rename("%s/dcs/%s.time_tracker.txt" % (d, sid), tt)
            status["sc"], _ = _run_script("singleton_correction.py", ["--singleton", sing] + bed)
            moved = {}
            for nm in ("sscs.correction", "singleton.correction", "uncorrected"):
                src = "%s/sscs/%s.%s.bam" % (d, sid, nm)
                dst = "%s/sscs_sc/%s.%s.bam" % (d, sid, nm)
                os.rename(src, dst)
                moved[nm] = sort_index(dst)
            sscs_sc = "%s/sscs_sc/%s.sscs.sc.bam" % (d, sid)
            merge(sscs_sc, sscs, moved["sscs.correction"], moved["singleton.correction"])
            sscs_sc = sort_index(sscs_sc)
            os.rename(stats, "%s/dcs_sc/%s.stats.txt" % (d, sid))
            os.rename(tt, "%s/dcs_sc/%s.time_tracker.txt" % (d, sid))
            dcs_sc = "%s/dcs_sc/%s.dcs.sc.bam" % (d, sid)
            status["dcs_sc"], _ = _run_script("DCS_maker.py",
                                              ["--infile", sscs_sc, "--outfile", dcs_sc] + bed)
            dcs_sc = sort_index(dcs_sc)
            sscs_sc_sing = sort_index("%s/dcs_sc/%s.sscs.sc.singleton.bam" % (d, sid))
            allu = "%s/dcs_sc/%s.all.unique.dcs.bam" % (d, sid)
            merge(allu, dcs_sc, sscs_sc_sing, moved["uncorrected"])
            sort_index(allu)
            stats_final = "%s/dcs_sc/%s.stats.txt" % (d, sid)
    files = {}
    for sub in ("sscs", "dcs", "sscs_sc", "dcs_sc"):
        for fn in sorted(os.listdir(os.path.join(d, sub))):
            if fn.endswith(".bam"):
                files[fn[len(sid) + 1:-4]] = _load(os.path.join(d, sub, fn))["records"]
    with open(stats_final) as fh:
        stats_txt = fh.read()
    with open("%s/sscs/%s.read_families.txt" % (d, sid)) as fh:
        fam_txt = fh.read()
    if not keep and workdir is None:
        shutil.rmtree(tmp, ignore_errors=True)
    return {"files": files, "stats": stats_txt, "read_families": fam_txt, "status": status}


def load_reference_functions():
    """The reference's pure functions, imported from its own modules (for KATs)."""
    out = {}
    with _reference_env():
        import importlib
        import warnings
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            ch = importlib.import_module("consensus_helper")
            for nm in ("which_read", "which_strand", "cigar_order", "sscs_qname", "unique_tag",
                       "duplex_tag", "read_mode", "consensus_flag"):
                out[nm] = getattr(ch, nm)
            import pysam
            out["AlignedSegment"] = pysam.AlignedSegment
            for mod, names in (("SSCS_maker", ("consensus_maker",)),
                               ("DCS_maker", ("duplex_consensus", "dcs_consensus_tag")),
                               ("singleton_correction", ("duplex_consensus",))):
                ns = runpy.run_path(os.path.join(REF_PKG, mod + ".py"), run_name="not_main")
                for n in names:
                    out[mod + "." + n] = ns[n]
            for k in ("SSCS_maker", "DCS_maker", "singleton_correction"):
                sys.modules.pop(k, None)
    return out
