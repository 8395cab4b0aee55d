"""CPU: barcode extraction (consensuscruncher_b200/extract.py) against outputs of the unmodified reference
`extract_barcodes.py` on the reference's own test FASTQ files (oracle/make_golden_extract.py)."""
import gzip
import json
import os

import pytest

from consensuscruncher_b200 import extract

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = ["extract_pattern_nnt", "extract_pattern_long", "extract_pattern_inner_spacer", "extract_list"]


def _run(tmp_path, g, gz):
    d = str(tmp_path)
    spec = g["spec"]
    for r in ("R1", "R2"):
        name = os.path.join(d, "%s.fastq%s" % (r, ".gz" if gz else ""))
        with (gzip.open(name, "wt") if gz else open(name, "w")) as fh:
            fh.write(g["inputs"][r])
    os.makedirs(os.path.join(d, "out"))
    argv = ["--read1", os.path.join(d, "R1.fastq" + (".gz" if gz else "")),
            "--read2", os.path.join(d, "R2.fastq" + (".gz" if gz else "")), "--outfile", os.path.join(d, "out", "sample")]
    if "bpattern" in spec:
        argv += ["--bpattern", spec["bpattern"]]
    if "blist" in spec:
        with open(os.path.join(d, "blist.txt"), "w") as fh:
            fh.write("\n".join(spec["blist"]) + "\n")
        argv += ["--blist", os.path.join(d, "blist.txt")]
    assert extract.main(argv) == 0
    return d


@pytest.mark.parametrize("name", CASES)
@pytest.mark.parametrize("gz", [True, False])
def test_outputs_equal_the_reference(tmp_path, capsys, name, gz):
    g = json.load(gzip.open(os.path.join(GOLDEN, name + ".json.gz"), "rt"))
    d = _run(tmp_path, g, gz)
    for rel, text in g["files"].items():
        assert open(os.path.join(d, rel)).read() == text, rel
    made = {os.path.relpath(os.path.join(r, f), d) for r, _, fs in os.walk(d) for f in fs}
    assert made - {"R1.fastq", "R2.fastq", "R1.fastq.gz", "R2.fastq.gz", "blist.txt"} == set(g["files"])
    assert capsys.readouterr().err == g["stderr"]


def test_argument_checks(tmp_path):
    with pytest.raises(ValueError):
        extract.main(["--read1", "a", "--read2", "b", "--outfile", str(tmp_path / "o")])
    with pytest.raises(ValueError):
        extract.extract_pattern("a", "b", str(tmp_path / "o"), "NNX")
    with pytest.raises(ValueError):
        extract.extract_list("a", "b", str(tmp_path / "o"), ["ACT", "GGA"])
