"""TEST INFRASTRUCTURE - golden-vector files (tests/golden/*.json.gz) and the
conversions between the product's `Read` and the record tuples the reference
harness (oracle/run_reference.py, oracle/refshim/pysam.py) works on."""
import gzip
import json
import os

import numpy as np

from consensuscruncher_b200.records import Read

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def read_to_tuple(r):
    tags = dict(r.aux) if isinstance(r.aux, dict) else {}
    if r.rg is not None:
        tags["RG"] = r.rg
    return (r.qname, int(r.flag), int(r.tid), int(r.pos), int(r.mapq), r.cigar, int(r.mtid),
            int(r.mpos), int(r.tlen), r.seq, bytes(np.asarray(r.qual, np.uint8)), tags)


def tuple_to_read(t):
    tags = dict(t[11])
    rg = tags.pop("RG", None)
    qual = np.frombuffer(t[10], np.uint8).copy() if t[10] is not None else np.zeros(0, np.uint8)
    return Read(t[0], t[1], t[2], t[3], t[4], t[5] or "", t[6], t[7], t[8], t[9] or "", qual,
                rg=rg, aux=tags if tags else b"")


def _enc(t):
    return [t[0], t[1], t[2], t[3], t[4], t[5], t[6], t[7], t[8], t[9],
            None if t[10] is None else bytes(t[10]).hex(), t[11]]


def _dec(j):
    return (j[0], j[1], j[2], j[3], j[4], j[5], j[6], j[7], j[8], j[9],
            None if j[10] is None else bytes.fromhex(j[10]), j[11])


def save(name, params, header, records, out):
    blob = {"params": params, "header": header, "input": [_enc(t) for t in records],
            "files": {k: [_enc(t) for t in v] for k, v in out["files"].items()},
            "stats": out["stats"], "read_families": out["read_families"], "status": out["status"]}
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    path = os.path.join(GOLDEN_DIR, name + ".json.gz")
    with gzip.GzipFile(path, "wb", mtime=0) as fh:
        fh.write(json.dumps(blob, sort_keys=True).encode())
    return path


def load(name):
    with gzip.open(os.path.join(GOLDEN_DIR, name + ".json.gz"), "rb") as fh:
        blob = json.loads(fh.read().decode())
    blob["input"] = [_dec(j) for j in blob["input"]]
    blob["files"] = {k: [_dec(j) for j in v] for k, v in blob["files"].items()}
    return blob


def names():
    """Cases of the consensus path (extract_* belong to the barcode extraction step, tests/test_extract.py)."""
    return sorted(f[:-8] for f in os.listdir(GOLDEN_DIR) if f.endswith(".json.gz") and not f.startswith("extract_"))
