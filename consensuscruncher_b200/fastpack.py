"""ctypes face of libccpack.so: multi-threaded BGZF inflate and the packing rules of `pack.py`
in C++ (csrc/ccpack.cpp), for inputs where decoding every record into a Python object would
dominate the run.  `pack_reads` here returns the same `PackedReads` as `pack.pack_reads`
(tests/test_fastpack.py compares them byte for byte); records are decoded lazily, only when the
emit step asks for one (family templates, singletons, bad reads).
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import bam
from .pack import FIELDS_DT, DomainError, PackedReads

HERE = os.path.dirname(os.path.abspath(__file__))
_lib = None
COUNTER_NAMES = ("total", "unmapped", "unmapped_mate", "multiple_mapping", "bad_spacer")


class EmitIn(C.Structure):
    """struct ccp_emit_in of csrc/ccemit.cpp."""
    _fields_ = ([("n_reads", C.c_uint64), ("n_fam", C.c_uint64), ("lpad", C.c_uint32), ("stride", C.c_uint32),
                 ("src", C.c_void_p), ("fields", C.c_void_p),
                 ("barcodes", C.c_char_p), ("cigar_pairs", C.c_char_p), ("rgs", C.c_char_p),
                 ("n_bc", C.c_uint32), ("n_cig", C.c_uint32), ("n_rg", C.c_uint32),
                 ("bad", C.c_void_p), ("n_bad", C.c_uint64)] +
                [(k, C.c_void_p) for k in ("read_fam", "fam_size", "fam_tmpl", "fam_mapq", "fam_flag", "fam_tlen", "fam_rg",
                                           "partner", "sscs_row", "corr_row", "dcs_row", "dcssc_row",
                                           "sscs_rows", "corr_rows", "dcs_rows", "dcssc_rows",
                                           "dcs_mapq", "dcs_flag", "dcs_tlen", "dcs_rg",
                                           "dcssc_mapq", "dcssc_flag", "dcssc_tlen", "dcssc_rg")])


# output lists in the order of csrc/ccemit.cpp's F_* enum
FILES = ("sscs.sorted", "singleton.sorted", "dcs.sorted", "sscs.singleton.sorted", "sscs.correction.sorted",
         "singleton.correction.sorted", "uncorrected.sorted", "sscs.sc.sorted", "dcs.sc.sorted",
         "sscs.sc.singleton.sorted", "all.unique.dcs.sorted", "badReads")
_RESULT_DTYPES = dict(read_fam=np.uint32, fam_size=np.uint32, fam_tmpl=np.uint32, fam_mapq=np.uint8, fam_flag=np.uint16,
                      fam_tlen=np.int32, fam_rg=np.uint8, partner=np.int32, sscs_row=np.int32, corr_row=np.int32,
                      dcs_row=np.int32, dcssc_row=np.int32, sscs_rows=np.uint8, corr_rows=np.uint8, dcs_rows=np.uint8,
                      dcssc_rows=np.uint8, dcs_mapq=np.uint8, dcs_flag=np.uint16, dcs_tlen=np.int32, dcs_rg=np.uint8,
                      dcssc_mapq=np.uint8, dcssc_flag=np.uint16, dcssc_tlen=np.int32, dcssc_rg=np.uint8)


def load_library():
    """dlopen libccpack.so (built by `build.build_pack_lib`); raises if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    path = os.environ.get("CCPACK_LIB") or os.path.join(HERE, "libccpack.so")
    if not os.path.exists(path):
        from . import build
        path = build.build_pack_lib()
    lib = C.CDLL(path)
    P, u64p, u32p = C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_uint32)
    protos = {
        "ccp_bam_open": (P, [C.c_char_p, C.c_int, C.c_char_p, C.c_int]),
        "ccp_bam_close": (None, [P]),
        "ccp_bam_n_records": (C.c_uint64, [P]),
        "ccp_bam_data": (P, [P, u64p]),
        "ccp_bam_offsets": (P, [P]),
        "ccp_bam_header_text": (C.c_char_p, [P]),
        "ccp_bam_n_refs": (C.c_int, [P]),
        "ccp_bam_ref_name": (C.c_char_p, [P, C.c_int]),
        "ccp_bam_ref_len": (C.c_int32, [P, C.c_int]),
        "ccp_pack_run": (P, [P, C.c_int, C.c_char_p, P, C.c_int, C.c_int, C.c_char_p, C.c_int]),
        "ccp_pack_free": (None, [P]),
        "ccp_pack_n": (C.c_uint64, [P]),
        "ccp_pack_geometry": (None, [P, u32p, u32p]),
        "ccp_pack_fields": (P, [P]),
        "ccp_pack_payload": (P, [P]),
        "ccp_pack_src": (P, [P]),
        "ccp_pack_lens": (P, [P]),
        "ccp_pack_counters": (P, [P]),
        "ccp_pack_bad": (P, [P, u64p]),
        "ccp_pack_leftover": (P, [P, u64p]),
        "ccp_pack_dup_bc": (P, [P, u32p]),
        "ccp_pack_dict": (P, [P, C.c_int, u32p, u64p]),
        "ccp_emit": (C.c_int, [P, C.POINTER(EmitIn), C.c_int, C.POINTER(C.c_char_p), C.c_int,
                               C.POINTER(C.c_int64), C.c_char_p, C.c_int]),
    }
    for name, (res, args) in protos.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    _lib = lib
    return lib


SYMBOLS = ("ccp_bam_open", "ccp_bam_close", "ccp_bam_n_records", "ccp_bam_data", "ccp_bam_offsets",
           "ccp_bam_header_text", "ccp_bam_n_refs", "ccp_bam_ref_name", "ccp_bam_ref_len", "ccp_pack_run",
           "ccp_pack_free", "ccp_pack_n", "ccp_pack_geometry", "ccp_pack_fields", "ccp_pack_payload",
           "ccp_pack_src", "ccp_pack_lens", "ccp_pack_counters", "ccp_pack_bad", "ccp_pack_leftover",
           "ccp_pack_dup_bc", "ccp_pack_dict", "ccp_emit")


def _view(ptr, nbytes, dtype):
    """numpy view (no copy) of `nbytes` at `ptr`."""
    if not nbytes:
        return np.empty(0, dtype)
    return np.frombuffer((C.c_uint8 * nbytes).from_address(ptr), dtype=dtype)


class BamFile:
    """An inflated BAM held by the native library; indexing decodes one record (`records.Read`)."""

    def __init__(self, path, threads=None):
        lib = load_library()
        err = C.create_string_buffer(512)
        threads = threads or min(32, os.cpu_count() or 1)
        self._h = lib.ccp_bam_open(os.fsencode(path), int(threads), err, len(err))
        if not self._h:
            raise ValueError("%s: %s" % (path, err.value.decode()))
        self.threads = int(threads)
        n = C.c_uint64()
        ptr = lib.ccp_bam_data(self._h, C.byref(n))
        self._view = _view(ptr, n.value, np.uint8)
        self._data = memoryview(self._view)
        self._n = int(lib.ccp_bam_n_records(self._h))
        self._offsets = _view(lib.ccp_bam_offsets(self._h), 8 * self._n, np.uint64)
        refs = [(lib.ccp_bam_ref_name(self._h, i).decode(), int(lib.ccp_bam_ref_len(self._h, i)))
                for i in range(lib.ccp_bam_n_refs(self._h))]
        self.header = bam.BamHeader(lib.ccp_bam_header_text(self._h).decode(),
                                    [r[0] for r in refs], [r[1] for r in refs])
        self._cache = {}

    def __len__(self):
        return self._n

    def __getitem__(self, i):
        i = int(i)
        if i < 0:
            i += self._n
        r = self._cache.get(i)
        if r is None:
            r, _ = bam.decode_record(self._data, self._view, int(self._offsets[i]))
            self._cache[i] = r
        return r

    def __iter__(self):
        return (self[i] for i in range(self._n))

    def close(self):
        if self._h:
            self._view = self._data = self._offsets = None
            load_library().ccp_bam_close(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def open_bam(path, threads=None):
    """-> (BamHeader, BamFile); the BamFile stands in for the record list of `bam.read_bam`."""
    f = BamFile(path, threads)
    return f.header, f


def pack_reads(bamfile, stage="sscs", bdelim="|", regions=None):
    """`pack.pack_reads` over a `BamFile`, done in C++ (same arguments, same result)."""
    lib = load_library()
    err = C.create_string_buffer(512)
    if regions is None:
        reg, n_reg = None, -1
    else:
        arr = np.ascontiguousarray(np.asarray(list(regions), np.int32).reshape(-1, 3))
        reg, n_reg = arr.ctypes.data, len(arr)
    h = lib.ccp_pack_run(bamfile._h, 0 if stage == "sscs" else 1, bdelim.encode(), reg, n_reg,
                         bamfile.threads, err, len(err))
    if not h:
        msg = err.value.decode()
        if msg.startswith("domain: "):
            raise DomainError(msg[len("domain: "):])
        raise ValueError(msg)
    try:
        n = int(lib.ccp_pack_n(h))
        lpad, stride = C.c_uint32(), C.c_uint32()
        lib.ccp_pack_geometry(h, C.byref(lpad), C.byref(stride))
        lpad, stride = lpad.value, stride.value
        fields = _view(lib.ccp_pack_fields(h), 32 * n, FIELDS_DT).copy()
        payload = _view(lib.ccp_pack_payload(h), n * stride, np.uint8).reshape(n, stride).copy()
        src = _view(lib.ccp_pack_src(h), 8 * n, np.int64).copy()
        lens = _view(lib.ccp_pack_lens(h), 4 * n, np.int32).copy()
        ct = _view(lib.ccp_pack_counters(h), 40, np.int64)
        counters = {k: int(v) for k, v in zip(COUNTER_NAMES, ct)}
        m = C.c_uint64()
        p = lib.ccp_pack_bad(h, C.byref(m))
        bad = _view(p, 8 * m.value, np.int64).tolist()
        p = lib.ccp_pack_leftover(h, C.byref(m))
        leftover = _view(p, 8 * m.value, np.int64).tolist()
        nb = C.c_uint32()
        p = lib.ccp_pack_dup_bc(h, C.byref(nb))
        dup_bc = _view(p, 4 * nb.value, np.uint32).copy()
        dicts = []
        for which in range(3):
            cnt, nbytes = C.c_uint32(), C.c_uint64()
            p = lib.ccp_pack_dict(h, which, C.byref(cnt), C.byref(nbytes))
            raw = C.string_at(p, nbytes.value) if nbytes.value else b""
            names = raw.decode().split("\0")[:-1] if raw else []
            assert len(names) == cnt.value
            dicts.append(names)
    finally:
        lib.ccp_pack_free(h)
    return PackedReads(fields, payload, lpad, dicts[0], dup_bc, dicts[1], dicts[2], src=src, lens=lens,
                       counters=counters, bad=bad, leftover=leftover)


def emit_files(bamfile, packed, res, scorrect, paths, threads=None):
    """`emit.emit_all` + `bam.write_bam` of every list, in C++ (csrc/ccemit.cpp): writes the BAM
    (+ BAI for the sorted ones) of each entry of `paths` ({list name: path}, names as in `FILES`)
    and returns ({list name: record count}, stats text, read_families text)."""
    from . import emit
    lib = load_library()
    keep = []

    def arr(a, dt):
        a = np.ascontiguousarray(a, dt)
        keep.append(a)
        return a.ctypes.data if a.size else None

    def joined(names):
        return b"".join(n.encode() + b"\0" for n in names) + b"\0"

    e = EmitIn()
    e.n_reads, e.n_fam, e.lpad, e.stride = packed.n, res.n_fam, res.lpad, res.stride
    e.src = arr(packed.src, np.int64)
    e.fields = arr(packed.fields, FIELDS_DT)
    e.barcodes, e.cigar_pairs, e.rgs = joined(packed.barcodes), joined(packed.cigar_pairs), joined(packed.rgs)
    e.n_bc, e.n_cig, e.n_rg = len(packed.barcodes), len(packed.cigar_pairs), len(packed.rgs)
    e.bad = arr(np.asarray(packed.bad, np.int64), np.int64)
    e.n_bad = len(packed.bad)
    for k, dt in _RESULT_DTYPES.items():
        setattr(e, k, arr(getattr(res, k), dt))
    c_paths = (C.c_char_p * len(FILES))(*[os.fsencode(paths[k]) if paths.get(k) else None for k in FILES])
    counts = (C.c_int64 * len(FILES))()
    err = C.create_string_buffer(512)
    rc = lib.ccp_emit(bamfile._h, C.byref(e), 1 if scorrect else 0, c_paths, int(threads or bamfile.threads), counts,
                      err, len(err))
    if rc != 0:
        msg = err.value.decode()
        if msg.startswith("domain: "):
            raise DomainError(msg[len("domain: "):])
        raise ValueError(msg)
    n = {k: int(counts[i]) for i, k in enumerate(FILES)}
    return n, emit.stats_text(packed.counters, n, scorrect), emit.families_text(res)
