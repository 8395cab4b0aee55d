"""TEST INFRASTRUCTURE - ctypes wrapper around oracle/cc_oracle.c (the CPU
restatement of the reference's consensus path; see that file's header for the
reference lines it follows and for the parity status).  Used by tests/,
__graft_entry__.smoke() and bench.py's CPU-baseline legs only."""
import ctypes
import os
import subprocess

import numpy as np

from consensuscruncher_b200.result import ConsensusResult

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

_DTYPES = {
    "read_fam": np.uint32, "fam_size": np.uint32, "fam_tmpl": np.uint32, "fam_mapq": np.uint8,
    "fam_flag": np.uint16, "fam_tlen": np.int32, "fam_rg": np.uint8, "partner": np.int32,
    "sscs_row": np.int32, "corr_row": np.int32, "dcs_row": np.int32, "dcssc_row": np.int32,
    "dcs_mapq": np.uint8, "dcs_flag": np.uint16, "dcs_tlen": np.int32, "dcs_rg": np.uint8,
    "dcssc_mapq": np.uint8, "dcssc_flag": np.uint16, "dcssc_tlen": np.int32, "dcssc_rg": np.uint8,
}
_IDS = {"read_fam": 0, "fam_size": 1, "fam_tmpl": 2, "fam_mapq": 3, "fam_flag": 4, "fam_tlen": 5,
        "fam_rg": 6, "partner": 7, "sscs_row": 8, "corr_row": 9, "dcs_row": 10, "dcssc_row": 11,
        "sscs_rows": 12, "corr_rows": 13, "dcs_rows": 14, "dcssc_rows": 15,
        "dcs_mapq": 16, "dcs_flag": 17, "dcs_tlen": 18, "dcs_rg": 19,
        "dcssc_mapq": 20, "dcssc_flag": 21, "dcssc_tlen": 22, "dcssc_rg": 23, "stats": 24}


def build(force=False):
    so = os.path.join(_HERE, "libcc_oracle.so")
    src = os.path.join(_HERE, "cc_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        L = ctypes.CDLL(build())
        L.cco_create.restype = ctypes.c_void_p
        L.cco_create.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint32,
                                 ctypes.c_uint32, ctypes.c_void_p, ctypes.c_uint32, ctypes.c_double]
        L.cco_run.argtypes = [ctypes.c_void_p, ctypes.c_int]
        L.cco_count.restype = ctypes.c_uint64
        L.cco_count.argtypes = [ctypes.c_void_p, ctypes.c_int]
        L.cco_domain_faults.restype = ctypes.c_uint64
        L.cco_domain_faults.argtypes = [ctypes.c_void_p]
        L.cco_array.restype = ctypes.c_void_p
        L.cco_array.argtypes = [ctypes.c_void_p, ctypes.c_int]
        L.cco_destroy.argtypes = [ctypes.c_void_p]
        L.cco_consensus_maker.argtypes = [ctypes.c_void_p, ctypes.c_uint32, ctypes.c_uint32,
                                          ctypes.c_uint32, ctypes.c_double, ctypes.c_void_p]
        L.cco_duplex_consensus.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint32,
                                           ctypes.c_uint32, ctypes.c_int, ctypes.c_void_p]
        _LIB = L
    return _LIB


class OracleError(RuntimeError):
    pass


_ERR = {-1: "out of memory", -2: "N base with Q>=30 in a family (reference: IndexError, SSCS_maker.py:122)",
        -3: "base outside A/C/G/T/N (reference: ValueError, SSCS_maker.py:120)", -4: "bad geometry"}


def _copy(ptr, count, dtype):
    if count == 0 or not ptr:
        return np.zeros(0, dtype)
    buf = (ctypes.c_uint8 * (count * np.dtype(dtype).itemsize)).from_address(ptr)
    return np.frombuffer(buf, dtype=dtype, count=count).copy()


def run(packed, cutoff=0.7, scorrect=True, return_faults=False):
    """Full consensus path (group, SSCS, DCS, rescue, DCS after rescue) on the CPU."""
    L = lib()
    fields = np.ascontiguousarray(packed.fields)
    payload = np.ascontiguousarray(packed.payload)
    dup = np.ascontiguousarray(packed.dup_bc, dtype=np.uint32)
    ctx = L.cco_create(fields.ctypes.data, payload.ctypes.data, packed.n, packed.lpad,
                       max(packed.stride, 1), dup.ctypes.data, len(dup), float(cutoff))
    if not ctx:
        raise MemoryError
    try:
        rc = L.cco_run(ctx, 1 if scorrect else 0)
        if rc:
            raise OracleError(_ERR.get(rc, str(rc)))
        nf = L.cco_count(ctx, 0)
        counts = {"sscs_rows": L.cco_count(ctx, 1), "corr_rows": L.cco_count(ctx, 2),
                  "dcs_rows": L.cco_count(ctx, 3), "dcssc_rows": L.cco_count(ctx, 4)}
        arrays = {}
        for name, dt in _DTYPES.items():
            if name == "read_fam":
                cnt = packed.n
            elif name.startswith("dcssc_") and name != "dcssc_row":
                cnt = counts["dcssc_rows"]
            elif name.startswith("dcs_") and name != "dcs_row":
                cnt = counts["dcs_rows"]
            else:
                cnt = nf
            arrays[name] = _copy(L.cco_array(ctx, _IDS[name]), cnt, dt)
        for name, cnt in counts.items():
            arrays[name] = _copy(L.cco_array(ctx, _IDS[name]), cnt * packed.stride,
                                 np.uint8).reshape(cnt, packed.stride)
        arrays["stats"] = _copy(L.cco_array(ctx, 24), 16, np.int64)
        faults = L.cco_domain_faults(ctx)
    finally:
        L.cco_destroy(ctx)
    res = ConsensusResult(packed.lpad, packed.stride, scorrect, **arrays)
    return (res, int(faults)) if return_faults else res


def consensus_maker_rows(rows, lpad, cutoff):
    """One family: rows u8[n][stride] -> consensus row (SSCS_maker.py:75-157)."""
    rows = np.ascontiguousarray(rows, np.uint8)
    out = np.zeros(rows.shape[1], np.uint8)
    rc = lib().cco_consensus_maker(rows.ctypes.data, rows.shape[0], lpad, rows.shape[1],
                                   float(cutoff), out.ctypes.data)
    if rc:
        raise OracleError(_ERR.get(rc, str(rc)))
    return out


def duplex_consensus_rows(r1, r2, lpad, sc):
    """One pair: DCS_maker.py:99-121 (sc=False) or singleton_correction.py:61-84 (sc=True)."""
    r1 = np.ascontiguousarray(r1, np.uint8)
    r2 = np.ascontiguousarray(r2, np.uint8)
    out = np.zeros(r1.shape[0], np.uint8)
    lib().cco_duplex_consensus(r1.ctypes.data, r2.ctypes.data, lpad, r1.shape[0], 1 if sc else 0,
                               out.ctypes.data)
    return out
