"""ctypes binding of libccgpu.so (include/ccgpu.h) and the `GpuConsensus` engine.

There is no CPU fallback: importing this module is harmless, but creating an
engine without the built library or without an sm_100 GPU raises.
"""
from __future__ import annotations

import ctypes
import os

import numpy as np

from .pack import DomainError, PackedReads
from .result import ConsensusResult

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libccgpu.so")

# enum cc_array / element types (include/ccgpu.h)
ARRAYS = {
    "read_fam": (0, np.uint32), "fam_size": (1, np.uint32), "fam_tmpl": (2, np.uint32),
    "fam_mapq": (3, np.uint8), "fam_flag": (4, np.uint16), "fam_tlen": (5, np.int32),
    "fam_rg": (6, np.uint8), "partner": (7, np.int32), "sscs_row": (8, np.int32),
    "corr_row": (9, np.int32), "dcs_row": (10, np.int32), "dcssc_row": (11, np.int32),
    "sscs_rows": (12, np.uint8), "corr_rows": (13, np.uint8), "dcs_rows": (14, np.uint8),
    "dcssc_rows": (15, np.uint8), "dcs_mapq": (16, np.uint8), "dcs_flag": (17, np.uint16),
    "dcs_tlen": (18, np.int32), "dcs_rg": (19, np.uint8), "dcssc_mapq": (20, np.uint8),
    "dcssc_flag": (21, np.uint16), "dcssc_tlen": (22, np.int32), "dcssc_rg": (23, np.uint8),
    "stats": (24, np.int64), "fam_hist": (25, np.uint64),
    "sscs_planes": (26, np.uint8), "corr_planes": (27, np.uint8), "dcs_planes": (28, np.uint8), "dcssc_planes": (29, np.uint8),
}
PHASES = ("h2d", "hash", "sort", "segment", "join", "classify", "vote", "rescue", "dcs", "dcs_sc", "vote_aux")
CC_ERR_BASE, CC_ERR_DOMAIN = -4, -5

# every symbol include/ccgpu.h declares
SYMBOLS = ("cc_create", "cc_destroy", "cc_last_error", "cc_set_params", "cc_set_stream", "cc_load_reads",
           "cc_load_reads_device", "cc_group", "cc_sscs_vote", "cc_dcs_join", "cc_sc_rescue", "cc_run",
           "cc_run_host", "cc_counts", "cc_fetch", "cc_array_bytes", "cc_device_ptr", "cc_consensus_maker",
           "cc_duplex_consensus", "cc_profile", "cc_phase_ms", "cc_kernel_launches", "cc_device_bytes",
           "cc_consensus_maker_static", "cc_sort_coordinates", "cc_load_reads_q4", "cc_run_host_q4",
           "cc_load_reads_compact", "cc_run_host_compact", "cc_load_reads_classes", "cc_run_host_classes", "cc_pack_rows", "cc_join_records", "cc_fetch_many", "cc_comm_unique_id", "cc_comm_init", "cc_stats_allreduce", "cc_comm_destroy", "cc_comm_info")
COMM_ID_BYTES = 128


class CcParams(ctypes.Structure):
    _fields_ = [("cutoff", ctypes.c_double), ("scorrect", ctypes.c_int32)]


class GpuError(RuntimeError):
    pass


class BaseError(GpuError):
    """N / IUPAC base with Q>=30 in a family (reference: IndexError / ValueError)."""


_lib = None


def load_library(path=None):
    """dlopen libccgpu.so and declare its prototypes.  Raises if it was not built.
    CCGPU_LIB names an alternative build of the same library (kernel experiments)."""
    global _lib
    if _lib is not None:
        return _lib
    path = path or os.environ.get("CCGPU_LIB") or LIB_PATH
    if not os.path.exists(path):
        raise GpuError("libccgpu.so is missing (%s): build it with `python -m consensuscruncher_b200.build`; "
                       "this package has no CPU fallback" % path)
    L = ctypes.CDLL(path)
    vp, u64, u32, i32 = ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint32, ctypes.c_int
    L.cc_create.restype = vp
    L.cc_create.argtypes = [i32, ctypes.POINTER(CcParams)]
    L.cc_destroy.restype = None
    L.cc_destroy.argtypes = [vp]
    L.cc_last_error.restype = ctypes.c_char_p
    L.cc_last_error.argtypes = [vp]
    L.cc_set_params.argtypes = [vp, ctypes.POINTER(CcParams)]
    L.cc_set_stream.argtypes = [vp, vp]
    load_args = [vp, vp, vp, u64, u32, u32, vp, u32]
    L.cc_load_reads.argtypes = load_args
    L.cc_load_reads_device.argtypes = load_args
    L.cc_run_host.argtypes = load_args
    L.cc_load_reads_q4.argtypes = [vp, vp, vp, u64, u32, u32, vp, vp, u32]
    L.cc_run_host_q4.argtypes = [vp, vp, vp, u64, u32, u32, vp, vp, u32]
    for nm in ("cc_group", "cc_sscs_vote", "cc_dcs_join", "cc_sc_rescue", "cc_run"):
        getattr(L, nm).argtypes = [vp]
    L.cc_counts.argtypes = [vp, ctypes.POINTER(u64)]
    L.cc_fetch.argtypes = [vp, i32, vp, u64]
    L.cc_array_bytes.restype = u64
    L.cc_array_bytes.argtypes = [vp, i32]
    L.cc_device_ptr.restype = vp
    L.cc_device_ptr.argtypes = [vp, i32]
    L.cc_consensus_maker.argtypes = [vp, vp, u32, u32, u32, ctypes.c_double, vp]
    L.cc_duplex_consensus.argtypes = [vp, vp, vp, u64, u32, u32, i32, vp]
    L.cc_sort_coordinates.argtypes = [vp, vp, vp, vp, u64, vp]
    L.cc_consensus_maker_static.argtypes = [vp, u32, u32, u32, ctypes.c_double, vp, vp]
    L.cc_profile.argtypes = [vp, i32]
    L.cc_phase_ms.argtypes = [vp, i32, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(u64)]
    L.cc_kernel_launches.restype = u64
    L.cc_kernel_launches.argtypes = [vp]
    L.cc_device_bytes.restype = u64
    L.cc_device_bytes.argtypes = [vp]
    L.cc_load_reads_compact.argtypes = [vp, vp, vp, u64, u32, u32, u32, vp, vp, vp, u32]
    L.cc_run_host_compact.argtypes = [vp, vp, vp, u64, u32, u32, u32, vp, vp, vp, u32]
    L.cc_load_reads_classes.argtypes = [vp, vp, vp, u32, vp, u64, u32, u32, u32, vp, vp, vp, u32]
    L.cc_run_host_classes.argtypes = [vp, vp, vp, u32, vp, u64, u32, u32, u32, vp, vp, vp, u32]
    L.cc_fetch_many.argtypes = [vp, u32, vp, vp, vp]
    L.cc_pack_rows.argtypes = [vp, vp, vp, vp, vp]
    L.cc_join_records.argtypes = [vp, vp, u64, vp, u32, vp]
    L.cc_comm_unique_id.argtypes = [vp]
    L.cc_comm_init.argtypes = [vp, vp, i32, i32]
    L.cc_stats_allreduce.argtypes = [vp]
    L.cc_comm_destroy.argtypes = [vp]
    L.cc_comm_info.argtypes = [vp, ctypes.POINTER(i32), ctypes.POINTER(i32), ctypes.POINTER(i32)]
    _lib = L
    return L


class GpuConsensus:
    """One context on one GPU: group -> SSCS vote -> DCS join -> singleton rescue."""

    def __init__(self, device=0, cutoff=0.7, scorrect=True):
        self.lib = load_library()
        self.params = CcParams(float(cutoff), 1 if scorrect else 0)
        self.ctx = self.lib.cc_create(int(device), ctypes.byref(self.params))
        if not self.ctx:
            raise GpuError("cc_create(device=%d) failed: %s" %
                           (device, self.lib.cc_last_error(None).decode()))
        self.device = int(device)
        self.lpad = self.stride = 0
        self._keep = None
        self.packing = None

    # -- plumbing ----------------------------------------------------------
    def close(self):
        if getattr(self, "ctx", None):
            self.lib.cc_destroy(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _check(self, rc):
        if rc == 0:
            return
        msg = self.lib.cc_last_error(self.ctx).decode()
        if rc == CC_ERR_BASE:
            raise BaseError(msg)
        if rc == CC_ERR_DOMAIN:
            raise DomainError(msg)
        raise GpuError("libccgpu error %d: %s" % (rc, msg))

    def set_params(self, cutoff=None, scorrect=None):
        if cutoff is not None:
            self.params.cutoff = float(cutoff)
        if scorrect is not None:
            self.params.scorrect = 1 if scorrect else 0
        self._check(self.lib.cc_set_params(self.ctx, ctypes.byref(self.params)))

    def set_stream(self, cuda_stream_handle):
        self._check(self.lib.cc_set_stream(self.ctx, ctypes.c_void_p(cuda_stream_handle or 0)))

    # -- input ---------------------------------------------------------------
    def _host_args(self, packed: PackedReads):
        fields = np.ascontiguousarray(packed.fields)
        payload = np.ascontiguousarray(packed.payload)
        dup = np.ascontiguousarray(packed.dup_bc, dtype=np.uint32)
        self._keep = (fields, payload, dup)
        self.lpad, self.stride = packed.lpad, (packed.stride or 64)
        return (self.ctx, fields.ctypes.data, payload.ctypes.data, packed.n, packed.lpad,
                self.stride, dup.ctypes.data, len(dup))

    def load(self, packed: PackedReads):
        """Host SoA -> device (cc_load_reads)."""
        self._check(self.lib.cc_load_reads(*self._host_args(packed)))

    def load_q4(self, packed: PackedReads, q4=None):
        """Host SoA with 4-bit quality codes -> device (cc_load_reads_q4); raises if the batch has
        too many distinct qualities for the compact rows."""
        from .pack import row_geometry, to_q4
        q4 = q4 or to_q4(packed)
        if q4 is None:
            raise GpuError("more than 15 distinct non-zero quality values: use load()")
        payload4, table = q4
        fields = np.ascontiguousarray(packed.fields)
        dup = np.ascontiguousarray(packed.dup_bc, dtype=np.uint32)
        self._keep = (fields, payload4, dup, table)
        self.lpad, self.stride = packed.lpad, row_geometry(packed.lpad)[1]
        self._check(self.lib.cc_load_reads_q4(self.ctx, fields.ctypes.data, payload4.ctypes.data, packed.n, packed.lpad,
                                              payload4.shape[1], table.ctypes.data, dup.ctypes.data, len(dup)))

    def load_compact(self, packed: PackedReads, compact=None):
        """Pair records + bit-plane rows -> device (cc_load_reads_compact); raises if the batch does not fit
        the format (pairs that do not mirror, more than 16 quality / literal-base combinations)."""
        from .pack import row_geometry, to_compact
        compact = compact or to_compact(packed)
        if compact is None:
            raise GpuError("the batch does not fit the compact format: use load()")
        pairs, planes, qbits, qual_table, base_table = compact
        dup = np.ascontiguousarray(packed.dup_bc, dtype=np.uint32)
        self._keep = (pairs, planes, dup, qual_table, base_table)
        self.lpad, self.stride = packed.lpad, row_geometry(packed.lpad)[1]
        self._check(self.lib.cc_load_reads_compact(self.ctx, pairs.ctypes.data, planes.ctypes.data, packed.n, packed.lpad,
                                                   planes.shape[1], int(qbits), qual_table.ctypes.data,
                                                   base_table.ctypes.data, dup.ctypes.data, len(dup)))

    def load_classes(self, packed: PackedReads, compact=None):
        """16-byte pair records + class table + bit-plane rows -> device (cc_load_reads_classes); raises if the
        batch does not fit (see pack.to_classes)."""
        from .pack import row_geometry, to_classes, to_compact
        compact = compact or to_compact(packed)
        short = to_classes(compact[0]) if compact is not None else None
        if short is None:
            raise GpuError("the batch does not fit the short pair records: use load_compact() or load()")
        pairs, classes = short
        _, planes, qbits, qual_table, base_table = compact
        dup = np.ascontiguousarray(packed.dup_bc, dtype=np.uint32)
        self._keep = (pairs, classes, planes, dup, qual_table, base_table)
        self.lpad, self.stride = packed.lpad, row_geometry(packed.lpad)[1]
        self._check(self.lib.cc_load_reads_classes(self.ctx, pairs.ctypes.data, classes.ctypes.data, len(classes),
                                                   planes.ctypes.data, packed.n, packed.lpad, planes.shape[1], int(qbits),
                                                   qual_table.ctypes.data, base_table.ctypes.data, dup.ctypes.data, len(dup)))

    def load_device(self, fields_ptr, payload_ptr, n, lpad, stride, dup_ptr, n_bc):
        """Adopt buffers that already live in device memory (cc_load_reads_device)."""
        self.lpad, self.stride = int(lpad), int(stride)
        self._check(self.lib.cc_load_reads_device(self.ctx, fields_ptr, payload_ptr, int(n), int(lpad),
                                                  int(stride), dup_ptr, int(n_bc)))

    # -- stages --------------------------------------------------------------------
    def group(self):
        self._check(self.lib.cc_group(self.ctx))

    def sscs_vote(self):
        self._check(self.lib.cc_sscs_vote(self.ctx))

    def dcs_join(self):
        self._check(self.lib.cc_dcs_join(self.ctx))

    def sc_rescue(self):
        self._check(self.lib.cc_sc_rescue(self.ctx))

    def run(self):
        self._check(self.lib.cc_run(self.ctx))

    def run_host(self, packed: PackedReads):
        """cc_run_host: copy in and run every stage."""
        self._check(self.lib.cc_run_host(*self._host_args(packed)))

    # -- results --------------------------------------------------------------------
    def counts(self):
        c = (ctypes.c_uint64 * 8)()
        self._check(self.lib.cc_counts(self.ctx, c))
        return dict(reads=c[0], fam=c[1], sscs=c[2], corr=c[3], dcs=c[4], dcssc=c[5])

    def fetch(self, name):
        what, dt = ARRAYS[name]
        nbytes = self.lib.cc_array_bytes(self.ctx, what)
        out = np.zeros(nbytes // np.dtype(dt).itemsize, dt)
        if nbytes:
            self._check(self.lib.cc_fetch(self.ctx, what, out.ctypes.data, nbytes))
        if name.endswith("_rows"):
            out = out.reshape(-1, self.stride)
        if name.endswith("_planes") and self.packing:
            out = out.reshape(-1, self.packing[1])
        return out

    def device_ptr(self, name):
        return self.lib.cc_device_ptr(self.ctx, ARRAYS[name][0])

    def fetch_many(self, names):
        """Several result arrays with one wait for all the copies (cc_fetch_many)."""
        k = len(names)
        whats = (ctypes.c_int * k)(*[ARRAYS[nm][0] for nm in names])
        sizes = (ctypes.c_uint64 * k)(*[self.lib.cc_array_bytes(self.ctx, ARRAYS[nm][0]) for nm in names])
        outs = [np.zeros(sizes[j] // np.dtype(ARRAYS[nm][1]).itemsize, ARRAYS[nm][1]) for j, nm in enumerate(names)]
        dsts = (ctypes.c_void_p * k)(*[o.ctypes.data if o.size else None for o in outs])
        self._check(self.lib.cc_fetch_many(self.ctx, k, whats, dsts, sizes))
        shape = lambda nm, o: (o.reshape(-1, self.stride) if nm.endswith("_rows") else
                               o.reshape(-1, self.packing[1]) if nm.endswith("_planes") and self.packing else o)
        return {nm: shape(nm, o) for nm, o in zip(names, outs)}

    def pack_rows(self):
        """cc_pack_rows: the produced rows once more in bit planes.  (qbits, stride_c, qual_table, base_table), or
        None when the rows hold more than 16 (quality, literal base) combinations."""
        qbits, stride_c = ctypes.c_uint32(0), ctypes.c_uint32(0)
        qt, bt = np.zeros(16, np.uint8), np.zeros(16, np.uint8)
        self._check(self.lib.cc_pack_rows(self.ctx, ctypes.byref(qbits), ctypes.byref(stride_c), qt.ctypes.data, bt.ctypes.data))
        self.packing = (qbits.value, stride_c.value, qt, bt) if qbits.value else None
        return self.packing

    def result(self, planes=False) -> ConsensusResult:
        """Every result array.  planes=True reads the rows back in bit planes (half the bytes) when they fit and
        expands them on the host - same result."""
        from .result import ALL_ARRAYS
        names = list(ALL_ARRAYS)
        if planes and self.pack_rows():
            from .pack import planes_to_rows
            qbits, _, qt, bt = self.packing
            row_names = [nm for nm in names if nm.endswith("_rows")]
            arrays = self.fetch_many([nm for nm in names if nm not in row_names] + [nm[:-4] + "planes" for nm in row_names])
            for nm in row_names:
                arrays[nm] = planes_to_rows(arrays.pop(nm[:-4] + "planes"), qbits, qt, bt, self.lpad)
        else:
            arrays = self.fetch_many(names)
        return ConsensusResult(self.lpad, self.stride, bool(self.params.scorrect), **arrays)

    def consensus(self, packed: PackedReads, compact=False) -> ConsensusResult:
        """The whole path on one batch: host SoA in, family-level results out.
        compact=True sends the rows with 4-bit quality codes (cc_load_reads_q4)."""
        q4 = None
        if compact in ("auto", "planes", "classes"):      # the most compact form the batch allows
            from .pack import to_classes, to_compact, to_q4
            planes = to_compact(packed)
            if planes is not None:
                if compact != "planes" and to_classes(planes[0]) is not None:
                    self.load_classes(packed, planes)
                elif compact == "classes":
                    raise GpuError("the batch does not fit the short pair records")
                else:
                    self.load_compact(packed, planes)
                self.run()
                return self.result()
            if compact in ("planes", "classes"):
                raise GpuError("the batch does not fit the compact format")
            q4 = to_q4(packed)
            compact = q4 is not None
        if compact:
            self.load_q4(packed, q4)
            self.run()
        else:
            self.run_host(packed)
        return self.result()

    # -- stand-alone primitives ----------------------------------------------------
    def join_records(self, fields, dup_bc):
        """cc_join_records: index of the record with the complementary key, or -1, for every record."""
        fields = np.ascontiguousarray(fields)
        dup = np.ascontiguousarray(dup_bc, dtype=np.uint32)
        out = np.full(len(fields), -1, np.int32)
        if len(fields):
            self._check(self.lib.cc_join_records(self.ctx, fields.ctypes.data, len(fields), dup.ctypes.data if len(dup) else None,
                                                 len(dup), out.ctypes.data))
        return out

    def consensus_maker_rows(self, rows, lpad, cutoff):
        rows = np.ascontiguousarray(rows, np.uint8)
        out = np.zeros(rows.shape[1], np.uint8)
        self._check(self.lib.cc_consensus_maker(self.ctx, rows.ctypes.data, rows.shape[0], int(lpad),
                                                rows.shape[1], float(cutoff), out.ctypes.data))
        return out

    def duplex_consensus_rows(self, rows1, rows2, lpad, sc):
        rows1 = np.ascontiguousarray(np.atleast_2d(rows1), np.uint8)
        rows2 = np.ascontiguousarray(np.atleast_2d(rows2), np.uint8)
        out = np.zeros_like(rows1)
        self._check(self.lib.cc_duplex_consensus(self.ctx, rows1.ctypes.data, rows2.ctypes.data,
                                                 rows1.shape[0], int(lpad), rows1.shape[1],
                                                 1 if sc else 0, out.ctypes.data))
        return out

    def sort_coordinates(self, tid, pos, flag):
        """Stable coordinate order of records given as arrays: the permutation `sort_index` /
        `samtools merge` would apply (ConsensusCruncher.py:10-29,248).  -> uint32[n]."""
        tid = np.ascontiguousarray(tid, np.int32)
        pos = np.ascontiguousarray(pos, np.int32)
        flag = np.ascontiguousarray(flag, np.uint16)
        order = np.empty(len(tid), np.uint32)
        self._check(self.lib.cc_sort_coordinates(self.ctx, tid.ctypes.data, pos.ctypes.data, flag.ctypes.data,
                                                 len(tid), order.ctypes.data))
        return order

    def sort_records(self, records):
        """`records.sort_records` with the ordering done on the GPU."""
        n = len(records)
        if n < 2:
            return list(records)
        order = self.sort_coordinates(np.fromiter((r.tid for r in records), np.int32, n),
                                      np.fromiter((r.pos for r in records), np.int32, n),
                                      np.fromiter((r.flag for r in records), np.uint16, n))
        return [records[i] for i in order]

    # -- multi-GPU: one rank per context, NCCL called by the library itself ---------------
    @staticmethod
    def comm_unique_id():
        """128 bytes from ncclGetUniqueId; rank 0 makes them, the host hands them to every rank."""
        lib = load_library()
        buf = (ctypes.c_uint8 * COMM_ID_BYTES)()
        rc = lib.cc_comm_unique_id(buf)
        if rc:
            raise GpuError("cc_comm_unique_id: %s" % lib.cc_last_error(None).decode())
        return bytes(buf)

    def comm_init(self, unique_id, rank, nranks):
        """Collective: every rank calls it with the same id (ncclCommInitRank on this context's GPU)."""
        if len(unique_id) != COMM_ID_BYTES:
            raise ValueError("the communicator id is %d bytes" % COMM_ID_BYTES)
        buf = (ctypes.c_uint8 * COMM_ID_BYTES).from_buffer_copy(unique_id)
        self._check(self.lib.cc_comm_init(self.ctx, buf, int(rank), int(nranks)))

    def stats_allreduce(self):
        """`stats` and `fam_hist` become the sums over all ranks (one ncclAllReduce, in place, on the
        context's stream); a no-op without a communicator."""
        self._check(self.lib.cc_stats_allreduce(self.ctx))

    def comm_destroy(self):
        self._check(self.lib.cc_comm_destroy(self.ctx))

    def comm_info(self):
        r, n, v = ctypes.c_int(0), ctypes.c_int(1), ctypes.c_int(0)
        self._check(self.lib.cc_comm_info(self.ctx, ctypes.byref(r), ctypes.byref(n), ctypes.byref(v)))
        return {"rank": r.value, "nranks": n.value, "nccl_version": v.value}

    # -- timing ------------------------------------------------------------------------
    def profile(self, enable=True):
        self._check(self.lib.cc_profile(self.ctx, 1 if enable else 0))

    def phase_ms(self):
        out = {}
        for i, nm in enumerate(PHASES):
            ms, runs = ctypes.c_double(0), ctypes.c_uint64(0)
            self._check(self.lib.cc_phase_ms(self.ctx, i, ctypes.byref(ms), ctypes.byref(runs)))
            out[nm] = (ms.value, runs.value)
        return out

    def kernel_launches(self):
        return int(self.lib.cc_kernel_launches(self.ctx))

    def device_bytes(self):
        return int(self.lib.cc_device_bytes(self.ctx))
