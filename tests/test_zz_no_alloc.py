"""GPU: cc_consensus_maker_static, the consensus_maker entry point that allocates nothing.

Some callers of libccgpu.so may not allocate device memory while they run: code living
inside a framework whose pool already owns the GPU, hosts that forbid driver allocations
after start-up, or code that must not trigger the device-wide synchronisation a cudaMalloc
implies.  For them the library keeps the vote kernel's working memory in a fixed array in
its own device image; the caller brings the family's rows and the result row as device
pointers.  This test calls it through the C ABI twice in a row and checks each result,
byte for byte, against the allocating entry point cc_consensus_maker on the same family
and - because both run the same vote kernel - against the CPU oracle's consensus_maker too.

It is collected last (file name) so that a run stopping at the first failure has already
run everything else, and it sets up its own context so it also passes alone.
"""
import ctypes

import numpy as np
import pytest

from consensuscruncher_b200 import pack
from consensuscruncher_b200.engine import GpuConsensus, load_library
from oracle import oracle

pytestmark = pytest.mark.gpu

CC_ERR_BASE, CC_ERR_TOO_LARGE = -4, -6
MAX_READS = 4096


def _family(rng, n, lpad, n_with_q=False):
    stride = pack.row_geometry(lpad)[1]
    codes = np.array([1, 2, 4, 8], np.uint8)[rng.integers(0, 4, (n, lpad))]
    codes[:, : lpad * 3 // 4] = codes[0, : lpad * 3 // 4]          # mostly agreeing columns
    qual = rng.integers(5, 45, (n, lpad)).astype(np.uint8)
    low_n = rng.random((n, lpad)) < 0.02
    codes[low_n] = 15
    qual[low_n] = 2
    if n_with_q:
        codes[0, 3] = 15
        qual[0, 3] = 40
    rows = np.zeros((n, stride), np.uint8)
    for i in range(n):
        pack.pack_row(codes[i], qual[i], lpad, rows[i])
    return rows, stride


def _static_call(lib, torch, rows, lpad, cutoff):
    """Rows and result live in memory the CALLER allocated (here: torch)."""
    d_rows = torch.from_numpy(rows).cuda()
    d_out = torch.full((rows.shape[1],), 0xEE, dtype=torch.uint8, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    rc = lib.cc_consensus_maker_static(d_rows.data_ptr(), rows.shape[0], lpad, rows.shape[1],
                                       float(cutoff), d_out.data_ptr(), ctypes.c_void_p(stream))
    return rc, d_out.cpu().numpy()


def test_static_entry_point_matches_allocating_one_twice_in_a_row():
    torch = pytest.importorskip("torch")
    lib = load_library()
    torch.cuda.set_device(0)
    torch.zeros(1, device="cuda")                      # caller-side context; nothing of ours is allocated here
    rng = np.random.default_rng(2024)
    with GpuConsensus(0) as eng:
        # two calls in a row; the second family is smaller and narrower than the first, so
        # anything the first call left in the shared scratch would show
        for n, lpad, cutoff in ((300, 160, 0.7), (5, 32, 0.5), (2, 128, 1.0), (MAX_READS, 64, 0.7)):
            rows, stride = _family(rng, n, lpad)
            expected = eng.consensus_maker_rows(rows, lpad, cutoff)        # existing entry point
            rc, got = _static_call(lib, torch, rows, lpad, cutoff)
            assert rc == 0, lib.cc_last_error(None).decode()
            assert np.array_equal(got, expected), "n=%d lpad=%d" % (n, lpad)
            # both entry points share the kernel: the oracle is the independent check
            assert np.array_equal(got, oracle.consensus_maker_rows(rows, lpad, cutoff)), "oracle, n=%d lpad=%d" % (n, lpad)

        # an error raised by one call must not stick to the next one
        bad, _ = _family(rng, 4, 32, n_with_q=True)
        rc, _ = _static_call(lib, torch, bad, 32, 0.7)
        assert rc == CC_ERR_BASE
        good, _ = _family(rng, 4, 32)
        rc, got = _static_call(lib, torch, good, 32, 0.7)
        assert rc == 0 and np.array_equal(got, eng.consensus_maker_rows(good, 32, 0.7))
        assert np.array_equal(got, oracle.consensus_maker_rows(good, 32, 0.7))

        # the call synchronises, so a capturing stream is refused instead of breaking the capture
        side = torch.cuda.Stream()
        d_rows = torch.from_numpy(good).cuda()
        d_out = torch.zeros(good.shape[1], dtype=torch.uint8, device="cuda")
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.stream(side):
            graph.capture_begin()
            rc = lib.cc_consensus_maker_static(d_rows.data_ptr(), good.shape[0], 32, good.shape[1], 0.7,
                                               d_out.data_ptr(), ctypes.c_void_p(side.cuda_stream))
            graph.capture_end()
        assert rc == -2 and b"capturing" in lib.cc_last_error(None)

        # sizes beyond the fixed scratch are refused with a clear message, nothing is launched
        big, _ = _family(rng, 8, 32)
        d_rows = torch.from_numpy(big).cuda()
        d_out = torch.zeros(big.shape[1], dtype=torch.uint8, device="cuda")
        rc = lib.cc_consensus_maker_static(d_rows.data_ptr(), MAX_READS + 1, 32, big.shape[1], 0.7,
                                           d_out.data_ptr(), None)
        assert rc == CC_ERR_TOO_LARGE
        assert b"CC_STATIC_MAX_READS" in lib.cc_last_error(None)
