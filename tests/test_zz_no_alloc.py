"""aqgpu_partition_counts_noalloc: the allocation-free variant of aqgpu_partition_rows' counting step (static device
scratch, include/aqgpu.h).  Called through the C ABI and checked against the existing entry point's counts (bit-exact:
integer work) on two calls in a row that share the scratch array, plus the documented limits.  Collected last on
purpose (the file name)."""
import ctypes
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

pytestmark = pytest.mark.gpu

MAX_ROWS = 8192 * 2048          # NOALLOC_MAX_ROWS (csrc/partition.cu)


@pytest.fixture(scope="module")
def aq():
    from aquery2_b200 import aqgpu
    aqgpu.init(0)
    return aqgpu


def _counts_existing(aq, owner_dev, n, nparts):
    counts = (ctypes.c_uint32 * nparts)()
    aq.check(aq.lib().aqgpu_partition_rows(ctypes.c_void_p(owner_dev.ptr), ctypes.c_uint32(n), ctypes.c_uint32(nparts),
                                           None, None, 0, None, counts))
    return np.frombuffer(counts, dtype=np.uint32).copy()


def _counts_noalloc(aq, owner_dev, n, nparts, out_dev):
    aq.check(aq.lib().aqgpu_partition_counts_noalloc(ctypes.c_void_p(owner_dev.ptr), ctypes.c_uint32(n),
                                                     ctypes.c_uint32(nparts), ctypes.c_void_p(out_dev.ptr)))
    aq.sync()
    return out_dev.numpy(nparts)


def test_noalloc_counts_match_existing_entry_point_twice(aq):
    rng = np.random.default_rng(20261105)
    out = aq.DeviceColumn.from_numpy(np.full(64, 0xdeadbeef, np.uint32))      # caller-owned output, poisoned
    # two calls in a row on the shared scratch: a large skewed input first, then a smaller one with fewer owners
    # and a ragged last chunk, whose cells the first call has left non-zero
    cases = [(3_000_017, 64), (70_001, 5), (2048, 1), (1, 3)]
    for n, nparts in cases:
        owner = (rng.zipf(1.3, n) % nparts).astype(np.uint32) if n > 1000 else rng.integers(0, nparts, n, dtype=np.uint32)
        d = aq.DeviceColumn.from_numpy(owner)
        want = _counts_existing(aq, d, n, nparts)
        assert np.array_equal(want, np.bincount(owner, minlength=nparts).astype(np.uint32))
        got = _counts_noalloc(aq, d, n, nparts, out)
        assert np.array_equal(got, want), (n, nparts, got, want)
        got2 = _counts_noalloc(aq, d, n, nparts, out)                          # same input again: same answer
        assert np.array_equal(got2, want), (n, nparts, got2, want)
        d.free()
    out.free()


def test_noalloc_largest_size_and_empty_input(aq):
    nparts = 7
    owner = (np.arange(MAX_ROWS, dtype=np.uint32) * np.uint32(2654435761) >> np.uint32(7)) % np.uint32(nparts)
    d = aq.DeviceColumn.from_numpy(owner)
    out = aq.DeviceColumn.from_numpy(np.full(nparts, 77, np.uint32))
    want = _counts_existing(aq, d, MAX_ROWS, nparts)
    assert np.array_equal(_counts_noalloc(aq, d, MAX_ROWS, nparts, out), want)
    assert int(want.sum()) == MAX_ROWS
    assert np.array_equal(_counts_noalloc(aq, d, 0, nparts, out), np.zeros(nparts, np.uint32))
    d.free(); out.free()


def test_noalloc_rejects_sizes_beyond_the_static_scratch(aq):
    L = aq.lib()
    d = aq.DeviceColumn.from_numpy(np.zeros(16, np.uint32))
    out = aq.DeviceColumn.from_numpy(np.zeros(64, np.uint32))
    rc = L.aqgpu_partition_counts_noalloc(ctypes.c_void_p(d.ptr), ctypes.c_uint32(MAX_ROWS + 1), ctypes.c_uint32(4),
                                          ctypes.c_void_p(out.ptr))
    assert rc != 0
    msg = L.aqgpu_last_error().decode()
    assert "maximum" in msg and str(MAX_ROWS) in msg, msg
    rc = L.aqgpu_partition_counts_noalloc(ctypes.c_void_p(d.ptr), ctypes.c_uint32(16), ctypes.c_uint32(65),
                                          ctypes.c_void_p(out.ptr))
    assert rc != 0 and "owners" in L.aqgpu_last_error().decode()
    host = np.zeros(64, np.uint32)                                            # a host pointer is refused, not dereferenced
    rc = L.aqgpu_partition_counts_noalloc(ctypes.c_void_p(d.ptr), ctypes.c_uint32(16), ctypes.c_uint32(4),
                                          host.ctypes.data_as(ctypes.c_void_p))
    assert rc != 0 and "device" in L.aqgpu_last_error().decode()
    # the library still works after the refusals
    got = _counts_noalloc(aq, d, 16, 4, out)
    assert np.array_equal(got, np.array([16, 0, 0, 0], np.uint32))
    d.free(); out.free()
