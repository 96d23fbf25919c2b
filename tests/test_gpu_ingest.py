"""Ingest / write-back on the copy streams (include/aqgpu.h aqgpu_ingest_async ... aqgpu_writeback_wait):
columns arrive from pinned host memory on the ingest stream, results leave on the write-back stream, ordering is by device
events only.  The pipelined result must equal the synchronous one, step after step, with the host buffers changing under it."""
import ctypes

import numpy as np
import pytest

from aquery2_b200 import synth
from util import assert_same

pytestmark = pytest.mark.gpu


def test_pipelined_steps_equal_synchronous_ones(gpu, port):
    n = 2_000_000
    L = gpu.lib()
    cols = ["id1", "id3", "v1", "v3"]
    dev = {c: gpu.DeviceColumn(n, gpu.NP2CODE[np.dtype(np.float32 if c == "v3" else np.int32)]) for c in cols}
    pin = {c: gpu.PinnedArray(n, dev[c].code) for c in cols}
    q_small = gpu.AutoQuery([dev["id1"]], [("sum", dev["v1"]), ("avg", dev["v3"])], want_first_row=True)
    q_big = gpu.AutoQuery([dev["id3"]], [("sum", dev["v1"]), ("count",)], want_first_row=True)
    out = {}
    for step in range(3):                                        # every step: new table contents in the SAME pinned buffers
        t = synth.h2o_table(n, 100, seed=100 + step, columns=cols)
        if step:
            gpu.writeback_wait()                                 # the host may not touch pinned result buffers still in flight
        for c in cols:
            pin[c].array[:] = t[c]
        for c in ("id3", "v1"):
            gpu.ingest_async(dev[c], pin[c].ptr)
        gpu.writeback_fence()
        gpu.ingest_fence()
        ng_big = q_big.run()                                     # started without waiting for id1 / v3
        for c in ("id1", "v3"):
            gpu.ingest_async(dev[c], pin[c].ptr)
        res = []
        for q, ng in ((q_big, ng_big),):
            for i, col in enumerate(list(q.kouts) + list(q.outs)):
                key = ("big", i)
                if key not in out:
                    out[key] = gpu.PinnedArray(n, col.code)
                gpu.writeback_async(out[key].ptr, col.ptr, ng * col.itemsize)
        gpu.ingest_fence()
        ng_small = q_small.run()
        for i, col in enumerate(list(q_small.kouts) + list(q_small.outs)):
            key = ("small", i)
            if key not in out:
                out[key] = gpu.PinnedArray(n, col.code)
            gpu.writeback_async(out[key].ptr, col.ptr, ng_small * col.itemsize)
        gpu.writeback_wait()
        # the synchronous answer on the same contents
        want_big = gpu.groupby_agg([t["id3"]], [("sum", t["v1"]), ("count",)])
        want_small = gpu.groupby_agg([t["id1"]], [("sum", t["v1"]), ("avg", t["v3"])])
        assert ng_big == want_big.ngroups and ng_small == want_small.ngroups
        assert_same(out[("big", 0)].array[:ng_big], want_big.keys[0], f"step {step} key")
        assert_same(out[("big", 1)].array[:ng_big], want_big.aggs[0], f"step {step} sum")
        assert_same(out[("big", 2)].array[:ng_big], want_big.aggs[1], f"step {step} count")
        assert_same(out[("small", 0)].array[:ng_small], want_small.keys[0], f"step {step} key")
        assert_same(out[("small", 1)].array[:ng_small], want_small.aggs[0], f"step {step} sum")
        assert_same(out[("small", 2)].array[:ng_small], want_small.aggs[1], f"step {step} avg")
    ng, gid, off, perm = port.groupby([t["id1"]])
    assert ng == ng_small
    for p in list(pin.values()) + list(out.values()):
        p.free()
    q_small.free(); q_big.free()
