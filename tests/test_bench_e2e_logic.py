"""Host logic of bench.py's e2e arm (bench_h2o.run_e2e) with the device replaced by host memory: the order of ingest /
fence / run / write-back calls, the pinned-buffer bookkeeping, the cleanup on failure and the spot check that results
landed.  No GPU: every 'device' pointer is a numpy buffer and the copy entry points are memmoves that log themselves."""
import ctypes
import os
import sys
import types

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


class FakeCol:
    def __init__(self, a):
        self.a = np.ascontiguousarray(a)
        self.n, self.itemsize, self.nbytes, self.code = len(self.a), self.a.dtype.itemsize, self.a.nbytes, 0
        self.ptr = self.a.ctypes.data


class FakeAq:
    """the subset of aquery2_b200.aqgpu that run_e2e uses"""
    def __init__(self, fail_on=None):
        self.log, self.live, self.fail_on = [], 0, fail_on
        aq = self

        class Pinned:
            def __init__(self, n, code):
                self.n, self.code = int(n), code
                self.buf = np.zeros(max(self.n, 1) * 16, np.uint8)
                self.ptr = self.buf.ctypes.data
                aq.live += 1

            def free(self):
                if self.buf is not None:
                    self.buf = None
                    aq.live -= 1
        self.PinnedArray = Pinned
        L = types.SimpleNamespace()
        L.aqgpu_col_download = lambda h, d, n: (ctypes.memmove(h.value, d.value, n.value), self.log.append("download"), 0)[2]
        L.aqgpu_col_upload = lambda d, h, n: (ctypes.memmove(d.value, h.value, n.value), self.log.append("upload"), 0)[2]
        self._L = L

    def lib(self):
        return self._L

    def check(self, rc):
        assert rc == 0

    def _op(self, name):
        self.log.append(name)
        if self.fail_on == name and self.log.count(name) > 12:
            raise RuntimeError(f"injected failure in {name}")

    def ingest_async(self, dst, src_ptr, nbytes=None):
        ctypes.memmove(dst.ptr, src_ptr, dst.nbytes if nbytes is None else nbytes); self._op("ingest")

    def writeback_async(self, dst_ptr, src_ptr, nbytes):
        ctypes.memmove(dst_ptr, src_ptr, nbytes); self._op("writeback")

    def ingest_fence(self): self._op("ingest_fence")
    def writeback_fence(self): self._op("writeback_fence")
    def writeback_wait(self): self._op("writeback_wait")
    def sync(self): self._op("sync")
    def timer_begin(self): self._op("timer_begin")
    def timer_end(self): self._op("timer_end"); return 12.5


class FakeRunner:
    def __init__(self, aq, table):
        self.aq, self.table, self.cols = aq, table, {}

    def run(self, q):
        self.aq.log.append(f"run {q}")
        ng = 7 if q != "q10" else self.table["id1"].n
        self.cols[q] = [FakeCol(np.arange(ng, dtype=np.int32) * (3 if q == "q10" else 1) + int(self.table["id1"].a[0]))]
        return ng

    def result_columns(self, q):
        return self.cols[q]


def _setup(fail_on=None):
    import bench_h2o
    n = 1000
    table = {c: FakeCol(np.full(n, i + 1, np.int32 if c != "v3" else np.float32)) for i, c in
             enumerate(["id1", "id2", "id3", "id4", "id5", "id6", "v1", "v2", "v3"])}
    aq = FakeAq(fail_on)
    args = types.SimpleNamespace(e2e_steps=2)
    return bench_h2o, aq, args, table, FakeRunner(aq, table), n


def test_pipeline_call_order_and_cleanup(monkeypatch):
    monkeypatch.delenv("AQ_E2E_SERIAL", raising=False)
    bench_h2o, aq, args, table, runner, n = _setup()
    queries = ["q1", "q3", "q10"]
    out = bench_h2o.run_e2e(aq, args, table, runner, queries, n)
    assert out["results_landed"] and out["steps"] == 2 and out["h2d_bytes_per_step"] == sum(c.nbytes for c in table.values())
    assert out["d2h_bytes_per_step"] == 7 * 4 * 2 + n * 4
    assert aq.live == 0                                           # every pinned buffer was given back
    log = aq.log
    # per step: all nine columns are enqueued before the first query; q10 runs last; the write-back fence sits right before it
    step = log[log.index("timer_begin"):]
    first_run = step.index("run q1")
    assert step[:first_run].count("ingest") == 9 and "ingest_fence" in step[:first_run]
    i10 = step.index("run q10")
    assert step[i10 - 1] == "writeback_fence" and step.index("run q3") < i10
    assert "upload" not in step                                   # the pipelined path issues no synchronous copies of columns


def test_serial_path_and_failure_cleanup(monkeypatch):
    monkeypatch.setenv("AQ_E2E_SERIAL", "1")
    bench_h2o, aq, args, table, runner, n = _setup()
    out = bench_h2o.run_e2e(aq, args, table, runner, ["q1", "q10"], n)
    assert out["results_landed"] and "one synchronous copy after the other" in out["path"] and aq.live == 0
    assert "ingest" not in aq.log and aq.log.count("upload") == 9 * 3          # warm-up + two timed steps
    monkeypatch.delenv("AQ_E2E_SERIAL")
    bench_h2o, aq, args, table, runner, n = _setup(fail_on="writeback")
    with pytest.raises(RuntimeError, match="injected"):
        bench_h2o.run_e2e(aq, args, table, runner, ["q1", "q3", "q10"] * 3, n)
    assert aq.live == 0                                           # nothing pinned survives the exception
