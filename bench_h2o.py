"""bench_h2o.py -- bench.py's hooks for the h2o workload (aquery2_b200/h2o.py): verification of the timed
results (bench_verify.py), the CPU baseline leg (the reference build, passed in by bench.py) and the e2e arm.
Bench infrastructure, not product code."""
from __future__ import annotations

import ctypes
import time

from aquery2_b200.h2o import *          # noqa: F401,F403  (QUERIES, ORDER, ALG_BYTES, Plan, device_table, ...)
from aquery2_b200.h2o import QUERIES, ORDER, key_bound, host_table, table_name


# ------------------------------------------------------------------------------------------------
# hooks for bench.py
# ------------------------------------------------------------------------------------------------
def verify(aq, comm, runner, table, queries, args, n_total, ngs) -> dict:
    """size-independent checks of the results the timed loop left in HBM (bench_verify.py)"""
    if comm is not None:
        return runner.verify(table, queries, args, n_total, ngs)
    import torch
    import bench_verify as bv
    aq.trim_pool()
    report = {}
    t0 = time.perf_counter()
    for q in queries:
        keys, aggs, epi = QUERIES[q]
        ranges = [key_bound(c, n_total, args.k)[1] for c in keys]
        bv.verify_query(torch, aq, q, keys, aggs, epi, table, runner.prepared[q], runner.extra[q], int(ngs[q]), ranges, report)
        torch.cuda.synchronize()
        torch.cuda.empty_cache()
        aq.trim_pool()
    return {"status": "ok", "rows": n_total, "per_query": report, "seconds": round(time.perf_counter() - t0, 1),
            "checks": "independent distinct count; sum(count)=rows; first rows ascending; key==col[first_row]; column-total "
                      "identities; groups recomputed through row masks; epilogues"}


def cpu_step(impl, t, queries) -> float:
    """one pass of the queries through hashtable_push / ht_postproc / per-group aggregates; seconds
    inside the postproc bracket (aqo_time_groupby_reduce)"""
    total = 0.0
    for q in queries:
        keys, aggs, _ = QUERIES[q]
        specs = [(a[0],) + tuple(t[c] for c in a[1:]) for a in aggs]
        sec, _ = impl.time_groupby_reduce([t[c] for c in keys], specs)
        total += sec
    return total


def cpu_baseline(aq, args, queries, impl, note, check=True) -> dict:
    """the reference build timed on one host core over the cpu_baseline table; with check=True the very
    results it produced are compared with the GPU's on the same table (bench_verify.verify_vs_oracle)"""
    import os
    n = int(min(args.cpu_rows, args.rows))
    t = host_table(args, n, queries)
    out = {}
    if check:
        import bench_verify as bv
        rep = {}
        try:
            sec = bv.verify_vs_oracle(aq, impl, t, queries, QUERIES, rep)
            out["oracle_parity"] = {"status": f"ok ({len(queries)} queries, {n:.0e} rows, bit-exact ints/order, fp64 <= {bv.FP_RTOL_ORACLE})",
                                    "kind": impl.kind, "per_query": rep}
        except bv.VerifyError as e:
            out["oracle_parity"] = {"status": f"{e}"[:500], "kind": impl.kind}
            sec = cpu_step(impl, t, queries)
    else:
        sec = cpu_step(impl, t, queries)
    out.update({"value": len(queries) * n / sec, "unit": "rows/s", "cores": 1,
                "kind": "reference" if impl.kind == "reference" else "port",
                "sample": f"{table_name(args, n)} {queries[0]}-{queries[-1]} once ({sec:.1f} s), " + note,
                "host_cores": os.cpu_count()})
    return out


def run_e2e(aq, args, table, runner, queries, n_total) -> dict:
    """End to end with HOST buffers: the table lives in pinned host memory (the MonetDB-owned buffers of
    the reference, server/monetdb_conn.cpp:203-224); a step ingests it into the HBM column store
    (the 'C' cache message, server/server.cpp:552-583: one H2D copy per column), runs the queries on the
    store and writes every result column back to pinned host memory.  H2D and D2H are inside the timed
    region.  Default: the copy-stream pipeline (aqgpu_ingest_async / aqgpu_writeback_async): a query starts as soon
    as ITS columns have arrived, its results leave while later columns arrive and later queries run; the query with
    the largest result (q10, G ~ N) goes last, so that its write-back overlaps the next step's ingest.  AQ_E2E_SERIAL=1: one synchronous copy after the other."""
    import os
    from aquery2_b200.h2o import columns_of
    L = aq.lib()
    serial = os.environ.get("AQ_E2E_SERIAL") == "1"
    pinned, host = [], {}
    for c, d in table.items():
        pa = aq.PinnedArray(d.n, d.code)
        if d.nbytes:
            aq.check(L.aqgpu_col_download(ctypes.c_void_p(pa.ptr), ctypes.c_void_p(d.ptr), ctypes.c_size_t(d.nbytes)))
        pinned.append(pa)
        host[c] = pa
    res_host = {}
    # Copies of one direction execute in submission order across streams (measured, profiles/r02/e2e_pipeline_trace.txt): with
    # q10 (36 GB of results) first, q1's 100-byte flag read-back waited 645 ms behind q10's write-back.  So the query with the
    # largest result goes LAST: its write-back overlaps the NEXT step's ingest (the two directions do overlap, 45 + 45 GB/s).
    order = list(queries) if serial else sorted(queries, key=lambda q: 1 if q == "q10" else 0)
    need = {q: [c for c in columns_of([q]) if c in table] for q in queries}

    def result_buffer(q, i, col, ng):
        nb = ng * col.itemsize
        key = (q, i)
        if key not in res_host or res_host[key].n * res_host[key].itemsize < nb:
            if key in res_host:
                aq.writeback_wait()
                res_host[key].free()
            res_host[key] = aq.PinnedArray(max(ng, 1), col.code)
            res_host[key].itemsize = col.itemsize
            pinned.append(res_host[key])
        return res_host[key], nb

    trace = os.environ.get("AQ_E2E_TRACE") == "1"
    t_step = [0.0]

    def mark(what):
        if trace:
            import sys
            print(f"[e2e trace] +{(time.perf_counter() - t_step[0]) * 1e3:8.1f} ms host: {what}", file=sys.stderr, flush=True)

    def one_step():
        bo = 0
        ngs = {}
        t_step[0] = time.perf_counter()
        if serial:
            for c, d in table.items():                 # ingest
                aq.check(L.aqgpu_col_upload(ctypes.c_void_p(d.ptr), ctypes.c_void_p(host[c].ptr), ctypes.c_size_t(d.nbytes)))
        if not serial:
            # Every column of the step is enqueued on the ingest stream BEFORE the first query runs: a query blocks the host until
            # its group count is back, and that read-back queues behind the previous step's big write-back (same direction), so
            # anything enqueued later would start only when that write-back is over (measured: 1.33 s per step that way).
            arrived = []
            for q in order:
                arrived += [c for c in need[q] if c not in arrived]
            arrived += [c for c in table if c not in arrived]
            for c in arrived:                          # H2D, ordered after the queries already enqueued (they may still read the column)
                aq.ingest_async(table[c], host[c].ptr)
            aq.ingest_fence()                          # the queries below wait for the columns
        for qi, q in enumerate(order):
            if not serial and qi == len(order) - 1:
                aq.writeback_fence()                   # the last query's previous results (the big write-back) have left before it overwrites
            mark(f"{q}: enqueued")                     # them; the earlier queries waited only for the small write-backs before that fence
            ng = ngs[q] = runner.run(q)
            mark(f"{q}: done ({ng} groups)")
            for i, col in enumerate(runner.result_columns(q)):     # write-back
                buf, nb = result_buffer(q, i, col, ng)
                if nb and serial:
                    aq.check(L.aqgpu_col_download(ctypes.c_void_p(buf.ptr), ctypes.c_void_p(col.ptr), ctypes.c_size_t(nb)))
                elif nb:
                    aq.writeback_async(buf.ptr, col.ptr, nb)       # D2H on the write-back stream, after the query
                bo += nb
            mark(f"{q}: write-back enqueued")
        return ngs, bo

    try:
        return _run_e2e_timed(aq, L, args, table, runner, queries, n_total, serial, order, one_step, res_host)
    finally:                                       # also on an exception: drain the copy streams, give the pinned memory back
        try:
            aq.writeback_wait(); aq.ingest_fence(); aq.sync()
        except Exception:
            pass
        for pa in pinned:
            pa.free()


def _run_e2e_timed(aq, L, args, table, runner, queries, n_total, serial, order, one_step, res_host) -> dict:
    one_step()                                     # warm-up (also sizes the pinned result buffers)
    aq.writeback_wait()
    aq.ingest_fence()
    aq.sync()
    t0 = time.perf_counter()
    aq.timer_begin()
    for _ in range(args.e2e_steps):
        ngs, bo = one_step()
    aq.writeback_fence()                           # the timer's end event waits for the last result to land ...
    aq.ingest_fence()
    ms = aq.timer_end()
    aq.writeback_wait()
    wall = time.perf_counter() - t0
    # the results that arrived in host memory are the ones the HBM-resident run produced (spot check: first query's first column)
    q0 = order[-1]
    col0 = runner.result_columns(q0)[0]
    import numpy as np
    got = np.ctypeslib.as_array((ctypes.c_uint8 * (ngs[q0] * col0.itemsize)).from_address(res_host[(q0, 0)].ptr)).copy()
    want = np.zeros(ngs[q0] * col0.itemsize, np.uint8)
    if want.size:
        aq.check(L.aqgpu_col_download(want.ctypes.data_as(ctypes.c_void_p), ctypes.c_void_p(col0.ptr), ctypes.c_size_t(want.size)))
    landed = bool(np.array_equal(got, want))
    bi = sum(d.nbytes for d in table.values())
    out = {"value": len(queries) * n_total / (ms * 1e-3 / args.e2e_steps), "unit": "rows/s",
           "h2d_bytes_per_step": bi, "d2h_bytes_per_step": bo, "steps": args.e2e_steps,
           "ms_per_step": ms / args.e2e_steps, "wall_s_per_step": wall / args.e2e_steps,
           "host_memory": "pinned (aqgpu_host_alloc)", "results_landed": landed,
           "path": ("ingest every column once per step (H2D) -> queries on the HBM store -> every result column D2H, one synchronous copy after the other"
                    if serial else
                    "per step: every column H2D on the ingest stream, then the queries (q10 last), every result column D2H on the write-back "
                    "stream while later queries run; consecutive steps overlap on the "
                    "two copy streams (a step's ingest waits for the previous step's queries and small write-backs, its last query for the previous step's big write-back)")}
    return out
