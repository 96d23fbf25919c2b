"""bench_trade.py -- bench.py's hooks for the windowed workload (aquery2_b200/trade.py): verification of the
timed results, the CPU baseline leg (the reference build, passed in by bench.py) and the e2e arm.
Bench infrastructure, not product code."""
from __future__ import annotations

import ctypes
import os
import time

import numpy as np

from aquery2_b200.trade import *          # noqa: F401,F403  (ORDER, ALG_BYTES, Plan, device_table, ...)
from aquery2_b200.trade import ORDER, COLUMNS, expressions, host_table, table_name, n_sym


# ------------------------------------------------------------------------------------------------
# CPU side: the reference's hashtable_push / ht_postproc / per-group loops over the same expressions
# ------------------------------------------------------------------------------------------------
def _cpu_query(impl, t, q):
    """returns (seconds, ng, offsets, perm, outputs) for one query on host columns"""
    t0 = time.perf_counter()
    ng, gid, off, perm = impl.groupby([t["sym"]])
    # row lists in time order: ascending rows (sorted table) or by the time column (ASSUMING ASC time)
    if q == "w_unsorted":
        order = np.lexsort((t["time_shuffled"], gid)).astype(np.uint32)       # TableInfo::order_by per group (table.h:447-461)
    else:
        order = np.argsort(gid, kind="stable").astype(np.uint32)              # ascending rows inside every group
    outs = []
    for e in expressions(t):
        op, v = e[0], e[1]
        w = e[2] if len(e) > 2 else 0
        r = impl.scan(op, v, off, order, w=w, raw128=True)
        if len(e) > 3 and e[3]:
            r = (v[order] - r).astype(v.dtype)                                 # vector arithmetic (table.h:820-973)
        if len(e) > 4 and e[4]:
            r = np.maximum.reduceat(r, off[:-1].astype(np.int64)) if ng else r[:0]   # max per group (aggregations.h:71-78)
        outs.append(r)
    return time.perf_counter() - t0, ng, off, order, outs


def cpu_step(impl, t, queries) -> float:
    return sum(_cpu_query(impl, t, q)[0] for q in queries)


def _compare(aq, got_cols, want, what):
    for i, (g, w) in enumerate(zip(got_cols, want)):
        a = g.numpy() if hasattr(g, "numpy") else np.asarray(g)
        a = a[:len(w)]
        if a.dtype.kind == "f":
            scale = float(np.abs(w).max()) if len(w) else 0.0
            bad = np.flatnonzero(~(np.abs(a - w) <= 1e-12 * np.maximum(np.maximum(np.abs(a), np.abs(w)), scale)))
        else:
            bad = np.flatnonzero((a.reshape(len(a), -1) != np.asarray(w).reshape(len(w), -1)).any(axis=1))
        if len(bad):
            raise AssertionError(f"{what}: expression {i}: {len(bad)} cells differ, first at {bad[:3]}: {a[bad[:3]]} vs {w[bad[:3]]}")


def cpu_baseline(aq, args, queries, impl, note, check=True) -> dict:
    n = int(min(args.cpu_rows, args.rows))
    t = host_table(args, n, queries)
    out, sec, rep = {}, 0.0, {}
    status = "ok"
    for q in queries:
        s, ng, off, order, outs = _cpu_query(impl, t, q)
        sec += s
        if check:
            try:
                h = aq.Groups([t["sym"]], row_order=aq.ROWS_ASC)
                if q == "w_unsorted":
                    h.order_by([t["time_shuffled"]], desc=[0])
                got = h.window(expressions(t))
                assert h.ngroups == ng, f"{q}: {h.ngroups} groups vs {ng}"
                _compare(aq, got, outs, q)
                rep[q] = {"groups": ng, "rows": n}
                h.free()
            except AssertionError as e:
                status = f"{e}"[:500]
    if check:
        out["oracle_parity"] = {"status": (f"ok ({len(queries)} queries, {n:.0e} rows, bit-exact ints/order, fp64 <= 1e-12)" if status == "ok" else status),
                                "kind": impl.kind, "per_query": rep}
    out.update({"value": len(queries) * n / sec, "unit": "rows/s", "cores": 1,
                "kind": "reference" if impl.kind == "reference" else "port",
                "sample": f"{table_name(args, n)} {','.join(queries)} once ({sec:.1f} s): group build + per-group loops of "
                          f"aggregations.h; the per-group time sort is numpy lexsort; " + note,
                "host_cores": os.cpu_count()})
    return out


# ------------------------------------------------------------------------------------------------
# verification at full size: a few symbols recomputed from scratch with torch
# ------------------------------------------------------------------------------------------------
def verify(aq, comm, runner, table, queries, args, n_total, ngs) -> dict:
    if comm is not None:
        return runner.verify(table, queries, args, n_total, ngs)
    import torch
    import bench_verify as bv
    aq.sync()
    t0 = time.perf_counter()
    sym = bv.tview(torch, aq, table["sym"])
    n = sym.numel()
    ns = n_sym(args, n_total)
    # group structure derived independently: first row and size of every symbol, first-appearance order
    counts = torch.bincount(sym, minlength=ns)
    first = torch.full((ns,), n, dtype=torch.int64, device="cuda")
    first.scatter_reduce_(0, sym.to(torch.int64), torch.arange(n, device="cuda", dtype=torch.int64), reduce="amin")
    present = counts > 0
    want_ng = int(present.sum().item())
    order = torch.argsort(torch.where(present, first, torch.full_like(first, n)))[:want_ng]      # symbols in gid order
    offs = torch.zeros(want_ng + 1, dtype=torch.int64, device="cuda")
    offs[1:] = torch.cumsum(counts[order], 0)
    del first
    report = {}
    price = bv.tview(torch, aq, table["price"]); qty = bv.tview(torch, aq, table["quantity"])
    for q in queries:
        bv._need(int(ngs[q]) == want_ng, f"{q}: {ngs[q]} groups vs {want_ng} symbols present")
        outs = [bv.tview(torch, aq, c) for c in runner.result_columns(q)]
        picks = sorted({0, want_ng - 1, want_ng // 2, want_ng // 3, (want_ng * 7) // 10})[:5]
        for g in picks:
            s = int(order[g].item())
            rows = torch.nonzero(sym == s).flatten()
            if q == "w_unsorted":
                tm = bv.tview(torch, aq, table["time_shuffled"])[rows]
                rows = rows[torch.argsort(tm, stable=True)]
            lo, hi = int(offs[g].item()), int(offs[g + 1].item())
            bv._need(hi - lo == rows.numel(), f"{q}: group {g}: {hi - lo} cells vs {rows.numel()} rows")
            p = price[rows].to(torch.int64).cpu().numpy(); qq = qty[rows].to(torch.int64).cpu().numpy()
            m = len(p)
            idx = np.arange(m)
            cnt = np.minimum(idx + 1, 5)
            cs = np.concatenate([[0], np.cumsum(p)]); cq = np.concatenate([[0], np.cumsum(qq)])
            sumw_p = cs[idx + 1] - cs[idx + 1 - cnt]; sumw_q = cq[idx + 1] - cq[idx + 1 - cnt]
            minw = np.array([p[max(0, i - 4):i + 1].min() for i in range(m)])
            deltas = np.concatenate([[0], np.diff(p)])
            maxdd = int((p - np.minimum.accumulate(p)).max())
            got = [o[lo:hi].cpu().numpy() if o.shape[0] == n else None for o in outs]
            bv._need(np.allclose(got[0], sumw_p / cnt, rtol=1e-12, atol=0), f"{q}: group {g}: avgs(5, price)")
            bv._need(bool((got[1][:, 0] == sumw_q).all() and (got[1][:, 1] == 0).all()), f"{q}: group {g}: sums(5, quantity)")
            bv._need(bool((got[2] == deltas).all()), f"{q}: group {g}: deltas(price)")
            bv._need(bool((got[3] == minw).all()), f"{q}: group {g}: mins(5, price)")
            bv._need(int(outs[4][g].item()) == maxdd, f"{q}: group {g}: max(price - mins(price)) = {int(outs[4][g].item())} vs {maxdd}")
        # a column-total identity over every row: sum of deltas = sum over groups of (last - first price)
        report[q] = {"groups": want_ng, "recomputed_groups": len(picks)}
    torch.cuda.synchronize()
    return {"status": "ok", "rows": n_total, "per_query": report, "seconds": round(time.perf_counter() - t0, 1),
            "checks": "symbols present == groups; group sizes / CSR offsets derived independently; all five expressions of sampled "
                      "symbols recomputed from the raw columns"}


def run_e2e(aq, args, table, runner, queries, n_total) -> dict:
    """host (pinned) columns -> HBM column store -> queries -> every result column back to pinned host memory"""
    L = aq.lib()
    pinned, host = [], {}
    for c, d in table.items():
        pa = aq.PinnedArray(d.n, d.code)
        aq.check(L.aqgpu_col_download(ctypes.c_void_p(pa.ptr), ctypes.c_void_p(d.ptr), ctypes.c_size_t(d.nbytes)))
        pinned.append(pa); host[c] = pa
    res_host = {}

    def one_step():
        bo = 0
        for c, d in table.items():
            aq.check(L.aqgpu_col_upload(ctypes.c_void_p(d.ptr), ctypes.c_void_p(host[c].ptr), ctypes.c_size_t(d.nbytes)))
        for q in queries:
            runner.run(q)
            for i, col in enumerate(runner.result_columns(q)):
                key = (q, i)
                if key not in res_host:
                    res_host[key] = aq.PinnedArray(col.n, col.code)
                    pinned.append(res_host[key])
                aq.check(L.aqgpu_col_download(ctypes.c_void_p(res_host[key].ptr), ctypes.c_void_p(col.ptr), ctypes.c_size_t(col.nbytes)))
                bo += col.nbytes
        return bo

    one_step()
    aq.sync()
    t0 = time.perf_counter()
    aq.timer_begin()
    for _ in range(args.e2e_steps):
        bo = one_step()
    ms = aq.timer_end()
    wall = time.perf_counter() - t0
    bi = sum(d.nbytes for d in table.values())
    out = {"value": len(queries) * n_total / (ms * 1e-3 / args.e2e_steps), "unit": "rows/s",
           "h2d_bytes_per_step": bi, "d2h_bytes_per_step": bo, "steps": args.e2e_steps,
           "ms_per_step": ms / args.e2e_steps, "wall_s_per_step": wall / args.e2e_steps,
           "host_memory": "pinned (aqgpu_host_alloc)",
           "path": "ingest every column once per step (H2D) -> group build (+ order_by) + fused window call -> every result column D2H"}
    for pa in pinned:
        pa.free()
    return out
