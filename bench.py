#!/usr/bin/env python
"""bench.py -- h2o group-by q1-q10 throughput (rows/s) of the B200 post-processor path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--rows R]

A step = one pass of the hot path over the synthetic table G1_{R}_1e2_0_0: the ten queries of the
reference's benchmark/h2o/groupby.sql, each from "columns resident in HBM" to "result columns
resident in HBM" (the reference's postproc_time bracket, server/server.cpp:303-305).
value = (queries x rows) / time, whole job.  One JSON line on stdout (rank 0).

--impl reference times the reference's own CPU post-processor (oracle/_ref/libaqref.so = the
reference headers compiled here; else the C port) on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import ctypes
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "h2o groupby rows/sec (q1-q10)"
UNIT = "rows/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--rows", type=float, default=1e9, help="total rows of the table (all ranks)")
    ap.add_argument("--k", type=int, default=100, help="h2o K (id1,id2,id4,id5 cardinality)")
    ap.add_argument("--queries", default="q1,q2,q3,q4,q5,q6,q7,q8,q9,q10")
    ap.add_argument("--e2e-steps", type=int, default=2)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--cpu-rows", type=float, default=1e7, help="rows of the CPU-baseline sample table")
    ap.add_argument("--no-bounds", action="store_true", help="do not pass key statistics to the library")
    return ap.parse_args()


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms during the timed region"""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx, self.proc, self.path = gpu_index, None, f"/tmp/aq_clocks_{os.getpid()}.csv"

    def start(self):
        try:
            self.f = open(self.path, "w")
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.idx)], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.f.close()
        sm, mx, reasons = [], [], set()
        for line in open(self.path):
            p = [x.strip() for x in line.split(",")]
            if len(p) < 9:
                continue
            try:
                sm.append(float(p[1])); mx.append(float(p[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), p[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        try:
            os.remove(self.path)
        except OSError:
            pass
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------------
# CPU baseline: the reference's own post-processor (the ONLY place bench.py executes oracle/)
# ------------------------------------------------------------------------------------------------
def load_cpu_impl():
    from oracle import aqo
    ref = aqo.load_ref()
    return ref if ref is not None else aqo.load_port()


def cpu_step(impl, t, queries):
    """one pass of the ten queries through hashtable_push / ht_postproc / per-group aggregates;
    returns seconds inside the postproc bracket (aqo_time_groupby_reduce)"""
    from aquery2_b200 import h2o
    total = 0.0
    for q in queries:
        keys, aggs, _ = h2o.QUERIES[q]
        specs = [(a[0],) + tuple(t[c] for c in a[1:]) for a in aggs]
        sec, _ = impl.time_groupby_reduce([t[c] for c in keys], specs)
        total += sec
    return total


def cpu_baseline(args, queries, rows):
    from aquery2_b200 import h2o, synth
    impl = load_cpu_impl()
    n = int(rows)
    t = synth.h2o_table(n, args.k, columns=h2o.columns_of(queries))
    sec = cpu_step(impl, t, queries)
    return {"value": len(queries) * n / sec, "unit": UNIT, "cores": 1,
            "kind": "reference" if impl.kind == "reference" else "port",
            "sample": f"G1_{n:.0e}_{args.k:.0e}_0_0 {queries[0]}-{queries[-1]} once ({sec:.1f} s), single thread "
                      f"(the reference path has no threads); g++ -O3 -march=x86-64-v3 (no clang in the image)",
            "host_cores": os.cpu_count()}


def run_reference(args, queries):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from aquery2_b200 import h2o, synth
    impl = load_cpu_impl()
    total_steps = max(args.steps + args.warmup, 1)
    n = int(min(args.cpu_rows, max(1e6, 1.2e8 / total_steps)))      # ~10 s per step at 1e7 rows
    t = synth.h2o_table(n, args.k, columns=h2o.columns_of(queries))
    for _ in range(args.warmup):
        cpu_step(impl, t, queries)
    secs = [cpu_step(impl, t, queries) for _ in range(args.steps)]
    tot = sum(secs)
    value = args.steps * len(queries) * n / tot
    sample = f"G1_{n:.0e}_{args.k:.0e}_0_0 q1-q10 per step, single thread (reference path is single-threaded)"
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tot / args.steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "int64/f64",
            "data": "synthetic", "config": {"workload": f"h2o G1_{args.rows:.0e}_{args.k:.0e}_0_0 {','.join(queries)}",
                                            "sample_rows": n},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": 1,
                             "kind": "reference" if impl.kind == "reference" else "port", "sample": sample,
                             "host_cores": os.cpu_count()},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
# the B200 arm
# ------------------------------------------------------------------------------------------------
def run_b200(args, queries):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    from aquery2_b200 import aqgpu as aq, h2o
    aq.init(local_rank)                      # raises without the CUDA library / device: no fallback
    comm = None
    if world > 1:
        from aquery2_b200 import dist
        comm = dist.Comm(local_rank)         # torch.distributed over NCCL, one process per GPU

    n_total = int(args.rows)
    lo, hi = rank * n_total // world, (rank + 1) * n_total // world
    cols = h2o.columns_of(queries)
    table = h2o.device_table(aq, n_total, args.k, lo, hi, cols)
    aq.sync()
    if comm is None:
        plan = h2o.Plan(aq, table, queries, n_total, args.k, device_out=True, use_bounds=not args.no_bounds)
        runner = plan
    else:
        plan = None
        runner = dist.DistPlan(comm, table, queries, n_total, args.k, lo, use_bounds=not args.no_bounds)

    def barrier():
        if comm is not None:
            comm.barrier()
        aq.sync()

    for _ in range(args.warmup):
        runner.step()
    # ---- timed region: exactly K steps, CUDA events on the library stream, barrier + sync around
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    barrier()
    launches0 = aq.launch_count()
    aq.timer_begin()
    for _ in range(args.steps):
        ngs = runner.step()
    ms = aq.timer_end()
    barrier()
    launches = aq.launch_count() - launches0
    clocks = sampler.stop() if rank == 0 else None
    if comm is not None:
        ms = comm.max_float(ms)
        launches = int(comm.sum_float(float(launches)))
    ms_per_step = ms / args.steps
    value = len(queries) * n_total / (ms_per_step * 1e-3)
    exchange = None
    if comm is not None:
        ngs = runner.total_groups(ngs)
        exchange = {q: {"strategy": runner.dq[q].strategy, "bytes_in_per_rank": runner.dq[q].exchanged_bytes} for q in queries}

    # ---- per-kernel / per-query breakdown (separate pass, CUDA events around every launch)
    aq.profile_enable(True)
    aq.profile_report()
    prof_steps = 2
    per_query_ms = {}
    for _ in range(prof_steps):
        for q in queries:
            aq.profile_tag(q)
            aq.timer_begin()
            runner.run(q)
            per_query_ms[q] = per_query_ms.get(q, 0.0) + aq.timer_end() / prof_steps
    aq.profile_tag("")
    prof = aq.profile_report()
    aq.profile_enable(False)
    peak, peak_src = load_peaks()
    nrows_local = hi - lo
    top = max(prof.items(), key=lambda kv: kv[1][1]) if prof else None
    roofline = None
    kernels = {}
    for name, (calls, tot_ms) in sorted(prof.items(), key=lambda kv: -kv[1][1])[:80]:
        kernels[name] = {"calls": calls, "ms_per_launch": tot_ms / calls}
    # DRAM bytes per row of the kernels captured with `ncu --set full` (dram__bytes_read.sum + dram__bytes_write.sum
    # per launch / rows of that launch; profiles/r01/ncu_full_hc_scatter.csv), scaled to this launch's rows
    NCU_DRAM_BYTES_PER_ROW = {"k_hc_scatter": (28.133142e9 + 31.992585e9) / 1e9}
    if top:
        name, (calls, tot_ms) = top
        q = name.split(":")[0]
        alg = h2o.ALG_BYTES.get(q, 0) * nrows_local
        avg_ms = tot_ms / calls
        achieved = alg / (avg_ms * 1e-3) / 1e9
        roofline = {"bound": "hbm", "kernel": name, "achieved": achieved, "peak": peak, "unit": "GB/s",
                    "frac": achieved / peak,
                    "traffic": next((v * nrows_local for k, v in NCU_DRAM_BYTES_PER_ROW.items() if k in name), None),
                    "peak_source": peak_src,
                    "algorithmic_bytes_per_launch": alg, "ms_per_launch": avg_ms,
                    "share_of_step": tot_ms / prof_steps / max(sum(per_query_ms.values()), 1e-9)}
    per_query = {q: {"ms": per_query_ms[q], "rows_per_s": n_total / (per_query_ms[q] * 1e-3),
                     "hbm_frac": h2o.ALG_BYTES[q] * nrows_local / (per_query_ms[q] * 1e-3) / 1e9 / peak,
                     "groups": int(ngs[q])} for q in queries}

    # ---- e2e: the same step through the C ABI with HOST buffers (pinned), H2D + D2H inside
    e2e = None
    if not args.no_e2e and comm is None:
        e2e = run_e2e(args, aq, h2o, table, plan, queries, n_total)
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu = cpu_baseline(args, queries, args.cpu_rows)

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "int64/f64", "data": "synthetic",
                "config": {"workload": f"h2o G1_{n_total:.0e}_{args.k:.0e}_0_0 {','.join(queries)}",
                           "rows_total": n_total, "rows_per_gpu": nrows_local, "queries": len(queries),
                           "l2": "inputs (>= 8 B/row x rows) far larger than the 126 MB L2; no flush needed",
                           "parallelism": "1 GPU" if world == 1 else f"{world} GPUs: row-range shards, local partial aggregation, NCCL merge"},
                "clocks": clocks, "e2e": e2e, "gpu_launches": launches // max(args.steps, 1) * args.steps,
                "roofline": roofline, "cpu_baseline": cpu, "per_query": per_query, "kernels": kernels, "exchange": exchange}
        print(json.dumps(line), flush=True)
    if comm is not None:
        comm.close()


def run_e2e(args, aq, h2o, table, plan, queries, n_total):
    """Host-resident columns in pinned memory -> aqgpu_groupby_agg stages them itself (H2D), results
    land in pinned host columns (D2H): the call a reference user makes with MonetDB-owned buffers."""
    pinned, host = [], {}
    for c, d in table.items():
        pa = aq.PinnedArray(d.n, d.code)
        if d.nbytes:
            aq.check(aq.lib().aqgpu_col_download(ctypes.c_void_p(pa.ptr), ctypes.c_void_p(d.ptr), ctypes.c_size_t(d.nbytes)))
        pinned.append(pa)
        host[c] = pa.array
    del plan
    for d in table.values():
        d.free()                               # the e2e arm starts with nothing resident in HBM

    def host_alloc(code, n):
        pa = aq.PinnedArray(n, code)
        pinned.append(pa)
        return pa.array

    hplan = h2o.Plan(aq, host, queries, n_total, args.k, device_out=False, host_alloc=host_alloc,
                     use_bounds=not args.no_bounds)
    hplan.step()                               # warm-up
    aq.sync()
    t0 = time.perf_counter()
    aq.timer_begin()
    for _ in range(args.e2e_steps):
        ngs = hplan.step()
    ms = aq.timer_end()
    wall = time.perf_counter() - t0
    bi, bo = hplan.io_bytes(ngs)
    out = {"value": len(queries) * n_total / (ms * 1e-3 / args.e2e_steps), "unit": UNIT,
           "h2d_bytes_per_step": bi, "d2h_bytes_per_step": bo, "steps": args.e2e_steps,
           "ms_per_step": ms / args.e2e_steps, "wall_s_per_step": wall / args.e2e_steps,
           "host_memory": "pinned (aqgpu_host_alloc)"}
    for pa in pinned:
        pa.free()
    return out


def main():
    args = parse_args()
    queries = [q.strip() for q in args.queries.split(",") if q.strip()]
    if args.impl == "reference":
        run_reference(args, queries)
    else:
        run_b200(args, queries)


if __name__ == "__main__":
    main()
