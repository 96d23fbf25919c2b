#!/usr/bin/env python
"""bench.py -- throughput (rows/s) of the B200 post-processor path on the reference's own workloads.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
                    [--workload h2o|trade] [--rows R] [--no-verify] [--hints]

--workload h2o (default; BASELINE.json configs[2]): a step = one pass of the hot path over the synthetic
table G1_{R}_1e2_0_0: the ten queries of the reference's benchmark/h2o/groupby.sql, each from "columns
resident in HBM" to "result columns resident in HBM" (the reference's postproc_time bracket,
server/server.cpp:303-305).  value = (queries x rows) / time, whole job.
--workload trade (configs[3]): ASSUMING-ordered windowed aggregates by sym over the trade table of
datagen.cpp:90-107 (see aquery2_b200/trade.py, bench_trade.py).

By default the library is told NOTHING the reference's path would not know: no key bounds, no group-count
hints, no output capacities (it samples the key columns itself inside the timed call and allocates the
result columns once it knows the group count).  --hints passes the generator's exact statistics instead
(the round-1 behaviour), for comparison only.

After the last timed step the results left in HBM are VERIFIED (bench_verify.py): independent group counts,
sum(count) = rows, first-appearance order, column-total identities, groups recomputed from scratch, and
the same ten queries compared with the reference-built CPU oracle on the cpu_baseline table.  A failed
check prints "verify": "FAILED: ..." and exits non-zero.

--impl reference times the reference's own CPU post-processor (oracle/_ref/libaqref.so = the reference
headers compiled here; else the C port) on a bounded sample of the same workload.
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

UNIT = "rows/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="h2o", choices=["h2o", "trade"])
    ap.add_argument("--rows", type=float, default=1e9, help="total rows of the table (all ranks)")
    ap.add_argument("--k", type=int, default=100, help="h2o K (id1,id2,id4,id5 cardinality)")
    ap.add_argument("--nsym", type=float, default=1e6, help="trade: number of symbols (groups)")
    ap.add_argument("--queries", default="")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-verify", action="store_true")
    ap.add_argument("--cpu-rows", type=float, default=1e7, help="rows of the CPU-baseline sample table")
    ap.add_argument("--hints", action="store_true", help="pass the generator's key bounds / group counts to the library")
    return ap.parse_args()


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def load_ncu_traffic():
    """{kernel name: {"bytes_per_row": .., "capture": file}} from an `ncu --set full` capture of THIS round
    (profiles/r02/ncu_traffic.json); absent kernels report traffic null."""
    p = os.path.join(ROOT, "profiles", "r02", "ncu_traffic.json")
    try:
        return json.load(open(p))
    except Exception:
        return {}


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


def workload_module(args):
    if args.workload == "trade":
        import bench_trade as wl
    else:
        import bench_h2o as wl
    return wl


# ------------------------------------------------------------------------------------------------
# CPU baseline: the reference's own post-processor (the ONLY place bench.py executes oracle/)
# ------------------------------------------------------------------------------------------------
def load_cpu_impl():
    from oracle import aqo
    ref = aqo.load_ref()
    return ref if ref is not None else aqo.load_port()


def cpu_kind(impl):
    return "reference" if impl.kind == "reference" else "port"


CPU_NOTE = ("single thread (the reference path has no threads); g++ -O3 -march=x86-64-v3 (the reference wants "
            "clang -Ofast, absent from the image)")


def run_reference(args, wl, queries):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    impl = load_cpu_impl()
    total_steps = max(args.steps + args.warmup, 1)
    n = int(min(args.cpu_rows, args.rows, max(1e6, 1.2e8 / total_steps)))      # ~10 s per step at 1e7 rows
    t = wl.host_table(args, n)
    for _ in range(args.warmup):
        wl.cpu_step(impl, t, queries)
    secs = [wl.cpu_step(impl, t, queries) for _ in range(args.steps)]
    tot = sum(secs)
    value = args.steps * len(queries) * n / tot
    sample = f"{wl.table_name(args, n)} {queries[0]}-{queries[-1]} per step, " + CPU_NOTE
    line = {"impl": "reference", "metric": wl.METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tot / args.steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": wl.DTYPE,
            "data": "synthetic",
            "config": {"workload": f"{wl.table_name(args, n)} {','.join(queries)} -- a {n:.0e}-row SAMPLE of the b200 arm's "
                                   f"{wl.table_name(args, int(args.rows))} (CPU time bound); group cardinalities shrink with the sample",
                       "sample_rows": n, "same_config": False},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": 1, "kind": cpu_kind(impl), "sample": sample,
                             "host_cores": os.cpu_count()},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
# the B200 arm
# ------------------------------------------------------------------------------------------------
def run_b200(args, wl, queries):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    from aquery2_b200 import aqgpu as aq
    aq.init(local_rank)                      # raises without the CUDA library / device: no fallback
    comm = None
    if world > 1:
        from aquery2_b200 import dist
        comm = dist.Comm(local_rank)         # torch.distributed over NCCL, one process per GPU

    n_total = int(args.rows)
    lo, hi = rank * n_total // world, (rank + 1) * n_total // world
    table = wl.device_table(aq, args, n_total, lo, hi, queries)
    aq.sync()
    if comm is None:
        runner = wl.Plan(aq, table, queries, args, n_total, hints=args.hints)
    else:
        runner = wl.dist_plan(comm, table, queries, args, n_total, lo, hints=args.hints)

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
    step_wall = []
    for _ in range(args.steps):
        t_s = time.perf_counter()
        ngs = runner.step()
        step_wall.append(round(1e3 * (time.perf_counter() - t_s), 2))      # host view of each step (diagnostic; the metric is the event time)
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
        ngs_total = runner.total_groups(ngs)
        exchange = runner.exchange_report()
    else:
        ngs_total = ngs

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
    # ---- verification of what the timed loop left in HBM (outside the timed region)
    verify = None
    if not args.no_verify:
        try:
            verify = wl.verify(aq, comm, runner, table, queries, args, n_total, ngs)
        except Exception as e:      # VerifyError or anything the checker tripped over: never report a number as verified
            import traceback
            traceback.print_exc()
            verify = {"status": f"FAILED: {type(e).__name__}: {e}"[:600]}

    peak, peak_src = load_peaks()
    nrows_local = hi - lo
    top = max(prof.items(), key=lambda kv: kv[1][1]) if prof else None
    roofline = None
    kernels = {}
    for name, (calls, tot_ms) in sorted(prof.items(), key=lambda kv: -kv[1][1])[:80]:
        kernels[name] = {"calls": calls, "ms_per_launch": tot_ms / calls}
    ncu = load_ncu_traffic()
    if top:
        name, (calls, tot_ms) = top
        q = name.split(":")[0]
        alg = wl.ALG_BYTES.get(q, 0) * nrows_local
        alg_note = "SURVEY 8(d): every needed input column read once x rows of the launch"
        if "emit" in name and q in ngs:      # the kernel that writes the result of a groups ~ rows query: SURVEY 8(d) counts q10's output too
            ob = wl.out_bytes(q, int(ngs[q]))
            if ob * 4 > alg:
                alg += ob
                alg_note += " + the result columns it writes (G ~ N: 24 B keys + 8 B sum + 4 B count per group)"
        avg_ms = tot_ms / calls
        achieved = alg / (avg_ms * 1e-3) / 1e9
        kname = name.split(":", 1)[-1]
        cap = next((v for k, v in ncu.items() if k in kname), None)
        roofline = {"bound": "hbm", "kernel": name, "achieved": achieved, "peak": peak, "unit": "GB/s",
                    "frac": achieved / peak,
                    "traffic": cap["bytes_per_row"] * nrows_local if cap and world == 1 else None,
                    "traffic_source": (f"ncu --set full capture {cap['capture']}, dram bytes/row x rows" if cap and world == 1
                                       else "no capture of this kernel at this configuration"),
                    "peak_source": peak_src,
                    "algorithmic_bytes_per_launch": alg, "algorithmic_bytes_note": alg_note,
                    "frac_input_bytes_only": wl.ALG_BYTES.get(q, 0) * nrows_local / (avg_ms * 1e-3) / 1e9 / peak, "ms_per_launch": avg_ms,
                    "share_of_step": tot_ms / prof_steps / max(sum(per_query_ms.values()), 1e-9)}
    per_query = {q: {"ms": per_query_ms[q], "rows_per_s": n_total / (per_query_ms[q] * 1e-3),
                     "hbm_frac": wl.ALG_BYTES[q] * nrows_local / (per_query_ms[q] * 1e-3) / 1e9 / peak,
                     "hbm_frac_in_out": (wl.ALG_BYTES[q] * nrows_local + wl.out_bytes(q, int(ngs[q]))) / (per_query_ms[q] * 1e-3) / 1e9 / peak,
                     "groups": int(ngs_total[q])} for q in queries}
    step_alg = sum(wl.ALG_BYTES[q] for q in queries) * nrows_local
    step_frac = step_alg / (ms_per_step * 1e-3) / 1e9 / peak

    # ---- e2e: host-resident (pinned) columns -> HBM column store -> queries -> results back in pinned host memory
    e2e = None
    if not args.no_e2e and comm is None:
        try:
            e2e = wl.run_e2e(aq, args, table, runner, queries, n_total)
        except Exception as ex:                    # the copy-stream pipeline failed: measure the serial path and say so
            first = f"{type(ex).__name__}: {ex}"[:300]
            os.environ["AQ_E2E_SERIAL"] = "1"
            try:
                e2e = wl.run_e2e(aq, args, table, runner, queries, n_total)
                e2e["pipeline_error"] = first
            except Exception as ex2:
                e2e = {"error": first, "serial_error": f"{type(ex2).__name__}: {ex2}"[:300]}
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        try:
            cpu = wl.cpu_baseline(aq, args, queries, load_cpu_impl(), CPU_NOTE, check=not args.no_verify)
            if verify is not None and "oracle_parity" in cpu:
                verify["oracle_parity"] = cpu.pop("oracle_parity")
                if not verify["oracle_parity"]["status"].startswith("ok"):
                    verify["status"] = "FAILED: " + verify["oracle_parity"]["status"]
        except Exception as e:
            import traceback
            traceback.print_exc()
            if verify is not None:
                verify["status"] = f"FAILED: cpu parity leg: {type(e).__name__}: {e}"[:600]

    if rank == 0:
        line = {"metric": wl.METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": wl.DTYPE, "data": "synthetic",
                "config": {"workload": f"{wl.table_name(args, n_total)} {','.join(queries)}",
                           "rows_total": n_total, "rows_per_gpu": nrows_local, "queries": len(queries),
                           "statistics_given_to_library": "generator bounds + group counts (--hints)" if args.hints
                                                          else "none: the library samples keys inside the timed call",
                           "l2": "inputs (>= 8 B/row x rows) far larger than the 126 MB L2; no flush needed",
                           "parallelism": "1 GPU" if world == 1 else f"{world} GPUs: row-range shards, local partial aggregation, NCCL merge"},
                "clocks": clocks, "e2e": e2e, "gpu_launches": launches // max(args.steps, 1) * args.steps,
                "verify": verify["status"] if verify else "skipped (--no-verify)", "verify_detail": verify,
                "step_ms_host": step_wall, "roofline": roofline, "step_hbm_frac": step_frac, "cpu_baseline": cpu, "per_query": per_query,
                "kernels": kernels, "exchange": exchange}
        print(json.dumps(line), flush=True)
    if comm is not None:
        if hasattr(runner, "close"):
            runner.close()
        comm.close()
    if verify is not None and not str(verify.get("status", "")).startswith(("ok", "skipped")):
        sys.exit(3)


def main():
    args = parse_args()
    wl = workload_module(args)
    queries = [q.strip() for q in args.queries.split(",") if q.strip()] or list(wl.ORDER)
    if args.impl == "reference":
        run_reference(args, wl, queries)
    else:
        run_b200(args, wl, queries)


if __name__ == "__main__":
    main()
