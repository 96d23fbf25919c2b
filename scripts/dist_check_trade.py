"""multi-GPU parity check of the windowed (order-sensitive) path, launched with torchrun (one process per GPU):
    torchrun --nproc-per-node N scripts/dist_check_trade.py [rows] [nsym]
Every rank holds a row shard of the synthetic trade table; DistWindowPlan repartitions the rows by symbol owner
(aqgpu_repartition_rows) and runs the windowed queries locally.  Checked on every rank, against the WHOLE table:
  * the received columns equal table[column][global row ids], the ids ascend, every row arrived exactly once and
    every symbol lives on exactly one rank;
  * the GPU's windowed results equal the CPU oracle's on the received rows (bit-exact integers, fp64 <= 1e-12).
Test infrastructure (tests/test_dist_gpu.py)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402


def main():
    n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 400_000
    nsym = int(float(sys.argv[2])) if len(sys.argv) > 2 else 3000
    from aquery2_b200 import aqgpu as aq, dist, synth, trade
    import bench_trade
    from oracle import aqo
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    aq.init(local_rank)
    comm = dist.Comm(local_rank)
    port = aqo.load_port()
    queries = list(trade.ORDER)
    cols = trade.columns_of(queries)
    full = synth.trade_table(n, nsym, columns=cols)
    lo, hi = comm.rank * n // comm.world, (comm.rank + 1) * n // comm.world
    table = {c: aq.DeviceColumn.from_numpy(np.ascontiguousarray(a[lo:hi])) for c, a in full.items()}

    class Args:
        pass
    args = Args(); args.nsym = nsym
    plan = dist.DistWindowPlan(comm, table, queries, args, n, lo)
    bad = []
    for q in queries:
        try:
            ng = plan.run(q)
            t = {c: d.numpy() for c, d in plan.local[q].items()}
            rows = t["_row"].astype(np.int64)
            assert np.all(rows[1:] > rows[:-1]), "global row ids do not ascend"
            for c in trade.COLUMNS[q]:
                assert np.array_equal(t[c], full[c][rows]), f"received column {c} differs from the table at the received rows"
            parts = [None] * comm.world
            comm.dist.all_gather_object(parts, (rows, np.unique(t["sym"])))
            allrows = np.sort(np.concatenate([p[0] for p in parts]))
            assert np.array_equal(allrows, np.arange(n)), "rows lost or duplicated in the exchange"
            syms = np.concatenate([p[1] for p in parts])
            assert len(np.unique(syms)) == len(syms), "a symbol lives on two ranks"
            _, ng_o, off, order, want = bench_trade._cpu_query(port, t, q)
            assert ng == ng_o, f"{ng} groups, oracle {ng_o}"
            bench_trade._compare(aq, plan.results[q], want, q)
            print(f"[dist_check_trade] rank {comm.rank} {q}: ok ({ng} groups, {len(rows)} rows here)", flush=True)
        except AssertionError as e:
            bad.append(q)
            print(f"[dist_check_trade] rank {comm.rank} {q}: FAILED {e}", flush=True)
    nbad = int(comm.sum_float(float(len(bad))))
    plan.close()
    comm.close()
    if comm.rank == 0:
        print("[dist_check_trade] " + ("ALL OK" if not nbad else f"FAILED on {nbad} (rank, query) pairs"), flush=True)
    sys.exit(1 if nbad else 0)


if __name__ == "__main__":
    main()
