"""multi-GPU parity check, launched with torchrun (one process per GPU):
    torchrun --nproc-per-node N scripts/dist_check.py [rows] [k]
Every rank holds a row shard of the synthetic h2o table; the h2o queries run through the library's sharded entry
point (aqgpu_groupby_run_dist; DistQuery for the shapes it declines), the PRODUCT assembles the global result
(ShardPlan.gather_ordered) and rank 0 compares it with the CPU oracle on the whole table: bit-exact keys, first
rows, order and integer aggregates, fp64 within 1e-12.  Test infrastructure (tests/test_dist_gpu.py)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
os.environ.setdefault("AQGPU_MC_MIN_ROWS", "4096")

import numpy as np  # noqa: E402


def main():
    n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 2_000_000
    k = int(sys.argv[2]) if len(sys.argv) > 2 else 10
    queries = sys.argv[3].split(",") if len(sys.argv) > 3 else None
    from aquery2_b200 import aqgpu as aq, dist, h2o, synth
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    aq.init(local_rank)
    comm = dist.Comm(local_rank)
    queries = queries or list(h2o.ORDER)
    cols = h2o.columns_of(queries)
    full = synth.h2o_table(n, k, seed=42, columns=cols)
    ndup = int(os.environ.get("AQ_CHECK_DUPS", "0"))
    if ndup:                                   # rows that repeat another row's whole key tuple (across shards too): q10's rare path
        rng = np.random.default_rng(7)
        src, dst = rng.integers(0, n, ndup), rng.integers(0, n, ndup)
        for c in ("id1", "id2", "id3", "id4", "id5", "id6"):
            if c in full:
                full[c][dst] = full[c][src]
    lo, hi = comm.rank * n // comm.world, (comm.rank + 1) * n // comm.world
    table = {c: aq.DeviceColumn.from_numpy(np.ascontiguousarray(a[lo:hi])) for c, a in full.items()}
    plan = dist.ShardPlan(comm, table, queries, n, k, lo)
    bad = []
    port = None
    if comm.rank == 0:
        from oracle import aqo
        from util import assert_same
        port = aqo.load_port()
    for rep in range(2):
        for q in queries:
            plan.run(q)
            keys_g, aggs_g, first_g = plan.gather_ordered(q)
            if comm.rank != 0 or rep == 0:
                continue
            keys, aggs, _ = h2o.QUERIES[q]
            ng, gid, off, perm = port.groupby([full[c] for c in keys])
            first = perm[off[1:] - 1].astype(np.uint32)
            try:
                assert len(first_g) == ng, f"{len(first_g)} groups, oracle {ng}"
                assert_same(first_g.astype(np.uint32), first, "first_row")
                for i, c in enumerate(keys):
                    assert_same(keys_g[i], full[c][first], f"key {c}")
                for a, agg in enumerate(aggs):
                    want = port.reduce(agg[0], full[agg[1]] if len(agg) > 1 else None, off, perm,
                                       v2=full[agg[2]] if len(agg) > 2 else None, raw128=True)
                    assert_same(aggs_g[a], want, f"{q} {agg}")
                print(f"[dist_check] {q}: ok ({ng} groups, {plan.exchange_report()[q]['strategy']})", flush=True)
            except AssertionError as e:
                bad.append(q)
                print(f"[dist_check] {q}: FAILED {e}", flush=True)
    plan.close()
    comm.close()
    if comm.rank == 0:
        print("[dist_check] " + ("ALL OK" if not bad else f"FAILED: {bad}"), flush=True)
        sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
