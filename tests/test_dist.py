"""Multi-rank orchestration (aquery2_b200/dist.py) on CPU: world_size 2 and 3 under gloo with the
oracle-backed engine (tests/cpu_engine.py), both exchange strategies, checked against the oracle run on
the whole table.  The GPU engine itself is covered by test_gpu_parity.py::test_dist_*."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

QUERIES = [
    (["id1"], [("sum", "v1"), ("avg", "v3"), ("count",)]),
    (["id1", "id2"], [("stddev", "v3"), ("corr", "v1", "v2"), ("var", "v1")]),
    (["id3"], [("max", "v1"), ("min", "v2"), ("first", "v3"), ("second", "v3"), ("last", "v1")]),
    (["id6"], [("sum", "v1"), ("sum", "v3"), ("second", "v1")]),
]


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, n, k, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
    from aquery2_b200 import aqgpu as aq, dist, synth
    from cpu_engine import CpuEngine
    from oracle import aqo
    from util import assert_same
    comm = dist.Comm(0, backend="gloo")
    eng = CpuEngine(comm, aqo.load_port())
    lo, hi = rank * n // world, (rank + 1) * n // world
    if rank == world - 1 and world == 3:
        lo = hi = n                                  # an empty shard
    elif world == 3 and rank == 1:
        hi = n
    shard = synth.h2o_table(n, k, lo, hi, seed=21)
    full = synth.h2o_table(n, k, seed=21)
    codes = {c: aq.NP2CODE[a.dtype] for c, a in shard.items()}
    port_o = eng.o
    for keys, aggs in QUERIES:
        ng, gid, off, perm = port_o.groupby([full[c] for c in keys])
        first = perm[off[1:] - 1]
        want_keys = [full[c][first] for c in keys]
        want = [port_o.reduce(a[0], full[a[1]] if len(a) > 1 else None, off, perm, v2=full[a[2]] if len(a) > 2 else None, raw128=True)
                for a in aggs]
        # all-gather of partials, hash-partitioned all-to-all of partials, all-to-all of the rows themselves
        for limit, hint in ((10**9, 0), (0, 0), (0, n)):
            q = dist.DistQuery(comm, eng, shard, codes, keys, aggs, lo, max_groups=max(hi - lo, 1), hint=hint, gather_limit=limit)
            ngm, mkeys, res, fg = q.run()
            assert q.strategy == ("allgather" if limit else ("alltoall-rows" if hint else "alltoall"))
            cols = [np.asarray(x[:ngm]) for x in mkeys + res] + [np.asarray(fg[:ngm])]
            if limit == 0:                           # bring the owners' groups together, order by first global row
                counts = comm.all_counts(ngm)
                cols = [eng.from_bytes(comm.allgather_bytes(eng.bytes_of(c, ngm), counts, c.dtype.itemsize * (2 if c.ndim == 2 else 1)),
                                       sum(counts), aq.INT128 if c.ndim == 2 else aq.NP2CODE[c.dtype]) for c in cols]
                order = np.argsort(cols[-1], kind="stable")
                cols = [c[order] for c in cols]
            assert len(cols[-1]) == ng, (keys, limit, len(cols[-1]), ng)
            assert_same(cols[-1].astype(np.uint32), first.astype(np.uint32), "first global row")
            for i, c in enumerate(keys):
                assert_same(cols[i], want_keys[i], f"key {c}")
            for j, a in enumerate(aggs):
                assert_same(cols[len(keys) + j], want[j], f"{keys} {a} limit={limit}")
    comm.barrier()
    comm.close()
    open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")


@pytest.mark.parametrize("world", [2, 3])
def test_dist_query_gloo(world, tmp_path):
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_worker, args=(world, port, 60_000, 10, str(tmp_path)), nprocs=world, join=True)
    assert all(os.path.exists(tmp_path / f"ok{r}") for r in range(world))


def test_decompose_deduplicates():
    from aquery2_b200 import aqgpu as aq, dist
    p, m, f = dist.decompose([("avg", "a"), ("sum", "a"), ("count",), ("stddev", "a")], {"a": aq.INT32})
    assert [x[0] for x in p] == ["sum", "count", "sumsq"]            # one partial sum / count shared by all four
    assert len(f) == 4
    with pytest.raises(ValueError):
        dist.decompose([("median", "a")], {"a": aq.INT32})
