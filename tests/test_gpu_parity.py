"""Parity of the CUDA path (through the C ABI, include/aqgpu.h) with the CPU oracle and with the
fixtures produced by the reference's own headers.  Bit-exact for keys, group order, counts, row lists
and integer aggregates; FP64_RTOL (1e-12) for fp64 sums / averages / windowed means."""
import numpy as np
import pytest

from aquery2_b200 import synth
from oracle.gen_golden import H2O_QUERIES, SCAN_CASES, trade_values
from util import assert_same

pytestmark = pytest.mark.gpu


def run_query(gpu, t, keys, aggs, **kw):
    specs = []
    for agg in aggs:
        specs.append((agg[0],) + tuple(t[c] for c in agg[1:]))
    return gpu.groupby_agg([t[c] for c in keys], specs, **kw)


def check_query_vs_oracle(gpu, oracle, t, keys, aggs, **kw):
    res = run_query(gpu, t, keys, aggs, **kw)
    ng, gid, off, perm = oracle.groupby([t[c] for c in keys])
    assert res.ngroups == ng
    first = perm[off[1:] - 1] if ng else np.zeros(0, np.uint32)     # descending lists: last entry = first row
    assert_same(res.first_row, first.astype(np.uint32), "first_row")
    for k, c in enumerate(keys):
        assert_same(res.keys[k], t[c][first], f"key {c}")
    for a, agg in enumerate(aggs):
        v = t[agg[1]] if len(agg) > 1 else None
        v2 = t[agg[2]] if len(agg) > 2 else None
        assert_same(res.aggs[a], oracle.reduce(agg[0], v, off, perm, v2=v2, raw128=True), f"{keys} {agg}")
    return res


# ---- golden fixtures (reference-built) ---------------------------------------------------------
def test_h2o_golden(gpu, golden):
    g = golden("h2o_small.npz")
    t = synth.h2o_table(int(g["n"]), int(g["k"]), seed=int(g["seed"]))
    for q, (keys, aggs) in H2O_QUERIES.items():
        res = run_query(gpu, t, keys, aggs)
        assert res.ngroups == int(g[f"{q}_ngroups"]), q
        for a, agg in enumerate(aggs):
            assert_same(res.aggs[a], g[f"{q}_agg{a}"], f"{q} {agg}")
        h = gpu.Groups([t[c] for c in keys])
        assert h.ngroups == int(g[f"{q}_ngroups"])
        assert_same(h.gid(), g[f"{q}_gid"], f"{q} gid")
        off, perm = h.csr()
        assert_same(off, g[f"{q}_offsets"], f"{q} offsets")
        assert_same(perm, g[f"{q}_perm"], f"{q} perm")
        h.free()


def test_trade_golden(gpu, golden):
    g = golden("trade_small.npz")
    t = synth.trade_table(int(g["n"]), int(g["nsym"]), seed=int(g["seed"]))
    vals = trade_values(t)
    h = gpu.Groups([t["sym"]])
    assert h.ngroups == int(g["ngroups"])
    assert_same(h.gid(), g["gid"])
    off, perm = h.csr()
    assert_same(off, g["offsets"]); assert_same(perm, g["perm"])
    for vn, v in vals.items():
        for op in ("sum", "avg", "min", "max", "first", "last", "var", "stddev"):
            if vn == "big_i64" and op in ("var", "stddev"):
                continue      # arr[i]*arr[i] overflows int64 in the reference (UB), not a parity case
            assert_same(h.reduce(op, v), g[f"reduce_{op}_{vn}"], f"{op} {vn}")
        for op, w in SCAN_CASES:
            assert_same(h.scan(op, v, w=w), g[f"scan_{op}_{w}_{vn}"], f"{op}({w}) {vn}", scale=float(np.abs(v).max()))
    assert_same(h.reduce("count"), g["reduce_count"])
    assert_same(h.reduce("corr", vals["price_i32"], vals["qty_i32"]), g["reduce_corr_i32"])
    h.free()


def test_keys_golden(gpu, golden):
    g = golden("keys_small.npz")
    for name, keys in (("k64", [g["k64"]]), ("k64x2", [g["k64"], g["k64b"]]), ("kneg", [g["kneg"]])):
        h = gpu.Groups(keys)
        assert h.ngroups == int(g[f"{name}_ngroups"])
        assert_same(h.gid(), g[f"{name}_gid"])
        off, perm = h.csr()
        assert_same(off, g[f"{name}_offsets"]); assert_same(perm, g[f"{name}_perm"])
        h.free()


# ---- oracle on the same seeded inputs -----------------------------------------------------------
@pytest.mark.parametrize("q", list(H2O_QUERIES))
@pytest.mark.parametrize("n,k", [(100_000, 100), (1_000_000, 100), (300_000, 3)])
def test_h2o_vs_oracle(gpu, port, q, n, k):
    keys, aggs = H2O_QUERIES[q]
    cols = set(keys) | {c for a in aggs for c in a[1:]}
    t = synth.h2o_table(n, k, seed=42, columns=sorted(cols))
    check_query_vs_oracle(gpu, port, t, keys, aggs)


def test_h2o_device_columns_and_bounds(gpu, port):
    """device-resident columns (the HBM column store), key bounds (ColRef::stats), groups_hint"""
    n, k = 500_000, 100
    t = synth.h2o_table(n, k, seed=42, columns=["id1", "id2", "id4", "v1", "v3"])
    d = {c: gpu.DeviceColumn.from_numpy(a) for c, a in t.items()}
    ng, gid, off, perm = port.groupby([t["id1"], t["id2"]])
    want = port.reduce("sum", t["v1"], off, perm, raw128=True)
    for bounds in (None, [(1, k), (1, k)], [(1, k // 2), (1, k)]):      # last: wrong bounds -> silent retry
        res = gpu.groupby_agg([d["id1"], d["id2"]], [("sum", d["v1"]), ("avg", d["v3"])], key_bounds=bounds, groups_hint=k * k)
        assert res.ngroups == ng
        assert_same(res.aggs[0], want)
        assert_same(res.aggs[1], port.reduce("avg", t["v3"], off, perm))
    res = gpu.groupby_agg([d["id1"], d["id2"]], [("sum", d["v1"])], device_out=True)
    assert_same(res.aggs[0].numpy(res.ngroups), want)


def test_synth_device_matches_host(gpu):
    n, k = 100_003, 100
    for name, kind, kk in (("id1", 0, k), ("id3", 0, n // k), ("v1", 0, 5), ("v2", 0, 15)):
        dev = gpu.synth_column(kind, synth.stream_key(42, synth.STREAM[name]), n, kk)
        assert_same(dev.numpy(), synth.h2o_column(name, n, k))
    dev = gpu.synth_column(2, synth.stream_key(42, synth.STREAM["v3"]), n, 100_000_000, code=gpu.FLOAT)
    assert_same(dev.numpy(), synth.h2o_column("v3", n, k))
    # a shard [lo, hi) of the global table
    lo, hi = 70_000, 90_001
    dev = gpu.synth_column(0, synth.stream_key(42, synth.STREAM["id6"]), hi - lo, n // k, row_lo=lo)
    assert_same(dev.numpy(), synth.h2o_column("id6", n, k, lo, hi))
    for name, kind in (("sym", 1), ("quantity", 5)):
        kk, aux = (1000, 0) if name == "sym" else (9901, 100)
        dev = gpu.synth_column(kind, synth.stream_key(42, synth.STREAM[name]), n, kk, aux_k=aux)
        assert_same(dev.numpy(), synth.trade_column(name, n, 1000))
    dev = gpu.synth_column(4, synth.stream_key(42, synth.STREAM["price"]), n, 1000,
                           key1=synth.stream_key(42, synth.STREAM["sym"]), key2=synth.stream_key(42, synth.STREAM["price_base"]))
    assert_same(dev.numpy(), synth.trade_column("price", n, 1000))


@pytest.mark.parametrize("n,nsym", [(200_000, 1000), (50_000, 20_000), (5000, 1)])
def test_windowed_vs_oracle(gpu, port, n, nsym):
    t = synth.trade_table(n, nsym, seed=9)
    vals = trade_values(t)
    h = gpu.Groups([t["sym"]])
    ng, gid, off, perm = port.groupby([t["sym"]])
    assert h.ngroups == ng
    assert_same(h.gid(), gid)
    o2, p2 = h.csr()
    assert_same(o2, off); assert_same(p2, perm)
    for vn in ("price_i32", "price_f32", "big_i64", "qty_f64"):
        v = vals[vn]
        for op, w in SCAN_CASES + [("sumw", 100), ("avgw", 100), ("minw", 300), ("maxw", 10_000)]:
            assert_same(h.scan(op, v, w=w), port.scan(op, v, off, perm, w=w, raw128=True), f"{op}({w}) {vn}", scale=float(np.abs(v).max()))
    h.free()


def test_row_order_asc_and_order_by(gpu, port):
    """AQ_ROWS_ASC (legacy emplace_back order) and ASSUMING via order_by (common/groupby.py:54-59)."""
    n, nsym = 60_000, 300
    t = synth.trade_table(n, nsym, seed=5)
    rng = np.random.default_rng(0)
    tm = rng.permutation(n).astype(np.int32)          # unsorted time column
    h = gpu.Groups([t["sym"]], row_order=gpu.ROWS_ASC)
    ng, gid, off, perm = port.groupby([t["sym"]])
    o2, p2 = h.csr()
    assert_same(o2, off)
    asc = np.concatenate([perm[off[g]:off[g + 1]][::-1] for g in range(ng)]).astype(np.uint32)
    assert_same(p2, asc)
    h.order_by([tm], desc=[0])
    o3, p3 = h.csr()
    want = np.concatenate([np.sort(asc[off[g]:off[g + 1]], kind="stable")[np.argsort(tm[np.sort(asc[off[g]:off[g + 1]])], kind="stable")]
                           for g in range(ng)]).astype(np.uint32)
    assert_same(p3, want)
    got = h.scan("avgw", t["price"], w=5)
    assert_same(got, port.scan("avgw", t["price"], off, want, w=5))
    h.order_by([tm], desc=[1])
    _, p4 = h.csr()
    want_d = np.concatenate([want[off[g]:off[g + 1]][::-1] for g in range(ng)]).astype(np.uint32)
    assert_same(p4, want_d)
    h.free()


# ---- edge cases ----------------------------------------------------------------------------------
def test_edge_cases(gpu, port):
    # empty input
    res = gpu.groupby_agg([np.zeros(0, np.int32)], [("sum", np.zeros(0, np.int32))])
    assert res.ngroups == 0
    h = gpu.Groups([np.zeros(0, np.int32)])
    assert h.ngroups == 0 and h.csr()[0].tolist() == [0]
    # a single row, a single group, all-distinct keys, ragged tails (n % 4 != 0)
    for n, hi in ((1, 1), (2, 1), (3, 5), (5, 1), (1023, 7), (4097, 4097), (65537, 2)):
        rng = np.random.default_rng(n)
        key = rng.integers(0, hi, n).astype(np.int32) if hi < n else rng.permutation(n).astype(np.int32)
        v = rng.integers(-1000, 1000, n).astype(np.int32)
        t = {"k": key, "v": v, "f": v.astype(np.float32) / 3}
        check_query_vs_oracle(gpu, port, t, ["k"], [("sum", "v"), ("min", "v"), ("max", "v"), ("count",), ("avg", "f"),
                                                   ("first", "v"), ("last", "f")])
    # all-negative fp max: the reference's numeric_limits<T>::min() identity (aggregations.h:73)
    t = {"k": np.array([1, 1, 2], np.int32), "d": np.array([-1.0, -2.0, -3.0]), "f": np.array([-1.0, -2.0, -3.0], np.float32)}
    res = check_query_vs_oracle(gpu, port, t, ["k"], [("max", "d"), ("max", "f"), ("min", "d")])
    assert res.aggs[0].tolist() == [np.finfo(np.float64).tiny] * 2
    # extreme int32 values and sums beyond int32
    t = {"k": np.array([3, 3, 3, 4], np.int32), "v": np.array([2**31 - 1, 2**31 - 1, 2**31 - 1, -2**31], np.int32)}
    res = check_query_vs_oracle(gpu, port, t, ["k"], [("sum", "v"), ("avg", "v"), ("min", "v"), ("max", "v")])
    assert res.aggs[0][:, 0].tolist() == [3 * (2**31 - 1), -2**31]
    # capacity error
    with pytest.raises(gpu.AqgpuError, match="capacity"):
        gpu.groupby_agg([np.arange(100, dtype=np.int32)], [("count",)], max_groups=10)


def test_high_cardinality_and_collisions(gpu, port):
    """G ~ N (q10 regime), keys that differ only in high words, 64-bit keys."""
    n = 400_000
    rng = np.random.default_rng(3)
    a = rng.integers(0, 1 << 30, n).astype(np.int32)
    b = (rng.integers(0, 4, n) << 28).astype(np.int32)
    v = rng.integers(0, 100, n).astype(np.int32)
    t = {"a": a, "b": b, "v": v, "k64": (a.astype(np.int64) << 20) ^ b.astype(np.int64)}
    check_query_vs_oracle(gpu, port, t, ["a", "b"], [("sum", "v"), ("count",)])
    check_query_vs_oracle(gpu, port, t, ["k64"], [("sum", "v")])
    check_query_vs_oracle(gpu, port, t, ["b", "k64", "a"], [("max", "v")])


def test_wide_exact_keys_fused_emit(gpu, port):
    """q10 shape: six bounded key columns whose packed image needs 76 bits -> wide exact slots
    (k_probe_wide, 128-bit CAS) + the fused look-back rank/emit; right bounds, wrong bounds (silent
    retry without them), duplicates-heavy and all-distinct inputs, several tile counts."""
    for n, k in ((300_007, 100), (5_000, 3), (1023, 100), (1025, 2), (1_300_001, 100)):   # last: groupby_hc.cu path
        t = synth.h2o_table(n, k, seed=13, columns=["id1", "id2", "id3", "id4", "id5", "id6", "v1", "v3"])
        t["id3"] = (t["id3"].astype(np.int64) * 1_000 % 10_000_000 + 1).astype(np.int32)   # 24-bit fields
        t["id6"] = (t["id6"].astype(np.int64) * 777 % 10_000_000 + 1).astype(np.int32)
        keys = ["id1", "id2", "id3", "id4", "id5", "id6"]
        good = [(1, k), (1, k), (1, 10_000_000), (1, k), (1, k), (1, 10_000_000)]
        wrong = [(1, k), (1, k), (1, 10), (1, k), (1, k), (1, 10_000_000)]
        for bounds in (good, wrong, None):
            check_query_vs_oracle(gpu, port, t, keys, [("sum", "v3"), ("count",)], groups_hint=n, key_bounds=bounds)
        check_query_vs_oracle(gpu, port, t, keys[:4], [("sum", "v1"), ("avg", "v3"), ("max", "v1"), ("first", "v3"), ("second", "v1")],
                              groups_hint=n, key_bounds=good[:4])
    # groupby_hc.cu with a skewed key: one group holds a third of the rows -> its partition overflows ->
    # silent fallback to the global table; and a moderately duplicated one that stays on the path
    n = 1_500_000
    rng = np.random.default_rng(23)
    t = {c: rng.integers(0, 1 << 20, n).astype(np.int32) for c in ("a", "b", "c")}
    t["v"] = rng.integers(-50, 50, n).astype(np.int32); t["f"] = rng.random(n).astype(np.float32)
    b3 = [(0, (1 << 20) - 1)] * 3
    dup = rng.integers(0, n, n // 4)
    for c in ("a", "b", "c"):
        t[c][dup] = t[c][(dup * 7) % 1000]                    # ~1000 hot keys share a quarter of the rows
    check_query_vs_oracle(gpu, port, t, ["a", "b", "c"], [("sum", "v"), ("avg", "f"), ("count",), ("min", "v"), ("first", "f")],
                          key_bounds=b3, groups_hint=n)
    for c in ("a", "b", "c"):
        t[c][::3] = 5
    check_query_vs_oracle(gpu, port, t, ["a", "b", "c"], [("sum", "v"), ("count",)], key_bounds=b3, groups_hint=n)
    # few groups over many rows: most rows are NOT first rows; full-range fields fill the 64-bit word
    n = 200_000
    rng = np.random.default_rng(17)
    t = {"a": rng.integers(-2**31, 2**31 - 1, 50).astype(np.int32)[rng.integers(0, 50, n)],
         "b": rng.integers(0, 2**31 - 1, 40).astype(np.int32)[rng.integers(0, 40, n)],
         "c": rng.integers(0, 1 << 20, n).astype(np.int32) >> 12, "v": rng.integers(-9, 9, n).astype(np.int32)}
    b3 = [(-2**31, 2**31 - 1), (0, 2**31 - 1), (0, 255)]
    check_query_vs_oracle(gpu, port, t, ["a", "b", "c"], [("sum", "v"), ("count",), ("min", "v")], key_bounds=b3)
    check_query_vs_oracle(gpu, port, t, ["c", "a", "b"], [("sum", "v"), ("last", "v")], key_bounds=[b3[2], b3[0], b3[1]], groups_hint=1000)


# ---- partitioned path (groupby_part.cu): groups beyond one SM's shared memory ---------------------
ALL_AGGS = [("sum", "v1"), ("avg", "v3"), ("min", "v2"), ("max", "v3"), ("count",), ("first", "v3"), ("last", "v1"),
            ("second", "v3")]


@pytest.mark.parametrize("n,k", [(1_000_000, 20), (300_003, 3)])
def test_partitioned_one_key(gpu, port, n, k):
    """id3 has n/k distinct values: with a groups_hint the library hash-partitions the rows
    (HASH32 tables without bounds, directly indexed tables with bounds); all three must agree."""
    t = synth.h2o_table(n, k, seed=7, columns=["id3", "id6", "v1", "v2", "v3"])
    g = n // k
    for bounds in (None, [(1, g)], [(1, g // 2)]):            # last: wrong bounds -> silent retry without them
        for aggs in (ALL_AGGS[:2], ALL_AGGS[2:5], ALL_AGGS[5:], [("sum", "v1"), ("sum", "v2"), ("sum", "v3")],
                     [("var", "v3"), ("stddev", "v1")]):
            check_query_vs_oracle(gpu, port, t, ["id3"], aggs, groups_hint=g, key_bounds=bounds)


def test_partitioned_two_keys(gpu, port):
    """two key columns: bit-packed tag with bounds (q6 / q9 shape), 64-bit shared-memory keys without"""
    n, k = 600_000, 200
    t = synth.h2o_table(n, k, seed=11, columns=["id2", "id4", "v1", "v2", "v3"])
    rng = np.random.default_rng(1)
    t["wide"] = (rng.integers(0, 30_000, n).astype(np.int64) * 70_001 - 1_000_000_000).astype(np.int32)
    for keys, bounds, hint in ((["id2", "id4"], [(1, k), (1, k)], k * k), (["id2", "id4"], None, k * k),
                               (["wide", "id4"], None, 200_000)):
        check_query_vs_oracle(gpu, port, t, keys, [("corr", "v1", "v2"), ("stddev", "v3")], groups_hint=hint, key_bounds=bounds)
        check_query_vs_oracle(gpu, port, t, keys, [("sum", "v1"), ("max", "v2"), ("count",), ("first", "v3")], groups_hint=hint, key_bounds=bounds)


def test_partitioned_hint_too_small_and_special_keys(gpu, port):
    """a hint far below the real group count (tables overflow into the global table) and keys equal
    to the EMPTY markers of the shared-memory tables"""
    n = 500_000
    rng = np.random.default_rng(5)
    key = rng.integers(0, 200_000, n).astype(np.int32)
    key[::1000] = -1                                          # 0xffffffff == EMPTY32
    t = {"k": key, "k2": np.where(key == -1, -1, key % 7).astype(np.int32), "v": rng.integers(-5, 50, n).astype(np.int32),
         "f": rng.random(n).astype(np.float32)}
    check_query_vs_oracle(gpu, port, t, ["k"], [("sum", "v"), ("avg", "f"), ("count",)], groups_hint=20_000)
    check_query_vs_oracle(gpu, port, t, ["k", "k2"], [("sum", "v"), ("min", "f")], groups_hint=20_000)
    check_query_vs_oracle(gpu, port, t, ["k", "k2"], [("sum", "v"), ("min", "f")], groups_hint=400_000)


def test_host_header_selftest(gpu):
    """emitter-shaped C++ (unchanged per-group loop, AQHashTable generation, whole-column form) compiled
    against the drop-in headers and run on the device: aquery2_b200/host/selftest.cpp"""
    import os
    import subprocess
    exe = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "aquery2_b200", "host", "selftest")
    assert os.path.exists(exe), "run __graft_entry__.build() first"
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "selftest ok" in r.stdout


# ---- multi-GPU path on one device: the GPU engine of aquery2_b200/dist.py ---------------------------
@pytest.fixture(scope="module")
def comm1(gpu):
    import os
    import socket
    from aquery2_b200 import dist
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK="0", WORLD_SIZE="1")
    c = dist.Comm(0)
    yield c
    gpu.set_stream(None)
    c.close()


def test_dist_engine_single_rank(gpu, port, comm1):
    """local partial aggregation -> exchange (world size 1) -> merge -> finalize on the device, both
    exchange strategies, against the oracle on the same table"""
    from aquery2_b200 import dist
    n, k = 300_000, 50
    t = synth.h2o_table(n, k, seed=33)
    d = {c: gpu.DeviceColumn.from_numpy(a) for c, a in t.items()}
    codes = {c: x.code for c, x in d.items()}
    eng = dist.GpuEngine(comm1)
    for keys, aggs in ((["id1"], [("sum", "v1"), ("avg", "v3"), ("count",)]),
                       (["id1", "id2"], [("stddev", "v3"), ("corr", "v1", "v2"), ("var", "v1")]),
                       (["id3"], [("max", "v1"), ("min", "v2"), ("first", "v3"), ("second", "v3"), ("last", "v1")]),
                       (["id6"], [("sum", "v1"), ("sum", "v3"), ("second", "v1")])):
        ng, gid, off, perm = port.groupby([t[c] for c in keys])
        first = perm[off[1:] - 1]
        for limit, rows in ((10**9, False), (0, False), (0, True)):
            q = dist.DistQuery(comm1, eng, d, codes, keys, aggs, 0, max_groups=n, hint=n if rows else 0, gather_limit=limit)
            q.by_rows = rows                     # world size 1 never chooses the row exchange by itself
            ngm, mkeys, res, fg = q.run()
            assert ngm == ng and q.strategy == ("allgather" if limit else ("alltoall-rows" if rows else "alltoall"))
            assert_same(fg.numpy(ngm).astype(np.uint32), first.astype(np.uint32), "first global row")
            for i, c in enumerate(keys):
                assert_same(mkeys[i].numpy(ngm), t[c][first], f"key {c}")
            for j, a in enumerate(aggs):
                want = port.reduce(a[0], t[a[1]] if len(a) > 1 else None, off, perm, v2=t[a[2]] if len(a) > 2 else None, raw128=True)
                assert_same(res[j].numpy(ngm), want, f"{keys} {a} limit={limit}")


def test_dist_partition_by_owner(gpu, comm1):
    """rows grouped by owner rank with their order kept (the send buffers of the all-to-all)"""
    from aquery2_b200 import dist
    n, parts = 200_001, 4
    rng = np.random.default_rng(2)
    key = rng.integers(0, 50_000, n).astype(np.int32)
    val = rng.random(n)
    idx = np.arange(n, dtype=np.int64)
    eng = dist.GpuEngine(comm1)
    cols = [gpu.DeviceColumn.from_numpy(a) for a in (key, val, idx)]
    counts, out = eng.partition([cols[0]], n, parts, cols)
    assert sum(counts) == n and len(counts) == parts and min(counts) > 0
    k2, v2, i2 = (c.numpy(n) for c in out)
    assert_same(k2, key[i2]); assert_same(v2, val[i2])
    bounds = np.concatenate([[0], np.cumsum(counts)])
    owners = []
    for r in range(parts):
        blk = i2[bounds[r]:bounds[r + 1]]
        assert np.all(np.diff(blk) > 0)                           # stable inside an owner
        owners.append(set(np.unique(k2[bounds[r]:bounds[r + 1]]).tolist()))
    for a in range(parts):
        for b in range(a + 1, parts):
            assert not (owners[a] & owners[b])                    # a key has exactly one owner


# ---- round 2: planned paths (plan.cu -> lc.cu / hcfp.cu), no bounds, no hints ---------------------------
def _lc_table(n, seed, ranges, neg=False):
    rng = np.random.default_rng(seed)
    t = {f"k{i}": (rng.integers(0, r, n) + (-7 if neg else 3)).astype(np.int32) for i, r in enumerate(ranges)}
    t["v1"] = rng.integers(-50 if neg else 1, 60, n).astype(np.int32)
    t["v2"] = rng.integers(1, 16, n).astype(np.int32)
    t["f"] = np.round(rng.random(n) * 100, 6).astype(np.float32)
    t["g"] = (rng.standard_normal(n) * 1000).astype(np.float32)        # signed floats, wide exponent spread
    return t


@pytest.mark.parametrize("n,ranges", [(4096, (7,)), (100_003, (100,)), (300_001, (100, 100)), (250_000, (13, 9, 11)),
                                      (400_000, (150, 150)), (2_500_000, (100, 37))])
def test_low_cardinality_direct_table(gpu, port, n, ranges):
    """lc.cu: bulk-copy ring + direct shared-memory table, cluster split for fat tables, ragged tails"""
    t = _lc_table(n, n, ranges, neg=(n % 2 == 1))
    keys = [f"k{i}" for i in range(len(ranges))]
    for aggs in ([("sum", "v1")], [("avg", "v1"), ("avg", "v2"), ("avg", "f")], [("stddev", "f")], [("corr", "v1", "v2")],
                 [("sum", "g"), ("min", "g"), ("max", "g"), ("min", "v1"), ("max", "v1"), ("count",), ("var", "v1")],
                 [("sum", "v1"), ("first", "f"), ("last", "v2")]):          # first / last: generic path
        check_query_vs_oracle(gpu, port, t, keys, aggs)


def test_low_cardinality_speculation_fails(gpu, port):
    """a key the sample cannot see and a float the fixed point cannot hold: the scan notices, the query re-runs"""
    n = 3_000_000
    t = _lc_table(n, 5, (50,))
    t["k0"][2_999_999] = 1_000_000                 # outside every sampled chunk's range
    check_query_vs_oracle(gpu, port, t, ["k0"], [("sum", "v1"), ("avg", "f")])
    t = _lc_table(n, 6, (50,))
    t["f"][1_234_567] = np.float32(1e-30)          # 2^-100: below the fixed-point unit chosen from the sample
    check_query_vs_oracle(gpu, port, t, ["k0"], [("sum", "f"), ("count",)])
    t["f"][7] = np.float32(np.inf)
    res = run_query(gpu, t, ["k0"], [("sum", "f")])
    assert np.isinf(res.aggs[0]).sum() == 1
    # one-word integer cells are chosen from the sampled magnitudes: an unsampled outlier must not wrap a cell
    t = _lc_table(n, 7, (50,))
    t["v1"][2_999_998] = 2_000_000_000
    t["v1"][1_999_998] = -2_000_000_000
    check_query_vs_oracle(gpu, port, t, ["k0"], [("sum", "v1"), ("avg", "v1"), ("max", "v1")])   # (no squares: int32 overflow is UB in the reference)
    # sum only (no count requested): empty slots are told apart by the seen marker, ragged tail rows included
    t = _lc_table(n - 3, 9, (40, 3))
    check_query_vs_oracle(gpu, port, t, ["k0", "k1"], [("sum", "v1")])
    check_query_vs_oracle(gpu, port, t, ["k1", "k0"], [("max", "v1"), ("min", "f")])


def test_large_integer_sums_low_cardinality(gpu, port):
    """values near the int32 limits and big products: spill words / two-word cells stay exact"""
    n = 200_000
    rng = np.random.default_rng(8)
    t = {"k": rng.integers(0, 5, n).astype(np.int32), "v": rng.integers(-2**31, 2**31 - 1, n).astype(np.int32),
         "w": rng.integers(-40_000, 40_000, n).astype(np.int32)}
    check_query_vs_oracle(gpu, port, t, ["k"], [("sum", "v"), ("avg", "v"), ("min", "v"), ("max", "v")])
    check_query_vs_oracle(gpu, port, t, ["k"], [("var", "w"), ("sum", "w")])


@pytest.mark.parametrize("n", [2_200_000, 5_000_000])
def test_groups_like_rows_fingerprint_path(gpu, port, n):
    """hcfp.cu: all-distinct key tuples, a sprinkling of duplicates (records), 64-bit keys, one level and two"""
    rng = np.random.default_rng(n)
    t = {c: rng.integers(0, 1 << 30, n).astype(np.int32) for c in ("a", "b", "c")}
    t["v"] = rng.integers(-50, 50, n).astype(np.int32); t["f"] = np.round(rng.random(n) * 100, 6).astype(np.float32)
    t["k64"] = (t["a"].astype(np.int64) << 31) | t["b"].astype(np.int64)
    aggs = [("sum", "f"), ("count",)]
    check_query_vs_oracle(gpu, port, t, ["a", "b", "c"], aggs)
    dup = rng.integers(0, n, n // 200)                                    # ~0.5 % of the rows repeat another row's key
    for c in ("a", "b", "c"):
        t[c][dup] = t[c][(dup * 13 + 5) % n]
    t["k64"] = (t["a"].astype(np.int64) << 31) | t["b"].astype(np.int64)
    check_query_vs_oracle(gpu, port, t, ["a", "b", "c"], aggs + [("min", "v"), ("first", "f"), ("sum", "v"), ("second", "v")])
    check_query_vs_oracle(gpu, port, t, ["k64", "c"], [("sum", "v"), ("avg", "f"), ("max", "v")])


def test_result_set_api(gpu, port):
    """aqgpu_groupby_run: the library allocates the result columns; no capacities, bounds or hints"""
    n, k = 700_000, 50
    t = synth.h2o_table(n, k, seed=3, columns=["id1", "id3", "v1", "v3"])
    d = {c: gpu.DeviceColumn.from_numpy(a) for c, a in t.items()}
    for keys in (["id1"], ["id3"], ["id1", "id3"]):
        q = gpu.AutoQuery([d[c] for c in keys], [("sum", d["v1"]), ("avg", d["v3"])], want_first_row=True)
        for _ in range(2):
            ng = q.run()
            ngo, gid, off, perm = port.groupby([t[c] for c in keys])
            assert ng == ngo
            first = perm[off[1:] - 1]
            assert_same(q.fr.numpy(ng), first.astype(np.uint32))
            for i, c in enumerate(keys):
                assert_same(q.kouts[i].numpy(ng), t[c][first])
            assert_same(q.outs[0].numpy(ng), port.reduce("sum", t["v1"], off, perm, raw128=True))
            assert_same(q.outs[1].numpy(ng), port.reduce("avg", t["v3"], off, perm))
        q.free()
    q = gpu.AutoQuery([np.zeros(0, np.int32)], [("count",)])
    assert q.run() == 0


# ---- BASELINE.json sizes against the C port ----------------------------------------------------------------
def _check_h2o(gpu, port, n, k, queries):
    from aquery2_b200 import h2o
    cols = h2o.columns_of(queries)
    t = synth.h2o_table(n, k, seed=42, columns=cols)
    d = {c: gpu.DeviceColumn.from_numpy(a) for c, a in t.items()}
    for q in queries:
        keys, aggs, _ = h2o.QUERIES[q]
        res = gpu.groupby_agg([d[c] for c in keys], [(a[0],) + tuple(d[c] for c in a[1:]) for a in aggs])
        ng, gid, off, perm = port.groupby([t[c] for c in keys])
        assert res.ngroups == ng, q
        first = perm[off[1:] - 1]
        assert_same(res.first_row, first.astype(np.uint32), f"{q} first_row")
        for i, c in enumerate(keys):
            assert_same(res.keys[i], t[c][first], f"{q} key {c}")
        for a, agg in enumerate(aggs):
            want = port.reduce(agg[0], t[agg[1]] if len(agg) > 1 else None, off, perm, v2=t[agg[2]] if len(agg) > 2 else None, raw128=True)
            assert_same(res.aggs[a], want, f"{q} {agg}")
    for c in d.values():
        c.free()


def test_h2o_config1_1e7_all_queries(gpu, port):
    """BASELINE.json configs[0]: G1_1e7_1e2, q1-q10, the library on its own (no bounds, no hints)"""
    from aquery2_b200 import h2o
    _check_h2o(gpu, port, 10_000_000, 100, h2o.ORDER)


def test_h2o_config2_1e8_size_dependent_paths(gpu, port):
    """BASELINE.json configs[1] size: G1_1e8_1e2 -- one query per size-dependent path: direct table in a
    cluster (q9), partitioned with 1e6 groups (q3), fingerprint-partitioned with two levels (q10)"""
    _check_h2o(gpu, port, 100_000_000, 100, ["q9", "q3", "q10"])


def test_generic_paths_at_1e8_without_planner(gpu, port, monkeypatch):
    """the round-1 paths at BASELINE size, forced by switching the planner off: hash-partitioned aggregation
    with bounds + hint (q5 shape), and the global-table path with nothing (q3 shape)"""
    monkeypatch.setenv("AQGPU_DISABLE", "plan")
    n, k = 100_000_000, 100
    t = synth.h2o_table(n, k, seed=42, columns=["id6", "v1", "v3"])
    d = {c: gpu.DeviceColumn.from_numpy(a) for c, a in t.items()}
    ng, gid, off, perm = port.groupby([t["id6"]])
    first = perm[off[1:] - 1]
    want = [port.reduce("sum", t["v1"], off, perm, raw128=True), port.reduce("sum", t["v3"], off, perm), port.reduce("count", None, off, perm)]
    for kw in (dict(key_bounds=[(1, n // k)], groups_hint=n // k), dict()):
        res = gpu.groupby_agg([d["id6"]], [("sum", d["v1"]), ("sum", d["v3"]), ("count",)], **kw)
        assert res.ngroups == ng
        assert_same(res.first_row, first.astype(np.uint32))
        for a in range(3):
            assert_same(res.aggs[a], want[a], f"agg {a} {kw}")


# ---- A4 / A10 helpers against numpy ------------------------------------------------------------------------
def test_gather_and_elementwise_helpers(gpu, port):
    n, nsym = 120_000, 700
    t = synth.trade_table(n, nsym, seed=4)
    h = gpu.Groups([t["sym"]])
    ng, gid, off, perm = port.groupby([t["sym"]])
    import ctypes
    L = gpu.lib()
    for v in (t["price"], t["price"].astype(np.float64) / 3, t["quantity"].astype(np.int64) * 1_000_003):
        out = np.empty_like(v)
        gpu.check(L.aqgpu_group_gather(h.h, gpu.NP2CODE[v.dtype], v.ctypes.data_as(ctypes.c_void_p), out.ctypes.data_as(ctypes.c_void_p)))
        assert_same(out, v[perm], "gather col[vecs[g]]")                  # server/table.h:184-189
    h.free()
    assert gpu.col_minmax(t["price"]) == (int(t["price"].min()), int(t["price"].max()))
    big = t["quantity"].astype(np.int64) * -3_000_000_007
    assert gpu.col_minmax(big) == (int(big.min()), int(big.max()))
    a, b = t["price"], t["quantity"]
    for x, y in ((a, b), (a.astype(np.float32), b.astype(np.float32)), (a.astype(np.float64), b.astype(np.float64)), (a.astype(np.int64), b.astype(np.int64))):
        o = np.empty_like(x)
        gpu.vec_sub(gpu.DeviceColumn.from_numpy(x), gpu.DeviceColumn.from_numpy(y), dc := gpu.DeviceColumn(n, gpu.NP2CODE[x.dtype]), n, gpu.NP2CODE[x.dtype])
        assert_same(dc.numpy(), x - y, "vec_sub")
    x = np.linspace(-1, 1, 1001)
    dx, do = gpu.DeviceColumn.from_numpy(x), gpu.DeviceColumn(1001, gpu.DOUBLE)
    gpu.vec_pow(dx, 2.0, do, 1001)
    assert_same(do.numpy(), np.power(x, 2.0), "vec_pow")
    s = np.arange(1, 1001, dtype=np.float64) * 1.5
    cnt = np.arange(1, 1001, dtype=np.uint32)
    o = np.empty(1000)
    gpu.check(L.aqgpu_avg_finalize(gpu.DOUBLE, s.ctypes.data_as(ctypes.c_void_p), cnt.ctypes.data_as(ctypes.c_void_p),
                                   o.ctypes.data_as(ctypes.c_void_p), ctypes.c_uint32(1000)))
    assert_same(o, s / cnt, "avg_finalize")


# ------------------------------------------------------------------------------------------------
# several expressions in one call (aqgpu_group_window): fused sort + window kernel vs the oracle
# ------------------------------------------------------------------------------------------------
def _window_exprs(t):
    p, q, f = t["price"], t["quantity"], t["price"].astype(np.float32) * np.float32(0.37)
    return [("avgw", p, 5), ("sumw", q, 5), ("deltas", p), ("minw", p, 5), ("mins", p, 0, True, "max"),
            ("maxs", f), ("sums", q), ("avgs", f)], f


def _check_window(gpu, port, h, t, off, perm, ng):
    exprs, f = _window_exprs(t)
    for lo in (0, 5):                                         # at most 3 value columns and 2 running expressions per fused call
        part = exprs[lo:lo + 5] if lo == 0 else exprs[5:]
        got = h.window(part)
        for e, g in zip(part, got):
            op, v = e[0], e[1]
            w = e[2] if len(e) > 2 else 0
            want = port.scan(op, v, off, perm, w=w, raw128=True)
            if len(e) > 3 and e[3]:
                want = (v[perm] - want).astype(v.dtype)
            if len(e) > 4 and e[4]:
                ident = np.iinfo(np.int32).min if e[4] == "max" else np.iinfo(np.int32).max
                red = np.full(ng, ident, dtype=v.dtype)
                fn = np.maximum if e[4] == "max" else np.minimum
                fn.at(red, np.repeat(np.arange(ng), np.diff(off)), want)
                want = red
            assert_same(g, want, f"window {op}({w})", scale=float(np.abs(v).max()))


@pytest.mark.parametrize("n,nsym", [(300_000, 1000), (200_000, 70_000), (50_000, 3), (9_000, 1), (1_000_003, 5000)])
def test_fused_window_vs_oracle(gpu, port, n, nsym):
    """window.cu: one or two radix passes (nsym below / above 2048), groups spanning many tiles (nsym 1 and 3), ragged tails"""
    t = synth.trade_table(n, nsym, seed=11)
    h = gpu.Groups([t["sym"]])
    ng, gid, off, perm = port.groupby([t["sym"]])
    assert h.ngroups == ng
    _check_window(gpu, port, h, t, off, perm, ng)
    h.free()
    h = gpu.Groups([t["sym"]], row_order=gpu.ROWS_ASC)       # time-ascending row lists (presorted table)
    asc = np.concatenate([perm[off[g]:off[g + 1]][::-1] for g in range(ng)]).astype(np.uint32)
    _check_window(gpu, port, h, t, off, asc, ng)
    h.free()


def test_fused_window_after_order_by(gpu, port):
    """unsorted table + ASSUMING ASC time: the handle's own row lists feed the same kernel"""
    n, nsym = 150_000, 800
    t = synth.trade_table(n, nsym, seed=12)
    tm = np.random.default_rng(3).permutation(n).astype(np.int32)
    h = gpu.Groups([t["sym"]], row_order=gpu.ROWS_ASC)
    ng, gid, off, perm = port.groupby([t["sym"]])
    h.order_by([tm], desc=[0])
    _, p3 = h.csr()
    _check_window(gpu, port, h, t, off, p3, ng)
    h.free()


def test_window_unfused_fallback(gpu, port):
    """expressions the fused kernel does not serve (64-bit values, long windows, ratios) take the per-op path"""
    n, nsym = 40_000, 50
    t = synth.trade_table(n, nsym, seed=13)
    big = t["price"].astype(np.int64) * 3_000_000_007
    h = gpu.Groups([t["sym"]])
    ng, gid, off, perm = port.groupby([t["sym"]])
    got = h.window([("sumw", big, 7), ("minw", t["price"], 500), ("mins", big, 0, True, "max"), ("ratiow", t["price"], 2)])
    assert_same(got[0], port.scan("sumw", big, off, perm, w=7, raw128=True), "sumw i64")
    assert_same(got[1], port.scan("minw", t["price"], off, perm, w=500), "minw 500")
    e = big[perm] - port.scan("mins", big, off, perm)
    want = np.array([e[off[g]:off[g + 1]].max() for g in range(ng)], dtype=np.int64)
    assert_same(got[2], want, "max(v - mins(v)) i64")
    assert_same(got[3], port.scan("ratiow", t["price"], off, perm, w=2), "ratiow")
    h.free()


def test_direct_group_build(gpu, port):
    """groups.cu build_direct: one bounded 32-bit key, >= 2^20 rows; a key outside the sampled range falls back"""
    n = 1_300_007
    rng = np.random.default_rng(21)
    k = (rng.integers(0, 50_000, n) - 20_000).astype(np.int32)            # negative keys, ~50k groups
    v = rng.integers(-9, 9, n).astype(np.int32)
    for outlier in (False, True):
        if outlier:
            k[n - 5] = 2_000_000_000
        h = gpu.Groups([k])
        ng, gid, off, perm = port.groupby([k])
        assert h.ngroups == ng
        assert_same(h.gid(), gid)
        assert_same(h.keys(0), k[perm[off[:-1]]] if False else h.keys(0))   # keys checked through first rows below
        assert_same(h.first_row(), np.array([perm[off[g + 1] - 1] for g in range(ng)], dtype=np.uint32))
        o2, p2 = h.csr()
        assert_same(o2, off); assert_same(p2, perm)
        assert_same(h.keys(0), k[h.first_row()])
        assert_same(h.reduce("sum", v), port.reduce("sum", v, off, perm, raw128=True))
        assert_same(h.scan("sums", v), port.scan("sums", v, off, perm, raw128=True))
        h.free()


# ---- range-partitioned direct tables (mc.cu): groups beyond one SM, bounded keys ---------------------------
def _took_mc(gpu, fn):
    """runs fn() with per-kernel profiling on; returns (result, True if k_mc_agg ran)"""
    gpu.profile_enable(True); gpu.profile_report()
    try:
        res = fn()
    finally:
        rep = gpu.profile_report(); gpu.profile_enable(False)
    return res, any("k_mc_agg" in k for k in rep)


MC_AGG_SETS = ([("sum", "v1"), ("avg", "v3")], [("sum", "v1"), ("sum", "v2"), ("sum", "v3")], [("max", "v1"), ("min", "v2")],
               [("first", "v3"), ("second", "v3")], [("count",)], [("min", "v3"), ("max", "v3"), ("last", "v1")],
               [("var", "v3"), ("stddev", "v1")], [("corr", "v1", "v2")])


@pytest.mark.parametrize("n,k", [(1_000_000, 8), (777_777, 5)])
def test_range_partitioned_one_key(gpu, port, monkeypatch, n, k):
    """mc.cu on its own statistics (sampled bounds), every cell kind, ragged row counts; and with wrong bounds"""
    monkeypatch.setenv("AQGPU_MC_MIN_ROWS", "4096")
    t = synth.h2o_table(n, k, seed=3, columns=["id3", "v1", "v2", "v3"])
    for aggs in MC_AGG_SETS:
        _, took = _took_mc(gpu, lambda: check_query_vs_oracle(gpu, port, t, ["id3"], aggs))
        assert took, f"mc path not taken for {aggs}"
    check_query_vs_oracle(gpu, port, t, ["id3"], MC_AGG_SETS[0], key_bounds=[(1, n // k // 2)])       # wrong bounds: silent retry


def test_range_partitioned_two_keys_skew_and_row_order(gpu, port, monkeypatch):
    monkeypatch.setenv("AQGPU_MC_MIN_ROWS", "4096")
    n = 1_500_000
    rng = np.random.default_rng(17)
    t = {"a": rng.integers(-200, 200, n).astype(np.int32), "b": rng.integers(1000, 1300, n).astype(np.int32),
         "v1": rng.integers(-50, 50, n).astype(np.int32), "v2": rng.integers(1, 16, n).astype(np.int32),
         "v3": np.round(rng.random(n) * 100, 6).astype(np.float32)}
    for aggs in (MC_AGG_SETS[0], MC_AGG_SETS[2], MC_AGG_SETS[3], MC_AGG_SETS[7]):
        _, took = _took_mc(gpu, lambda: check_query_vs_oracle(gpu, port, t, ["a", "b"], aggs))
        assert took
    # ascending row lists (PerfectHashTable / legacy order): first = value at the smallest row, second at the next one
    res, took = _took_mc(gpu, lambda: run_query(gpu, t, ["b", "a"], MC_AGG_SETS[3], row_order=gpu.ROWS_ASC))
    assert took
    ng, gid, off, perm = port.groupby([t["b"], t["a"]])
    assert res.ngroups == ng
    r1 = perm[off[1:] - 1]
    two = (off[1:] - off[:-1]) >= 2
    r2 = perm[np.maximum(off[1:].astype(np.int64) - 2, 0)]
    assert_same(res.aggs[0], t["v3"][r1], "asc first")
    assert_same(res.aggs[1], np.where(two, t["v3"][r2], np.float32(0)).astype(np.float32), "asc second")
    # a hot key: one partition holds a third of the rows (exact per-partition positions, no capacity speculation)
    hot = rng.random(n) < 0.33
    t["a"][hot] = 7; t["b"][hot] = 1111
    _, took = _took_mc(gpu, lambda: check_query_vs_oracle(gpu, port, t, ["a", "b"], [("sum", "v1"), ("avg", "v3"), ("max", "v2")]))
    assert took


def test_range_partitioned_speculation_fails(gpu, port, monkeypatch):
    """unsampled outliers: a key outside the sampled range, a float below the fixed-point unit, integers that would
    wrap a 32-bit cell; and a sparse key domain (declined)"""
    monkeypatch.setenv("AQGPU_MC_MIN_ROWS", "4096")
    n = 3_000_000
    rng = np.random.default_rng(23)
    def table():
        return {"k": rng.integers(0, 300_000, n).astype(np.int32), "v": rng.integers(0, 10, n).astype(np.int32),
                "f": np.round(rng.random(n) * 100, 6).astype(np.float32)}
    t = table(); t["k"][2_999_999] = 5_000_000
    check_query_vs_oracle(gpu, port, t, ["k"], [("sum", "v"), ("avg", "f")])
    t = table(); t["f"][1_234_567] = np.float32(1e-30)
    check_query_vs_oracle(gpu, port, t, ["k"], [("sum", "f"), ("count",)])
    t = table(); t["v"][2_999_998] = 2_000_000_000; t["v"][1_999_998] = -2_000_000_000
    t["k"][2_999_998] = t["k"][5]; t["k"][1_999_998] = t["k"][5]
    check_query_vs_oracle(gpu, port, t, ["k"], [("sum", "v"), ("max", "v")])
    t = table(); t["k"] = (t["k"].astype(np.int64) * 5003 - 700_000_000).astype(np.int32)
    _, took = _took_mc(gpu, lambda: check_query_vs_oracle(gpu, port, t, ["k"], [("sum", "v"), ("avg", "f")]))
    assert not took


# ---- A9 second wave: running / windowed variance and standard deviation -------------------------------------
@pytest.mark.parametrize("n,nsym", [(200_000, 1000), (30_000, 7)])
def test_moment_scans_vs_oracle(gpu, port, n, nsym):
    """vars / stddevs (aggregations.h:350-381, 419-425) and varw / stddevw against the oracle; the GPU evaluates
    exact prefix sums of x and x*x instead of Welford's recurrence: fp64 results within 1e-12 of the value scale^2"""
    t = synth.trade_table(n, nsym, seed=13)
    vals = trade_values(t)
    h = gpu.Groups([t["sym"]])
    ng, gid, off, perm = port.groupby([t["sym"]])
    for vn in ("price_i32", "price_f32", "signed_i32", "qty_f64"):
        v = vals[vn]
        sc = float(np.abs(v.astype(np.float64)).max())
        for op, w in (("vars", 0), ("stddevs", 0), ("varw", 1), ("varw", 3), ("varw", 64), ("stddevw", 3), ("stddevw", 50)):
            got, want = h.scan(op, v, w=w), port.scan(op, v, off, perm, w=w)
            if op.startswith("stddev"):            # the tolerance is stated on the variance: sqrt amplifies rounding noise near 0
                got, want = got * got, want * want
            assert_same(got, want, f"{op}({w}) {vn}", scale=sc * sc)
    h.free()


def test_group_median_vs_numpy(gpu, port):
    """h2o q6's median(v3) by id4, id5 (benchmark/h2o/groupby.sql:11-12).  The reference library has no median: the
    check is numpy's, per group, on the oracle's row lists -- NOT a reference result."""
    n, k = 300_000, 40
    t = synth.h2o_table(n, k, seed=21, columns=["id4", "id5", "v1", "v3"])
    h = gpu.Groups([t["id4"], t["id5"]])
    ng, gid, off, perm = port.groupby([t["id4"], t["id5"]])
    assert h.ngroups == ng
    for col in ("v3", "v1"):
        got = h.median(t[col])
        want = np.array([np.median(t[col][perm[off[g]:off[g + 1]]].astype(np.float64)) for g in range(ng)])
        assert_same(got, want, f"median({col})")
    o2, p2 = h.csr()
    assert_same(p2, perm)                      # the handle's own row lists are untouched
    h.free()


@pytest.mark.parametrize("n,nsym,desc", [(300_000, 400, False), (120_000, 9, True), (50_000, 30_000, False)])
def test_window_ordered_inside_the_call(gpu, port, n, nsym, desc):
    """aqgpu_group_window_ordered (ASSUMING order inside the call: key carried through the scatter, groups sorted in shared
    memory; 120000 / 9 = groups longer than 8192 rows -> the two-call route) against order_by + window on a second handle,
    and against the oracle over numpy's lexsort"""
    t = synth.trade_table(n, nsym, seed=31, columns=["sym", "price", "quantity", "time_shuffled"])
    exprs = [("avgw", t["price"], 5), ("sumw", t["quantity"], 5), ("deltas", t["price"]), ("minw", t["price"], 5), ("mins", t["price"], 0, True, "max")]
    for key in (t["time_shuffled"], (t["time_shuffled"] % 1000).astype(np.int32)):          # unique keys / many ties (stability)
        h1 = gpu.Groups([t["sym"]], row_order=gpu.ROWS_ASC)
        got = h1.window(exprs, order_by=key, desc=desc)
        h2 = gpu.Groups([t["sym"]], row_order=gpu.ROWS_ASC)
        h2.order_by([key], desc=[int(desc)])
        want = h2.window(exprs)
        for a, b in zip(got, want):
            assert_same(np.asarray(a), np.asarray(b), "ordered window vs order_by + window")
        ng, gid, off, perm = port.groupby([t["sym"]])
        k64 = key.astype(np.int64)
        order = np.lexsort((np.arange(n), -k64 if desc else k64, gid)).astype(np.uint32)
        ref = port.scan("avgw", t["price"], off, order, w=5)
        assert_same(np.asarray(got[0]), ref, "ordered window vs oracle", scale=float(np.abs(t["price"]).max()))
        h1.free(); h2.free()
