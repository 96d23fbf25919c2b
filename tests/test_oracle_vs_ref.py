"""Pins the plain-C restatement (kind "port") against the reference's own headers (kind "reference",
oracle/_ref/libaqref.so) on seeded random inputs that the golden fixtures do not cover: ragged group
sizes, w larger than the group, single-row groups, many keys."""
import numpy as np
import pytest

from aquery2_b200 import synth
from util import assert_same

REDUCE_OPS = ("sum", "avg", "min", "max", "first", "last", "var", "stddev")
SCAN_OPS = [("sums", 0), ("avgs", 0), ("mins", 0), ("maxs", 0), ("sumw", 2), ("sumw", 7), ("avgw", 2), ("avgw", 100),
            ("minw", 3), ("maxw", 4), ("minw", 1000), ("deltas", 0), ("prev", 0), ("next", 0), ("ratiow", 1), ("ratiow", 4)]


def test_kinds(port, ref):
    assert port.kind == "port" and ref.kind == "reference"


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_groupby_random(port, ref, seed):
    rng = np.random.default_rng(seed)
    n = int(rng.integers(1, 5000))
    for nkeys in (1, 2, 3, 6):
        keys = [rng.integers(-3, 4 + 2 * seed, n).astype(np.int32) for _ in range(nkeys)]
        a, b = port.groupby(keys), ref.groupby(keys)
        assert a[0] == b[0]
        for x, y in zip(a[1:], b[1:]):
            assert_same(x, y)


@pytest.mark.parametrize("dtype", [np.int32, np.int64, np.float32, np.float64])
def test_aggregates_random(port, ref, dtype):
    rng = np.random.default_rng(17)
    n = 3000
    # ragged: zipf-like group sizes incl. many singletons
    key = np.minimum(rng.zipf(1.3, n), 400).astype(np.int32)
    ng, gid, off, perm = ref.groupby([key])
    if np.issubdtype(dtype, np.integer):
        v = rng.integers(1, 1000, n).astype(dtype)          # ratiow divides by elements: keep them non-zero
        v[::7] *= -1
    else:
        v = (rng.random(n) * 200 - 100).astype(dtype)
    for op in REDUCE_OPS:
        assert_same(port.reduce(op, v, off, perm, raw128=True), ref.reduce(op, v, off, perm, raw128=True), op)
    for op, w in SCAN_OPS:
        assert_same(port.scan(op, v, off, perm, w=w, raw128=True), ref.scan(op, v, off, perm, w=w, raw128=True), f"{op}({w})", scale=float(np.abs(v).max()))
    assert_same(port.reduce("count", None, off, perm), ref.reduce("count", None, off, perm))


def test_corr_int(port, ref):
    t = synth.h2o_table(5000, 10, seed=3)
    ng, gid, off, perm = ref.groupby([t["id2"], t["id4"]])
    assert_same(port.reduce("corr", t["v1"], off, perm, v2=t["v2"]), ref.reduce("corr", t["v1"], off, perm, v2=t["v2"]))


def test_h2o_1e5(port, ref):
    t = synth.h2o_table(100_000, 100, seed=42)
    for keys in (["id1"], ["id3"], ["id1", "id2", "id3", "id4", "id5", "id6"]):
        a, b = port.groupby([t[c] for c in keys]), ref.groupby([t[c] for c in keys])
        assert a[0] == b[0]
        for x, y in zip(a[1:], b[1:]):
            assert_same(x, y)


def test_moment_scans_port_vs_reference_and_definition():
    """vars / stddevs: the C port against the reference's own Welford loops (aggregations.h:350-381, 419-425; the
    harness parks stdout because the reference printf()s every element).  varw / stddevw have no defined reference
    result (arr[i-w-1] at i == w, aggregations.h:311): the port states the windowed population variance, checked
    here against numpy."""
    from oracle import aqo
    ref = aqo.load_ref()
    if ref is None:
        pytest.skip("reference build not available")
    port = aqo.load_port()
    rng = np.random.default_rng(1)
    n = 3000
    k = rng.integers(0, 9, n).astype(np.int32)
    ng, gid, off, perm = port.groupby([k])
    for dt in (np.int32, np.int64, np.float32, np.float64):
        v = rng.integers(-500, 500, n).astype(dt) if np.issubdtype(dt, np.integer) else (rng.random(n) * 100).astype(dt)
        for op in ("vars", "stddevs"):
            a, b = port.scan(op, v, off, perm), ref.scan(op, v, off, perm)
            assert np.all(np.abs(a - b) <= 1e-12 * np.maximum(np.abs(b), 1.0)), (dt, op)
        for w in (1, 3, 50):
            got = port.scan("varw", v, off, perm, w=w)
            sd = port.scan("stddevw", v, off, perm, w=w)
            want = np.zeros(n)
            for g in range(ng):
                x = v[perm[off[g]:off[g + 1]]].astype(np.float64)
                for i in range(len(x)):
                    c = min(w, i + 1)
                    want[off[g] + i] = np.var(x[i + 1 - c:i + 1])
            assert np.all(np.abs(got - want) <= 1e-9 * np.maximum(np.abs(want), 1.0)), (dt, w)
            assert np.all(np.abs(sd - np.sqrt(want)) <= 1e-9 * np.maximum(np.sqrt(want), 1.0)), (dt, w)


@pytest.mark.parametrize("seed", [11, 12])
def test_groupby_special_keys_port_vs_reference(port, ref, seed):
    """string / date / time / timestamp keys: the restatement's equality rules against AQHashTable with
    the reference's own hashes (server/hasher.h:97-132) and operator== (types.h:299-309, libaquery.cpp:112-203)"""
    import keydata
    rng = np.random.default_rng(seed)
    n = int(rng.integers(2000, 6000))
    cols = {"str": keydata.strings(rng, n, 300, long_tail=True), "date": keydata.dates(rng, n, 900),
            "time": keydata.times(rng, n, 64), "timestamp": keydata.timestamps(rng, n, 64)}
    for name, c in cols.items():
        a, b = port.groupby([c]), ref.groupby([c])
        assert a[0] == b[0] and 1 < a[0] < n, name
        for x, y in zip(a[1:], b[1:]):
            assert_same(x, y, name)
    k2 = rng.integers(0, 4, n).astype(np.int32)
    a, b = port.groupby([cols["str"], k2]), ref.groupby([cols["str"], k2])
    assert a[0] == b[0]
    for x, y in zip(a[1:], b[1:]):
        assert_same(x, y, "(str, int32)")
    # the padding byte of time_t is not part of the key: same cells with other padding give the same groups
    t2 = cols["time"].copy(); t2["pad"] = 0
    assert_same(ref.groupby([t2])[1], ref.groupby([cols["time"]])[1], "time padding")
    # a single row, and all rows equal
    one = keydata.strings(rng, 1, 5)
    assert port.groupby([one])[0] == ref.groupby([one])[0] == 1


def test_vector_ops_port_vs_reference(port, ref):
    """element-wise operators, sqrt, truncate: the C restatement against the reference's own templates, result
    types included, for 16 type pairs x every operand form the reference can evaluate"""
    import itertools
    from oracle.gen_golden import VEC_DTYPES, VEC_OPS, vec_operand
    rng = np.random.default_rng(31)
    n, forms_checked = 3001, 0
    for (na, ta), (nb, tb) in itertools.product(VEC_DTYPES.items(), VEC_DTYPES.items()):
        a, b = vec_operand(rng, ta, n), vec_operand(rng, tb, n)
        b[::3] = a[::3].astype(tb)
        for opn, op in VEC_OPS.items():
            forms = [(a, b)] + ([(a, b[1])] if op <= 3 or op == 7 else []) + ([(a[1], b)] if op <= 3 else [])
            for x, y in forms:
                (r, rc), (p, pc) = ref.vec_binary(op, x, y), port.vec_binary(op, x, y)
                assert rc == pc and r.tobytes() == p.tobytes(), (opn, na, nb, np.ndim(x), np.ndim(y))
                forms_checked += 1
    assert forms_checked == 16 * 19
    for ta in VEC_DTYPES.values():
        x = vec_operand(rng, ta, n)
        assert ref.vec_unary(0, np.abs(x)).tobytes() == port.vec_unary(0, np.abs(x)).tobytes()
        for prec in (0, 1, 3, 6, 7, 15, 16, 20):
            assert ref.vec_unary(1, x, prec).tobytes() == port.vec_unary(1, x, prec).tobytes(), (ta, prec)
    with pytest.raises(RuntimeError):
        port.vec_binary(8, np.zeros(3, np.int32), np.int32(1))          # `vector < scalar`: not a form the reference evaluates as written
