"""Element-wise vector arithmetic / compare (include/aqgpu.h aqgpu_vec_binary) against the reference's
own operators (server/vector_type.hpp:372-396, server/table.h:826-973) compiled into oracle/_ref, cell for cell and type
for type: bit-exact, floating point included (one IEEE operation per cell, the same on both sides).  Forms the reference
cannot evaluate (its `vector < scalar` evaluates `>`) are checked against the C++ semantics restated in numpy."""
import itertools

import numpy as np
import pytest

from util import assert_same

DT = {"i32": np.int32, "i64": np.int64, "f32": np.float32, "f64": np.float64}


def operands(rng, dt, n):
    if np.dtype(dt).kind == "i":
        v = rng.integers(-2000, 2000, n).astype(dt)
        v[v == 0] = 7                                   # integer division by zero traps in the reference
        if dt == np.int64:
            v[::5] *= 3_000_000_007                     # beyond 32 bits, beyond the 24-bit float mantissa
        return v
    v = (rng.standard_normal(n) * 1000).astype(dt)
    v[::7] = np.round(v[::7])
    return v


@pytest.fixture(scope="module")
def refop(port):
    """the reference's operators where oracle/_ref is built, else the C restatement that tests/test_oracle_vs_ref.py pins to them"""
    from oracle import aqo
    return aqo.load_ref() or port


def test_result_types_follow_the_reference_templates(refop):
    """aqgpu_vec_out_type (host code, no device needed) == the element type the reference's templates deduce"""
    from aquery2_b200 import aqgpu as aq
    rng = np.random.default_rng(0)
    for (na, ta), (nb, tb) in itertools.product(DT.items(), DT.items()):
        a, b = operands(rng, ta, 8), operands(rng, tb, 8)
        for op in (aq.V_ADD, aq.V_SUB, aq.V_MUL, aq.V_DIV, aq.V_GT):
            forms = [(a, b), (a, b[0])] + ([(a[0], b)] if op != aq.V_GT else [])
            for x, y in forms:
                _, code = refop.vec_binary(op, x, y)
                assert aq.vec_out_type(op, x, y) == code, (op, na, nb, np.ndim(x), np.ndim(y))
        for op in (aq.V_LT, aq.V_GTE, aq.V_LTE, aq.V_EQ, aq.V_NEQ):
            assert aq.vec_out_type(op, a, b) == refop.vec_binary(op, a, b)[1] == aq.BOOL


@pytest.mark.gpu
@pytest.mark.parametrize("na,nb", list(itertools.product(DT, DT)))
def test_cells_equal_the_reference_operators(gpu, refop, na, nb):
    rng = np.random.default_rng(len(na) * 7 + len(nb))
    n = 20_011
    a, b = operands(rng, DT[na], n), operands(rng, DT[nb], n)
    b[::3] = a[::3].astype(DT[nb])                      # equal cells for == / != / >=
    for op in (gpu.V_ADD, gpu.V_SUB, gpu.V_MUL, gpu.V_DIV, gpu.V_GT):
        forms = [(a, b), (a, b[1])] + ([(a[1], b)] if op != gpu.V_GT else [])
        for x, y in forms:
            want, code = refop.vec_binary(op, x, y)
            got = gpu.vec_binary(op, x, y)
            assert got.dtype == want.dtype, (op, na, nb, got.dtype, want.dtype, code)
            assert got.tobytes() == want.tobytes(), f"op {op} {na} {nb} forms {np.ndim(x)} {np.ndim(y)}: " \
                f"first difference at {np.flatnonzero((got != want).reshape(len(got), -1).any(axis=1))[:5]}"
    for op in (gpu.V_LT, gpu.V_GTE, gpu.V_LTE, gpu.V_EQ, gpu.V_NEQ):
        want, _ = refop.vec_binary(op, a, b)            # aqop_lt ... (table.h:949-973)
        assert_same(gpu.vec_binary(op, a, b), want, f"aqop {op} {na} {nb}")


@pytest.mark.gpu
def test_comparisons_with_a_scalar_mean_what_they_say(gpu):
    """waived quirk: the reference evaluates `>` for every `vector OP scalar` comparison (vector_type.hpp:372-385)"""
    rng = np.random.default_rng(3)
    a = operands(rng, np.int32, 5000)
    f = operands(rng, np.float64, 5000)
    import operator as o
    for op, fn in ((gpu.V_GT, o.gt), (gpu.V_LT, o.lt), (gpu.V_GTE, o.ge), (gpu.V_LTE, o.le), (gpu.V_EQ, o.eq), (gpu.V_NEQ, o.ne)):
        assert_same(gpu.vec_binary(op, a, np.int32(120)), fn(a, 120).astype(np.uint8), f"int32 op {op} 120")
        assert_same(gpu.vec_binary(op, f, np.float64(a[3])), fn(f, float(a[3])).astype(np.uint8), f"f64 op {op}")
        assert_same(gpu.vec_binary(op, np.int64(5), a), fn(5, a).astype(np.uint8), f"5 op {op} int32")


@pytest.mark.gpu
def test_explicit_result_types_bitwise_and_bool_operands(gpu):
    rng = np.random.default_rng(4)
    n = 10_000
    a, b = operands(rng, np.int32, n), operands(rng, np.int32, n)
    # aqop_x(lhs, rhs, ret): the caller's Ret decides the cell type (table.h:949-973)
    assert_same(gpu.vec_binary(gpu.V_GT, a, b, out_type=gpu.INT32), (a > b).astype(np.int32), "gt -> int32")
    assert_same(gpu.vec_binary(gpu.V_ADD, a, b, out_type=gpu.INT64), (a + b).astype(np.int64), "add -> int64")
    assert_same(gpu.vec_binary(gpu.V_MUL, a, b, out_type=gpu.DOUBLE), (a * b).astype(np.float64), "mul -> f64")
    assert_same(gpu.vec_binary(gpu.V_AND, a, b), a & b, "and"); assert_same(gpu.vec_binary(gpu.V_OR, a, b), a | b, "or")
    assert_same(gpu.vec_binary(gpu.V_XOR, a, b.astype(np.int64)), a ^ b.astype(np.int64), "xor")
    p, q = a > 0, b > 0                                   # bool cells promote to int like C++
    assert_same(gpu.vec_binary(gpu.V_AND, p, q, out_type=gpu.BOOL), (p & q).astype(np.uint8), "bool and")
    assert_same(gpu.vec_binary(gpu.V_ADD, p, np.int32(0), out_type=gpu.INT32), p.astype(np.int32), "bool -> int32")
    # int32 wraps, integer division truncates toward zero, x / 0 = 0 here, INT_MIN / -1 wraps
    big = np.array([2**31 - 1, -2**31, 7, -7, 5], np.int32)
    d = np.array([1, -1, 2, 2, 0], np.int32)
    assert_same(gpu.vec_binary(gpu.V_ADD, big, np.int32(1)), np.array([-2**31, -2**31 + 1, 8, -6, 6], np.int32), "wrap")
    assert_same(gpu.vec_binary(gpu.V_DIV, big, d), np.array([2**31 - 1, -2**31, 3, -3, 0], np.int32), "int division")
    # errors
    import ctypes
    L = gpu.lib()
    f = np.zeros(4, np.float32)
    assert L.aqgpu_vec_binary(gpu.V_AND, gpu.FLOAT, f.ctypes.data, 0, gpu.FLOAT, f.ctypes.data, 0, -1, f.ctypes.data, 4) != 0
    assert L.aqgpu_vec_binary(99, gpu.FLOAT, f.ctypes.data, 0, gpu.FLOAT, f.ctypes.data, 0, -1, f.ctypes.data, 4) != 0
    assert L.aqgpu_vec_binary(gpu.V_ADD, gpu.FLOAT, f.ctypes.data, 1, gpu.FLOAT, f.ctypes.data, 1, -1, f.ctypes.data, 4) != 0


@pytest.mark.gpu
def test_network_a_session_key(gpu, port):
    """tests/network.a:8-11: GROUP BY src, dst, sums(deltas(_time) > 120) -- string keys + a key computed by vector ops"""
    import keydata
    rng = np.random.default_rng(8)
    n = 60_000
    src, dst = keydata.strings(rng, n, 12), keydata.strings(rng, n, 9)
    t = np.cumsum(rng.integers(1, 200, n)).astype(np.int32)
    ln = rng.integers(40, 1500, n).astype(np.int32)
    one = gpu.Groups([np.zeros(n, np.uint32)], row_order=gpu.ROWS_ASC)           # non-grouped deltas: the table is one group
    deltas = one.scan("deltas", t)
    gap = gpu.vec_binary(gpu.V_GT, deltas, np.int32(120), out_type=gpu.INT32)     # aqop_gt(deltas, 120, ret<int>)
    sess = gpu.i128_to_int64(one.scan("sums", gap))
    want_sess = np.cumsum(np.diff(t, prepend=t[0]) > 120).astype(np.int64)
    assert_same(sess, want_sess, "sums(deltas(_time) > 120)")
    cs, _ = gpu.string_dict(*src.packed()); cd, _ = gpu.string_dict(*dst.packed())
    res = gpu.groupby_agg([cs, cd, sess], [("count",), ("avg", ln)])
    ng, gid, off, perm = port.groupby([src, dst, want_sess])
    assert res.ngroups == ng
    assert_same(res.first_row, perm[off[1:] - 1].astype(np.uint32), "first rows")
    assert_same(res.aggs[1], port.reduce("avg", ln, off, perm), "avg(len)")


@pytest.mark.gpu
def test_take_is_the_gather_of_table_h(gpu):
    """out[j] = col[idx[j]] (server/table.h:184-189; TableInfo::materialize :429-440) for every cell width"""
    import keydata
    rng = np.random.default_rng(9)
    n, m = 50_000, 80_000
    idx = rng.integers(0, n, m).astype(np.uint32)
    cols = [rng.integers(-9, 9, n).astype(np.int32), rng.standard_normal(n), rng.integers(0, 2, n).astype(np.uint8),
            rng.integers(0, 1 << 15, n).astype(np.int16), keydata.timestamps(rng, n, 64), keydata.times(rng, n, 64),
            rng.integers(-5, 5, (n, 2)).astype(np.int64).view([("lo", "<i8"), ("hi", "<i8")]).reshape(n)]
    for c in cols:
        got = gpu.take(c, idx)
        assert got.tobytes() == c[idx].tobytes(), c.dtype
    assert len(gpu.take(cols[0], np.zeros(0, np.uint32))) == 0
    with pytest.raises(gpu.AqgpuError, match="beyond"):
        gpu.take(cols[0], np.array([0, n], np.uint32))


@pytest.mark.gpu
def test_cells_equal_the_golden_fixture(gpu, golden):
    """every operand form against the cells the reference's operators produced (tests/golden/special_small.npz): the same
    check as test_cells_equal_the_reference_operators for machines without the reference build"""
    from oracle.gen_golden import VEC_DTYPES, VEC_OPS
    g = golden("special_small.npz")
    checked = 0
    for na in VEC_DTYPES:
        for nb in VEC_DTYPES:
            a, b = g[f"vec_a_{na}"], g[f"vec_b_{nb}"]
            for opn, op in VEC_OPS.items():
                for fn, (x, y) in (("vv", (a, b)), ("vs", (a, b[1])), ("sv", (a[1], b))):
                    key = f"vec_{opn}_{na}_{nb}_{fn}"
                    if key not in g.files:
                        continue
                    if fn == "vs" and op >= gpu.V_LT:
                        continue                          # the reference evaluates `>` there (waived quirk): not stored
                    got = gpu.vec_binary(op, x, y)
                    assert np.ascontiguousarray(got).view(np.uint8).reshape(-1).tobytes() == g[key].tobytes(), key
                    checked += 1
    assert checked == 16 * (10 + 5 + 4)


@pytest.mark.gpu
def test_sqrt_and_truncate_equal_the_reference(gpu, golden):
    """server/aggregations.h:34-69 through aqgpu_vec_unary against the reference's cells (tests/golden/special_small.npz) and,
    where oracle/_ref is built, against the reference's templates on larger seeded inputs; bit-exact (IEEE sqrt / mul / div)"""
    from oracle import aqo
    from oracle.gen_golden import VEC_DTYPES, UNARY_PRECISIONS
    g = golden("special_small.npz")
    for na, ta in VEC_DTYPES.items():
        x = g[f"una_x_{na}"]
        assert gpu.vec_unary(gpu.U_SQRT, np.abs(x)).tobytes() == g[f"una_sqrt_{na}"].tobytes(), f"sqrt {na}"
        for prec in UNARY_PRECISIONS:
            xin = (x / 7).astype(ta) if np.dtype(ta).kind == "f" else x
            assert gpu.vec_unary(gpu.U_TRUNCATE, xin, prec).tobytes() == g[f"una_trunc{prec}_{na}"].tobytes(), f"truncate {na} {prec}"
    ref = aqo.load_ref()
    if ref is not None:
        rng = np.random.default_rng(12)
        for ta in VEC_DTYPES.values():
            x = (rng.standard_normal(50_000) * 1e4).astype(ta)
            x[:4] = [0, 1, 2, 3]
            assert gpu.vec_unary(gpu.U_SQRT, np.abs(x)).tobytes() == ref.vec_unary(0, np.abs(x)).tobytes()
            for prec in (0, 1, 3, 6, 7, 15, 16, 20):
                assert gpu.vec_unary(gpu.U_TRUNCATE, x, prec).tobytes() == ref.vec_unary(1, x, prec).tobytes(), (ta, prec)
    big = np.array([3.0e38, -3.0e38, 1.5, -2.5, 2.5], np.float32)          # beyond max / 10^p: left as they are
    assert_same(gpu.vec_unary(gpu.U_TRUNCATE, big, 3), np.array([3.0e38, -3.0e38, 1.5, -2.5, 2.5], np.float32), "truncate near max")
    assert_same(gpu.vec_unary(gpu.U_TRUNCATE, big, 0), np.array([3.0e38, -3.0e38, 2.0, -3.0, 3.0], np.float32), "half away from zero")
