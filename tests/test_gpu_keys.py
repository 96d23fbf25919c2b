"""String / date / time / timestamp key columns on the GPU (include/aqgpu.h: aqgpu_string_dict,
aqgpu_key_split, aqgpu_key_join) against the CPU oracle -- the reference's own AQHashTable with its hashes and
operator== where oracle/_ref is built, else the C restatement that tests/test_oracle_vs_ref.py pins to it.
Everything here is integer / index work: bit-exact."""
import ctypes

import numpy as np
import pytest

import keydata
from util import assert_same

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def oracle(port):
    from oracle import aqo
    return aqo.load_ref() or port


def first_rows(off, perm, ng):
    return perm[off[1:] - 1].astype(np.uint32) if ng else np.zeros(0, np.uint32)      # descending lists: last = first row


@pytest.mark.parametrize("n,nd,long_tail", [(1, 1, False), (37, 10, False), (5000, 300, True), (300_000, 20_000, True),
                                            (200_000, 3, False)])
def test_string_dict_is_the_reference_group_id(gpu, oracle, n, nd, long_tail):
    rng = np.random.default_rng(n + nd)
    col = keydata.strings(rng, n, nd, long_tail)
    data, off = col.packed()
    codes, first = gpu.string_dict(data, off)
    ng, gid, goff, perm = oracle.groupby([col])
    assert len(first) == ng
    assert_same(codes, gid, "codes = reversemap of AQHashTable<astring_view>")
    assert_same(first, first_rows(goff, perm, ng), "first row of every distinct string")


def test_string_key_with_an_integer_key_and_aggregates(gpu, oracle):
    rng = np.random.default_rng(5)
    n = 250_000
    col = keydata.strings(rng, n, 2000)
    k2 = rng.integers(0, 7, n).astype(np.int32)
    v = rng.integers(-1000, 1000, n).astype(np.int32)
    x = rng.random(n).astype(np.float32)
    data, off = col.packed()
    codes, dict_first = gpu.string_dict(data, off)
    res = gpu.groupby_agg([codes, k2], [("sum", v), ("avg", x), ("count",), ("max", v)])
    ng, gid, goff, perm = oracle.groupby([col, k2])
    assert res.ngroups == ng
    first = first_rows(goff, perm, ng)
    assert_same(res.first_row, first, "first rows")
    # result key: code -> the string of the dictionary's first row
    got = [col.strings[dict_first[c]] for c in res.keys[0]]
    assert got == [col.strings[r] for r in first]
    assert_same(res.keys[1], k2[first], "int key")
    assert_same(res.aggs[0], oracle.reduce("sum", v, goff, perm, raw128=True), "sum")
    assert_same(res.aggs[1], oracle.reduce("avg", x, goff, perm), "avg")
    assert_same(res.aggs[3], oracle.reduce("max", v, goff, perm), "max")
    # the group handle route (row lists): same ids and lists
    h = gpu.Groups([codes, k2])
    o2, p2 = h.csr()
    assert h.ngroups == ng
    assert_same(h.gid(), gid, "gid"); assert_same(o2, goff, "offsets"); assert_same(p2, perm, "row lists")


@pytest.mark.parametrize("kind", ["date", "time", "timestamp"])
@pytest.mark.parametrize("n", [1, 1000, 400_000])
def test_date_time_keys(gpu, oracle, kind, n):
    rng = np.random.default_rng(n + len(kind))
    col = {"date": keydata.dates, "time": keydata.times, "timestamp": keydata.timestamps}[kind](rng, n, 64)
    v = rng.integers(0, 100, n).astype(np.int32)
    lowered = gpu.key_split(col)
    res = gpu.groupby_agg(lowered, [("sum", v), ("count",)])
    ng, gid, goff, perm = oracle.groupby([col])
    assert res.ngroups == ng
    first = first_rows(goff, perm, ng)
    assert_same(res.first_row, first, "first rows")
    assert_same(res.aggs[0], oracle.reduce("sum", v, goff, perm, raw128=True), "sum")
    back = gpu.key_join(col.dtype, res.keys[: len(lowered)])
    want = col[first].copy()
    if kind == "time":
        want["pad"] = 0                       # padding comes back as zero; the fields are the reference's
    if kind == "timestamp":
        want["time"]["pad"] = 0
    assert back.tobytes() == want.tobytes()
    h = gpu.Groups(lowered)
    assert_same(h.gid(), gid, "gid")


def test_time_image_is_order_preserving(gpu):
    """the int64 image of time_t sorts like (hours, minutes, seconds, ms): usable as an ASSUMING order column"""
    rng = np.random.default_rng(3)
    t = keydata.times(rng, 5000, 64)
    img = gpu.key_split(t)[0]
    want = np.lexsort((t["ms"], t["seconds"], t["minutes"], t["hours"]))
    got = np.argsort(img, kind="stable")
    key = lambda idx: np.stack([t["hours"][idx], t["minutes"][idx], t["seconds"][idx], t["ms"][idx]])
    assert np.array_equal(key(got), key(want))
    assert img.min() >= 0


def test_string_dict_errors(gpu):
    L = gpu.lib()
    data = np.frombuffer(b"abcdef", np.uint8).copy()
    nd = ctypes.c_uint32(0)
    codes = np.zeros(4, np.uint32)
    first = np.zeros(4, np.uint32)

    def call(off, cap):
        off = np.asarray(off, np.uint64)
        return L.aqgpu_string_dict(data.ctypes.data_as(ctypes.c_void_p), ctypes.c_uint64(data.size), off.ctypes.data_as(ctypes.c_void_p),
                                   ctypes.c_uint32(len(off) - 1), codes.ctypes.data_as(ctypes.c_void_p), ctypes.c_uint32(cap),
                                   first.ctypes.data_as(ctypes.c_void_p), ctypes.byref(nd))
    assert call([0, 2, 1, 6], 4) != 0 and "offsets" in L.aqgpu_last_error().decode()        # descending
    assert call([0, 2, 4, 9], 4) != 0 and "offsets" in L.aqgpu_last_error().decode()        # past the buffer
    assert call([0, 2, 4, 6], 2) != 0 and "capacity" in L.aqgpu_last_error().decode()       # 3 distinct, room for 2
    assert call([0, 2, 4, 6], 4) == 0 and nd.value == 3
    assert call([0, 0, 0, 0], 4) == 0 and nd.value == 1 and list(codes[:3]) == [0, 0, 0]    # three empty strings
    assert call([0], 4) == 0 and nd.value == 0                                              # no rows


def test_special_keys_golden(gpu, golden):
    """the GPU against the fixture the reference's own AQHashTable produced (tests/golden/special_small.npz)"""
    g = golden("special_small.npz")
    cols = keydata.from_golden(g)
    codes, first = gpu.string_dict(*cols["str"].packed())
    assert len(first) == int(g["str_ngroups"])
    assert_same(codes, g["str_gid"], "string codes")
    for name in ("date", "time", "timestamp"):
        h = gpu.Groups(gpu.key_split(cols[name]))
        o, p = h.csr()
        assert h.ngroups == int(g[f"{name}_ngroups"]), name
        assert_same(h.gid(), g[f"{name}_gid"], name); assert_same(o, g[f"{name}_offsets"], name); assert_same(p, g[f"{name}_perm"], name)
    h = gpu.Groups([codes, cols["k2"]])
    o, p = h.csr()
    assert h.ngroups == int(g["str_k2_ngroups"])
    assert_same(h.gid(), g["str_k2_gid"], "(str, int32)"); assert_same(o, g["str_k2_offsets"]); assert_same(p, g["str_k2_perm"])
