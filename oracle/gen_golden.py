"""Generates tests/golden/*.npz with the REFERENCE-BUILT oracle (oracle/_ref/libaqref.so).

TEST INFRASTRUCTURE ONLY.  Run in the container that has /root/reference mounted:
    make -C oracle ref && python oracle/gen_golden.py
Inputs are regenerated from aquery2_b200/synth.py by the tests (only seeds / shapes are stored);
outputs of the reference's own headers are stored verbatim.  The fixtures pin the plain-C restatement
(oracle/aq_oracle.c) and the CUDA path on any machine, including the GPU box where the reference
tree is absent.
"""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from aquery2_b200 import synth  # noqa: E402
from oracle import aqo  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")

# h2o queries of the reference (benchmark/h2o/groupby.sql); q6's median has no reference implementation
H2O_QUERIES = {
    "q1": (["id1"], [("sum", "v1")]),
    "q2": (["id1", "id2"], [("sum", "v1")]),
    "q3": (["id3"], [("sum", "v1"), ("avg", "v3")]),
    "q4": (["id4"], [("avg", "v1"), ("avg", "v2"), ("avg", "v3")]),
    "q5": (["id6"], [("sum", "v1"), ("sum", "v2"), ("sum", "v3")]),
    "q6": (["id4", "id5"], [("stddev", "v3")]),
    "q7": (["id3"], [("max", "v1"), ("min", "v2")]),
    "q8": (["id6"], [("first", "v3"), ("second", "v3")]),   # subvec(v3, 0, 2)
    "q9": (["id2", "id4"], [("corr", "v1", "v2")]),
    "q10": (["id1", "id2", "id3", "id4", "id5", "id6"], [("sum", "v3"), ("count",)]),
}
SCAN_CASES = [("sums", 0), ("avgs", 0), ("mins", 0), ("maxs", 0), ("sumw", 1), ("sumw", 3), ("sumw", 5),
              ("avgw", 1), ("avgw", 3), ("avgw", 5), ("minw", 2), ("minw", 5), ("maxw", 2), ("maxw", 5),
              ("deltas", 0), ("prev", 0), ("next", 0), ("ratiow", 1), ("ratiow", 3)]


def trade_values(t):
    """value columns exercised by the scan fixtures, derived from the trade table"""
    price, qty = t["price"], t["quantity"]
    return {
        "price_i32": price,
        "qty_i32": qty,
        "signed_i32": (price - 275).astype(np.int32),
        "big_i64": qty.astype(np.int64) * 1_000_003 - 5_000_000_000,
        "price_f32": (price.astype(np.float32) * np.float32(0.37) - np.float32(60.0)),
        "qty_f64": qty.astype(np.float64) / 7.0 - 500.0,
    }


def main():
    ref = aqo.load_ref()
    if ref is None:
        raise SystemExit("oracle/_ref/libaqref.so missing: make -C oracle ref (needs /root/reference)")
    assert ref.kind == "reference"
    os.makedirs(GOLD, exist_ok=True)

    # ---- h2o group-by + scalar aggregates ------------------------------------------------------
    n, k, seed = 6000, 20, 7
    t = synth.h2o_table(n, k, seed=seed)
    out = {"n": n, "k": k, "seed": seed}
    for q, (keys, aggs) in H2O_QUERIES.items():
        ng, gid, off, perm = ref.groupby([t[c] for c in keys])
        out[f"{q}_ngroups"] = ng
        out[f"{q}_gid"] = gid
        out[f"{q}_offsets"] = off
        out[f"{q}_perm"] = perm
        for a, agg in enumerate(aggs):
            v = t[agg[1]] if len(agg) > 1 else None
            v2 = t[agg[2]] if len(agg) > 2 else None
            out[f"{q}_agg{a}"] = ref.reduce(agg[0], v, off, perm, v2=v2, raw128=True)
    np.savez_compressed(os.path.join(GOLD, "h2o_small.npz"), **out)

    # ---- every scalar aggregate on every value dtype --------------------------------------------
    n, nsym, seed = 4000, 37, 11
    t = synth.trade_table(n, nsym, seed=seed)
    vals = trade_values(t)
    ng, gid, off, perm = ref.groupby([t["sym"]])
    out = {"n": n, "nsym": nsym, "seed": seed, "ngroups": ng, "gid": gid, "offsets": off, "perm": perm}
    for vn, v in vals.items():
        for op in ("sum", "avg", "min", "max", "first", "last", "var", "stddev"):
            out[f"reduce_{op}_{vn}"] = ref.reduce(op, v, off, perm, raw128=True)
    out["reduce_count"] = ref.reduce("count", None, off, perm)
    out["reduce_corr_i32"] = ref.reduce("corr", vals["price_i32"], off, perm, v2=vals["qty_i32"])
    out["reduce_corr_i64"] = ref.reduce("corr", vals["big_i64"] // 1000, off, perm, v2=vals["qty_i32"].astype(np.int64))
    for vn, v in vals.items():
        for op, w in SCAN_CASES:
            out[f"scan_{op}_{w}_{vn}"] = ref.scan(op, v, off, perm, w=w, raw128=True)
    np.savez_compressed(os.path.join(GOLD, "trade_small.npz"), **out)

    # ---- int64 keys, multi-key, negative keys ---------------------------------------------------
    rng = np.random.default_rng(5)
    n = 3000
    k64 = rng.integers(-5, 6, n).astype(np.int64) * 4_000_000_007
    k64b = rng.integers(0, 3, n).astype(np.int64)
    kneg = rng.integers(-40, 40, n).astype(np.int32)
    out = {"k64": k64, "k64b": k64b, "kneg": kneg}
    for name, keys in (("k64", [k64]), ("k64x2", [k64, k64b]), ("kneg", [kneg])):
        ng, gid, off, perm = ref.groupby(keys)
        out[f"{name}_ngroups"] = ng
        out[f"{name}_gid"] = gid
        out[f"{name}_offsets"] = off
        out[f"{name}_perm"] = perm
    np.savez_compressed(os.path.join(GOLD, "keys_small.npz"), **out)
    for f in sorted(os.listdir(GOLD)):
        print(f, os.path.getsize(os.path.join(GOLD, f)))


VEC_DTYPES = {"i32": np.int32, "i64": np.int64, "f32": np.float32, "f64": np.float64}
UNARY_PRECISIONS = (0, 2, 5, 7, 16)
VEC_OPS = {"add": 0, "sub": 1, "mul": 2, "div": 3, "gt": 7, "lt": 8, "gte": 9, "lte": 10, "eq": 11, "neq": 12}


def vec_operand(rng, dt, n):
    if np.dtype(dt).kind == "i":
        v = rng.integers(-2000, 2000, n).astype(dt)
        v[v == 0] = 7                                   # integer division by zero traps in the reference
        if dt == np.int64:
            v[::5] *= 3_000_007                         # beyond 32 bits and beyond the float mantissa, products stay below 2^63
        return v
    v = (rng.standard_normal(n) * 1000).astype(dt)
    v[::7] = np.round(v[::7])
    return v


def special_keys(seed=21, n=1500):
    """the key columns of special_small.npz, rebuilt by the tests from the stored bytes"""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import keydata
    rng = np.random.default_rng(seed)
    return {"str": keydata.strings(rng, n, 120, long_tail=True), "date": keydata.dates(rng, n, 900),
            "time": keydata.times(rng, n, 64), "timestamp": keydata.timestamps(rng, n, 64),
            "k2": rng.integers(0, 4, n).astype(np.int32)}


def gen_special():
    """string / date / time / timestamp keys through the reference's AQHashTable (server/hasher.h:97-132) and the reference's
    element-wise operators (server/vector_type.hpp:372-396, server/table.h:826-973): inputs AND outputs stored"""
    ref = aqo.load_ref()
    if ref is None:
        raise SystemExit("oracle/_ref/libaqref.so missing: make -C oracle ref (needs /root/reference)")
    cols = special_keys()
    data, off = cols["str"].packed()
    out = {"strcol_bytes": data, "strcol_offsets": off, "k2": cols["k2"]}
    for name in ("date", "time", "timestamp"):
        out[f"{name}_cells"] = cols[name].view(np.uint8).reshape(len(cols[name]), -1)
    for name, keys in (("str", [cols["str"]]), ("date", [cols["date"]]), ("time", [cols["time"]]), ("timestamp", [cols["timestamp"]]),
                       ("str_k2", [cols["str"], cols["k2"]])):
        ng, gid, offs, perm = ref.groupby(keys)
        out[f"{name}_ngroups"] = ng; out[f"{name}_gid"] = gid; out[f"{name}_offsets"] = offs; out[f"{name}_perm"] = perm
    rng = np.random.default_rng(22)
    n = 257
    ops = {}
    for na, ta in VEC_DTYPES.items():
        out[f"vec_a_{na}"] = vec_operand(rng, ta, n)
        out[f"vec_b_{na}"] = vec_operand(rng, ta, n)
    for (na, ta), (nb, tb) in [(x, y) for x in VEC_DTYPES.items() for y in VEC_DTYPES.items()]:
        a, b = out[f"vec_a_{na}"], out[f"vec_b_{nb}"]
        for opn, op in VEC_OPS.items():
            forms = {"vv": (a, b)}
            if op <= 3 or op == 7:
                forms["vs"] = (a, b[1])
            if op <= 3:
                forms["sv"] = (a[1], b)
            for fn, (x, y) in forms.items():
                cells, code = ref.vec_binary(op, x, y)
                out[f"vec_{opn}_{na}_{nb}_{fn}"] = np.ascontiguousarray(cells).view(np.uint8).reshape(-1)
                out[f"vect_{opn}_{na}_{nb}_{fn}"] = code
    # sqrt / truncate (server/aggregations.h:34-69)
    for na, ta in VEC_DTYPES.items():
        x = np.abs(out[f"vec_a_{na}"]) if np.dtype(ta).kind == "f" else out[f"vec_a_{na}"]
        out[f"una_x_{na}"] = x
        out[f"una_sqrt_{na}"] = ref.vec_unary(0, np.abs(x))
        for prec in UNARY_PRECISIONS:
            out[f"una_trunc{prec}_{na}"] = ref.vec_unary(1, (x / 7).astype(ta) if np.dtype(ta).kind == "f" else x, prec)
    np.savez_compressed(os.path.join(GOLD, "special_small.npz"), **out)
    print("special_small.npz", os.path.getsize(os.path.join(GOLD, "special_small.npz")))


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "special":
        gen_special()
    else:
        main()
        gen_special()
