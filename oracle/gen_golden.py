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


if __name__ == "__main__":
    main()
