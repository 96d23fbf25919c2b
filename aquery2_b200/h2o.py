"""The reference's h2o group-by workload (benchmark/h2o/groupby.sql, schema
benchmark/h2o/groupby_prep.sql:1: source(id1..id6 int, v1 int, v2 int, v3 float)) as library calls.

Each query = one fused group-by + aggregate call (+ an element-wise epilogue for q7 / q9), i.e. what the
generated post-processor does between "columns resident" and "result columns produced"
(server/server.cpp:303-305).  ALG_BYTES is SURVEY.md section 8(d)'s algorithmic bytes per row: every
needed input column read exactly once.

Table generation and the query plan only; bench.py's hooks (verification, CPU baseline, e2e arm) live in
bench_h2o.py at the repository root.
"""
from __future__ import annotations

from . import synth

METRIC = "h2o groupby rows/sec (q1-q10)"
DTYPE = "int64/f64"

# name -> (key columns, [(op, value column[, second column])], epilogue)
QUERIES = {
    "q1": (["id1"], [("sum", "v1")], None),                                   # groupby.sql:1
    "q2": (["id1", "id2"], [("sum", "v1")], None),                            # :3
    "q3": (["id3"], [("sum", "v1"), ("avg", "v3")], None),                    # :5
    "q4": (["id4"], [("avg", "v1"), ("avg", "v2"), ("avg", "v3")], None),     # :7
    "q5": (["id6"], [("sum", "v1"), ("sum", "v2"), ("sum", "v3")], None),     # :9
    "q6": (["id4", "id5"], [("stddev", "v3")], None),                         # :11-12 (median: no reference implementation)
    "q7": (["id3"], [("max", "v1"), ("min", "v2")], "sub"),                   # :14  max(v1) - min(v2)
    "q8": (["id6"], [("first", "v3"), ("second", "v3")], None),               # :17  subvec(v3, 0, 2)
    "q9": (["id2", "id4"], [("corr", "v1", "v2")], "pow2"),                   # :20  pow(corr(v1, v2), 2)
    "q10": (["id1", "id2", "id3", "id4", "id5", "id6"], [("sum", "v3"), ("count",)], None),   # :23
}
ALG_BYTES = {"q1": 8, "q2": 12, "q3": 12, "q4": 16, "q5": 16, "q6": 12, "q7": 12, "q8": 8, "q9": 16, "q10": 28}
ORDER = ["q1", "q2", "q3", "q4", "q5", "q6", "q7", "q8", "q9", "q10"]
_OUT_CELL = {"sum_i": 16, "sum_f": 8, "avg": 8, "min": 4, "max": 4, "first": 4, "second": 4, "last": 4,
             "count": 4, "stddev": 8, "var": 8, "corr": 8}

# device generator kinds (csrc/synth.cu) per column: (kind, k as a function of (N, K))
_GEN = {
    "id1": (0, lambda n, k: k), "id2": (0, lambda n, k: k), "id3": (0, lambda n, k: max(n // k, 1)),
    "id4": (0, lambda n, k: k), "id5": (0, lambda n, k: k), "id6": (0, lambda n, k: max(n // k, 1)),
    "v1": (0, lambda n, k: 5), "v2": (0, lambda n, k: 15), "v3": (2, lambda n, k: 100_000_000),
}


def key_bound(col: str, n_total: int, k: int) -> tuple[int, int]:
    """inclusive value range of a generated key column (the statistics ColRef::stats would hold)"""
    return (1, _GEN[col][1](n_total, k))


def max_groups(q: str, n_total: int, k: int, nrows: int) -> int:
    g = 1
    for c in QUERIES[q][0]:
        g *= key_bound(c, n_total, k)[1]
        if g >= nrows:
            return nrows
    return min(g, nrows)


def out_bytes(q: str, ng: int) -> int:
    """bytes of the result columns of one query (keys + aggregates), the other half of its HBM traffic"""
    keys, aggs, epi = QUERIES[q]
    b = 4 * len(keys)
    for a in aggs:
        if a[0] == "sum":
            b += _OUT_CELL["sum_f" if a[1] == "v3" else "sum_i"]
        else:
            b += _OUT_CELL[a[0]]
    if epi == "sub":
        b += 4
    elif epi == "pow2":
        b += 8
    return b * ng


def columns_of(queries) -> list[str]:
    cols = []
    for q in queries:
        keys, aggs, _ = QUERIES[q]
        for c in list(keys) + [c for a in aggs for c in a[1:]]:
            if c not in cols:
                cols.append(c)
    return cols


def table_name(args, n: int) -> str:
    return f"h2o G1_{n:.0e}_{args.k:.0e}_0_0"


def host_table(args, n: int, queries=ORDER) -> dict:
    return synth.h2o_table(n, args.k, columns=columns_of(queries))


def device_table(aq, args, n_total: int, lo: int, hi: int, queries, seed: int = 42) -> dict:
    """rows [lo, hi) of G1_{n_total}_{k}_0_0 generated in HBM (bit-identical to synth.h2o_table)"""
    out = {}
    for c in columns_of(queries):
        kind, kf = _GEN[c]
        out[c] = aq.synth_column(kind, synth.stream_key(seed, synth.STREAM[c]), hi - lo, kf(n_total, args.k),
                                 row_lo=lo, code=aq.FLOAT if c == "v3" else aq.INT32)
    return out


class Plan:
    """Prepared queries over one device-resident table.  hints=False: aqgpu_groupby_run, the library plans
    and allocates by itself; hints=True: aqgpu_groupby_agg with the generator's bounds / group counts and
    caller-allocated result columns."""

    def __init__(self, aq, table: dict, queries, args, n_total: int, hints: bool = False):
        self.aq, self.queries, self.table = aq, list(queries), table
        self.prepared, self.epilogue, self.extra = {}, {}, {}
        self.nrows = next(iter(table.values())).n
        for q in self.queries:
            keys, aggs, epi = QUERIES[q]
            specs = [(a[0],) + tuple(table[c] for c in a[1:]) for a in aggs]
            if hints:
                mg = max_groups(q, n_total, args.k, self.nrows)
                bounds = [key_bound(c, n_total, args.k) for c in keys]
                self.prepared[q] = aq.PreparedQuery([table[c] for c in keys], specs, max_groups=mg, groups_hint=mg,
                                                    key_bounds=bounds, device_out=True)
            else:
                self.prepared[q] = aq.AutoQuery([table[c] for c in keys], specs)
            self.epilogue[q] = epi
            self.extra[q] = None

    def run(self, q: str) -> int:
        p = self.prepared[q]
        ng = p.run()
        epi = self.epilogue[q]
        if epi:
            code = self.aq.INT32 if epi == "sub" else self.aq.DOUBLE
            e = self.extra[q]
            if e is None or e.n < ng:
                if e is not None:
                    e.free()
                e = self.extra[q] = self.aq.DeviceColumn(max(ng, 1), code)
            if epi == "sub":              # q7: max(v1) - min(v2)   (server/table.h:820-973 vector arithmetic)
                self.aq.vec_sub(p.outs[0], p.outs[1], e, ng, code)
            else:                         # q9: pow(corr, 2)        (server/aggregations.h:409-411)
                self.aq.vec_pow(p.outs[0], 2.0, e, ng)
        return ng

    def step(self) -> dict:
        return {q: self.run(q) for q in self.queries}

    def result_columns(self, q: str):
        p = self.prepared[q]
        cols = list(p.kouts) + list(p.outs)
        if self.extra[q] is not None:
            cols.append(self.extra[q])
        return cols


def dist_plan(comm, table, queries, args, n_total, lo, hints=False):
    from . import dist
    return dist.ShardPlan(comm, table, queries, n_total, args.k, lo, use_bounds=hints)
