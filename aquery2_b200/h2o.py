"""The reference's h2o group-by workload (benchmark/h2o/groupby.sql, schema
benchmark/h2o/groupby_prep.sql:1: source(id1..id6 int, v1 int, v2 int, v3 float)) as library calls.

Each query = one fused aqgpu_groupby_agg call (+ an element-wise epilogue for q7 / q9), i.e. what the
generated post-processor does between "columns resident" and "result columns produced"
(server/server.cpp:303-305).  ALG_BYTES is SURVEY.md section 8(d)'s algorithmic bytes per row: every
needed input column read exactly once.
"""
from __future__ import annotations

from . import synth

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


def columns_of(queries) -> list[str]:
    cols = []
    for q in queries:
        keys, aggs, _ = QUERIES[q]
        for c in list(keys) + [c for a in aggs for c in a[1:]]:
            if c not in cols:
                cols.append(c)
    return cols


def device_table(aq, n_total: int, k: int, lo: int, hi: int, cols, seed: int = 42) -> dict:
    """rows [lo, hi) of G1_{n_total}_{k}_0_0 generated in HBM (bit-identical to synth.h2o_table)"""
    out = {}
    for c in cols:
        kind, kf = _GEN[c]
        out[c] = aq.synth_column(kind, synth.stream_key(seed, synth.STREAM[c]), hi - lo, kf(n_total, k),
                                 row_lo=lo, code=aq.FLOAT if c == "v3" else aq.INT32)
    return out


class Plan:
    """PreparedQuery objects for a set of h2o queries over one table (device or pinned-host columns)."""

    def __init__(self, aq, table: dict, queries, n_total: int, k: int, device_out: bool = True,
                 host_alloc=None, use_bounds: bool = True):
        self.aq, self.queries = aq, list(queries)
        self.prepared, self.epilogue, self.extra = {}, {}, {}
        any_col = next(iter(table.values()))
        self.nrows = any_col.n if isinstance(any_col, aq.DeviceColumn) else len(any_col)
        for q in self.queries:
            keys, aggs, epi = QUERIES[q]
            mg = max_groups(q, n_total, k, self.nrows)
            bounds = [key_bound(c, n_total, k) for c in keys] if use_bounds else None
            specs = [(a[0],) + tuple(table[c] for c in a[1:]) for a in aggs]
            self.prepared[q] = aq.PreparedQuery([table[c] for c in keys], specs, max_groups=mg, groups_hint=mg,
                                                key_bounds=bounds, device_out=device_out, host_alloc=host_alloc)
            self.epilogue[q] = epi
            if epi == "sub":
                self.extra[q] = aq.DeviceColumn(mg, aq.INT32) if device_out else host_alloc(aq.INT32, mg)
            elif epi == "pow2":
                self.extra[q] = aq.DeviceColumn(mg, aq.DOUBLE) if device_out else host_alloc(aq.DOUBLE, mg)

    def run(self, q: str) -> int:
        p = self.prepared[q]
        ng = p.run()
        epi = self.epilogue[q]
        if epi == "sub":                  # q7: max(v1) - min(v2)   (server/table.h:820-973 vector arithmetic)
            self.aq.vec_sub(p.outs[0], p.outs[1], self.extra[q], ng, self.aq.INT32)
        elif epi == "pow2":               # q9: pow(corr, 2)        (server/aggregations.h:409-411)
            self.aq.vec_pow(p.outs[0], 2.0, self.extra[q], ng)
        return ng

    def step(self, tag: bool = False) -> dict:
        ngs = {}
        for q in self.queries:
            if tag:
                self.aq.profile_tag(q)
            ngs[q] = self.run(q)
        if tag:
            self.aq.profile_tag("")
        return ngs

    def io_bytes(self, ngs: dict) -> tuple[int, int]:
        """(input bytes, output bytes) one step moves across the C ABI"""
        bi = sum(self.prepared[q].in_bytes for q in self.queries)
        bo = sum(self.prepared[q].out_bytes(ngs[q]) for q in self.queries)
        return bi, bo
