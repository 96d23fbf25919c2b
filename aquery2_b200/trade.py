"""trade.py -- the windowed workload (BASELINE.json configs[3]): per-symbol moving aggregates over the
trade table of the reference's datagen.cpp:90-107, in the shape of benchmark/quries/Aquery/q7.a / q10.a,
tests/moving_avg.a and tests/stock.a:

    SELECT sym, avgs(5, price), sums(5, quantity), deltas(price), mins(5, price), max(price - mins(price))
    FROM trade [ASSUMING ASC time] GROUP BY sym

w_sorted    the table is in time order (the hybrid engine sorts in SQL first): group build with
            ascending row lists + one aqgpu_group_window call.
w_unsorted  time is a permutation of 1..N: group build + the per-group sort of TableInfo::order_by
            (server/table.h:447-461, common/groupby.py:54-59) + the same window call over the ordered lists.

Host-side mirror only: every number is produced by libaqgpu through aquery2_b200.aqgpu."""
from __future__ import annotations

import os

from . import synth

METRIC = "windowed aggregation rows/sec (trade: avgs sums deltas mins max by sym)"
DTYPE = "int64/f64"
ORDER = ["w_sorted", "w_unsorted"]
# per row: sym 4 + price 4 + quantity 4 (+ time 4) read; avgs 8 + sums 16 (int128) + deltas 4 + mins 4 written
ALG_BYTES = {"w_sorted": 12 + 32, "w_unsorted": 16 + 32}
COLUMNS = {"w_sorted": ["sym", "quantity", "price"], "w_unsorted": ["sym", "quantity", "price", "time_shuffled"]}


def expressions(table):
    """(op, value column, w, minus_from_v, reduce) of aqgpu_group_window"""
    p, q = table["price"], table["quantity"]
    return [("avgw", p, 5), ("sumw", q, 5), ("deltas", p), ("minw", p, 5), ("mins", p, 0, True, "max")]


def out_bytes(q: str, ng: int) -> int:
    return 0                        # outputs are per-row and already counted in ALG_BYTES


def columns_of(queries):
    out = []
    for q in queries:
        for c in COLUMNS[q]:
            if c not in out:
                out.append(c)
    return out


def n_sym(args, n: int) -> int:
    return max(1, min(int(args.nsym), n))


def table_name(args, n: int) -> str:
    return f"trade {n:.0e} rows x {n_sym(args, n):.0e} syms"


def host_table(args, n: int, queries=ORDER) -> dict:
    return synth.trade_table(n, n_sym(args, n), columns=columns_of(queries))


_GEN = {"sym": 1, "time": 3, "quantity": 5, "price": 4, "time_shuffled": 6}


def device_table(aq, args, n_total: int, lo: int, hi: int, queries, seed: int = 42) -> dict:
    """rows [lo, hi) of the trade table generated in HBM (bit-identical to synth.trade_table)"""
    ns = n_sym(args, n_total)
    key = lambda c: synth.stream_key(seed, synth.STREAM[c])
    out = {}
    for c in columns_of(queries):
        if c == "sym":
            out[c] = aq.synth_column(1, key("sym"), hi - lo, ns, row_lo=lo)
        elif c == "quantity":
            out[c] = aq.synth_column(5, key("quantity"), hi - lo, 9901, row_lo=lo, aux_k=100)
        elif c == "price":
            out[c] = aq.synth_column(4, key("price"), hi - lo, ns, row_lo=lo, key1=key("sym"), key2=key("price_base"))
        elif c == "time":
            out[c] = aq.synth_column(3, 0, hi - lo, 0, row_lo=lo)
        elif c == "time_shuffled":
            out[c] = aq.synth_column(6, 0, hi - lo, n_total, row_lo=lo)
    return out


class Plan:
    """One group build (+ order_by) + one fused window call per query; results stay in HBM."""

    def __init__(self, aq, table: dict, queries, args, n_total: int, hints: bool = False):
        self.aq, self.queries, self.table = aq, list(queries), table
        self.nrows = table["sym"].n
        self.bounds = [(0, n_sym(args, n_total) - 1)] if hints else None
        self.hint = n_sym(args, n_total) if hints else 0
        self.results, self.handles = {}, {}

    def run(self, q: str) -> int:
        aq, t = self.aq, self.table
        old = self.handles.pop(q, None)
        if old is not None:
            old.free()
        for col in self.results.pop(q, []):
            col.free()
        h = aq.Groups([t["sym"]], row_order=aq.ROWS_ASC, groups_hint=self.hint, key_bounds=self.bounds)
        # ASSUMING ASC time: ordered inside the window call (aqgpu_group_window_ordered; AQ_TRADE_TWO_CALLS=1: order_by + window)
        if q == "w_unsorted" and os.environ.get("AQ_TRADE_TWO_CALLS"):
            h.order_by([t["time_shuffled"]], desc=[0])
            self.results[q] = h.window(expressions(t), device_out=True)
        else:
            self.results[q] = h.window(expressions(t), device_out=True, order_by=t["time_shuffled"] if q == "w_unsorted" else None)
        self.handles[q] = h
        return h.ngroups

    def step(self) -> dict:
        return {q: self.run(q) for q in self.queries}

    def result_columns(self, q: str):
        return list(self.results[q])


def dist_plan(comm, table, queries, args, n_total, lo, hints=False):
    from . import dist
    return dist.DistWindowPlan(comm, table, queries, args, n_total, lo)
