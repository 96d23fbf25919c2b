"""codegen.py -- the whole-column (GPU) variant of the reference's group-by emitter.

The reference turns `SELECT ... GROUP BY ...` into a C++ translation unit whose post-processor loops over the
groups (`groupby_c.produce / finalize`, engine/ast.py:628-793; legacy template common/groupby.py:18-70):

    auto g = HashTableFactory<...>::get<...>(k0, k1);            for (i < g.size) { col1[i] = sum(v1[vecs[i]]); ... }

This module restates that emitter for the form a GPU wants: one statement per OUTPUT COLUMN over all groups,

    auto g = aq::group_by(k0, k1);      col0 = g.key<0>();      col1 = sum(v1[g]);      wb.add<double>(AQ_AVGW, 5, price);

against the drop-in headers of aquery2_b200/host (same `./server/*.h` include lines, same `extern "C" int
name(Context*)` entry, same `server->getCol(i, type)` column fetch and `new TableInfo<...>` result as
engine/ast.py:244-330,629,681-716 emit).  The SQL front end (`mo_parsing`) is not in this image, so the input
is the parsed query as a plain dict -- the shape `aquery_parser` produces for the select list, reduced to what
the emitter reads:

    {"name": "q3", "table": "source", "schema": [("id1", "int"), ...], "group_by": ["id3"],
     "select": [("id3", {"col": "id3"}), ("sum_v1", {"fn": "sum", "args": [{"col": "v1"}]}), ...]}

Expression nodes: {"col": name} | {"lit": number} | {"fn": name, "args": [...]} | {"op": "-", "l": node, "r": node}.
`python -m aquery2_b200.codegen q3` prints the translation unit of one of the sixteen built-in queries.
"""
from __future__ import annotations

CTYPE = {"int": "int", "float": "float", "double": "double", "long": "long long", "str": "astring_view",
         "date": "types::date_t", "time": "types::time_t", "timestamp": "types::timestamp_t"}
AQTYPE = {"int": "types::AINT32", "float": "types::AFLOAT", "double": "types::ADOUBLE", "long": "types::AINT64", "str": "types::ASTR",
          "date": "types::ADATE", "time": "types::ATIME", "timestamp": "types::ATIMESTAMP"}
HOST_ONLY = {"str"}        # cells that are host pointers: lowered (packed + dictionary-encoded) by the group-by itself, never uploaded as is

H2O_SCHEMA = [("id1", "int"), ("id2", "int"), ("id3", "int"), ("id4", "int"), ("id5", "int"), ("id6", "int"),
              ("v1", "int"), ("v2", "int"), ("v3", "float")]                   # benchmark/h2o/groupby_prep.sql:1
TRADE_SCHEMA = [("sym", "int"), ("time", "int"), ("quantity", "int"), ("price", "int")]   # datagen.cpp:90-107


def col(c):
    return {"col": c}


def fn(name, *args):
    return {"fn": name, "args": [a if isinstance(a, dict) else {"lit": a} for a in args]}


def _h2o(name, keys, select):
    return {"name": name, "table": "source", "schema": H2O_SCHEMA, "group_by": keys,
            "select": [(k, col(k)) for k in keys] + select}


def _trade(name, select, assuming=None):
    q = {"name": name, "table": "trade", "schema": TRADE_SCHEMA, "group_by": ["sym"], "row_order": "asc",
         "select": [("sym", col("sym"))] + select}
    if assuming:
        q["assuming"] = assuming           # (column, "asc" | "desc"): FROM trade ASSUMING ASC time (aquery_parser/parser.py:300-301)
    return q


# benchmark/h2o/groupby.sql:1-23 and the windowed shapes of benchmark/quries/Aquery/q7.a, q10.a, tests/moving_avg.a
QUERIES = {
    "q1": _h2o("q1", ["id1"], [("v1", fn("sum", col("v1")))]),
    "q2": _h2o("q2", ["id1", "id2"], [("v1", fn("sum", col("v1")))]),
    "q3": _h2o("q3", ["id3"], [("v1", fn("sum", col("v1"))), ("v3", fn("avg", col("v3")))]),
    "q4": _h2o("q4", ["id4"], [("v1", fn("avg", col("v1"))), ("v2", fn("avg", col("v2"))), ("v3", fn("avg", col("v3")))]),
    "q5": _h2o("q5", ["id6"], [("v1", fn("sum", col("v1"))), ("v2", fn("sum", col("v2"))), ("v3", fn("sum", col("v3")))]),
    "q6": _h2o("q6", ["id4", "id5"], [("sd_v3", fn("stddev", col("v3")))]),
    "q7": _h2o("q7", ["id3"], [("range_v1_v2", {"op": "-", "l": fn("max", col("v1")), "r": fn("min", col("v2"))})]),
    "q8": _h2o("q8", ["id6"], [("first_v3", fn("first", col("v3"))), ("second_v3", fn("second", col("v3")))]),
    "q9": _h2o("q9", ["id2", "id4"], [("r2", fn("pow", fn("corr", col("v1"), col("v2")), 2))]),
    "q10": _h2o("q10", ["id1", "id2", "id3", "id4", "id5", "id6"], [("v3", fn("sum", col("v3"))), ("cnt", fn("count", {"col": "*"}))]),
    "w1": _trade("w1", [("avg5", fn("avgs", 5, col("price")))]),
    "w2": _trade("w2", [("sum5", fn("sums", 5, col("quantity")))]),
    "w3": _trade("w3", [("d", fn("deltas", col("price")))]),
    "w4": _trade("w4", [("min5", fn("mins", 5, col("price")))]),
    "w5": _trade("w5", [("avg5", fn("avgs", 5, col("price"))), ("sum5", fn("sums", 5, col("quantity"))), ("d", fn("deltas", col("price"))),
                        ("min5", fn("mins", 5, col("price"))),
                        ("drawup", fn("max", {"op": "-", "l": col("price"), "r": fn("mins", col("price"))}))]),
    # the same over an UNSORTED table: ASSUMING ASC time orders every group's rows inside the window call
    "w6": _trade("w6", [("avg5", fn("avgs", 5, col("price"))), ("d", fn("deltas", col("price"))),
                        ("drawup", fn("max", {"op": "-", "l": col("price"), "r": fn("mins", col("price"))}))], assuming=("time", "asc")),
}

# Queries beyond integer keys, emitted and compiled against the drop-in headers by tests/test_codegen.py; their
# hand-written twins run on the device in aquery2_b200/host/selftest.cpp.  tests/network.a:8-11:
#   SELECT src, dst, count(*), avg(len) FROM network ASSUMING ASC src, ASC dst, ASC _time GROUP BY src, dst, sums(deltas(_time) > 120)
NETWORK_SCHEMA = [("src", "str"), ("dst", "str"), ("len", "int"), ("_time", "int")]
EXTRA_QUERIES = {
    "network": {"name": "network", "table": "network", "schema": NETWORK_SCHEMA,
                "group_by": ["src", "dst", {"name": "session", "expr": {"fn": "sums", "args": [{"op": ">", "l": {"fn": "deltas", "args": [{"col": "_time"}]}, "r": {"lit": 120}}]}}],
                "select": [("src", {"col": "src"}), ("dst", {"col": "dst"}), ("cnt", {"fn": "count", "args": [{"col": "*"}]}),
                           ("avg_len", {"fn": "avg", "args": [{"col": "len"}]})]},
    # benchmark/quries/Aquery/load_data.a:1 `stocksymbol STRING`: max(price) - min(price) by symbol and trade date
    "symrange": {"name": "symrange", "table": "ticks",
                 "schema": [("stocksymbol", "str"), ("tradedate", "date"), ("price", "int")],
                 "group_by": ["stocksymbol", "tradedate"],
                 "select": [("stocksymbol", {"col": "stocksymbol"}), ("tradedate", {"col": "tradedate"}),
                            ("range", {"op": "-", "l": {"fn": "max", "args": [{"col": "price"}]}, "r": {"fn": "min", "args": [{"col": "price"}]}})]},
}

SCALAR_AGGS = {"sum", "avg", "min", "max", "first", "last", "second", "stddev", "var"}
# SQL spelling -> (scan op without window, scan op with window, result type as a C++ type expression of the value type T)
VECTOR_AGGS = {"sums": ("AQ_SUMS", "AQ_SUMW", "types::GetLongType<{T}>"), "avgs": ("AQ_AVGS", "AQ_AVGW", "double"),
               "mins": ("AQ_MINS", "AQ_MINW", "{T}"), "maxs": ("AQ_MAXS", "AQ_MAXW", "{T}"),
               "deltas": ("AQ_DELTAS", None, "{T}"), "prev": ("AQ_PREV", None, "{T}"), "next": ("AQ_NEXT", None, "{T}"),
               "ratios": (None, "AQ_RATIOW", "types::GetFPType<{T}>"),
               "vars": ("AQ_VARS", "AQ_VARW", "double"), "stddevs": ("AQ_STDDEVS", "AQ_STDDEVW", "double")}


class Emitter:
    """One query -> one translation unit.  Mirrors the statement order of groupby_c.produce / finalize
    (engine/ast.py:628-793): fetch the columns, build the groups, size the result table, fill its columns."""

    def __init__(self, q: dict, device: bool = True, dump: bool = False):
        self.q, self.device, self.dump = q, device, dump
        self.schema = dict(q["schema"])
        self.colpos = {c: i for i, (c, _) in enumerate(q["schema"])}
        self.lines, self.used, self.tmp = [], [], 0
        self.batch = []                      # queued WindowBatch statements

    # ---- naming (engine/ast.py gives every column a suffixed C identifier: id1_a, v1_b, ...)
    def cname(self, c: str) -> str:
        if c not in self.used:
            self.used.append(c)
        return f"{c}_{chr(ord('a') + self.used.index(c))}"

    def fresh(self, stem: str) -> str:
        self.tmp += 1
        return f"{stem}{self.tmp}"

    def emit(self, line: str = ""):
        self.lines.append("    " + line if line else "")

    def key_names(self):
        return [k if isinstance(k, str) else k["name"] for k in self.q["group_by"]]

    # ---- expressions over WHOLE columns outside any group (a computed group key such as sums(deltas(_time) > 120)): the
    # element-wise operators and the non-grouped vector functions of the drop-in headers (table.h:826-973, aggregations.h)
    def plain(self, e: dict) -> str:
        if "lit" in e:
            return repr(e["lit"])
        if "col" in e:
            return self.cname(e["col"])
        if "op" in e:
            if e["op"] not in ("+", "-", "*", "/", ">", "<", ">=", "<=", "==", "!="):
                raise NotImplementedError(e["op"])
            return f"({self.plain(e['l'])} {e['op']} {self.plain(e['r'])})"
        f, args = e["fn"], e["args"]
        if f in VECTOR_AGGS or f in ("sqrt", "truncate"):
            return f"{f}({', '.join(self.plain(a) for a in args)})"
        raise NotImplementedError(f"{f} outside a group")

    # ---- expressions: returns a C++ expression whose value is a result column (ColRef / VecCol)
    def expr(self, e: dict) -> str:
        if "lit" in e:
            return repr(e["lit"])
        if "col" in e:
            if e["col"] in self.key_names():
                return f"g.key<{self.key_names().index(e['col'])}>()"
            raise ValueError(f"column {e['col']} is neither aggregated nor a group key (engine/ast.py:742-746 rejects it too)")
        if "op" in e:
            if e["op"] != "-":
                raise NotImplementedError(e["op"])
            return f"({self.expr(e['l'])} - {self.expr(e['r'])})"
        f, args = e["fn"], e["args"]
        if f == "count":
            return "g.counts()"                                            # count(*) is the row list's size (engine/ast.py:763)
        if f == "pow":
            return f"pow({self.expr(args[0])}, (double){args[1]['lit']})"
        if f == "corr":
            return f"corr({self.cname(args[0]['col'])}[g], {self.cname(args[1]['col'])}[g])"
        if f in ("max", "min") and ("op" in args[0] or args[0].get("fn") in VECTOR_AGGS):
            return self.vector(args[0], reduce=f)                          # max(price - mins(price)): folded into the batch
        if f in SCALAR_AGGS:
            return f"{f}({self.cname(args[0]['col'])}[g])"
        if f in VECTOR_AGGS:
            return self.vector(e)
        raise NotImplementedError(f)

    def vector(self, e: dict, reduce: str | None = None) -> str:
        """queue one vector-valued expression on the query's WindowBatch; returns the name of its result column"""
        minus = False
        if "op" in e:                                                      # v - scan(v)
            if e["op"] != "-" or "col" not in e["l"] or e["r"].get("fn") not in VECTOR_AGGS or e["r"]["args"][-1].get("col") != e["l"]["col"]:
                raise NotImplementedError("only `col - scan(col)` is folded into the window call")
            minus, e = True, e["r"]
        f, args = e["fn"], e["args"]
        w = args[0]["lit"] if len(args) == 2 else 0
        c = args[-1]["col"]
        op0, opw, rt = VECTOR_AGGS[f]
        op = opw if len(args) == 2 or op0 is None else op0
        if f == "ratios" and len(args) == 1:
            w = 1
        T = CTYPE[self.schema[c]]
        name = self.fresh("w")
        if reduce:
            self.batch.append(f"auto& {name} = *new auto(wb.add_reduced<{T}>({op}, {w}, {self.cname(c)}, {str(minus).lower()}, AQ_{reduce.upper()}));")
        else:
            self.batch.append(f"auto& {name} = *new auto(wb.add<{rt.format(T=T)}>({op}, {w}, {self.cname(c)}, {str(minus).lower()}));")
        return name

    def generate(self) -> str:
        q = self.q
        # the select list first: it decides which columns are fetched
        body = []
        for i, (name, e) in enumerate(q["select"]):
            body.append((name, self.expr(e)))
        computed = []                        # (C++ name, expression) of group keys that are expressions over whole columns
        key_vars = []
        for k in q["group_by"]:
            if isinstance(k, str):
                key_vars.append(self.cname(k))
            else:
                computed.append((f"key_{k['name']}", self.plain(k["expr"])))
                key_vars.append(f"key_{k['name']}")
        if q.get("assuming"):
            self.cname(q["assuming"][0])
        out = ['#include "./server/libaquery.h"', '#include "./server/monetdb_conn.h"', '#include "./server/aggregations.h"',
               '#include "./server/hasher.h"', "",
               f"// {q['name']}: whole-column post-processor generated by aquery2_b200/codegen.py (engine/ast.py:628-793 restated)",
               f'extern "C" int dll_{q["name"]}(Context* cxt) {{']
        self.emit("using namespace std;")
        self.emit("using namespace types;")
        self.emit("auto server = static_cast<DataSource*>(cxt->curr_server);")
        self.emit("auto len = server->cnt;")
        for c in self.used:
            ty = self.schema[c]
            dev = ".to_device()" if self.device and ty not in HOST_ONLY else ""
            self.emit(f"auto {self.cname(c)} = ColRef<{CTYPE[ty]}>(len, server->getCol({self.colpos[c]}, {AQTYPE[ty]})){dev};")
        for name, ex in computed:
            self.emit(f"auto {name} = {ex};")
        keys = ", ".join(key_vars)
        if q.get("row_order") == "asc":
            note = "ties of the ASSUMING order keep table order" if q.get("assuming") else "table sorted by time: row lists in time order"
            self.emit(f"auto g = aq::group_by_rows_ascending({keys});     // {note}")
        else:
            self.emit(f"auto g = aq::group_by({keys});")
        if self.batch:
            self.emit("aq::WindowBatch wb(g);")
            if q.get("assuming"):
                oc, direction = q["assuming"]
                self.emit(f"wb.assuming({self.cname(oc)}, {str(direction == 'desc').lower()});     // ASSUMING {direction.upper()} {oc}: TableInfo::order_by per group, inside the call")
            for b in self.batch:
                self.emit(b)
            self.emit("wb.run();")
        cols = []
        for i, (name, ex) in enumerate(body):
            v = f"c{i}"
            self.emit(f"auto& {v} = *new auto({ex});" if not ex.startswith("w") or not ex[1:].isdigit() else f"auto& {v} = {ex};")
            cols.append(v)
        # vector-valued columns become vector_type cells over the flat CSR buffer (engine/ast.py:696-718)
        cells = []
        for v in cols:
            self.emit(f"auto& {v}_cells = *new auto(aq::table_cells({v}));")
            cells.append(f"{v}_cells")
        names = ", ".join(f'"{n}"' for n, _ in body)
        self.emit(f"const char* names[] = {{{names}}};")
        types_ = ", ".join(f"typename decays<decltype({c})>::value_t" for c in cells)
        self.emit(f'auto out = new TableInfo<{types_}>("out_{q["name"]}", names);')
        for i, c in enumerate(cells):
            self.emit(f"out->template get_col<{i}>().initfrom({c}, names[{i}]);")
        self.emit('out->print(",", "\\n", 8);' if not self.dump else 'out->print(",", "\\n", 8); aq::dump_table(*out, "out");')
        self.emit(f'cxt->tables["out_{q["name"]}"] = out;')
        self.emit("return 0;")
        return "\n".join(out + self.lines + ["}", ""])


def generate(name_or_query, device: bool = True, dump: bool = False) -> str:
    q = (QUERIES.get(name_or_query) or EXTRA_QUERIES[name_or_query]) if isinstance(name_or_query, str) else name_or_query
    return Emitter(q, device=device, dump=dump).generate()


if __name__ == "__main__":
    import sys
    for n in sys.argv[1:] or list(QUERIES):
        print(generate(n))
