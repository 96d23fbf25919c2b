"""ctypes binding of the CPU oracles declared in oracle/aqo.h.

TEST INFRASTRUCTURE ONLY.  Import this from tests/, from __graft_entry__.smoke() and from the
cpu_baseline / ``--impl reference`` legs of bench.py -- never from aquery2_b200/ (the product has no
CPU path and fails loudly when its CUDA library is missing).

Two libraries export the same symbols:
  * ``oracle/_ref/libaqref.so``  -- kind "reference": the reference's own C++ headers
    (server/hasher.h, server/table.h, server/aggregations.h) compiled by oracle/Makefile;
  * ``oracle/libaqoracle.so``    -- kind "port": the plain-C restatement aq_oracle.c.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_LIB = os.path.join(HERE, "_ref", "libaqref.so")
PORT_LIB = os.path.join(HERE, "libaqoracle.so")

# dtype codes = reference types::Type_t (server/aquery_types.h:1-5)
INT32, FLOAT, DOUBLE, INT64, INT128, UINT32 = 0, 1, 3, 5, 6, 11
# aqo_reduce_op
SUM, AVG, MIN, MAX, COUNT, FIRST, LAST, VAR, STDDEV, CORR, SECOND = range(11)
# aqo_scan_op
SUMS, AVGS, MINS, MAXS, SUMW, AVGW, MINW, MAXW, DELTAS, PREV, NEXT, RATIOW, VARS, STDDEVS, VARW, STDDEVW = range(16)

REDUCE_NAMES = {"sum": SUM, "avg": AVG, "min": MIN, "max": MAX, "count": COUNT, "first": FIRST,
                "last": LAST, "var": VAR, "stddev": STDDEV, "corr": CORR, "second": SECOND}
SCAN_NAMES = {"sums": SUMS, "avgs": AVGS, "mins": MINS, "maxs": MAXS, "sumw": SUMW, "avgw": AVGW,
              "minw": MINW, "maxw": MAXW, "deltas": DELTAS, "prev": PREV, "next": NEXT,
              "ratiow": RATIOW, "vars": VARS, "stddevs": STDDEVS, "varw": VARW, "stddevw": STDDEVW}

_NP2CODE = {np.dtype(np.int32): INT32, np.dtype(np.float32): FLOAT, np.dtype(np.float64): DOUBLE,
            np.dtype(np.int64): INT64, np.dtype(np.uint32): UINT32}
_CODE2NP = {INT32: np.int32, FLOAT: np.float32, DOUBLE: np.float64, INT64: np.int64,
            UINT32: np.uint32}


# key-only cell types (aqo_groupby): the reference's structs as numpy records (server/types.h:82-86,105-109,128-130)
STR, DATE, TIME, TIMESTAMP = 2, 8, 9, 18
DATE_DT = np.dtype([("day", "u1"), ("month", "u1"), ("year", "<i2")])
TIME_DT = np.dtype([("ms", "<u4"), ("seconds", "u1"), ("minutes", "u1"), ("hours", "u1"), ("pad", "u1")])
TIMESTAMP_DT = np.dtype([("date", DATE_DT), ("time", TIME_DT)])
assert DATE_DT.itemsize == 4 and TIME_DT.itemsize == 8 and TIMESTAMP_DT.itemsize == 12
_KEY_DT2CODE = {DATE_DT: DATE, TIME_DT: TIME, TIMESTAMP_DT: TIMESTAMP}


class StrColumn:
    """A string key column the way the reference holds it: an array of char pointers (ColRef<astring_view>)."""

    def __init__(self, strings):
        self.strings = [x if isinstance(x, bytes) else str(x).encode() for x in strings]
        self.bufs = [ctypes.create_string_buffer(x) for x in self.strings]      # NUL-terminated, distinct addresses
        self.ptrs = (ctypes.c_void_p * max(len(self.bufs), 1))(*[ctypes.addressof(b) for b in self.bufs])

    def __len__(self):
        return len(self.strings)

    def packed(self):
        """(bytes uint8[], offsets uint64[n + 1]): the Arrow-style layout aqgpu_string_dict takes"""
        lens = np.fromiter((len(x) for x in self.strings), np.uint64, len(self.strings))
        off = np.zeros(len(self.strings) + 1, np.uint64)
        np.cumsum(lens, out=off[1:])
        return np.frombuffer(b"".join(self.strings), np.uint8).copy(), off


def build(which: str = "all") -> None:
    """Run oracle/Makefile (``port``, ``ref`` or ``all``).  Building the checker is not using it."""
    subprocess.run(["make", "-s", "-C", HERE, which], check=True)


def i128_to_int64(cells: np.ndarray) -> np.ndarray:
    """(n,2) int64 [lo, hi] little-endian __int128 cells -> int64, asserting they fit."""
    cells = cells.reshape(-1, 2)
    lo, hi = cells[:, 0], cells[:, 1]
    if not np.array_equal(hi, lo >> 63):
        raise OverflowError("__int128 result does not fit in int64")
    return lo.copy()


def _ptr(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


class Oracle:
    def __init__(self, path: str):
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        self.path = path
        L = self.lib = ctypes.CDLL(path)
        L.aqo_kind.restype = ctypes.c_char_p
        L.aqo_groupby.restype = ctypes.c_int
        L.aqo_reduce.restype = ctypes.c_int
        L.aqo_scan.restype = ctypes.c_int
        L.aqo_reduce_out_type.restype = ctypes.c_int
        L.aqo_scan_out_type.restype = ctypes.c_int
        L.aqo_time_groupby_reduce.restype = ctypes.c_int64
        self.kind = L.aqo_kind().decode()

    # -- A1 + A2 -----------------------------------------------------------------------------
    def groupby(self, keys):
        keys = [k if isinstance(k, StrColumn) else np.ascontiguousarray(k) for k in keys]
        n = len(keys[0])
        kp = (ctypes.c_void_p * len(keys))(*[ctypes.addressof(k.ptrs) if isinstance(k, StrColumn) else k.ctypes.data for k in keys])
        kt = (ctypes.c_int * len(keys))(*[STR if isinstance(k, StrColumn) else _KEY_DT2CODE.get(k.dtype, None) or _NP2CODE[k.dtype]
                                          for k in keys])
        ng = ctypes.c_uint32(0)
        gid = np.empty(n, np.uint32)
        offsets = np.empty(n + 1, np.uint32)
        perm = np.empty(n, np.uint32)
        rc = self.lib.aqo_groupby(kp, kt, len(keys), ctypes.c_uint32(n), ctypes.byref(ng),
                                  _ptr(gid), _ptr(offsets), _ptr(perm))
        if rc != 0:
            raise RuntimeError(f"aqo_groupby rc={rc}")
        return ng.value, gid, offsets[: ng.value + 1].copy(), perm

    # -- A10 -------------------------------------------------------------------------------------
    def vec_binary(self, op: int, a, b):
        """a op b through the reference's own operators (server/vector_type.hpp:372-396, server/table.h:826-973);
        returns (cells, Type_t code of the result cells).  0-d operands are scalars."""
        a, b = np.asarray(a), np.asarray(b)
        sa, sb = int(a.ndim == 0), int(b.ndim == 0)
        a, b = np.ascontiguousarray(a.reshape(1) if sa else a), np.ascontiguousarray(b.reshape(1) if sb else b)
        n = len(b) if sa else len(a)
        out = np.zeros(2 * n + 2, np.int64)
        ot = ctypes.c_int(-1)
        sz = self.lib.aqo_vec_binary(op, _NP2CODE[a.dtype], _ptr(a), sa, _NP2CODE[b.dtype], _ptr(b), sb, ctypes.c_uint32(n),
                                     _ptr(out), ctypes.byref(ot))
        if sz < 0:
            raise RuntimeError(f"aqo_vec_binary rc={sz}")
        raw = out.view(np.uint8)[: sz * n]
        code = ot.value
        if code == INT128:
            return raw.view(np.int64).reshape(n, 2).copy(), code
        return raw.view({16: np.uint8, **_CODE2NP}[code]).copy(), code

    def vec_unary(self, op: int, x, param: int = 0):
        """sqrt (op 0) / truncate (op 1): server/aggregations.h:34-69"""
        x = np.ascontiguousarray(x)
        out = np.zeros(len(x), np.float64 if op == 0 else x.dtype)
        code = self.lib.aqo_vec_unary(op, _NP2CODE[x.dtype], _ptr(x), ctypes.c_uint32(param), ctypes.c_uint32(len(x)), _ptr(out))
        if code < 0:
            raise RuntimeError(f"aqo_vec_unary rc={code}")
        assert _CODE2NP[code] == out.dtype.type, (code, out.dtype)
        return out

    # -- A5 / A9 -----------------------------------------------------------------------------
    def reduce(self, op, v, offsets, perm, v2=None, raw128=False):
        if isinstance(op, str):
            op = REDUCE_NAMES[op]
        ng = len(offsets) - 1
        if op == COUNT:
            vt = INT32
        else:
            v = np.ascontiguousarray(v)
            vt = _NP2CODE[v.dtype]
        ot = self.lib.aqo_reduce_out_type(op, vt)
        if ot < 0:
            raise RuntimeError(f"unsupported reduce op={op} vtype={vt}")
        out = np.zeros((ng, 2), np.int64) if ot == INT128 else np.zeros(ng, _CODE2NP[ot])
        if v2 is not None:
            v2 = np.ascontiguousarray(v2)
        rc = self.lib.aqo_reduce(op, vt, _ptr(v) if op != COUNT else None, _ptr(v2),
                                 ctypes.c_uint32(ng), _ptr(offsets), _ptr(perm), _ptr(out))
        if rc != 0:
            raise RuntimeError(f"aqo_reduce rc={rc}")
        if ot == INT128 and not raw128:
            return i128_to_int64(out)
        return out

    # -- A6 / A7 / A8 ------------------------------------------------------------------------
    def scan(self, op, v, offsets, perm, w=0, raw128=False):
        if isinstance(op, str):
            op = SCAN_NAMES[op]
        v = np.ascontiguousarray(v)
        vt = _NP2CODE[v.dtype]
        ng = len(offsets) - 1
        n = int(offsets[-1]) if ng >= 0 and len(offsets) else 0
        ot = self.lib.aqo_scan_out_type(op, vt)
        if ot < 0:
            raise RuntimeError(f"unsupported scan op={op} vtype={vt}")
        out = np.zeros((n, 2), np.int64) if ot == INT128 else np.zeros(n, _CODE2NP[ot])
        rc = self.lib.aqo_scan(op, ctypes.c_uint32(w), vt, _ptr(v), ctypes.c_uint32(ng),
                               _ptr(offsets), _ptr(perm), _ptr(out))
        if rc != 0:
            raise RuntimeError(f"aqo_scan rc={rc}")
        if ot == INT128 and not raw128:
            return i128_to_int64(out)
        return out

    # -- the postproc_time bracket -----------------------------------------------------------
    def time_groupby_reduce(self, keys, aggs, want_outs=False):
        """aggs: list of (op, v[, v2]).  Returns (seconds, ngroups)."""
        keys = [np.ascontiguousarray(k) for k in keys]
        n = len(keys[0])
        kp = (ctypes.c_void_p * len(keys))(*[k.ctypes.data for k in keys])
        kt = (ctypes.c_int * len(keys))(*[_NP2CODE[k.dtype] for k in keys])
        na = len(aggs)
        ops = (ctypes.c_int * max(na, 1))()
        vts = (ctypes.c_int * max(na, 1))()
        vals = (ctypes.c_void_p * max(na, 1))()
        vals2 = (ctypes.c_void_p * max(na, 1))()
        keep = []
        for a, agg in enumerate(aggs):
            op = REDUCE_NAMES[agg[0]] if isinstance(agg[0], str) else agg[0]
            ops[a] = op
            if op == COUNT:
                vts[a] = INT32
                continue
            v = np.ascontiguousarray(agg[1]); keep.append(v)
            vts[a] = _NP2CODE[v.dtype]
            vals[a] = v.ctypes.data
            if len(agg) > 2 and agg[2] is not None:
                v2 = np.ascontiguousarray(agg[2]); keep.append(v2)
                vals2[a] = v2.ctypes.data
        outs_p, outs = None, []
        if want_outs:
            outs_p = (ctypes.c_void_p * max(na, 1))()
            for a in range(na):
                ot = self.lib.aqo_reduce_out_type(ops[a], vts[a])
                if ot < 0:
                    raise RuntimeError(f"unsupported reduce op={ops[a]} vtype={vts[a]}")
                o = np.zeros((n, 2), np.int64) if ot == INT128 else np.zeros(n, _CODE2NP[ot])
                outs.append(o)
                outs_p[a] = o.ctypes.data
        ng = ctypes.c_uint32(0)
        ns = self.lib.aqo_time_groupby_reduce(kp, kt, len(keys), ctypes.c_uint32(n), na, ops, vts,
                                              vals, vals2, outs_p, ctypes.byref(ng))
        if ns < 0:
            raise RuntimeError(f"aqo_time_groupby_reduce rc={ns}")
        if want_outs:
            return ns * 1e-9, ng.value, [o[: ng.value] for o in outs]
        return ns * 1e-9, ng.value

    def time_groupby_reduce_out(self, keys, aggs):
        """as time_groupby_reduce, also returning the aggregate columns (first-appearance group order)"""
        return self.time_groupby_reduce(keys, aggs, want_outs=True)


def load_port(build_if_missing: bool = True) -> Oracle:
    if not os.path.exists(PORT_LIB) and build_if_missing:
        build("port")
    return Oracle(PORT_LIB)


def load_ref() -> Oracle | None:
    """The reference-built oracle, or None when it has not been built (no /root/reference)."""
    return Oracle(REF_LIB) if os.path.exists(REF_LIB) else None
