"""ctypes binding of libaqgpu.so (include/aqgpu.h) -- plumbing for tests, bench.py and smoke().

The product is the C-ABI library plus the C++ drop-in headers under aquery2_b200/host/; this module
only marshals numpy arrays / device columns into the plain-pointer calls.  There is NO CPU fallback:
importing works anywhere, but every compute call fails loudly when the library or a CUDA device is
missing.  Nothing here imports or calls oracle/.
"""
from __future__ import annotations

import ctypes
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libaqgpu.so")

# dtype codes = reference types::Type_t (server/aquery_types.h:1-5)
INT32, FLOAT, DOUBLE, INT64, INT128, UINT32 = 0, 1, 3, 5, 6, 11
SUM, AVG, MIN, MAX, COUNT, FIRST, LAST, VAR, STDDEV, CORR, SECOND, SUMSQ, SUMXY = range(13)
SUMS, AVGS, MINS, MAXS, SUMW, AVGW, MINW, MAXW, DELTAS, PREV, NEXT, RATIOW, VARS, STDDEVS, VARW, STDDEVW = range(16)
ROWS_DESC, ROWS_ASC = 0, 1

REDUCE_NAMES = {"sum": SUM, "avg": AVG, "min": MIN, "max": MAX, "count": COUNT, "first": FIRST,
                "last": LAST, "var": VAR, "stddev": STDDEV, "corr": CORR, "second": SECOND, "sumsq": SUMSQ,
                "sumxy": SUMXY}
SCAN_NAMES = {"sums": SUMS, "avgs": AVGS, "mins": MINS, "maxs": MAXS, "sumw": SUMW, "avgw": AVGW,
              "minw": MINW, "maxw": MAXW, "deltas": DELTAS, "prev": PREV, "next": NEXT,
              "ratiow": RATIOW, "vars": VARS, "stddevs": STDDEVS, "varw": VARW, "stddevw": STDDEVW}

NP2CODE = {np.dtype(np.int32): INT32, np.dtype(np.float32): FLOAT, np.dtype(np.float64): DOUBLE,
           np.dtype(np.int64): INT64, np.dtype(np.uint32): UINT32}
CODE2NP = {INT32: np.dtype(np.int32), FLOAT: np.dtype(np.float32), DOUBLE: np.dtype(np.float64),
           INT64: np.dtype(np.int64), UINT32: np.dtype(np.uint32)}
CODE_SIZE = {INT32: 4, FLOAT: 4, DOUBLE: 8, INT64: 8, INT128: 16, UINT32: 4}

EXPORTS = [
    "aqgpu_abi_version", "aqgpu_init", "aqgpu_shutdown", "aqgpu_last_error", "aqgpu_set_stream",
    "aqgpu_sync", "aqgpu_launch_count", "aqgpu_timer_begin", "aqgpu_timer_end",
    "aqgpu_profile_enable", "aqgpu_profile_tag", "aqgpu_profile_report",
    "aqgpu_col_alloc", "aqgpu_col_free", "aqgpu_col_upload", "aqgpu_col_download",
    "aqgpu_host_alloc", "aqgpu_host_free", "aqgpu_ptr_is_device", "aqgpu_group_gather", "aqgpu_group_window",
    "aqgpu_reduce_out_type", "aqgpu_scan_out_type", "aqgpu_dtype_size",
    "aqgpu_groupby_agg", "aqgpu_groupby_build", "aqgpu_col_minmax", "aqgpu_groups_free",
    "aqgpu_groups_size", "aqgpu_groups_keys", "aqgpu_groups_first_row", "aqgpu_groups_gid",
    "aqgpu_groups_csr", "aqgpu_groups_device_views", "aqgpu_group_reduce", "aqgpu_group_scan",
    "aqgpu_groups_order_by", "aqgpu_vec_sub", "aqgpu_avg_finalize", "aqgpu_vec_pow", "aqgpu_finalize",
    "aqgpu_rows_to_global", "aqgpu_key_owner", "aqgpu_second_merge", "aqgpu_partition_rows",
    "aqgpu_synth_column", "aqgpu_groupby_run", "aqgpu_result_size", "aqgpu_result_key", "aqgpu_result_agg",
    "aqgpu_result_first_row", "aqgpu_result_free", "aqgpu_trim_pool", "aqgpu_col_upload_async",
    "aqgpu_col_download_async",
    "aqgpu_comm_unique_id", "aqgpu_comm_init", "aqgpu_comm_destroy", "aqgpu_comm_size", "aqgpu_groupby_run_dist",
    "aqgpu_result_replicated", "aqgpu_group_median", "aqgpu_repartition_rows", "aqgpu_group_window_ordered",
    "aqgpu_partition_counts_noalloc", "aqgpu_key_cell_size", "aqgpu_key_split", "aqgpu_key_join", "aqgpu_string_dict",
    "aqgpu_vec_out_type", "aqgpu_vec_binary", "aqgpu_take", "aqgpu_vec_unary",
    "aqgpu_ingest_async", "aqgpu_ingest_fence", "aqgpu_writeback_async", "aqgpu_writeback_fence", "aqgpu_writeback_wait",
]


class AqgpuError(RuntimeError):
    pass


class WindowSpec(ctypes.Structure):
    """struct aqgpu_window (include/aqgpu.h)"""
    _fields_ = [("op", ctypes.c_int), ("w", ctypes.c_uint32), ("vtype", ctypes.c_int), ("v", ctypes.c_void_p),
                ("minus_from_v", ctypes.c_int), ("reduce", ctypes.c_int), ("out_type", ctypes.c_int), ("out", ctypes.c_void_p)]


class AggSpec(ctypes.Structure):
    """struct aqgpu_agg (include/aqgpu.h)."""
    _fields_ = [("op", ctypes.c_int), ("vtype", ctypes.c_int), ("v", ctypes.c_void_p),
                ("v2", ctypes.c_void_p), ("out_type", ctypes.c_int), ("out", ctypes.c_void_p)]


_lib = None


def lib() -> ctypes.CDLL:
    """The loaded C-ABI library.  Raises if it has not been built (no fallback of any kind)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise AqgpuError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                             "(or make -C aquery2_b200/csrc). libaqgpu has no CPU fallback.")
        L = ctypes.CDLL(LIB_PATH)
        L.aqgpu_last_error.restype = ctypes.c_char_p
        L.aqgpu_launch_count.restype = ctypes.c_uint64
        L.aqgpu_dtype_size.restype = ctypes.c_size_t
        for name in EXPORTS:
            getattr(L, name)              # AttributeError here = header/library mismatch
        _lib = L
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        raise AqgpuError(lib().aqgpu_last_error().decode(errors="replace"))


def init(device: int = -1) -> None:
    check(lib().aqgpu_init(device))


def sync() -> None:
    check(lib().aqgpu_sync())


def trim_pool() -> None:
    """give the unused memory of the library's stream-ordered pool back to the device (synchronises)"""
    check(lib().aqgpu_trim_pool())


def launch_count() -> int:
    return int(lib().aqgpu_launch_count())


def set_stream(cuda_stream_handle: int | None) -> None:
    check(lib().aqgpu_set_stream(ctypes.c_void_p(cuda_stream_handle or 0)))


def timer_begin() -> None:
    check(lib().aqgpu_timer_begin())


def timer_end() -> float:
    ms = ctypes.c_float(0)
    check(lib().aqgpu_timer_end(ctypes.byref(ms)))
    return float(ms.value)


def profile_enable(on: bool) -> None:
    check(lib().aqgpu_profile_enable(int(on)))


def profile_tag(tag: str) -> None:
    check(lib().aqgpu_profile_tag(tag.encode()))


def profile_report() -> dict[str, tuple[int, float]]:
    """{"tag:kernel": (calls, total_ms)} since the last report (synchronises the stream)."""
    buf = ctypes.create_string_buffer(1 << 16)
    check(lib().aqgpu_profile_report(buf, ctypes.c_size_t(len(buf))))
    out = {}
    for line in buf.value.decode().splitlines():
        name, calls, ms = line.rsplit("\t", 2)
        out[name] = (int(calls), float(ms))
    return out


class DeviceColumn:
    """One HBM-resident column of the column store (what a device ColRef::container points at)."""

    def __init__(self, n: int, code: int):
        self.n, self.code = int(n), int(code)
        self.itemsize = CODE_SIZE[code]
        p = ctypes.c_void_p(0)
        check(lib().aqgpu_col_alloc(ctypes.c_size_t(max(self.nbytes, 16)), ctypes.byref(p)))
        self.ptr = p.value

    @property
    def nbytes(self) -> int:
        return self.n * self.itemsize

    @classmethod
    def from_numpy(cls, a: np.ndarray) -> "DeviceColumn":
        a = np.ascontiguousarray(a)
        if a.ndim == 2 and a.shape[1] == 2 and a.dtype == np.int64:
            c = cls(a.shape[0], INT128)
        else:
            c = cls(a.size, NP2CODE[a.dtype])
        if c.nbytes:
            check(lib().aqgpu_col_upload(ctypes.c_void_p(c.ptr), a.ctypes.data_as(ctypes.c_void_p),
                                         ctypes.c_size_t(c.nbytes)))
        return c

    def upload(self, a: np.ndarray) -> None:
        a = np.ascontiguousarray(a)
        assert a.nbytes <= self.nbytes
        check(lib().aqgpu_col_upload(ctypes.c_void_p(self.ptr), a.ctypes.data_as(ctypes.c_void_p),
                                     ctypes.c_size_t(a.nbytes)))

    def numpy(self, n: int | None = None) -> np.ndarray:
        n = self.n if n is None else int(n)
        out = np.empty((n, 2), np.int64) if self.code == INT128 else np.empty(n, CODE2NP[self.code])
        if n:
            check(lib().aqgpu_col_download(out.ctypes.data_as(ctypes.c_void_p), ctypes.c_void_p(self.ptr),
                                           ctypes.c_size_t(n * self.itemsize)))
        return out

    def free(self) -> None:
        if getattr(self, "ptr", None):
            lib().aqgpu_col_free(ctypes.c_void_p(self.ptr))
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def _col(x):
    """(pointer, dtype code, keepalive) of a numpy array or DeviceColumn."""
    if isinstance(x, DeviceColumn):
        return x.ptr, x.code, x
    a = np.ascontiguousarray(x)
    return a.ctypes.data, NP2CODE[a.dtype], a


def _out_buf(code: int, n: int, device: bool):
    if device:
        return DeviceColumn(n, code)
    return np.zeros((n, 2), np.int64) if code == INT128 else np.zeros(n, CODE2NP[code])


def _out_ptr(b):
    return b.ptr if isinstance(b, DeviceColumn) else b.ctypes.data


def i128_to_int64(cells: np.ndarray) -> np.ndarray:
    cells = cells.reshape(-1, 2)
    lo, hi = cells[:, 0], cells[:, 1]
    if not np.array_equal(hi, lo >> 63):
        raise OverflowError("__int128 result does not fit in int64")
    return lo.copy()


def reduce_out_type(op: int, vtype: int) -> int:
    return int(lib().aqgpu_reduce_out_type(op, vtype))


def scan_out_type(op: int, vtype: int) -> int:
    return int(lib().aqgpu_scan_out_type(op, vtype))


def col_minmax(col) -> tuple[int, int]:
    p, code, keep = _col(col)
    n = col.n if isinstance(col, DeviceColumn) else keep.size
    lo, hi = ctypes.c_int64(0), ctypes.c_int64(0)
    check(lib().aqgpu_col_minmax(code, ctypes.c_void_p(p), ctypes.c_uint32(n), ctypes.byref(lo), ctypes.byref(hi)))
    return lo.value, hi.value


class GroupByResult:
    def __init__(self, ngroups, keys, first_row, aggs):
        self.ngroups, self.keys, self.first_row, self.aggs = ngroups, keys, first_row, aggs


def groupby_agg(keys, aggs, row_order: int = ROWS_DESC, groups_hint: int = 0, max_groups: int = 0,
                key_bounds=None, device_out: bool = False, want_keys: bool = True,
                want_first_row: bool = True, out_types=None) -> GroupByResult:
    """aqgpu_groupby_agg.  keys: list of columns; aggs: list of (op, v[, v2]) with op a name or code.

    Columns are numpy arrays (host pointers: the call stages them itself) or DeviceColumn objects.
    Results are numpy arrays (or DeviceColumn when device_out) sized max_groups, trimmed to ngroups
    for numpy outputs.  __int128 cells come back as (n, 2) int64 [lo, hi].
    """
    L = lib()
    nk = len(keys)
    kc = [_col(k) for k in keys]
    nrows = keys[0].n if isinstance(keys[0], DeviceColumn) else kc[0][2].size
    kp = (ctypes.c_void_p * nk)(*[c[0] for c in kc])
    kt = (ctypes.c_int * nk)(*[c[1] for c in kc])
    cap = max_groups or nrows
    na = len(aggs)
    specs = (AggSpec * max(na, 1))()
    keep, outs = [], []
    for a, agg in enumerate(aggs):
        op = REDUCE_NAMES[agg[0]] if isinstance(agg[0], str) else int(agg[0])
        v = agg[1] if len(agg) > 1 else None
        v2 = agg[2] if len(agg) > 2 else None
        vt, vp, v2p = INT32, None, None
        if op != COUNT:
            vp, vt, k1 = _col(v)
            keep.append(k1)
            if v2 is not None:
                v2p, _, k2 = _col(v2)
                keep.append(k2)
        ot = (out_types[a] if out_types and out_types[a] else 0) or reduce_out_type(op, vt)
        if ot < 0:
            raise AqgpuError(f"unsupported aggregate op={op} on dtype={vt}")
        ob = _out_buf(ot, cap, device_out)
        outs.append(ob)
        specs[a].op, specs[a].vtype, specs[a].v, specs[a].v2 = op, vt, vp, v2p
        specs[a].out_type, specs[a].out = ot, _out_ptr(ob)
    kouts = [_out_buf(kc[k][1], cap, device_out) for k in range(nk)] if want_keys else []
    okp = (ctypes.c_void_p * nk)(*[_out_ptr(b) for b in kouts]) if want_keys else None
    fr = _out_buf(UINT32, cap, device_out) if want_first_row else None
    kmin = kmax = None
    if key_bounds is not None:
        kmin = (ctypes.c_int64 * nk)(*[int(b[0]) for b in key_bounds])
        kmax = (ctypes.c_int64 * nk)(*[int(b[1]) for b in key_bounds])
    ng = ctypes.c_uint32(0)
    check(L.aqgpu_groupby_agg(kp, kt, nk, ctypes.c_uint32(nrows), kmin, kmax, row_order,
                              ctypes.c_uint32(groups_hint), specs, na, ctypes.c_uint32(cap), okp,
                              ctypes.c_void_p(_out_ptr(fr)) if fr is not None else None, ctypes.byref(ng)))
    n = ng.value
    if not device_out:
        outs = [o[:n] for o in outs]
        kouts = [o[:n] for o in kouts]
        fr = fr[:n] if fr is not None else None
    return GroupByResult(n, kouts, fr, outs)


class Groups:
    """aqgpu_groups_t handle == the reference's HashTableComponents {size, keys, values, offsets}
    (server/hasher.h:201-207), resident in HBM."""

    def __init__(self, keys, row_order: int = ROWS_DESC, groups_hint: int = 0, key_bounds=None):
        L = lib()
        nk = len(keys)
        kc = [_col(k) for k in keys]
        self.nrows = keys[0].n if isinstance(keys[0], DeviceColumn) else kc[0][2].size
        self.key_codes = [c[1] for c in kc]
        kp = (ctypes.c_void_p * nk)(*[c[0] for c in kc])
        kt = (ctypes.c_int * nk)(*self.key_codes)
        kmin = kmax = None
        if key_bounds is not None:
            kmin = (ctypes.c_int64 * nk)(*[int(b[0]) for b in key_bounds])
            kmax = (ctypes.c_int64 * nk)(*[int(b[1]) for b in key_bounds])
        h = ctypes.c_void_p(0)
        check(L.aqgpu_groupby_build(kp, kt, nk, ctypes.c_uint32(self.nrows), kmin, kmax, row_order,
                                    ctypes.c_uint32(groups_hint), ctypes.byref(h)))
        self.h = h
        ng = ctypes.c_uint32(0)
        check(L.aqgpu_groups_size(self.h, ctypes.byref(ng)))
        self.ngroups = ng.value

    def free(self):
        if getattr(self, "h", None):
            lib().aqgpu_groups_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass

    def keys(self, k: int) -> np.ndarray:
        out = np.empty(self.ngroups, CODE2NP[self.key_codes[k]])
        if self.ngroups:
            check(lib().aqgpu_groups_keys(self.h, k, out.ctypes.data_as(ctypes.c_void_p)))
        return out

    def first_row(self) -> np.ndarray:
        out = np.empty(self.ngroups, np.uint32)
        if self.ngroups:
            check(lib().aqgpu_groups_first_row(self.h, out.ctypes.data_as(ctypes.c_void_p)))
        return out

    def gid(self) -> np.ndarray:
        out = np.empty(self.nrows, np.uint32)
        if self.nrows:
            check(lib().aqgpu_groups_gid(self.h, out.ctypes.data_as(ctypes.c_void_p)))
        return out

    def csr(self):
        off = np.zeros(self.ngroups + 1, np.uint32)
        perm = np.empty(self.nrows, np.uint32)
        check(lib().aqgpu_groups_csr(self.h, off.ctypes.data_as(ctypes.c_void_p),
                                     perm.ctypes.data_as(ctypes.c_void_p) if self.nrows else None))
        return off, perm

    def reduce(self, op, v=None, v2=None, out_type: int = 0, device_out: bool = False):
        op = REDUCE_NAMES[op] if isinstance(op, str) else int(op)
        vt, vp, v2p, keep = INT32, None, None, []
        if op != COUNT:
            vp, vt, k1 = _col(v)
            keep.append(k1)
            if v2 is not None:
                v2p, _, k2 = _col(v2)
                keep.append(k2)
        ot = out_type or reduce_out_type(op, vt)
        if ot < 0:
            raise AqgpuError(f"unsupported aggregate op={op} on dtype={vt}")
        ob = _out_buf(ot, self.ngroups, device_out)
        check(lib().aqgpu_group_reduce(self.h, op, vt, ctypes.c_void_p(vp), ctypes.c_void_p(v2p), ot,
                                       ctypes.c_void_p(_out_ptr(ob))))
        return ob

    def scan(self, op, v, w: int = 0, out_type: int = 0, device_out: bool = False):
        op = SCAN_NAMES[op] if isinstance(op, str) else int(op)
        vp, vt, keep = _col(v)
        ot = out_type or scan_out_type(op, vt)
        if ot < 0:
            raise AqgpuError(f"unsupported scan op={op} on dtype={vt}")
        ob = _out_buf(ot, self.nrows, device_out)
        check(lib().aqgpu_group_scan(self.h, op, ctypes.c_uint32(w), vt, ctypes.c_void_p(vp), ot,
                                     ctypes.c_void_p(_out_ptr(ob))))
        return ob

    def median(self, v):
        """aqgpu_group_median: fp64 median of v per group (not a reference aggregate: checked against numpy)"""
        vp, vt, keep = _col(v)
        out = np.zeros(self.ngroups, np.float64)
        check(lib().aqgpu_group_median(self.h, vt, ctypes.c_void_p(vp), out.ctypes.data_as(ctypes.c_void_p)))
        return out

    def window(self, exprs, device_out: bool = False, order_by=None, desc: bool = False):
        """Several vector-valued expressions in one call (aqgpu_group_window).  exprs: tuples
        (op, v[, w[, minus_from_v[, reduce]]]) with op a scan name, reduce None / "max" / "min".
        Returns one output per expression: nrows cells in CSR order, or ngroups cells when reduced."""
        n = len(exprs)
        arr = (WindowSpec * n)()
        keep, outs = [], []
        for i, e in enumerate(exprs):
            op = SCAN_NAMES[e[0]] if isinstance(e[0], str) else int(e[0])
            vp, vt, k = _col(e[1])
            keep.append(k)
            w = int(e[2]) if len(e) > 2 and e[2] else 0
            minus = int(bool(e[3])) if len(e) > 3 else 0
            red = e[4] if len(e) > 4 else None
            red = -1 if red is None else (REDUCE_NAMES[red] if isinstance(red, str) else int(red))
            ot = scan_out_type(op, vt)
            if ot < 0:
                raise AqgpuError(f"unsupported scan op={op} on dtype={vt}")
            ob = _out_buf(ot, self.ngroups if red >= 0 else self.nrows, device_out)
            outs.append(ob)
            arr[i] = WindowSpec(op, w, vt, vp, minus, red, ot, _out_ptr(ob))
        if order_by is None:
            check(lib().aqgpu_group_window(self.h, arr, n))
        else:                               # ASSUMING order evaluated inside the call (aqgpu_group_window_ordered)
            op_, ot_, k_ = _col(order_by)
            keep.append(k_)
            check(lib().aqgpu_group_window_ordered(self.h, ctypes.c_void_p(op_), ot_, int(bool(desc)), arr, n))
        return outs

    def order_by(self, cols, desc=None):
        nc = len(cols)
        cc = [_col(c) for c in cols]
        cp = (ctypes.c_void_p * nc)(*[c[0] for c in cc])
        ct = (ctypes.c_int * nc)(*[c[1] for c in cc])
        d = (ctypes.c_int * nc)(*[int(bool(x)) for x in (desc or [0] * nc)])
        check(lib().aqgpu_groups_order_by(self.h, cp, ct, d, nc))


def synth_column(kind: int, key0: int, n: int, k: int, row_lo: int = 0, key1: int = 0, key2: int = 0,
                 aux_k: int = 0, code: int = INT32) -> DeviceColumn:
    out = DeviceColumn(n, code)
    check(lib().aqgpu_synth_column(kind, ctypes.c_uint64(key0), ctypes.c_uint64(key1), ctypes.c_uint64(key2),
                                   ctypes.c_uint32(k), ctypes.c_uint32(aux_k), ctypes.c_uint64(row_lo),
                                   ctypes.c_uint32(n), ctypes.c_void_p(out.ptr)))
    return out


class PreparedQuery:
    """One aqgpu_groupby_agg call with its argument block and output columns built once, so that a
    timed loop does nothing but the C-ABI call.  Inputs/outputs may be DeviceColumn (HBM-resident
    column store) or numpy arrays (host pointers: the call stages them itself, H2D + D2H)."""

    def __init__(self, keys, aggs, max_groups: int, groups_hint: int = 0, key_bounds=None,
                 row_order: int = ROWS_DESC, device_out: bool = True, out_types=None,
                 want_keys: bool = True, want_first_row: bool = False, host_alloc=None):
        self.nk = nk = len(keys)
        self._keep = []
        kc = [_col(k) for k in keys]
        self._keep += [c[2] for c in kc]
        self.nrows = keys[0].n if isinstance(keys[0], DeviceColumn) else kc[0][2].size
        self.kp = (ctypes.c_void_p * nk)(*[c[0] for c in kc])
        self.kt = (ctypes.c_int * nk)(*[c[1] for c in kc])
        self.cap = int(max_groups)
        self.groups_hint = int(groups_hint)
        self.row_order = row_order
        self.na = na = len(aggs)
        self.specs = (AggSpec * max(na, 1))()
        self.outs = []
        self.in_bytes = sum(CODE_SIZE[c[1]] for c in kc) * self.nrows
        mk = (lambda code, n: _out_buf(code, n, True)) if device_out else \
             (host_alloc or (lambda code, n: _out_buf(code, n, False)))
        seen = {c[0] for c in kc}
        for a, agg in enumerate(aggs):
            op = REDUCE_NAMES[agg[0]] if isinstance(agg[0], str) else int(agg[0])
            vt, vp, v2p = INT32, None, None
            if op != COUNT:
                vp, vt, k1 = _col(agg[1])
                self._keep.append(k1)
                if vp not in seen:
                    seen.add(vp); self.in_bytes += CODE_SIZE[vt] * self.nrows
                if len(agg) > 2 and agg[2] is not None:
                    v2p, _, k2 = _col(agg[2])
                    self._keep.append(k2)
                    if v2p not in seen:
                        seen.add(v2p); self.in_bytes += CODE_SIZE[vt] * self.nrows
            ot = (out_types[a] if out_types and out_types[a] else 0) or reduce_out_type(op, vt)
            ob = mk(ot, self.cap)
            self.outs.append(ob)
            s = self.specs[a]
            s.op, s.vtype, s.v, s.v2, s.out_type, s.out = op, vt, vp, v2p, ot, _out_ptr(ob)
        self.kouts = [mk(kc[k][1], self.cap) for k in range(nk)] if want_keys else []
        self.okp = (ctypes.c_void_p * nk)(*[_out_ptr(b) for b in self.kouts]) if want_keys else None
        self.fr = mk(UINT32, self.cap) if want_first_row else None
        self.kmin = self.kmax = None
        if key_bounds is not None:
            self.kmin = (ctypes.c_int64 * nk)(*[int(b[0]) for b in key_bounds])
            self.kmax = (ctypes.c_int64 * nk)(*[int(b[1]) for b in key_bounds])
        self.ng = ctypes.c_uint32(0)
        self._fn = lib().aqgpu_groupby_agg

    def out_bytes(self, ngroups: int) -> int:
        b = sum(CODE_SIZE[self.specs[a].out_type] for a in range(self.na))
        b += sum(CODE_SIZE[self.kt[k]] for k in range(self.nk)) if self.kouts else 0
        b += 4 if self.fr is not None else 0
        return b * ngroups

    def run(self) -> int:
        rc = self._fn(self.kp, self.kt, self.nk, ctypes.c_uint32(self.nrows), self.kmin, self.kmax,
                      self.row_order, ctypes.c_uint32(self.groups_hint), self.specs, self.na,
                      ctypes.c_uint32(self.cap), self.okp,
                      ctypes.c_void_p(_out_ptr(self.fr)) if self.fr is not None else None,
                      ctypes.byref(self.ng))
        check(rc)
        return self.ng.value


class ColumnView(DeviceColumn):
    """non-owning DeviceColumn over memory that belongs to something else (a result set)"""

    def __init__(self, ptr: int, n: int, code: int, keep=None):   # noqa: super().__init__ allocates; this must not
        self.n, self.code, self.itemsize = int(n), int(code), CODE_SIZE[code]
        self.ptr, self.keep = ptr, keep

    def free(self):
        self.ptr, self.keep = None, None


class AutoQuery:
    """aqgpu_groupby_run with its argument block built once: no bounds, no hints, no capacities -- the library
    samples the key columns inside the call, picks a path and allocates the result columns itself.  After
    run(), ``kouts`` / ``outs`` / ``fr`` are views of the result set (valid until the next run / free)."""

    def __init__(self, keys, aggs, row_order: int = ROWS_DESC, want_first_row: bool = False, key_bounds=None,
                 groups_hint: int = 0):
        self.nk = nk = len(keys)
        self._keep = []
        kc = [_col(k) for k in keys]
        self._keep += [c[2] for c in kc]
        self.nrows = keys[0].n if isinstance(keys[0], DeviceColumn) else kc[0][2].size
        self.kp = (ctypes.c_void_p * nk)(*[c[0] for c in kc])
        self.kt = (ctypes.c_int * nk)(*[c[1] for c in kc])
        self.key_codes = [c[1] for c in kc]
        self.row_order, self.want_fr, self.groups_hint = row_order, int(want_first_row), int(groups_hint)
        self.na = na = len(aggs)
        self.specs = (AggSpec * max(na, 1))()
        for a, agg in enumerate(aggs):
            op = REDUCE_NAMES[agg[0]] if isinstance(agg[0], str) else int(agg[0])
            vt, vp, v2p = INT32, None, None
            if op != COUNT:
                vp, vt, k1 = _col(agg[1])
                self._keep.append(k1)
                if len(agg) > 2 and agg[2] is not None:
                    v2p, _, k2 = _col(agg[2])
                    self._keep.append(k2)
            s = self.specs[a]
            s.op, s.vtype, s.v, s.v2, s.out_type, s.out = op, vt, vp, v2p, 0, None
        self.kmin = self.kmax = None
        if key_bounds is not None:
            self.kmin = (ctypes.c_int64 * nk)(*[int(b[0]) for b in key_bounds])
            self.kmax = (ctypes.c_int64 * nk)(*[int(b[1]) for b in key_bounds])
        self.h = None
        self.kouts, self.outs, self.fr, self.ngroups = [], [], None, 0
        self._fn = lib().aqgpu_groupby_run

    def free(self):
        if self.h is not None:
            lib().aqgpu_result_free(self.h)
            self.h = None
        self.kouts, self.outs, self.fr = [], [], None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass

    def run(self) -> int:
        self.free()
        h = ctypes.c_void_p(0)
        check(self._fn(self.kp, self.kt, self.nk, ctypes.c_uint32(self.nrows), self.kmin, self.kmax, self.row_order,
                       ctypes.c_uint32(self.groups_hint), self.specs, self.na, self.want_fr, ctypes.byref(h)))
        self.h = h
        L = lib()
        ng = ctypes.c_uint32(0)
        check(L.aqgpu_result_size(h, ctypes.byref(ng)))
        n = self.ngroups = ng.value
        p, t = ctypes.c_void_p(0), ctypes.c_int(0)
        for k in range(self.nk):
            check(L.aqgpu_result_key(h, k, ctypes.byref(p), ctypes.byref(t)))
            self.kouts.append(ColumnView(p.value, n, t.value))
        for a in range(self.na):
            check(L.aqgpu_result_agg(h, a, ctypes.byref(p), ctypes.byref(t)))
            self.outs.append(ColumnView(p.value, n, t.value))
        if self.want_fr:
            check(L.aqgpu_result_first_row(h, ctypes.byref(p)))
            self.fr = ColumnView(p.value, n, UINT32)
        return n


# ---- multi-GPU (one process per GPU): NCCL communicator inside the library, sharded group-by ------------------
def comm_unique_id() -> bytes:
    buf = ctypes.create_string_buffer(128)
    check(lib().aqgpu_comm_unique_id(buf))
    return buf.raw


def comm_init(nranks: int, rank: int, unique_id: bytes) -> None:
    check(lib().aqgpu_comm_init(int(nranks), int(rank), ctypes.c_char_p(unique_id)))


def comm_destroy() -> None:
    check(lib().aqgpu_comm_destroy())


def repartition_rows(keys, cols, row_base: int):
    """aqgpu_repartition_rows (collective): rows of this rank's shard -> the rank that owns their key tuple.  keys / cols are
    DeviceColumns; returns (the columns on their new rank..., the rows' global ids) as owning DeviceColumns."""
    nk, nc = len(keys), len(cols)
    kp = (ctypes.c_void_p * nk)(*[k.ptr for k in keys])
    kt = (ctypes.c_int * nk)(*[k.code for k in keys])
    cp = (ctypes.c_void_p * nc)(*[c.ptr for c in cols])
    eb = (ctypes.c_int * nc)(*[c.itemsize for c in cols])
    out = (ctypes.c_void_p * (nc + 1))()
    n = ctypes.c_uint32(0)
    check(lib().aqgpu_repartition_rows(kp, kt, nk, ctypes.c_uint32(keys[0].n), ctypes.c_uint64(int(row_base)), cp, eb, nc, out, ctypes.byref(n)))
    res = []
    for i in range(nc + 1):
        d = DeviceColumn.__new__(DeviceColumn)
        code = cols[i].code if i < nc else UINT32
        d.n, d.code, d.itemsize, d.ptr = int(n.value), code, CODE_SIZE[code], out[i]
        res.append(d)
    return res


class ShardQuery(AutoQuery):
    """aqgpu_groupby_run_dist over this rank's row shard (rows [row_base, row_base + n) of rows_total): called
    collectively.  run() returns the number of groups this rank holds, or None when the library declines the
    shape (the caller then exchanges rows / partials itself, aquery2_b200/dist.py).  ``replicated`` tells
    whether every rank holds all groups."""

    def __init__(self, keys, aggs, row_base: int, rows_total: int, row_order: int = ROWS_DESC, want_first_row: bool = True):
        super().__init__(keys, aggs, row_order=row_order, want_first_row=want_first_row)
        self.row_base, self.rows_total, self.replicated, self.spread = int(row_base), int(rows_total), True, 1

    def run(self):
        self.free()
        h, used = ctypes.c_void_p(0), ctypes.c_int(0)
        check(lib().aqgpu_groupby_run_dist(self.kp, self.kt, self.nk, ctypes.c_uint32(self.nrows), ctypes.c_uint64(self.row_base),
                                           ctypes.c_uint64(self.rows_total), self.row_order, self.specs, self.na, self.want_fr,
                                           ctypes.byref(h), ctypes.byref(used)))
        if not used.value:
            return None
        self.h = h
        L = lib()
        ng, rep = ctypes.c_uint32(0), ctypes.c_int(0)
        check(L.aqgpu_result_size(h, ctypes.byref(ng)))
        check(L.aqgpu_result_replicated(h, ctypes.byref(rep)))
        self.replicated, self.spread = rep.value == 1, rep.value
        n = self.ngroups = ng.value
        p, t = ctypes.c_void_p(0), ctypes.c_int(0)
        for k in range(self.nk):
            check(L.aqgpu_result_key(h, k, ctypes.byref(p), ctypes.byref(t)))
            self.kouts.append(ColumnView(p.value, n, t.value))
        for a in range(self.na):
            check(L.aqgpu_result_agg(h, a, ctypes.byref(p), ctypes.byref(t)))
            self.outs.append(ColumnView(p.value, n, t.value))
        if self.want_fr:
            check(L.aqgpu_result_first_row(h, ctypes.byref(p)))
            self.fr = ColumnView(p.value, n, UINT32)
        return n


def vec_sub(a, b, out, n: int, code: int) -> None:
    check(lib().aqgpu_vec_sub(code, ctypes.c_void_p(_out_ptr(a)), ctypes.c_void_p(_out_ptr(b)),
                              ctypes.c_void_p(_out_ptr(out)), ctypes.c_uint32(n)))


def vec_pow(x, p: float, out, n: int) -> None:
    check(lib().aqgpu_vec_pow(ctypes.c_void_p(_out_ptr(x)), ctypes.c_double(p), ctypes.c_void_p(_out_ptr(out)),
                              ctypes.c_uint32(n)))


class PinnedArray:
    """numpy view of pinned host memory from aqgpu_host_alloc (host side of the e2e arm)."""

    def __init__(self, n: int, code: int):
        self.n, self.code = int(n), code
        nbytes = max(self.n * CODE_SIZE[code], 16)
        p = ctypes.c_void_p(0)
        check(lib().aqgpu_host_alloc(ctypes.c_size_t(nbytes), ctypes.byref(p)))
        self.ptr = p.value
        buf = (ctypes.c_char * nbytes).from_address(self.ptr)
        if code == INT128:
            self.array = np.frombuffer(buf, np.int64, self.n * 2).reshape(self.n, 2)
        else:
            self.array = np.frombuffer(buf, CODE2NP[code], self.n)

    def free(self):
        if getattr(self, "ptr", None):
            self.array = None
            lib().aqgpu_host_free(ctypes.c_void_p(self.ptr))
            self.ptr = None


# ---- key columns that are not plain integers (include/aqgpu.h: aqgpu_key_split / aqgpu_key_join / aqgpu_string_dict) ----
KEY_DATE, KEY_TIME, KEY_TIMESTAMP = 1, 2, 3
# the reference's cells as numpy records (server/types.h:82-86,105-109,128-130)
DATE_DT = np.dtype([("day", "u1"), ("month", "u1"), ("year", "<i2")])
TIME_DT = np.dtype([("ms", "<u4"), ("seconds", "u1"), ("minutes", "u1"), ("hours", "u1"), ("pad", "u1")])
TIMESTAMP_DT = np.dtype([("date", DATE_DT), ("time", TIME_DT)])
_KEY_KIND = {DATE_DT: KEY_DATE, TIME_DT: KEY_TIME, TIMESTAMP_DT: KEY_TIMESTAMP}


def key_split(col: np.ndarray):
    """date / time / timestamp cells -> the integer key columns that stand for them: [uint32], [int64] or [uint32, int64]"""
    col = np.ascontiguousarray(col)
    kind, n = _KEY_KIND[col.dtype], len(col)
    a = np.zeros(n, np.int64 if kind == KEY_TIME else np.uint32)
    b = np.zeros(n, np.int64) if kind == KEY_TIMESTAMP else None
    check(lib().aqgpu_key_split(kind, col.ctypes.data_as(ctypes.c_void_p), ctypes.c_uint32(n), a.ctypes.data_as(ctypes.c_void_p),
                                b.ctypes.data_as(ctypes.c_void_p) if b is not None else None))
    return [a] if b is None else [a, b]


def key_join(kind_dtype: np.dtype, cols) -> np.ndarray:
    """inverse of key_split for the key columns of a result"""
    kind, n = _KEY_KIND[np.dtype(kind_dtype)], len(cols[0])
    out = np.zeros(n, kind_dtype)
    a = np.ascontiguousarray(cols[0])
    b = np.ascontiguousarray(cols[1]) if len(cols) > 1 else None
    check(lib().aqgpu_key_join(kind, a.ctypes.data_as(ctypes.c_void_p), b.ctypes.data_as(ctypes.c_void_p) if b is not None else None,
                               ctypes.c_uint32(n), out.ctypes.data_as(ctypes.c_void_p)))
    return out


def string_dict(data: np.ndarray, offsets: np.ndarray, max_distinct: int | None = None):
    """Arrow-style strings (uint8 bytes, uint64 offsets[n + 1]) -> (codes uint32[n] in first-appearance order,
    first_row uint32[ndistinct])."""
    data = np.ascontiguousarray(data, np.uint8)
    offsets = np.ascontiguousarray(offsets, np.uint64)
    n = len(offsets) - 1
    cap = n if max_distinct is None else int(max_distinct)
    codes = np.zeros(max(n, 1), np.uint32)
    first = np.zeros(max(cap, 1), np.uint32)
    nd = ctypes.c_uint32(0)
    check(lib().aqgpu_string_dict(data.ctypes.data_as(ctypes.c_void_p), ctypes.c_uint64(data.size), offsets.ctypes.data_as(ctypes.c_void_p),
                                  ctypes.c_uint32(n), codes.ctypes.data_as(ctypes.c_void_p), ctypes.c_uint32(cap),
                                  first.ctypes.data_as(ctypes.c_void_p), ctypes.byref(nd)))
    return codes[:n], first[: nd.value]


# ---- element-wise vector arithmetic / compare (aqgpu_vec_binary) -------------------------------------------------------
BOOL = 16
V_ADD, V_SUB, V_MUL, V_DIV, V_AND, V_OR, V_XOR, V_GT, V_LT, V_GTE, V_LTE, V_EQ, V_NEQ = range(13)
_VEC_NP = {INT32: np.int32, INT64: np.int64, FLOAT: np.float32, DOUBLE: np.float64, BOOL: np.uint8}


def _vec_operand(x):
    """(array of cells, dtype code, is_scalar): numpy arrays are vectors, 0-d arrays / numpy scalars are scalar operands"""
    x = np.asarray(x)
    scalar = x.ndim == 0
    x = np.ascontiguousarray(x.reshape(1) if scalar else x)
    if x.dtype == np.bool_:
        x = x.view(np.uint8)
    return x, (BOOL if x.dtype == np.uint8 else NP2CODE[x.dtype]), int(scalar)


def vec_out_type(op: int, a, b) -> int:
    (_, ta, sa), (_, tb, sb) = _vec_operand(a), _vec_operand(b)
    return int(lib().aqgpu_vec_out_type(op, ta, sa, tb, sb))


def vec_binary(op: int, a, b, out_type: int = -1) -> np.ndarray:
    """out = a op b, element-wise.  __int128 results come back as (n, 2) int64 [lo, hi], AQ_BOOL as uint8."""
    (a, ta, sa), (b, tb, sb) = _vec_operand(a), _vec_operand(b)
    n = len(b) if sa else len(a)
    ot = out_type if out_type >= 0 else int(lib().aqgpu_vec_out_type(op, ta, sa, tb, sb))
    if ot < 0:
        raise AqgpuError(f"vec_binary: no result type for op {op} on dtypes {ta}, {tb}")
    out = np.zeros((n, 2), np.int64) if ot == INT128 else np.zeros(n, _VEC_NP[ot])
    check(lib().aqgpu_vec_binary(op, ta, a.ctypes.data_as(ctypes.c_void_p), sa, tb, b.ctypes.data_as(ctypes.c_void_p), sb,
                                 ot, out.ctypes.data_as(ctypes.c_void_p), ctypes.c_uint32(n)))
    return out


def take(col: np.ndarray, idx: np.ndarray) -> np.ndarray:
    """out[j] = col[idx[j]] (aqgpu_take); col: any 1/2/4/8/12/16-byte cells"""
    col, idx = np.ascontiguousarray(col), np.ascontiguousarray(idx, np.uint32)
    out = np.zeros(len(idx), col.dtype)
    check(lib().aqgpu_take(col.dtype.itemsize, col.ctypes.data_as(ctypes.c_void_p), ctypes.c_uint32(len(col)),
                           idx.ctypes.data_as(ctypes.c_void_p), ctypes.c_uint32(len(idx)), out.ctypes.data_as(ctypes.c_void_p)))
    return out


# ---- ingest / write-back on the copy streams (aqgpu_ingest_async ...) ------------------------------------------------------
def ingest_async(dst: DeviceColumn, pinned_ptr: int, nbytes: int | None = None) -> None:
    check(lib().aqgpu_ingest_async(ctypes.c_void_p(dst.ptr), ctypes.c_void_p(pinned_ptr), ctypes.c_size_t(dst.nbytes if nbytes is None else nbytes)))


def ingest_fence() -> None:
    check(lib().aqgpu_ingest_fence())


def writeback_async(pinned_ptr: int, src_ptr: int, nbytes: int) -> None:
    check(lib().aqgpu_writeback_async(ctypes.c_void_p(pinned_ptr), ctypes.c_void_p(src_ptr), ctypes.c_size_t(nbytes)))


def writeback_fence() -> None:
    check(lib().aqgpu_writeback_fence())


def writeback_wait() -> None:
    check(lib().aqgpu_writeback_wait())


U_SQRT, U_TRUNCATE = 0, 1


def vec_unary(op: int, x: np.ndarray, param: int = 0) -> np.ndarray:
    """sqrt(x) (float64 cells) or truncate(x, param) (cells of x's type), aqgpu_vec_unary"""
    x = np.ascontiguousarray(x)
    out = np.zeros(len(x), np.float64 if op == U_SQRT else x.dtype)
    check(lib().aqgpu_vec_unary(op, NP2CODE[x.dtype], x.ctypes.data_as(ctypes.c_void_p), ctypes.c_uint32(param),
                                out.ctypes.data_as(ctypes.c_void_p), ctypes.c_uint32(len(x))))
    return out
