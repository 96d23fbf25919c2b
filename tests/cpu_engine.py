"""Oracle-backed stand-in for aquery2_b200.dist.GpuEngine (TEST INFRASTRUCTURE): the same methods over numpy
columns, so that the orchestration of dist.DistQuery (decomposition, exchange, merge, finalisation) runs
under gloo on CPU with world_size > 1.  __int128 cells are (n, 2) int64 arrays [lo, hi]."""
import numpy as np

from aquery2_b200 import aqgpu as aq

NP = {aq.INT32: np.int32, aq.FLOAT: np.float32, aq.DOUBLE: np.float64, aq.INT64: np.int64, aq.UINT32: np.uint32}


def _to128(x):
    x = np.asarray(x, np.int64)
    return np.stack([x, x >> 63], axis=1)


def _as_i64(col):
    col = np.asarray(col)
    return col[:, 0].copy() if col.ndim == 2 else col.astype(np.int64)


class _Local:
    def __init__(self, eng, keys, partials, cols):
        self.eng, self.keys, self.partials, self.cols = eng, keys, partials, cols

    def run(self):
        specs = [(p[0], self.cols[p[1]] if p[1] else None, self.cols[p[2]] if p[2] else None, p[3]) for p in self.partials]
        ng, self.kouts, self.outs, first = self.eng._groupby(self.keys, specs)
        self.fr = first.astype(np.uint32)
        return ng


class CpuEngine:
    def __init__(self, comm, oracle):
        self.comm, self.o, self.torch = comm, oracle, comm.torch

    def _reduce(self, op, v, v2, out, off, perm):
        o = self.o
        if op == "sumsq":
            v = (v.astype(np.int64) * v) if v.dtype.kind == "i" else (v * v).astype(v.dtype)
            op = "sum"
        elif op == "sumxy":
            v = (v.astype(np.int64) * v2) if v.dtype.kind == "i" else (v * v2).astype(v.dtype)
            v2, op = None, "sum"
        r = o.reduce(op, v, off, perm, v2=v2, raw128=True)
        if out == aq.INT64:
            r = _as_i64(r)
        elif out == aq.INT128 and r.ndim == 1:
            r = _to128(r)
        return r

    def _groupby(self, keys, specs):
        keys = [np.ascontiguousarray(k) for k in keys]
        if len(keys[0]) == 0:
            return 0, [k[:0] for k in keys], [np.zeros(0) for _ in specs], np.zeros(0, np.uint32)
        ng, gid, off, perm = self.o.groupby(keys)
        first = perm[off[1:] - 1]
        outs = [self._reduce(op, v, v2, out, off, perm) for op, v, v2, out in specs]
        return ng, [k[first] for k in keys], outs, first

    def local_prepare(self, keys, partials, cols, max_groups, hint, bounds):
        return _Local(self, keys, partials, cols)

    def rows_to_global(self, fr, n, add):
        return fr[:n].astype(np.int64) + add

    def bytes_of(self, col, n):
        a = np.ascontiguousarray(col[:n])
        return self.torch.from_numpy(a.view(np.uint8).reshape(-1).copy())

    def from_bytes(self, t, n, code):
        b = t.numpy()
        if code == aq.INT128:
            return b.view(np.int64).reshape(n, 2)
        return b.view(NP[code])[:n]

    def groupby(self, keys, specs, hint, bounds, max_groups):
        ng, k, outs, _ = self._groupby(keys, specs)
        return ng, k, outs

    def finalize(self, op, n, count, s_code, s_cols):
        cnt = np.asarray(count[:n], np.int64)
        if op == aq.COUNT:
            return cnt.astype(np.uint32)

        def big(c):          # exact Python integers of an int64 / __int128 column
            c = np.asarray(c)
            return [(int(hi) << 64) + (int(lo) & ((1 << 64) - 1)) for lo, hi in c[:n]] if c.ndim == 2 else [int(x) for x in c[:n]]

        def f64(c):
            c = np.asarray(c)
            return np.array([float(x) for x in big(c)]) if c.ndim == 2 else c[:n].astype(np.float64)

        if op == aq.AVG:
            return f64(s_cols[0]) / cnt.astype(np.float64)
        if op in (aq.VAR, aq.STDDEV):
            np1 = (cnt + 1).astype(np.float64)
            if s_code == aq.DOUBLE:
                s, q = f64(s_cols[0]), f64(s_cols[1])
                r = (q - s * s / np1) / np1
            else:
                s, q = big(s_cols[0]), f64(s_cols[1])
                r = (q - np.array([float(x * x) for x in s]) / np1) / np1
            return np.sqrt(r) if op == aq.STDDEV else r
        if op == aq.CORR:
            sx, sy, sxy, sx2, sy2 = (big(c) for c in s_cols)
            out = np.empty(n)
            for i in range(n):
                ln = int(cnt[i])
                num = float(ln * sxy[i]) - float(sx[i] * sy[i])
                dx = float(ln * sx2[i]) - float(sx[i] * sx[i])
                dy = float(ln * sy2[i]) - float(sy[i] * sy[i])
                with np.errstate(all="ignore"):
                    out[i] = np.float64(num) / np.sqrt(np.float64(dx) * np.float64(dy))
            return out
        raise ValueError(op)

    def second_merge(self, code, n, cnt_top, nparts, second_top, first_prev):
        zero = np.zeros(n, NP[code])
        return np.where(np.asarray(cnt_top[:n]) >= 2, second_top[:n], np.where(np.asarray(nparts[:n]) >= 2, first_prev[:n], zero))

    def partition(self, keys, n, nparts, cols):
        h = np.zeros(n, np.uint64)
        for k in keys:
            h = h * np.uint64(1000003) + np.asarray(k[:n]).astype(np.int64).astype(np.uint64)
        owner = (h % np.uint64(nparts)).astype(np.int64)
        order = np.argsort(owner, kind="stable")
        counts = np.bincount(owner, minlength=nparts).tolist()
        return counts, [np.ascontiguousarray(np.asarray(c[:n])[order]) for c in cols]

    def groupby_rows(self, keys, aggs, hint, bounds, max_groups):
        specs = [(a[0], a[1] if len(a) > 1 else None, a[2] if len(a) > 2 else None, 0) for a in aggs]
        ng, k, outs, first = self._groupby(keys, specs)
        return ng, k, outs, first.astype(np.uint32)

    def take(self, col, idx, n):
        return np.asarray(col)[np.asarray(idx[:n], np.int64)]

    def global_rows(self, lo, n):
        return np.arange(lo, lo + n, dtype=np.int64)

    def to_numpy(self, col, n):
        return np.asarray(col[:n])
