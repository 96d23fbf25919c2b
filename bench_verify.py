"""bench_verify.py -- result checks behind ``bench.py --verify`` (on by default).

Runs OUTSIDE the timed loop, after the last timed step, on the very result columns the timed step left
in HBM.  Nothing here is on the product path: torch is used as an independent calculator on device views
of the library's columns (``__cuda_array_interface__``), and the CPU leg compares the GPU result with
the reference-built oracle on the cpu_baseline table.  Three layers:

  1. size-independent properties at the FULL bench size, per query:
       * ngroups == an independently counted number of distinct key tuples (torch.unique of the packed
         key; two independent 64-bit hashes when the tuple does not fit 63 bits: h2o q10);
       * a second call with the same keys and (count, first_row): sum(count) == rows, first rows
         strictly ascending (= first-appearance group order), key[g] == key_col[first_row[g]];
       * every timed aggregate column: sum of sums == column total, sum(avg * count) == column total,
         max of maxes == column max, ...;
       * SAMPLE_GROUPS groups per query recomputed from scratch through a row mask (all rows whose key
         tuple equals the group's) with the reference's formulas in float64 / int64, compared bit-exact
         (integers) or to 1e-9 (fp; the recomputation itself sums in another order);
       * the q7 / q9 element-wise epilogues.
  2. the same ten queries on the cpu_baseline table (G1_1e7_1e2): GPU result vs oracle/_ref result,
     bit-exact for counts / integer aggregates / group order, 1e-12 for fp64.
Any failure raises VerifyError; bench.py then prints "verify": "FAILED: ..." and exits non-zero.
"""
from __future__ import annotations

import numpy as np

SAMPLE_GROUPS = 6
FP_RTOL_RECOMPUTE = 1e-9       # torch recomputation vs library (different summation orders at 1e7 rows/group)
FP_RTOL_ORACLE = 1e-12         # north_star tolerance vs the reference's own C++


class VerifyError(AssertionError):
    pass


def _need(cond, msg):
    if not cond:
        raise VerifyError(msg)


class _CudaBytes:
    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


def tview(torch, aq, col, n=None):
    """torch view of the first n cells of a DeviceColumn (INT128 -> (n, 2) int64, UINT32 -> int64 copy)"""
    n = col.n if n is None else int(n)
    if n == 0:
        return torch.empty(0, dtype=torch.int64, device="cuda")
    raw = torch.as_tensor(_CudaBytes(col.ptr, n * col.itemsize), device="cuda")
    if col.code == aq.INT32:
        return raw.view(torch.int32)
    if col.code == aq.UINT32:
        return raw.view(torch.int32).to(torch.int64) & 0xFFFFFFFF
    if col.code == aq.FLOAT:
        return raw.view(torch.float32)
    if col.code == aq.DOUBLE:
        return raw.view(torch.float64)
    if col.code == aq.INT64:
        return raw.view(torch.int64)
    if col.code == aq.INT128:
        return raw.view(torch.int64).view(n, 2)
    raise VerifyError(f"dtype code {col.code}")


def _i128_sum(torch, cells):
    """exact python int sum of (n, 2) int64 [lo, hi] cells whose hi word is the sign extension of lo"""
    lo, hi = cells[:, 0], cells[:, 1]
    _need(bool(torch.equal(hi, lo >> 63)), "__int128 cell does not fit int64 (unexpected for this workload)")
    return int(lo.sum().item())


def _close(a, b, rtol, scale=0.0):
    return abs(a - b) <= rtol * max(abs(a), abs(b), scale)


def distinct_count(torch, key_tensors, ranges):
    """independent number of distinct key tuples: exact packing when it fits 62 bits, else the larger of
    two 64-bit-hash distinct counts (a lower bound that is exact unless BOTH hashes collide)"""
    n = key_tensors[0].numel()
    bits = sum(int(r).bit_length() for r in ranges)
    if bits <= 62:
        packed = torch.zeros(n, dtype=torch.int64, device="cuda")
        for k, r in zip(key_tensors, ranges):
            packed = packed * int(r + 1) + k.to(torch.int64)
        return int(torch.unique(packed).numel())
    best = 0
    for mul, add in ((0x9E3779B97F4A7C15, 0x632BE59BD9B4E019), (0xC2B2AE3D27D4EB4F, 0x165667B19E3779F9)):
        h = torch.full((n,), add - (1 << 64) if add >= (1 << 63) else add, dtype=torch.int64, device="cuda")
        m = mul - (1 << 64) if mul >= (1 << 63) else mul
        for k in key_tensors:
            h = (h ^ k.to(torch.int64)) * m
            h = h ^ ((h >> 29) & ((1 << 35) - 1))            # logical shift: keeps the step a bijection
        best = max(best, int(torch.unique(h).numel()))
        del h
    return best


def _recompute(torch, op, v, v2, rows_desc):
    """the reference's aggregate (server/aggregations.h) over the group's rows, float64 / python ints.
    rows_desc: row ids of the group in DESCENDING order (the reference's row-list order)"""
    if op == "count":
        return int(rows_desc.numel())
    x = v[rows_desc]
    isint = not x.dtype.is_floating_point
    if op == "sum":
        return int(x.to(torch.int64).sum().item()) if isint else float(x.to(torch.float64).sum().item())
    if op == "avg":
        return float(x.to(torch.float64).sum().item()) / x.numel()
    if op == "min":
        return x.min().item()
    if op == "max":
        m = x.max().item()
        if not isint:
            tiny = float(np.finfo(np.float32 if x.dtype == torch.float32 else np.float64).tiny)
            m = m if m > tiny else tiny          # aggregations.h:73 identity numeric_limits<T>::min()
        return m
    if op == "first":
        return x[0].item()
    if op == "last":
        return x[-1].item()
    if op == "second":
        return x[1].item() if x.numel() > 1 else 0
    if op in ("var", "stddev"):
        n1 = x.numel() + 1
        if isint:
            s = int(x.to(torch.int64).sum().item()); q = int((x.to(torch.int64) ** 2).sum().item())
            r = (float(q) - float(s * s) / n1) / n1
        else:
            s = float(x.to(torch.float64).sum().item()); q = float((x * x).to(torch.float64).sum().item())
            r = (q - s * s / n1) / n1
        return r ** 0.5 if op == "stddev" else r
    if op == "corr":
        y = v2[rows_desc].to(torch.int64); xi = x.to(torch.int64)
        n = xi.numel()
        sx, sy = int(xi.sum().item()), int(y.sum().item())
        sxy, sxx, syy = int((xi * y).sum().item()), int((xi * xi).sum().item()), int((y * y).sum().item())
        num = float(n * sxy) - float(sx * sy)
        den = (float(n * sxx) - float(sx * sx)) * (float(n * syy) - float(sy * sy))
        return num / den ** 0.5 if den > 0 else float("nan")
    raise VerifyError(f"no recomputation for {op}")


def _cell(torch, t, g):
    """python value of result cell g (INT128 -> int)"""
    if t.dim() == 2:
        lo, hi = int(t[g, 0].item()), int(t[g, 1].item())
        return (hi << 64) + (lo & ((1 << 64) - 1))
    return t[g].item()


def verify_query(torch, aq, q, keys, aggs, epilogue, table, prepared, extra, ng, key_ranges, report):
    """table: {name: DeviceColumn}; prepared: the PreparedQuery the timed loop ran (its outs / kouts hold
    the last step's result); ng: its group count."""
    nrows = table[keys[0]].n
    aq.sync()
    kt = [tview(torch, aq, table[c]) for c in keys]
    # ---- 1. independent group count
    want_ng = distinct_count(torch, kt, key_ranges)
    _need(ng == want_ng, f"{q}: ngroups {ng} != independently counted {want_ng}")
    # ---- 2. same keys, (count, first_row): grouping machinery
    chk = aq.groupby_agg([table[c] for c in keys], [("count",)], device_out=True, want_keys=True, want_first_row=True,
                         max_groups=max(ng, 1))
    aq.sync()      # the library runs on its own stream: torch (another stream) must not read before it is done
    _need(chk.ngroups == ng, f"{q}: count/first_row call saw {chk.ngroups} groups, timed call {ng}")
    cnt = tview(torch, aq, chk.aggs[0], ng)
    fr = tview(torch, aq, chk.first_row, ng)
    _need(int(cnt.sum().item()) == nrows, f"{q}: sum(count) = {int(cnt.sum().item())} != rows {nrows}")
    if ng > 1:
        _need(bool((fr[1:] > fr[:-1]).all().item()), f"{q}: first rows are not strictly ascending (group order)")
    _need(ng == 0 or (int(fr[0].item()) == 0), f"{q}: first group does not start at row 0")
    for k, c in enumerate(keys):
        ko_chk = tview(torch, aq, chk.keys[k], ng)
        ko = tview(torch, aq, prepared.kouts[k], ng)
        src = kt[k][fr]
        _need(bool(torch.equal(ko_chk.to(torch.int64), src.to(torch.int64))), f"{q}: key {c} != key_col[first_row] (check call)")
        _need(bool(torch.equal(ko.to(torch.int64), src.to(torch.int64))), f"{q}: key {c} != key_col[first_row] (timed call)")
    # ---- 3. column-total identities of the timed aggregates
    outs = [tview(torch, aq, prepared.outs[a], ng) for a in range(len(aggs))]
    cntf = cnt.to(torch.float64)
    for a, agg in enumerate(aggs):
        op = agg[0]
        v = tview(torch, aq, table[agg[1]]) if len(agg) > 1 else None
        o = outs[a]
        if op == "sum":
            if o.dim() == 2:
                _need(_i128_sum(torch, o) == int(v.to(torch.int64).sum().item()), f"{q}: sum of sum({agg[1]}) != column total")
            else:
                tot = float(v.to(torch.float64).sum().item())
                _need(_close(float(o.sum().item()), tot, 1e-10), f"{q}: sum of sum({agg[1]}) = {float(o.sum().item())!r} vs column total {tot!r}")
        elif op == "avg":
            tot = float(v.to(torch.float64).sum().item())
            got = float((o * cntf).sum().item())
            _need(_close(got, tot, 1e-10), f"{q}: sum(avg({agg[1]}) * count) = {got!r} vs column total {tot!r}")
        elif op == "count":
            _need(bool(torch.equal(o.to(torch.int64), cnt.to(torch.int64))), f"{q}: count column differs from the check call")
        elif op == "max":
            _need(o.max().item() == v.max().item(), f"{q}: max of max({agg[1]}) != column max")
        elif op == "min":
            _need(o.min().item() == v.min().item(), f"{q}: min of min({agg[1]}) != column min")
        elif op in ("stddev", "var", "corr"):
            _need(bool(torch.isfinite(o).all().item()) or op == "corr", f"{q}: non-finite {op}")
    # ---- 4. a few groups recomputed from scratch through a row mask
    picks = sorted({0, ng - 1, ng // 2, ng // 3, (ng * 7) // 10, min(ng - 1, 12345)} if ng else set())[:SAMPLE_GROUPS]
    vt = {c: tview(torch, aq, table[c]) for agg in aggs for c in agg[1:]}
    for g in picks:
        mask = None
        for k in range(len(keys)):
            m = kt[k] == kt[k][fr[g]]
            mask = m if mask is None else (mask & m)
        rows = torch.nonzero(mask).flatten()
        _need(int(rows.numel()) == int(cnt[g].item()), f"{q}: group {g}: count {int(cnt[g].item())} vs {int(rows.numel())} matching rows")
        _need(int(rows[0].item()) == int(fr[g].item()), f"{q}: group {g}: first_row {int(fr[g].item())} vs smallest matching row {int(rows[0].item())}")
        rows_desc = torch.flip(rows, [0])
        for a, agg in enumerate(aggs):
            want = _recompute(torch, agg[0], vt.get(agg[1]) if len(agg) > 1 else None, vt.get(agg[2]) if len(agg) > 2 else None, rows_desc)
            got = _cell(torch, outs[a], g)
            if isinstance(want, int) and not isinstance(got, float):
                _need(int(got) == want, f"{q}: group {g} {agg}: {got} != recomputed {want}")
            else:
                ok = (want != want and got != got) or _close(float(got), float(want), FP_RTOL_RECOMPUTE)
                _need(ok, f"{q}: group {g} {agg}: {got!r} vs recomputed {want!r}")
        del mask, rows, rows_desc
    # ---- 5. epilogues
    if epilogue == "sub":
        e = tview(torch, aq, extra, ng)
        _need(bool(torch.equal(e, outs[0] - outs[1])), f"{q}: epilogue max - min differs")
    elif epilogue == "pow2":
        e = tview(torch, aq, extra, ng)
        w = outs[0] * outs[0]
        bad = ~((e == w) | ((e - w).abs() <= 1e-14 * w.abs()) | (torch.isnan(e) & torch.isnan(w)))
        _need(not bool(bad.any().item()), f"{q}: epilogue pow(corr, 2) differs")
    report[q] = {"groups": ng, "recomputed_groups": len(picks)}
    del chk


def verify_vs_oracle(aq, impl, t, queries, QUERIES, report):
    """GPU vs the reference-built oracle on the cpu_baseline table (host numpy columns): q1-q10,
    bit-exact counts / integer aggregates / order, 1e-12 fp.  Returns seconds of CPU time spent in the
    oracle's post-processor bracket (that is the cpu_baseline measurement itself)."""
    total = 0.0
    n = len(next(iter(t.values())))
    for q in queries:
        keys, aggs, _ = QUERIES[q]
        specs = [(a[0],) + tuple(t[c] for c in a[1:]) for a in aggs]
        sec, ng_cpu, outs_cpu = impl.time_groupby_reduce_out([t[c] for c in keys], specs)
        total += sec
        res = aq.groupby_agg([t[c] for c in keys], specs, want_first_row=True)
        _need(res.ngroups == ng_cpu, f"oracle {q}: ngroups {res.ngroups} vs CPU {ng_cpu}")
        fr = res.first_row.astype(np.int64)
        _need(bool(np.all(np.diff(fr) > 0)), f"oracle {q}: first rows not ascending")
        for k, c in enumerate(keys):
            _need(bool(np.array_equal(res.keys[k], t[c][fr])), f"oracle {q}: key {c} order")
        for a, agg in enumerate(aggs):
            got, want = np.asarray(res.aggs[a]), np.asarray(outs_cpu[a])[: ng_cpu]
            if got.dtype.kind in "iu":
                _need(bool(np.array_equal(got, want.reshape(got.shape))), f"oracle {q} {agg}: integer cells differ")
            else:
                with np.errstate(invalid="ignore"):
                    ok = (np.abs(got - want) <= FP_RTOL_ORACLE * np.maximum(np.abs(got), np.abs(want))) | (np.isnan(got) & np.isnan(want))
                _need(bool(ok.all()), f"oracle {q} {agg}: {int((~ok).sum())} fp cells beyond {FP_RTOL_ORACLE}, e.g. "
                                      f"{got[~ok][:2]} vs {want[~ok][:2]}")
        report[q] = {"groups": ng_cpu, "rows": n}
    return total
