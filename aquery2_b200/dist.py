"""Multi-GPU group-by: row shards, local partial aggregation, NCCL exchange, merge.

One process per GPU (``torch.distributed``, NCCL over NVLink; gloo for the CPU tests).  Rank r owns the
contiguous global rows [lo_r, hi_r).  A query runs as

  1. local:    ``aqgpu_groupby_agg`` over the shard with DECOMPOSABLE partial aggregates (``decompose``)
  2. exchange: the partial tables (key columns, partial states, global first rows) are all-gathered when
               small, or hash-partitioned by key (``aqgpu_key_owner``) and sent to the owning rank with one
               all-to-all when large
  3. merge:    a second ``aqgpu_groupby_agg`` over the concatenation of the partial tables in rank order.
               Shards are contiguous row ranges, so first appearance in that concatenation IS the
               reference's first-appearance group order (server/unordered_dense.h:1142-1146), and the
               largest row of the concatenation is the largest global row (FIRST / LAST / SECOND keep
               their meaning, server/hasher.h:193-196)
  4. finalize: avg / var / stddev / corr / count / second from the merged states (``aqgpu_finalize``)

The reference is single-process (no distributed backend); this file has no
counterpart there.  Compute goes through an *engine* (``GpuEngine`` = libaqgpu; the CPU tests plug in an
oracle-backed engine) and bytes through ``Comm``; nothing here touches oracle/.
"""
from __future__ import annotations

import os
import sys
import time

import numpy as np

from . import aqgpu as aq

INT_CODES = (aq.INT32, aq.INT64)


# ------------------------------------------------------------------------------------------------
# algebra: requested aggregates -> partial aggregates, merge aggregates, finalisation
# ------------------------------------------------------------------------------------------------
def decompose(aggs, codes):
    """aggs: [(op name, column name[, second column])]; codes: {column name: dtype code}.

    Returns (partials, merges, finals):
      partials  [(op, col, col2, out_code)]               run on the shard (deduplicated)
      merges    [(op, ("p", i) | ("grow",), out_code)]    run on the concatenated partial tables; ("p", i) =
                                                          output i of `partials`
      finals    [(kind, ...)] one per requested aggregate, indices into `merges`
    """
    partials, merges = [], []

    def part(op, col=None, col2=None, out=0):
        e = (op, col, col2, out)
        if e not in partials:
            partials.append(e)
        return partials.index(e)

    def merge(op, src, out=0):
        e = (op, src, out)
        if e not in merges:
            merges.append(e)
        return merges.index(e)

    def sum_of(op, col, col2=None):
        isint = codes[col] in INT_CODES
        p = part(op, col, col2, aq.INT64 if isint else aq.DOUBLE)
        return merge("sum", ("p", p), aq.INT128 if isint else aq.DOUBLE), (aq.INT128 if isint else aq.DOUBLE)

    def count_of():
        return merge("sum", ("p", part("count", out=aq.INT64)), aq.INT64)

    finals = []
    for a in aggs:
        op = a[0]
        if op == "sum":
            finals.append(("col", sum_of("sum", a[1])[0]))
        elif op == "count":
            finals.append(("fin", aq.COUNT, count_of(), 0, []))
        elif op == "avg":
            s, st = sum_of("sum", a[1])
            finals.append(("fin", aq.AVG, count_of(), st, [s]))
        elif op in ("min", "max", "first", "last"):
            finals.append(("col", merge(op, ("p", part(op, a[1])))))
        elif op in ("var", "stddev"):
            s, st = sum_of("sum", a[1])
            q, _ = sum_of("sumsq", a[1])
            finals.append(("fin", aq.VAR if op == "var" else aq.STDDEV, count_of(), st, [s, q]))
        elif op == "corr":
            x, y = a[1], a[2]
            cols = [sum_of("sum", x), sum_of("sum", y), sum_of("sumxy", x, y), sum_of("sumsq", x), sum_of("sumsq", y)]
            finals.append(("fin", aq.CORR, count_of(), cols[0][1], [c[0] for c in cols]))
        elif op == "second":
            pf, ps, pc = part("first", a[1]), part("second", a[1]), part("count", out=aq.INT64)
            finals.append(("second", codes[a[1]], merge("first", ("p", pc)), merge("count", None, aq.INT64),
                           merge("first", ("p", ps)), merge("second", ("p", pf))))
        else:
            raise ValueError(f"aggregate {op!r} has no multi-GPU decomposition")
    return partials, merges, finals


# ------------------------------------------------------------------------------------------------
# communication
# ------------------------------------------------------------------------------------------------
class Comm:
    """torch.distributed process group: NCCL when CUDA is there, gloo otherwise (CPU tests)."""
    CHUNK = 1 << 30            # bytes per peer and all-to-all call

    def __init__(self, local_rank: int = 0, backend: str | None = None):
        # the exchange buffers are multi-GB torch tensors next to the library's own cudaMallocAsync pool:
        # expandable segments keep torch's cache from fragmenting into cudaFree / cudaMalloc retries
        os.environ.setdefault("PYTORCH_CUDA_ALLOC_CONF", "expandable_segments:True")
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        cuda = torch.cuda.is_available()
        backend = backend or ("nccl" if cuda else "gloo")
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if cuda:
            torch.cuda.set_device(local_rank)
        self.device = torch.device("cuda", local_rank) if cuda else torch.device("cpu")
        if not dist.is_initialized():
            import datetime
            kw = {"device_id": self.device} if backend == "nccl" else {}
            dist.init_process_group(backend, timeout=datetime.timedelta(seconds=int(os.environ.get("AQ_DIST_TIMEOUT", "180"))), **kw)
        self.rank, self.world = dist.get_rank(), dist.get_world_size()
        self.stream = None
        if cuda:
            # ONE stream for the library's kernels, torch's indexing and the NCCL collectives.  torch's default
            # stream has handle 0, which aqgpu_set_stream reads as "use your own stream", so a real stream
            # is made current instead (kept alive by this object).
            self.stream = torch.cuda.Stream(device=self.device)
            torch.cuda.set_stream(self.stream)
            aq.set_stream(self.stream.cuda_stream)

    def barrier(self):
        self.dist.barrier()

    def _reduce(self, x: float, op):
        t = self.torch.tensor([x], dtype=self.torch.float64, device=self.device)
        self.dist.all_reduce(t, op=op)
        return float(t.item())

    def max_float(self, x: float) -> float:
        return self._reduce(x, self.dist.ReduceOp.MAX)

    def sum_float(self, x: float) -> float:
        return self._reduce(x, self.dist.ReduceOp.SUM)

    def all_counts(self, n: int) -> list[int]:
        t = self.torch.zeros(self.world, dtype=self.torch.int64, device=self.device)
        t[self.rank] = n
        self.dist.all_reduce(t)
        return [int(x) for x in t.tolist()]

    def count_matrix(self, row: list[int]) -> list[list[int]]:
        """m[src][dst] from every rank's row of per-destination counts"""
        t = self.torch.zeros(self.world, self.world, dtype=self.torch.int64, device=self.device)
        t[self.rank] = self.torch.tensor(row, dtype=self.torch.int64)
        self.dist.all_reduce(t)
        return [[int(x) for x in r] for r in t.tolist()]

    def _wide(self, t, itemsize: int):
        """the same bytes as 8- or 4-byte elements where the cell size allows: element counts of NCCL
        calls stay far below 2^31 (a 4e9-element uint8 all-to-all hung on B200 / NCCL 2.28)"""
        if itemsize % 8 == 0 and t.data_ptr() % 8 == 0:
            return t.view(self.torch.int64), 8
        if itemsize % 4 == 0 and t.data_ptr() % 4 == 0:
            return t.view(self.torch.int32), 4
        return t, 1

    def allgather_bytes(self, buf, counts: list[int], itemsize: int):
        """buf: uint8 tensor of counts[rank] * itemsize bytes -> concatenation over ranks (rank order)"""
        out = self.torch.empty(sum(counts) * itemsize, dtype=self.torch.uint8, device=self.device)
        off = 0
        for r, c in enumerate(counts):
            sl = out[off * itemsize:(off + c) * itemsize]
            if r == self.rank and c:
                sl.copy_(buf[:c * itemsize])
            if c:
                self.dist.broadcast(self._wide(sl, itemsize)[0], src=r)
            off += c
        return out

    def alltoall_bytes(self, buf, send_counts: list[int], recv_counts: list[int], itemsize: int, max_count: int | None = None):
        """buf holds the rows for rank 0, 1, ... back to back; returns what the others sent, in rank order"""
        out = self.torch.empty(sum(recv_counts) * itemsize, dtype=self.torch.uint8, device=self.device)
        ins = [c * itemsize for c in send_counts]
        outs = [c * itemsize for c in recv_counts]
        if self.dist.get_backend() == "gloo":      # gloo has no all_to_all_single on every build: pairwise
            offs_in = np.concatenate([[0], np.cumsum(ins)]).astype(int)
            offs_out = np.concatenate([[0], np.cumsum(outs)]).astype(int)
            for step in range(self.world):
                dst, src = (self.rank + step) % self.world, (self.rank - step) % self.world
                send = buf[offs_in[dst]:offs_in[dst + 1]]
                recv = out[offs_out[src]:offs_out[src + 1]]
                if dst == self.rank:
                    recv.copy_(send)
                    continue
                reqs = []
                if len(send):
                    reqs.append(self.dist.isend(send.contiguous(), dst))
                if len(recv):
                    reqs.append(self.dist.irecv(recv, src))
                for rq in reqs:
                    rq.wait()
        else:
            # pieces of <= CHUNK bytes per peer: one 4 GB all_to_all_single (2 GB per peer) never completed on
            # B200 / NCCL 2.28, 1 GB per peer did (and 1 GiB pieces move at 430 GB/s against 325 for 256 MiB)
            offs_in = np.concatenate([[0], np.cumsum(ins)]).astype(np.int64)
            offs_out = np.concatenate([[0], np.cumsum(outs)]).astype(np.int64)
            step = max(self.CHUNK // itemsize, 1) * itemsize
            if max_count is None:       # every rank must issue the same number of calls
                max_count = int(self.max_float(float(max(send_counts + recv_counts, default=0))))
            dt = self.torch.int64 if itemsize % 8 == 0 else self.torch.int32 if itemsize % 4 == 0 else self.torch.uint8
            done = 0
            while done < max_count * itemsize:
                send = [buf[offs_in[d] + min(done, ins[d]):offs_in[d] + min(done + step, ins[d])].view(dt) for d in range(self.world)]
                recv = [out[offs_out[r] + min(done, outs[r]):offs_out[r] + min(done + step, outs[r])].view(dt) for r in range(self.world)]
                self.dist.all_to_all(recv, send)
                done += step
        return out

    def close(self):
        if self.stream is not None:
            self.stream.synchronize()
            aq.set_stream(None)
            self.torch.cuda.set_stream(self.torch.cuda.default_stream(self.device))
            self.stream = None
        if self.dist.is_initialized():
            self.dist.destroy_process_group()


# ------------------------------------------------------------------------------------------------
# compute engine over libaqgpu (the CPU tests substitute an oracle-backed engine with the same methods)
# ------------------------------------------------------------------------------------------------
class _CudaBytes:
    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


class WrappedColumn(aq.DeviceColumn):
    """non-owning DeviceColumn over memory of a torch tensor (exchange buffers)"""

    def __init__(self, tensor, n: int, code: int):       # noqa: super().__init__ allocates; this one must not
        self.n, self.code, self.itemsize = int(n), int(code), aq.CODE_SIZE[code]
        self.keep, self.ptr = tensor, tensor.data_ptr()

    def free(self):
        self.ptr, self.keep = None, None


class GpuEngine:
    def __init__(self, comm: Comm):
        self.comm, self.torch = comm, comm.torch

    def local_prepare(self, keys, partials, cols, max_groups, hint, bounds):
        specs = [(p[0],) + tuple(cols[c] for c in p[1:3] if c is not None) for p in partials]
        return aq.PreparedQuery(keys, specs, max_groups=max_groups, groups_hint=hint, key_bounds=bounds, device_out=True,
                                out_types=[p[3] for p in partials], want_keys=True, want_first_row=True)

    def rows_to_global(self, fr, n, add):
        out = aq.DeviceColumn(max(n, 1), aq.INT64)
        aq.check(aq.lib().aqgpu_rows_to_global(aq.ctypes.c_void_p(fr.ptr), aq.ctypes.c_int64(add), aq.ctypes.c_void_p(out.ptr),
                                               aq.ctypes.c_uint32(n)))
        return out

    def bytes_of(self, col, n):
        nb = n * col.itemsize
        if nb == 0:
            return self.torch.empty(0, dtype=self.torch.uint8, device=self.comm.device)
        return self.torch.as_tensor(_CudaBytes(col.ptr, nb), device=self.comm.device)

    def from_bytes(self, t, n, code):
        return WrappedColumn(t, n, code)

    def groupby(self, keys, specs, hint, bounds, max_groups):
        res = aq.groupby_agg(keys, [(s[0],) + tuple(c for c in s[1:3] if c is not None) for s in specs], groups_hint=hint,
                             max_groups=max_groups, key_bounds=bounds, device_out=True, want_first_row=False,
                             out_types=[s[3] for s in specs])
        return res.ngroups, res.keys, res.aggs

    def finalize(self, op, n, count, s_code, s_cols):
        out = aq.DeviceColumn(max(n, 1), aq.UINT32 if op == aq.COUNT else aq.DOUBLE)
        sp = (aq.ctypes.c_void_p * 5)(*([c.ptr for c in s_cols] + [None] * (5 - len(s_cols))))
        aq.check(aq.lib().aqgpu_finalize(op, aq.ctypes.c_uint32(n), aq.ctypes.c_void_p(count.ptr), s_code, sp, aq.ctypes.c_void_p(out.ptr)))
        return out

    def second_merge(self, code, n, cnt_top, nparts, second_top, first_prev):
        out = aq.DeviceColumn(max(n, 1), code)
        aq.check(aq.lib().aqgpu_second_merge(code, aq.ctypes.c_void_p(cnt_top.ptr), aq.ctypes.c_void_p(nparts.ptr),
                                             aq.ctypes.c_void_p(second_top.ptr), aq.ctypes.c_void_p(first_prev.ptr),
                                             aq.ctypes.c_void_p(out.ptr), aq.ctypes.c_uint32(n)))
        return out

    def partition(self, keys, n, nparts, cols):
        """rows grouped by owner rank, original order kept inside an owner: (rows per owner, the columns in that order)"""
        if n == 0:
            return [0] * nparts, cols
        owner = aq.DeviceColumn(n, aq.UINT32)
        kp = (aq.ctypes.c_void_p * len(keys))(*[k.ptr for k in keys])
        kt = (aq.ctypes.c_int * len(keys))(*[k.code for k in keys])
        aq.check(aq.lib().aqgpu_key_owner(kp, kt, len(keys), aq.ctypes.c_uint32(n), aq.ctypes.c_uint32(nparts), aq.ctypes.c_void_p(owner.ptr)))
        out = [aq.DeviceColumn(n, c.code) for c in cols]
        cp = (aq.ctypes.c_void_p * len(cols))(*[c.ptr for c in cols])
        op = (aq.ctypes.c_void_p * len(cols))(*[c.ptr for c in out])
        eb = (aq.ctypes.c_int * len(cols))(*[c.itemsize for c in cols])
        counts = (aq.ctypes.c_uint32 * nparts)()
        aq.check(aq.lib().aqgpu_partition_rows(aq.ctypes.c_void_p(owner.ptr), aq.ctypes.c_uint32(n), aq.ctypes.c_uint32(nparts),
                                               cp, eb, len(cols), op, counts))
        owner.free()
        return [int(x) for x in counts], out

    def groupby_rows(self, keys, aggs, hint, bounds, max_groups):
        """whole groups on this rank (row exchange): the requested aggregates themselves, plus first rows"""
        res = aq.groupby_agg(keys, aggs, groups_hint=hint, max_groups=max_groups, key_bounds=bounds, device_out=True,
                             want_first_row=True)
        return res.ngroups, res.keys, res.aggs, res.first_row

    def take(self, col, idx, n):
        """col[idx[0..n)]: global row ids of the groups' first rows (n is G-sized; torch indexing is plumbing)"""
        t = self.torch
        src = self.bytes_of(col, col.n).view(t.int64 if col.itemsize == 8 else t.int32)
        ix = self.bytes_of(idx, n).view(t.int32).to(t.int64) & 0xffffffff
        return self.from_bytes(src[ix].contiguous().view(t.uint8), n, col.code)

    def global_rows(self, lo, n):
        """int64 column lo, lo + 1, ..., lo + n - 1"""
        t = self.torch
        return self.from_bytes(t.arange(lo, lo + n, dtype=t.int64, device=self.comm.device).view(t.uint8), n, aq.INT64)

    @staticmethod
    def _offsets(g):
        off = np.zeros(g.ngroups + 1, np.uint32)
        aq.check(aq.lib().aqgpu_groups_csr(g.h, off.ctypes.data_as(aq.ctypes.c_void_p), None))
        return off

    def _gather(self, g, src, dst, n):
        if src.itemsize in (4, 8):
            code = aq.INT32 if src.itemsize == 4 else aq.INT64
            aq.check(aq.lib().aqgpu_group_gather(g.h, code, aq.ctypes.c_void_p(src.ptr), aq.ctypes.c_void_p(dst.ptr)))
        else:           # 16-byte cells: two 8-byte halves through torch views would need strides; not exchanged today
            raise aq.AqgpuError("16-byte partial cells are not exchanged (partials are <= 8 bytes wide)")

    def _reorder_blocks(self, col, off, order, n):
        t = self.bytes_of(col, n)
        out = self.torch.empty_like(t)
        pos = 0
        for i in order:
            a, b = int(off[i]) * col.itemsize, int(off[i + 1]) * col.itemsize
            out[pos:pos + b - a].copy_(t[a:b])
            pos += b - a
        return self.from_bytes(out, n, col.code)

    def to_numpy(self, col, n):
        return col.numpy(n)


# ------------------------------------------------------------------------------------------------
# one distributed query
# ------------------------------------------------------------------------------------------------
GATHER_LIMIT = 4_000_000        # partial rows over all ranks up to which every rank merges everything


class DistQuery:
    """keys / aggs name columns of `cols` (this rank's shard, rows [lo, lo + n) of the global table)."""

    def __init__(self, comm: Comm, engine, cols: dict, codes: dict, keys, aggs, lo: int, max_groups: int,
                 hint: int = 0, bounds=None, gather_limit: int = GATHER_LIMIT, global_hint: int | None = None):
        self.comm, self.eng, self.lo = comm, engine, int(lo)
        self.key_names, self.aggs = list(keys), list(aggs)
        self.key_codes = [codes[k] for k in keys]
        self.hint, self.bounds, self.gather_limit = hint, bounds, gather_limit
        self.partials, self.merges, self.finals = decompose(aggs, codes)
        self.pcodes = [p[3] or aq.reduce_out_type(aq.REDUCE_NAMES[p[0]], codes[p[1]] if p[1] else aq.INT32) for p in self.partials]
        self.local = engine.local_prepare([cols[k] for k in keys], self.partials, cols, max_groups, hint, bounds)
        self.strategy = None
        self.exchanged_bytes = 0
        # groups ~ rows (h2o q10): local aggregation cannot shrink anything, so the ROWS are exchanged by key
        # owner and every group is aggregated whole on its owner (any aggregate, no merge algebra)
        self.cols = cols
        k0 = cols[keys[0]]
        self.nrows = int(k0.n) if hasattr(k0, "n") else len(k0)
        total_rows = sum(comm.all_counts(self.nrows))
        # decided from the GLOBAL group estimate against the GLOBAL row count (hint itself is capped at the shard size)
        ghint = hint if global_hint is None else global_hint
        self.by_rows = comm.world > 1 and ghint > 0 and 2 * ghint >= total_rows and total_rows > gather_limit

    def run_rows(self):
        c, e = self.comm, self.eng
        trace = os.environ.get("AQ_DIST_TRACE") and c.rank == 0
        t0 = [time.perf_counter()]

        def mark(what):
            if trace:
                if c.stream is not None:
                    c.stream.synchronize()
                t1 = time.perf_counter()
                print(f"[dist rows] {what}: {1e3 * (t1 - t0[0]):.1f} ms", file=sys.stderr, flush=True)
                t0[0] = t1
        nk = len(self.key_names)
        vnames = []
        for a in self.aggs:
            for name in a[1:]:
                if name not in vnames:
                    vnames.append(name)
        grow = e.global_rows(self.lo, self.nrows)
        table = [self.cols[k] for k in self.key_names] + [self.cols[v] for v in vnames] + [grow]
        tcodes = self.key_codes + [self.cols[v].code if hasattr(self.cols[v], "code") else aq.NP2CODE[self.cols[v].dtype] for v in vnames] + [aq.INT64]
        self.strategy = "alltoall-rows"
        mark("global rows")
        send, parted = e.partition(table[:nk], self.nrows, c.world, table)
        del grow, table
        mark("partition")
        m = c.count_matrix(send)
        rcounts = [m[src][c.rank] for src in range(c.world)]
        mmax = max(max(r) for r in m)
        nmerge = sum(rcounts)
        recv = []
        self.exchanged_bytes = 0
        for i, code in enumerate(tcodes):
            sz = aq.CODE_SIZE[code]
            recv.append(e.from_bytes(c.alltoall_bytes(e.bytes_of(parted[i], self.nrows), send, rcounts, sz, mmax), nmerge, code))
            parted[i] = None                                  # the send copy of this column is released
            self.exchanged_bytes += (self.nrows - send[c.rank]) * sz
        mark("all-to-all")
        rkeys, rgrow = recv[:nk], recv[-1]
        rv = dict(zip(vnames, recv[nk:-1]))
        aggs = [(a[0],) + tuple(rv[x] for x in a[1:]) for a in self.aggs]
        hint = min(nmerge, self.hint) if self.hint else 0
        ngm, mkeys, results, fr = e.groupby_rows(rkeys, aggs, hint, self.bounds, max(nmerge, 1))
        mark("group-by")
        fg = e.take(rgrow, fr, ngm)
        mark("take")
        return ngm, mkeys, results, fg

    def run(self):
        """returns (ngroups on this rank, key columns, result columns, first global rows); with the gather
        strategy every rank holds all groups in the reference's order, with the partitioned strategy
        rank r holds the groups it owns, each rank's in ascending first-global-row order"""
        c, e = self.comm, self.eng
        if self.by_rows:
            return self.run_rows()
        ng = self.local.run()
        grow = e.rows_to_global(self.local.fr, ng, self.lo)
        table = list(self.local.kouts) + list(self.local.outs) + [grow]
        tcodes = self.key_codes + self.pcodes + [aq.INT64]
        counts = c.all_counts(ng)
        total = sum(counts)
        self.exchanged_bytes = 0
        if total <= self.gather_limit:
            self.strategy = "allgather"
            recv = []
            for col, code in zip(table, tcodes):
                sz = aq.CODE_SIZE[code]
                recv.append(e.from_bytes(c.allgather_bytes(e.bytes_of(col, ng), counts, sz), total, code))
                self.exchanged_bytes += (total - ng) * sz
            nmerge = total
        else:
            self.strategy = "alltoall"
            send, parted = e.partition(table[:len(self.key_names)], ng, c.world, table)
            m = c.count_matrix(send)
            rcounts = [m[src][c.rank] for src in range(c.world)]
            mmax = max(max(r) for r in m)
            nmerge = sum(rcounts)
            recv = []
            for col, code in zip(parted, tcodes):
                sz = aq.CODE_SIZE[code]
                recv.append(e.from_bytes(c.alltoall_bytes(e.bytes_of(col, ng), send, rcounts, sz, mmax), nmerge, code))
                self.exchanged_bytes += (ng - send[c.rank]) * sz
        nk = len(self.key_names)
        rkeys, rparts, rgrow = recv[:nk], recv[nk:-1], recv[-1]
        specs = []
        for op, src, out in self.merges:
            specs.append((op, rparts[src[1]] if src else None, None, out))
        specs.append(("min", rgrow, None, aq.INT64))
        hint = min(nmerge, self.hint) if self.hint else 0
        ngm, mkeys, mouts = e.groupby(rkeys, specs, hint, self.bounds, max(nmerge, 1))
        first_global = mouts[-1]
        results = []
        for f in self.finals:
            if f[0] == "col":
                results.append(mouts[f[1]])
            elif f[0] == "fin":
                _, op, cnt, st, ss = f
                results.append(e.finalize(op, ngm, mouts[cnt], st, [mouts[i] for i in ss]))
            else:
                _, code, cnt_top, nparts, second_top, first_prev = f
                results.append(e.second_merge(code, ngm, mouts[cnt_top], mouts[nparts], mouts[second_top], mouts[first_prev]))
        return ngm, mkeys, results, first_global


class DistPlan:
    """the h2o queries of bench.py over row shards (one DistQuery each); mirrors h2o.Plan"""

    def __init__(self, comm: Comm, table: dict, queries, n_total: int, k: int, lo: int, use_bounds: bool = True):
        from . import h2o
        self.comm, self.queries = comm, list(queries)
        self.eng = GpuEngine(comm)
        codes = {c: d.code for c, d in table.items()}
        nrows = next(iter(table.values())).n
        self.dq, self.epilogue, self.extra = {}, {}, {}
        for q in self.queries:
            keys, aggs, epi = h2o.QUERIES[q]
            mg = h2o.max_groups(q, n_total, k, nrows)
            bounds = [h2o.key_bound(c, n_total, k) for c in keys] if use_bounds else None
            self.dq[q] = DistQuery(comm, self.eng, table, codes, keys, aggs, lo, mg, hint=mg, bounds=bounds,
                                   global_hint=h2o.max_groups(q, n_total, k, n_total))
            self.epilogue[q] = epi

    def run(self, q: str) -> int:
        ng, keys, res, fg = self.dq[q].run()
        epi = self.epilogue[q]
        if epi == "sub":
            out = aq.DeviceColumn(max(ng, 1), aq.INT32)
            aq.vec_sub(res[0], res[1], out, ng, aq.INT32)
        elif epi == "pow2":
            out = aq.DeviceColumn(max(ng, 1), aq.DOUBLE)
            aq.vec_pow(res[0], 2.0, out, ng)
        self.last = (ng, keys, res)
        self.last_first_global = fg
        if self.comm.stream is not None:
            # one query at a time on the device (the reference's engine thread does the same, server.cpp:236-238):
            # letting the host run ahead into the next query's collectives made steps 2-3x slower at 2 and 4 GPUs
            self.comm.stream.synchronize()
        return ng

    def step(self) -> dict:
        if not (os.environ.get("AQ_DIST_TRACE") and self.comm.rank == 0):
            return {q: self.run(q) for q in self.queries}
        out, line = {}, []
        for q in self.queries:                                # per-query wall time inside a step (debug aid)
            self.comm.stream.synchronize()
            t0 = time.perf_counter()
            out[q] = self.run(q)
            self.comm.stream.synchronize()
            line.append(f"{q} {1e3 * (time.perf_counter() - t0):.1f}")
        t = self.comm.torch
        print("[dist step] " + "  ".join(line) + f"  | torch reserved {t.cuda.memory_reserved() / 2**30:.1f} GiB", file=sys.stderr, flush=True)
        return out

    def total_groups(self, ngs: dict) -> dict:
        """groups over all ranks: the gather strategy replicates them, the partitioned one splits them"""
        out = {}
        for q, ng in ngs.items():
            out[q] = ng if self.dq[q].strategy == "allgather" else int(self.comm.sum_float(float(ng)))
        return out

    def exchange_report(self) -> dict:
        return {q: {"strategy": self.dq[q].strategy, "bytes_in_per_rank": self.dq[q].exchanged_bytes} for q in self.queries}

    def verify(self, table, queries, args, n_total, ngs) -> dict:
        return {"status": "skipped (multi-GPU result checks: tests/test_dist_gpu.py)"}


class ShardPlan(DistPlan):
    """the h2o queries over row shards through the library's own multi-GPU entry point (aqgpu_groupby_run_dist:
    NCCL communicator inside libaqgpu, dense record tables combined over NVLink); a query whose shape the library
    declines (groups ~ rows, first / second) runs through DistQuery (row / partial exchange driven from here)."""

    def __init__(self, comm: Comm, table: dict, queries, n_total: int, k: int, lo: int, use_bounds: bool = False):
        from . import h2o
        self.comm, self.queries = comm, list(queries)
        self.eng = GpuEngine(comm)
        self.table, self.n_total, self.k, self.lo, self.use_bounds = table, n_total, k, lo, use_bounds
        self.nrows = next(iter(table.values())).n
        uid = [aq.comm_unique_id() if comm.rank == 0 else None]
        comm.dist.broadcast_object_list(uid, src=0)
        aq.comm_init(comm.world, comm.rank, uid[0])
        self.sq, self.dq, self.epilogue, self.extra, self.native = {}, {}, {}, {}, {}
        for q in self.queries:
            keys, aggs, epi = h2o.QUERIES[q]
            specs = [(a[0],) + tuple(table[c] for c in a[1:]) for a in aggs]
            self.sq[q] = aq.ShardQuery([table[c] for c in keys], specs, row_base=lo, rows_total=n_total)
            self.epilogue[q] = epi
            self.native[q] = None
        self.exbytes = {}

    def _fallback(self, q: str):
        from . import h2o
        if q not in self.dq:
            keys, aggs, _ = h2o.QUERIES[q]
            codes = {c: d.code for c, d in self.table.items()}
            mg = h2o.max_groups(q, self.n_total, self.k, self.nrows)
            bounds = [h2o.key_bound(c, self.n_total, self.k) for c in keys] if self.use_bounds else None
            self.dq[q] = DistQuery(self.comm, self.eng, self.table, codes, keys, aggs, self.lo, mg, hint=mg, bounds=bounds,
                                   global_hint=h2o.max_groups(q, self.n_total, self.k, self.n_total))
        return self.dq[q]

    def run(self, q: str) -> int:
        if self.native[q] is not False:
            ng = self.sq[q].run()
            self.native[q] = ng is not None
            if ng is not None:
                res = self.sq[q].outs
                epi = self.epilogue[q]
                if epi:
                    code = aq.INT32 if epi == "sub" else aq.DOUBLE
                    e = self.extra.get(q)
                    if e is None or e.n < ng:
                        if e is not None:
                            e.free()
                        e = self.extra[q] = aq.DeviceColumn(max(ng, 1), code)
                    if epi == "sub":
                        aq.vec_sub(res[0], res[1], e, ng, code)
                    else:
                        aq.vec_pow(res[0], 2.0, e, ng)
                if self.comm.stream is not None:
                    self.comm.stream.synchronize()      # one query at a time on the device (as DistPlan.run does)
                return ng
        self._fallback(q)
        return DistPlan.run(self, q)

    def total_groups(self, ngs: dict) -> dict:
        out = {}
        for q, ng in ngs.items():
            if self.native[q]:
                out[q] = ng if self.sq[q].replicated else int(self.comm.sum_float(float(ng)))
            else:
                out[q] = ng if self.dq[q].strategy == "allgather" else int(self.comm.sum_float(float(ng)))
        return out

    def exchange_report(self) -> dict:
        rep = {}
        for q in self.queries:
            if self.native[q]:
                rep[q] = {"strategy": {1: "record tables: all-gather (replicated result)",
                                       0: "record tables: all-to-all of slot ranges (owner emits)",
                                       2: "fingerprints: all-to-all of (row, fingerprint) tuples (home of the first row emits)"}[self.sq[q].spread],
                          "native": True}
            else:
                rep[q] = {"strategy": self.dq[q].strategy, "bytes_in_per_rank": self.dq[q].exchanged_bytes, "native": False}
        return rep

    def gather_ordered(self, q: str):
        """the result of the last run(q) assembled in the reference's order (global first appearance) on every
        rank: (key columns, aggregate columns, first rows) as numpy arrays.  Replicated results are already in
        that order; partitioned ones (each rank: the groups it owns, ascending first row) are gathered and merged
        by first row.  For result checking and small results -- a consumer that keeps working on the GPUs uses
        the partitioned form."""
        import numpy as np
        if self.native[q]:
            sq = self.sq[q]
            ng = sq.ngroups
            keys = [c.numpy(ng) for c in sq.kouts]
            aggs = [c.numpy(ng) for c in sq.outs]
            first = sq.fr.numpy(ng).astype(np.int64)
            replicated = sq.replicated
        else:
            ng, kcols, res = self.last
            keys = [self.eng.to_numpy(c, ng) for c in kcols]
            aggs = [self.eng.to_numpy(c, ng) for c in res]
            first = self.eng.to_numpy(self.last_first_global, ng).astype(np.int64)
            replicated = self.dq[q].strategy == "allgather"
        if replicated:
            return keys, aggs, first
        parts = [None] * self.comm.world
        self.comm.dist.all_gather_object(parts, (keys, aggs, first))
        first = np.concatenate([p[2] for p in parts])
        order = np.argsort(first, kind="stable")
        keys = [np.concatenate([p[0][i] for p in parts])[order] for i in range(len(keys))]
        aggs = [np.concatenate([p[1][i] for p in parts])[order] for i in range(len(aggs))]
        return keys, aggs, first[order]

    def verify(self, table, queries, args, n_total, ngs) -> dict:
        """size-independent checks of the results the timed loop left in HBM, over all ranks: first rows strictly
        ascending on every rank; the key columns of first rows that lie in this rank's shard equal the table's key
        columns there; sum over all groups of every integer SUM / COUNT = the column total / the row count
        (all-reduced).  Bit-exact parity with the oracle at small sizes: tests/test_dist_gpu.py."""
        from . import h2o
        T, comm = self.comm.torch, self.comm
        dev = comm.device
        rep = {}

        def view(col, n, dt):
            return self.eng.bytes_of(col, n).view(dt) if n else T.zeros(0, dtype=dt, device=dev)

        for q in self.queries:
            if not self.native[q]:
                rep[q] = "not checked (row / partial exchange path)"
                continue
            sq = self.sq[q]
            keys, aggs, _ = h2o.QUERIES[q]
            ng = sq.ngroups
            fr = view(sq.fr, ng, T.int32).to(T.int64) & 0xffffffff
            if ng > 1 and not bool((fr[1:] > fr[:-1]).all()):
                raise RuntimeError(f"{q}: first rows not strictly ascending on rank {comm.rank}")
            local = (fr >= self.lo) & (fr < self.lo + self.nrows)
            idx = (fr[local] - self.lo)
            for i, c in enumerate(keys):
                col = view(table[c], self.nrows, T.int32)
                if not bool((view(sq.kouts[i], ng, T.int32)[local] == col[idx]).all()):
                    raise RuntimeError(f"{q}: key column {c} differs from the table at the groups' first rows")
            share = 1.0 if not sq.replicated else 1.0 / comm.world        # replicated results count once
            checked = []
            for a, agg in enumerate(aggs):
                if agg[0] == "count":
                    got = float(view(sq.outs[a], ng, T.int32).sum(dtype=T.int64)) * share
                    want = float(self.nrows)
                elif agg[0] == "sum" and table[agg[1]].code == aq.INT32:
                    got = float(view(sq.outs[a], ng, T.int64)[0::2].sum(dtype=T.int64)) * share      # low words of the int128 cells
                    want = float(view(table[agg[1]], self.nrows, T.int32).sum(dtype=T.int64))
                else:
                    continue
                got, want = comm.sum_float(got), comm.sum_float(want)
                if got != want:
                    raise RuntimeError(f"{q}: sum over groups of {agg} = {got}, column total {want}")
                checked.append(agg[0])
            total = int(ngs[q]) if sq.replicated else int(comm.sum_float(float(ng)))
            rep[q] = {"groups": total, "totals_checked": checked, "spread": sq.spread}
        return {"status": "ok", "rows": n_total, "per_query": rep,
                "checks": "per rank: first rows ascending, keys == table[first_row] for local first rows; over all ranks: "
                          "sum(count) = rows, sum of integer sums = column totals; oracle parity at small sizes: tests/test_dist_gpu.py"}

    def close(self):
        aq.comm_destroy()


class DistWindowPlan:
    """the windowed trade queries (aquery2_b200/trade.py) over row shards: the rows are repartitioned by the owner of
    their symbol (aqgpu_repartition_rows: stable split + one grouped NCCL exchange per column, rows arrive in global
    row order), then every rank runs the single-GPU pipeline on the symbols it owns -- group build (+ order_by) + one
    fused window call.  Results stay partitioned by symbol owner (order-sensitive aggregates)."""

    def __init__(self, comm: Comm, table: dict, queries, args, n_total: int, lo: int):
        self.comm, self.table, self.queries, self.lo, self.n_total = comm, table, list(queries), lo, n_total
        uid = [aq.comm_unique_id() if comm.rank == 0 else None]
        comm.dist.broadcast_object_list(uid, src=0)
        aq.comm_init(comm.world, comm.rank, uid[0])
        self.local, self.handles, self.results, self.moved = {}, {}, {}, {}

    def _free(self, q):
        h = self.handles.pop(q, None)
        if h is not None:
            h.free()
        for c in self.results.pop(q, []):
            c.free()
        for c in self.local.pop(q, {}).values():
            c.free()

    def run(self, q: str) -> int:
        from . import trade
        trace = os.environ.get("AQ_DIST_TRACE") and self.comm.rank == 0
        t0 = [time.perf_counter()]

        def mark(what):
            if trace:
                aq.sync()
                t1 = time.perf_counter()
                print(f"[dist window] {q} {what}: {1e3 * (t1 - t0[0]):.1f} ms", file=sys.stderr, flush=True)
                t0[0] = t1
        self._free(q)
        mark("free")
        names = trade.COLUMNS[q]
        cols = [self.table[c] for c in names]
        out = aq.repartition_rows([self.table["sym"]], cols, self.lo)
        mark("repartition")
        t = dict(zip(names, out[:-1]))
        t["_row"] = out[-1]
        self.local[q] = t
        self.moved[q] = sum(c.itemsize for c in cols) + 4
        h = aq.Groups([t["sym"]], row_order=aq.ROWS_ASC)
        mark("group build")
        self.results[q] = h.window(trade.expressions(t), device_out=True, order_by=t["time_shuffled"] if q == "w_unsorted" else None)
        mark("window")
        self.handles[q] = h
        aq.sync()             # one query at a time on the device: a host that runs ahead into the next query's collectives made steps slower
        return h.ngroups

    def step(self) -> dict:
        return {q: self.run(q) for q in self.queries}

    def total_groups(self, ngs: dict) -> dict:
        return {q: int(self.comm.sum_float(float(ng))) for q, ng in ngs.items()}

    def exchange_report(self) -> dict:
        return {q: {"strategy": "rows repartitioned by hash(sym) mod ranks (stable split + grouped send / receive per column)",
                    "bytes_per_row": self.moved.get(q), "native": True} for q in self.queries}

    def verify(self, table, queries, args, n_total, ngs) -> dict:
        """every row arrived exactly once (sum of local rows = table rows, sum of global row ids = n (n - 1) / 2), the global
        row ids ascend on every rank; oracle parity of the windowed results at small sizes: tests/test_dist_gpu.py"""
        T = self.comm.torch
        rep = {}
        for q in self.queries:
            rows = self.local[q]["_row"]
            r = T.as_tensor(_CudaBytes(rows.ptr, rows.n * 4), device=self.comm.device).view(T.int32).to(T.int64) & 0xffffffff if rows.n else T.zeros(0, dtype=T.int64, device=self.comm.device)
            if rows.n > 1 and not bool((r[1:] > r[:-1]).all()):
                raise RuntimeError(f"{q}: received rows are not in ascending global row order on rank {self.comm.rank}")
            tot_n = self.comm.sum_float(float(rows.n))
            tot_s = self.comm.sum_float(float(r.sum()))
            if tot_n != float(n_total) or tot_s != float(n_total) * float(n_total - 1) / 2.0:
                raise RuntimeError(f"{q}: {tot_n} rows arrived (sum of ids {tot_s}), table has {n_total}")
            rep[q] = {"rows_here": rows.n, "groups": int(self.comm.sum_float(float(ngs[q])))}
        return {"status": "ok", "rows": n_total, "per_query": rep,
                "checks": "every row arrives exactly once, in ascending global row order; oracle parity at small sizes: tests/test_dist_gpu.py"}

    def close(self):
        for q in list(self.local):
            self._free(q)
        aq.comm_destroy()
