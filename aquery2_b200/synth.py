"""Deterministic, counter-based synthetic inputs shared by the CPU oracle and every GPU shard.

Value of (seed, column stream, GLOBAL row) = f(splitmix64(stream_key + row)), so the host (numpy, this
file) and the device generator (csrc/synth.cu, ``aqgpu_synth_*``) produce bit-identical columns with no
transfer, and a rank that owns rows [lo, hi) of a table generates exactly that slice.

Shapes follow the reference's own benchmark files:
  * h2o "G1_N_K_0_0", numeric-id variant: ``source(id1..id6 int, v1 int, v2 int, v3 float)``
    (benchmark/h2o/groupby_prep.sql:1,4): id1,id2,id4,id5 = 1+U[0,K), id3,id6 = 1+U[0,N/K),
    v1 = 1+U[0,5), v2 = 1+U[0,15), v3 = round(U(0,100), 6) stored as 32-bit REAL.
  * trade table ``trade(sym int, time int, quantity int, price int)`` after datagen.cpp:9,90-107
    (quantity U[100,10000], price in [50,500]); time = row + 1 so the table is ASSUMING-ASC-time
    sorted, as the hybrid engine delivers it.
"""
from __future__ import annotations

import numpy as np

MASK64 = (1 << 64) - 1
GAMMA = 0x9E3779B97F4A7C15
H2O_COLUMNS = ("id1", "id2", "id3", "id4", "id5", "id6", "v1", "v2", "v3")
TRADE_COLUMNS = ("sym", "time", "quantity", "price")
# column -> stream id (also used by csrc/synth.cu)
STREAM = {name: i for i, name in enumerate(H2O_COLUMNS)}
STREAM.update({"sym": 16, "time": 17, "quantity": 18, "price": 19, "price_base": 20})


def _splitmix64_scalar(z: int) -> int:
    z = (z + GAMMA) & MASK64
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & MASK64
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & MASK64
    return z ^ (z >> 31)


def stream_key(seed: int, stream: int) -> int:
    """Per-(seed, column) 64-bit key; the device generator receives this value from the host."""
    return _splitmix64_scalar((seed * 0x100000001B3 + (stream + 1) * 0xD1B54A32D192ED03) & MASK64)


def _splitmix64(z: np.ndarray) -> np.ndarray:
    with np.errstate(over="ignore"):
        z = z + np.uint64(GAMMA)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def raw(seed: int, stream: int, lo: int, hi: int) -> np.ndarray:
    """64 random bits for global rows [lo, hi) of one column stream."""
    rows = np.arange(lo, hi, dtype=np.uint64)
    with np.errstate(over="ignore"):
        return _splitmix64(rows + np.uint64(stream_key(seed, stream)))


def _uniform(x: np.ndarray, k: int) -> np.ndarray:
    """U[0, k) from the high 32 bits by multiply-high (k < 2**32)."""
    return ((x >> np.uint64(32)) * np.uint64(k)) >> np.uint64(32)


def h2o_column(name: str, n_total: int, k: int, lo: int = 0, hi: int | None = None,
               seed: int = 42) -> np.ndarray:
    hi = n_total if hi is None else hi
    x = raw(seed, STREAM[name], lo, hi)
    big = max(n_total // k, 1)
    if name in ("id1", "id2", "id4", "id5"):
        return (1 + _uniform(x, k)).astype(np.int32)
    if name in ("id3", "id6"):
        return (1 + _uniform(x, big)).astype(np.int32)
    if name == "v1":
        return (1 + _uniform(x, 5)).astype(np.int32)
    if name == "v2":
        return (1 + _uniform(x, 15)).astype(np.int32)
    if name == "v3":
        return (_uniform(x, 100_000_000).astype(np.float64) / 1e6).astype(np.float32)
    raise KeyError(name)


def h2o_table(n_total: int, k: int, lo: int = 0, hi: int | None = None, seed: int = 42,
              columns=H2O_COLUMNS) -> dict[str, np.ndarray]:
    return {c: h2o_column(c, n_total, k, lo, hi, seed) for c in columns}


def trade_column(name: str, n_total: int, n_sym: int, lo: int = 0, hi: int | None = None,
                 seed: int = 42) -> np.ndarray:
    hi = n_total if hi is None else hi
    if name == "time":
        return np.arange(lo + 1, hi + 1, dtype=np.int64).astype(np.int32)
    if name == "time_shuffled":      # a permutation of 1..n_total (n_total not a multiple of 7): the unsorted variant's time
        r = np.arange(lo, hi, dtype=np.uint64)
        return (1 + (r * np.uint64(40353607) + np.uint64(12345)) % np.uint64(n_total)).astype(np.int64).astype(np.int32)
    if name == "sym":
        return _uniform(raw(seed, STREAM["sym"], lo, hi), n_sym).astype(np.int32)
    if name == "quantity":
        return (100 + _uniform(raw(seed, STREAM["quantity"], lo, hi), 9901)).astype(np.int32)
    if name == "price":
        sym = _uniform(raw(seed, STREAM["sym"], lo, hi), n_sym)
        with np.errstate(over="ignore"):
            base = _uniform(_splitmix64(sym + np.uint64(stream_key(seed, STREAM["price_base"]))), 441)
        noise = _uniform(raw(seed, STREAM["price"], lo, hi), 11)
        return (50 + base + noise).astype(np.int32)          # in [50, 500]
    raise KeyError(name)


def trade_table(n_total: int, n_sym: int, lo: int = 0, hi: int | None = None, seed: int = 42,
                columns=TRADE_COLUMNS) -> dict[str, np.ndarray]:
    return {c: trade_column(c, n_total, n_sym, lo, hi, seed) for c in columns}
