"""seeded key columns that are not plain integers: strings, date / time / timestamp cells"""
import numpy as np

from oracle import aqo


def strings(rng, n, ndistinct, long_tail=False):
    """n strings drawn (skewed) from a vocabulary with the awkward cases: the empty string, one-character strings,
    strings that are prefixes of each other, multi-byte characters, and (long_tail) a few strings of 100..400 bytes"""
    vocab = [b"", b"A", b"AA", b"AAA", b"IBM", b"IBM.N", b"a\xc3\xa9", b"\x01\x02", b" ", b"  "]
    alphabet = np.frombuffer(b"ABCDEFGHIJKLMNOPQRSTUVWXYZ.", np.uint8)
    seen = set(vocab)
    while len(vocab) < ndistinct:
        ln = int(rng.integers(100, 400)) if long_tail and len(vocab) % 97 == 0 else int(rng.integers(1, 9))
        w = rng.choice(alphabet, ln).tobytes()
        if w not in seen:
            seen.add(w); vocab.append(w)
    vocab = vocab[:ndistinct]
    pick = np.minimum(rng.zipf(1.2, n) - 1, len(vocab) - 1) if n > 50 else rng.integers(0, len(vocab), n)
    pick = rng.permutation(len(vocab))[pick]
    return aqo.StrColumn([vocab[i] for i in pick])


def times(rng, n, nd):
    t = np.zeros(n, aqo.TIME_DT)
    t["ms"] = rng.integers(0, max(nd // 8, 1), n) * 125
    t["seconds"] = rng.integers(0, 2, n) * 59
    t["minutes"] = rng.integers(0, 2, n) * 30
    t["hours"] = rng.integers(0, 2, n) * 23
    t["pad"] = rng.integers(0, 256, n)                  # garbage the reference's operator== never looks at
    return t


def dates(rng, n, nd):
    d = np.zeros(n, aqo.DATE_DT)
    d["day"] = rng.integers(1, 29, n)
    d["month"] = rng.integers(1, 13, n)
    d["year"] = rng.integers(-2, max(nd // 300, 1), n) + 2000 * rng.integers(0, 2, n)    # negative years too
    return d


def timestamps(rng, n, nd):
    ts = np.zeros(n, aqo.TIMESTAMP_DT)
    ts["date"] = dates(rng, n, 4)
    ts["date"]["day"] = rng.integers(1, 3, n)
    ts["date"]["month"] = 7
    ts["time"] = times(rng, n, nd)
    return ts


def from_golden(g):
    """the key columns stored in tests/golden/special_small.npz (oracle/gen_golden.py gen_special)"""
    data, off = g["strcol_bytes"], g["strcol_offsets"].astype(np.int64)
    raw = data.tobytes()
    cols = {"str": aqo.StrColumn([raw[off[i]:off[i + 1]] for i in range(len(off) - 1)]), "k2": g["k2"]}
    for name, dt in (("date", aqo.DATE_DT), ("time", aqo.TIME_DT), ("timestamp", aqo.TIMESTAMP_DT)):
        cols[name] = np.ascontiguousarray(g[f"{name}_cells"]).view(dt).reshape(-1)
    return cols
