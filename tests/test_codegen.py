"""aquery2_b200/codegen.py: the whole-column emitter for the ten h2o queries and the five windowed ones.
CPU: every generated translation unit compiles against the drop-in headers, laid out as the reference lays out
a snippet build (out.cpp next to server/).  GPU: each one is built into a program around tests/codegen/harness.cpp,
run on synthetic columns, and the table it produced is compared with the oracle."""
import os
import shutil
import subprocess

import numpy as np
import pytest

from aquery2_b200 import codegen, synth
from util import assert_same

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "aquery2_b200", "host")


def _workdir(tmp_path):
    shutil.copytree(os.path.join(HOST, "server"), tmp_path / "server")
    shutil.copy(os.path.join(HOST, "aqgpu_host.hpp"), tmp_path / "aqgpu_host.hpp")
    return tmp_path


def test_all_generated_units_compile(tmp_path):
    w = _workdir(tmp_path)
    for name in list(codegen.QUERIES) + list(codegen.EXTRA_QUERIES):          # EXTRA: string / date keys, a computed key (tests/network.a)
        (w / "out.cpp").write_text(codegen.generate(name))
        r = subprocess.run(["g++", "-std=c++17", "-O0", "-w", "-fsyntax-only", f"-I{os.path.join(ROOT, 'include')}", "out.cpp"],
                           cwd=w, capture_output=True, text=True)
        assert r.returncode == 0, f"{name}:\n" + r.stderr[-2500:]
    text = codegen.generate("w5")
    assert "wb.add_reduced<int>(AQ_MINS, 0, price_a, true, AQ_MAX)" in text and text.count("wb.run()") == 1
    assert 'extern "C" int dll_q9(Context* cxt)' in codegen.generate("q9")
    net = codegen.generate("network")
    assert "auto key_session = sums((deltas(_time_d) > 120));" in net and "aq::group_by(src_b, dst_c, key_session)" in net
    assert "ColRef<astring_view>(len, server->getCol(0, types::ASTR));" in net          # string cells are host pointers: not uploaded


def _oracle_columns(port, q, t):
    """expected result columns of a codegen query from the oracle's group-by / reduce / scan"""
    keys = q["group_by"]
    ng, gid, off, perm = port.groupby([t[k] for k in keys])
    order = perm
    if q.get("row_order") == "asc":
        order = np.concatenate([perm[off[g]:off[g + 1]][::-1] for g in range(ng)]).astype(np.uint32) if ng else perm
    if q.get("assuming"):                                          # TableInfo::order_by per group (table.h:447-461): stable, by the column
        oc, direction = q["assuming"]
        k = t[oc].astype(np.int64)
        order = np.lexsort((np.arange(len(gid)), -k if direction == "desc" else k, gid)).astype(np.uint32)
    first = np.array([perm[off[g + 1] - 1] for g in range(ng)], dtype=np.int64)

    def ev(e):
        if "col" in e:
            return t[e["col"]][first]
        if "op" in e:
            if "col" in e["l"]:                                    # col - scan(col): a vector
                return (t[e["l"]["col"]][order] - ev(e["r"])).astype(t[e["l"]["col"]].dtype)
            return ev(e["l"]) - ev(e["r"])
        f, a = e["fn"], e["args"]
        if f == "count":
            return np.diff(off).astype(np.uint32)
        if f == "pow":
            return ev(a[0]) ** a[1]["lit"]
        if f == "corr":
            return port.reduce("corr", t[a[0]["col"]], off, perm, v2=t[a[1]["col"]])
        if f in ("max", "min") and ("op" in a[0] or a[0].get("fn") in codegen.VECTOR_AGGS):
            v = ev(a[0])
            return (np.maximum if f == "max" else np.minimum).reduceat(v, off[:-1].astype(np.int64))
        if f in codegen.VECTOR_AGGS:
            w = a[0]["lit"] if len(a) == 2 else (1 if f == "ratios" else 0)
            op = {"sums": "sumw", "avgs": "avgw", "mins": "minw", "maxs": "maxw", "ratios": "ratiow"}.get(f, f) if len(a) == 2 or f == "ratios" else f
            return port.scan(op, t[a[-1]["col"]], off, order, w=w, raw128=True)
        return port.reduce(f, t[a[0]["col"]], off, perm, raw128=True)
    return [ev(e) for _, e in q["select"]]


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(codegen.QUERIES))
def test_generated_unit_runs_and_matches_the_oracle(gpu, port, tmp_path, name):
    q = codegen.QUERIES[name]
    n = 60_000
    if q["table"] == "source":
        t = synth.h2o_table(n, 10, seed=3)
        t["id3"] = (t["id3"] % 997).astype(np.int32); t["id6"] = (t["id6"] % 1013).astype(np.int32)
        if name == "q10":
            for c in ("id3", "id4", "id5", "id6"):
                t[c] = (t[c] % 3).astype(np.int32)
    else:
        t = synth.trade_table(n, 300, seed=3)
        if q.get("assuming"):                                      # an unsorted table: the time column is a permutation
            t["time"] = np.random.default_rng(5).permutation(n).astype(np.int32)
    w = _workdir(tmp_path)
    (w / "out.cpp").write_text(codegen.generate(name, dump=True))
    for i, (c, _) in enumerate(q["schema"]):
        t[c].tofile(w / f"col_{i}.bin")
    lib = os.path.join(ROOT, "aquery2_b200")
    r = subprocess.run(["g++", "-std=c++17", "-O1", "-w", f"-I{os.path.join(ROOT, 'include')}", '-DQUERY_TU="out.cpp"', f"-DQUERY_ENTRY=dll_{name}",
                        os.path.join(ROOT, "tests", "codegen", "harness.cpp"), "-I.", "-o", "prog", f"-L{lib}", "-laqgpu", f"-Wl,-rpath,{lib}"],
                       cwd=w, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2500:]
    run = subprocess.run(["./prog", str(n), str(len(q["schema"]))], cwd=w, capture_output=True, text=True, timeout=300)
    assert run.returncode == 0, run.stdout[-1500:] + run.stderr[-1500:]
    header = run.stdout.splitlines()[0].split(",")
    assert header == [nm for nm, _ in q["select"]], run.stdout[:300]          # the table the generated code printed
    want = _oracle_columns(port, q, t)
    for i, wcol in enumerate(want):
        wcol = np.ascontiguousarray(wcol)
        got = np.fromfile(w / f"out_{i}.bin", dtype=wcol.dtype).reshape(wcol.shape)
        assert_same(got, wcol, f"{name} column {i} ({header[i]})", scale=1.0)
