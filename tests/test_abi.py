"""The C-ABI library loads and exports every symbol include/aqgpu.h declares (no compute calls here:
this runs without a GPU), and refuses to compute without a CUDA device (no CPU fallback)."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "aqgpu.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(aqgpu_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported():
    from aquery2_b200 import aqgpu
    L = aqgpu.lib()
    syms = declared_symbols()
    assert len(syms) >= 30
    for s in syms:
        assert hasattr(L, s), f"libaqgpu.so does not export {s}"
    assert sorted(aqgpu.EXPORTS) == [s for s in syms if s in aqgpu.EXPORTS]
    assert set(aqgpu.EXPORTS) == set(syms), set(syms) ^ set(aqgpu.EXPORTS)
    assert L.aqgpu_abi_version() == 1


def test_result_types_follow_reference():
    """server/types.h:199-210 (GetLongType / GetFPType); no device needed for these entry points."""
    from aquery2_b200 import aqgpu as A
    assert A.reduce_out_type(A.SUM, A.INT32) == A.INT128
    assert A.reduce_out_type(A.SUM, A.INT64) == A.INT128
    assert A.reduce_out_type(A.SUM, A.FLOAT) == A.DOUBLE
    assert A.reduce_out_type(A.AVG, A.INT32) == A.DOUBLE
    assert A.reduce_out_type(A.MIN, A.FLOAT) == A.FLOAT
    assert A.reduce_out_type(A.COUNT, A.INT32) == A.UINT32
    assert A.scan_out_type(A.SUMS, A.INT32) == A.INT128
    assert A.scan_out_type(A.AVGW, A.INT32) == A.DOUBLE
    assert A.scan_out_type(A.RATIOW, A.INT32) == A.FLOAT
    assert A.scan_out_type(A.RATIOW, A.INT64) == A.DOUBLE
    assert A.scan_out_type(A.DELTAS, A.FLOAT) == A.FLOAT


def test_oracle_and_library_agree_on_types(port):
    from aquery2_b200 import aqgpu as A
    for vt in (A.INT32, A.INT64, A.FLOAT, A.DOUBLE):
        for op in range(10):
            assert A.reduce_out_type(op, vt) == port.lib.aqo_reduce_out_type(op, vt)
        for op in range(12):
            assert A.scan_out_type(op, vt) == port.lib.aqo_scan_out_type(op, vt)


def test_no_cpu_fallback():
    """Without a CUDA device every compute entry point must fail with a message, never compute."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from aquery2_b200 import aqgpu as A
    with pytest.raises(A.AqgpuError, match="CUDA device"):
        A.groupby_agg([np.array([1, 2, 1], np.int32)], [("sum", np.array([1, 2, 3], np.int32))])


def test_product_does_not_touch_oracle():
    """The product package never imports, links or loads anything under oracle/."""
    pkg = os.path.join(ROOT, "aquery2_b200")
    needles = ("import oracle", "from oracle", "libaqoracle", "libaqref", "aqo_", "aqo.")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp")) or f == "Makefile":
                txt = open(os.path.join(dp, f), errors="replace").read()
                for needle in needles:
                    assert needle not in txt, f"{os.path.join(dp, f)} mentions {needle}"


def test_host_headers_compile():
    """the drop-in C++ headers (aquery2_b200/host/server/*.h) compile emitter-shaped code against the C ABI"""
    import subprocess
    host = os.path.join(ROOT, "aquery2_b200", "host")
    subprocess.run(["make", "-s", "-C", host], check=True)
    assert os.path.exists(os.path.join(host, "selftest"))
    for h in ("vector_type.hpp", "table.h", "hasher.h", "aggregations.h", "types.h", "gc.h", "libaquery.h"):
        assert os.path.exists(os.path.join(host, "server", h)), h


def test_header_is_plain_c(tmp_path):
    """the boundary is a C ABI: include/aqgpu.h compiles as C99 / C11 with -pedantic, and a C program links against the library"""
    import subprocess
    src = os.path.join(ROOT, "tests", "native", "c_abi.c")
    for std in ("c99", "c11"):
        r = subprocess.run(["gcc", f"-std={std}", "-Wall", "-Wextra", "-pedantic", "-Werror", "-fsyntax-only", f"-I{os.path.join(ROOT, 'include')}", src],
                           capture_output=True, text=True)
        assert r.returncode == 0, r.stderr[-1500:]
    lib = os.path.join(ROOT, "aquery2_b200")
    exe = tmp_path / "c_abi"
    r = subprocess.run(["gcc", "-std=c11", f"-I{os.path.join(ROOT, 'include')}", src, "-o", str(exe), f"-L{lib}", "-laqgpu", f"-Wl,-rpath,{lib}"],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-1500:]
    assert subprocess.run([str(exe)]).returncode == 0           # aqgpu_abi_version() == AQGPU_ABI_VERSION
