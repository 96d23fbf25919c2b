"""SURVEY 8(b) "Threading": scratch-using entry points are serialised per process by a recursive lock that lives as long as
their ArenaScope (csrc/context.cu).  Host logic only -- no device needed: tests/native/scope_test.cpp opens nested scopes from
three threads and counts moments at which two threads were inside."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_scratch_scopes_are_mutually_exclusive_and_nest(tmp_path):
    lib = os.path.join(ROOT, "aquery2_b200")
    cuda_inc = os.path.join(os.environ.get("CUDA_HOME", "/usr/local/cuda"), "include")
    exe = tmp_path / "scope_test"
    r = subprocess.run(["g++", "-std=c++17", "-O1", f"-I{os.path.join(lib, 'csrc')}", f"-I{os.path.join(ROOT, 'include')}", f"-I{cuda_inc}",
                        os.path.join(ROOT, "tests", "native", "scope_test.cpp"), "-o", str(exe), f"-L{lib}", "-laqgpu", f"-Wl,-rpath,{lib}", "-lpthread"],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
    run = subprocess.run([str(exe)], capture_output=True, text=True, timeout=120)
    assert run.returncode == 0 and "violations 0" in run.stdout, run.stdout + run.stderr
