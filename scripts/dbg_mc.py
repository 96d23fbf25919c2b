"""debug driver: the mc.cu test cases one by one with timings (run under `timeout`)"""
import os, sys, time, faulthandler
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
faulthandler.dump_traceback_later(100, exit=True)
os.environ["AQGPU_MC_MIN_ROWS"] = "4096"
import numpy as np
from aquery2_b200 import aqgpu as gpu, synth
from oracle import aqo
gpu.init(0)
port = aqo.load_port()
n, k = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000, int(sys.argv[2]) if len(sys.argv) > 2 else 8
t = synth.h2o_table(n, k, seed=3, columns=["id3", "v1", "v2", "v3"])
SETS = ([("sum", "v1"), ("avg", "v3")], [("sum", "v1"), ("sum", "v2"), ("sum", "v3")], [("max", "v1"), ("min", "v2")],
        [("first", "v3"), ("second", "v3")], [("count",)], [("min", "v3"), ("max", "v3"), ("last", "v1")],
        [("var", "v3"), ("stddev", "v1")], [("corr", "v1", "v2")])
for aggs in SETS:
    t0 = time.time()
    print("run", aggs, flush=True)
    gpu.profile_enable(True); gpu.profile_report()
    res = gpu.groupby_agg([t["id3"]], [(a[0],) + tuple(t[c] for c in a[1:]) for a in aggs])
    rep = gpu.profile_report(); gpu.profile_enable(False)
    print("  ngroups", res.ngroups, "ms", round(1e3 * (time.time() - t0), 1), {k.split(":")[-1]: round(v[1], 3) for k, v in rep.items()}, flush=True)
print("done")
