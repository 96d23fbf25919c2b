"""debug aid: sum(count) of a no-hints group-by at full size under different AQGPU_DISABLE settings"""
import os, sys, subprocess, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
if len(sys.argv) > 1 and sys.argv[1] == "child":
    import torch
    from aquery2_b200 import aqgpu as aq, synth
    import bench_verify as bv
    n = int(float(sys.argv[2])); col = sys.argv[3]
    aq.init(0)
    kk = n // 100 if col in ("id3", "id6") else 100
    d = aq.synth_column(0, synth.stream_key(42, synth.STREAM[col]), n, kk)
    res = aq.groupby_agg([d], [("count",)], device_out=True, want_first_row=True)
    cnt = bv.tview(torch, aq, res.aggs[0], res.ngroups)
    print(json.dumps({"disable": os.environ.get("AQGPU_DISABLE", ""), "ngroups": res.ngroups, "sum_count": int(cnt.sum().item()), "rows": n}))
else:
    for dis in ("plan", "plan,smem", "plan,smem,part", ""):
        env = dict(os.environ, AQGPU_DISABLE=dis)
        r = subprocess.run([sys.executable, __file__, "child", sys.argv[1] if len(sys.argv) > 1 else "1e9", "id3"], env=env, capture_output=True, text=True)
        print(r.stdout.strip()[-300:], r.stderr.strip()[-300:])
