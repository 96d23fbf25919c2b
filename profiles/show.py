import json, sys
for f in sys.argv[1:]:
    try: d=json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e: print(f,"ERR",e); continue
    print(f, "value %.3e ms/step %.1f launches %s"%(d["value"], d["ms_per_step"], d.get("gpu_launches")))
    if d.get("e2e"): print("  e2e %.3e"%d["e2e"]["value"])
    if d.get("cpu_baseline"): print("  cpu %.3e"%d["cpu_baseline"]["value"])
    for q,v in d["per_query"].items(): print("  %-4s %9.2f ms %10.3e rows/s  hbm_frac %.3f groups %d"%(q, v["ms"], v["rows_per_s"], v["hbm_frac"], v["groups"]))
    for k,v in list(d["kernels"].items())[:14]: print("   ", k, v)
