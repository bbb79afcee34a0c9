#!/usr/bin/env python
"""Pretty-print bench.py JSON lines: python tools/show_bench.py file.json [...]"""
import json, sys
for f in sys.argv[1:]:
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e:
        print(f, "ERR", e, open(f).read()[:300]); continue
    print(f"{f}: {d['value']:.1f} {d['unit']}  {d['ms_per_step']:.3f} ms/step  e2e {d['e2e']['value']:.1f} ({d['e2e'].get('ms_per_step',0):.1f} ms)  launches {d.get('gpu_launches')}")
    r = d.get("roofline", {})
    if r: print("  roof:", r.get("kernel"), f"frac {r['frac']:.4f}", "whole", {k: (round(v,4) if isinstance(v,float) else v) for k,v in r.get("whole_build",{}).items()})
    for l in d["config"].get("levels", []): print("   ", l)
    for k, v in d.get("stage_ms", {}).items(): print("   ", k, v)
    if d.get("cpu_baseline"): print("  cpu", d["cpu_baseline"])
    print("  clocks", d.get("clocks"))
