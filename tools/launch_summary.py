#!/usr/bin/env python
"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel.
    python tools/launch_summary.py gpurun_out/launches.csv [first_n_launches_per_kernel]"""
import csv, sys, collections
rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if l.startswith('"'))]
hdr = rows[0]; ix = {h: i for i, h in enumerate(hdr)}
agg = collections.OrderedDict()
for r in rows[1:]:
    if r[ix["Metric Name"]] != "gpu__time_duration.sum": continue
    name = r[ix["Kernel Name"]].split("(")[0].replace("void ", "").replace("gmg::", "")
    v = float(r[ix["Metric Value"]].replace(",", "")); u = r[ix["Metric Unit"]]
    us = v * {"ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6}.get(u, 1e-3)
    agg.setdefault(name, []).append(us)
tot = sum(sum(v) for v in agg.values())
print(f"| kernel | launches | total us | share | first launch us (level 0) |\n|---|---:|---:|---:|---:|")
for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
    print(f"| `{k}` | {len(v)} | {sum(v):.1f} | {100*sum(v)/tot:.1f}% | {v[0]:.1f} |")
print(f"\nTotal {tot:.1f} us")
