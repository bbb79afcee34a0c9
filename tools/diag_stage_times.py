#!/usr/bin/env python
"""Per-build stage times of repeated hierarchy builds (diagnostic): python tools/diag_stage_times.py C3 exact 6"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gravomg_b200 import meshgen, capi
from gravomg_b200.capi import Context, Input, Hierarchy

name, sampler, reps = sys.argv[1], sys.argv[2], int(sys.argv[3])
scale = float(sys.argv[4]) if len(sys.argv) > 4 else 1.0
cfg = meshgen.make_config(name, scale)
V, A = cfg["V"], cfg["A"]
ctx = Context(0)
inp = Input(ctx, V, A.indptr, A.indices, None)
mode = capi.SAMPLER_EXACT if sampler == "exact" else capi.SAMPLER_PARALLEL
for r in range(reps):
    t0 = time.perf_counter()
    h = Hierarchy(ctx, inp, cfg["levels"], sampler_mode=mode)
    wall = 1e3 * (time.perf_counter() - t0)
    print(f"build {r}: wall {wall:.1f} ms, device {h.build_ms:.1f} ms")
    for l, i in enumerate(h.info):
        print("   L%d" % l, " ".join(f"{n[:6]}={v:.2f}" for n, v in zip(capi.STAGE_NAMES, i["stage_ms"])))
    h.close()
