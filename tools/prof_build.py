#!/usr/bin/env python
"""One hierarchy build for profiling (ncu / compute-sanitizer wrap this):
    python tools/prof_build.py <C1|C2|C3|C4> <scale> <exact|parallel|hashed> <galerkin 0|1> [builds]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gravomg_b200 import capi, meshgen  # noqa: E402

name, scale, sampler, galerkin = sys.argv[1], float(sys.argv[2]), sys.argv[3], int(sys.argv[4])
builds = int(sys.argv[5]) if len(sys.argv) > 5 else 1
cfg = meshgen.make_config(name, scale)
ctx = capi.Context(0)
inp = capi.Input(ctx, cfg["V"], cfg["A"].indptr, cfg["A"].indices, cfg["A"].data if galerkin else None)
mode = {"exact": capi.SAMPLER_EXACT, "parallel": capi.SAMPLER_PARALLEL, "hashed": capi.SAMPLER_HASHED}[sampler]
for _ in range(builds):
    h = capi.Hierarchy(ctx, inp, cfg["levels"], sampler_mode=mode, build_galerkin=bool(galerkin))
    print(name, scale, sampler, [(i["n"], i["m"]) for i in h.info], f"{h.build_ms:.3f} ms",
          [round(x, 3) for x in h.info[0]["stage_ms"]], [round(x, 3) for x in h.info[0]["galerkin_ms"]], flush=True)
    h.close()
