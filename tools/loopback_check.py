#!/usr/bin/env python
"""Loop-back (virtual ranks on one GPU) vs single-GPU build, with a report of what differs:
    python tools/loopback_check.py C3 0.05 2 parallel"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gravomg_b200 import capi, meshgen
from gravomg_b200.sharded import ShardedHierarchy

name, scale, world = sys.argv[1], float(sys.argv[2]), int(sys.argv[3])
mode = capi.SAMPLER_EXACT if sys.argv[4] == "exact" else capi.SAMPLER_PARALLEL
reps = int(sys.argv[5]) if len(sys.argv) > 5 else 1
cfg = meshgen.make_config(name, scale)
V, A = cfg["V"], cfg["A"]
ctx = capi.Context(0)
inp = capi.Input(ctx, V, A.indptr, A.indices, None)
for rep in range(reps):
    single = capi.Hierarchy(ctx, inp, cfg["levels"], sampler_mode=mode, seed=3)
    single2 = capi.Hierarchy(ctx, inp, cfg["levels"], sampler_mode=mode, seed=3)
    sh = ShardedHierarchy(ctx, inp, cfg["levels"], world, list(range(world)), comm=None, sampler_mode=mode, seed=3)
    for lvl in range(single.num_levels):
        w, w2, g = single.level_arrays(lvl), single2.level_arrays(lvl), sh.level_arrays(lvl)
        for k in "sel par ccp cri cval cpts prp pci pva".split():
            if not np.array_equal(w[k], w2[k]):
                d = np.flatnonzero(np.asarray(w[k]).ravel() != np.asarray(w2[k]).ravel())
                print(f"rep {rep} level {lvl} {k}: SINGLE-GPU build not reproducible at {len(d)} entries, first {d[:5]}")
            if w[k].shape != g[k].shape:
                print(f"rep {rep} level {lvl} {k}: shape {w[k].shape} vs {g[k].shape}")
                continue
            if not np.array_equal(w[k], g[k]):
                d = np.flatnonzero(np.asarray(w[k]).ravel() != np.asarray(g[k]).ravel())
                print(f"rep {rep} level {lvl} {k}: sharded differs at {len(d)} entries, first {d[:5]}: "
                      f"{np.asarray(w[k]).ravel()[d[:3]]} vs {np.asarray(g[k]).ravel()[d[:3]]}")
        for r in range(1, world):
            for key, which in (("cval", capi.H_COARSE_VAL), ("cri", capi.H_COARSE_ROWIDX)):
                x = sh.fetch(lvl, which, r)
                if not np.array_equal(x, w[key]):
                    d = np.flatnonzero(x != w[key])
                    print(f"rep {rep} level {lvl} vrank {r} {key}: differs at {len(d)} entries, first {d[:5]}: {w[key][d[:3]]} vs {x[d[:3]]}")
    print(f"rep {rep}: n={V.shape[0]} levels {[(i['n'], i['m']) for i in single.info]} done")
    sh.close(); single.close(); single2.close()
