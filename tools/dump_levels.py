#!/usr/bin/env python
"""Build a hierarchy on the GPU and dump every level to files (no viewer, no graphics stack):
    python tools/dump_levels.py C1 1.0 out_dir"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gravomg_b200 import capi, dump, meshgen

name = sys.argv[1] if len(sys.argv) > 1 else "C1"
scale = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
out = sys.argv[3] if len(sys.argv) > 3 else "gpurun_out/dump"
cfg = meshgen.make_config(name, scale)
ctx = capi.Context(0)
inp = capi.Input(ctx, cfg["V"], cfg["A"].indptr, cfg["A"].indices, None)
h = capi.Hierarchy(ctx, inp, cfg["levels"])
files = dump.dump_hierarchy(h, cfg["V"], out)
print(f"{len(files)} files in {out}: " + ", ".join(os.path.basename(f) for f in files[:9]) + " ...")
