#!/usr/bin/env python
"""Where the tile flow's time goes, from a per-rank trace (GMG_SAMPLER_TRACE=file makes the flow kernel stamp
%globaltimer when a rank is queued / popped / handed to the other-tile warps / finished):

    GMG_SAMPLER_TRACE=gpurun_out/trace.bin python tools/prof_build.py C3 0.25 exact 0 1     # on the GPU box
    python tools/flow_trace/analyse.py gpurun_out/trace.bin C3 0.25                          # anywhere

Reconstructs the critical chain from the measured pop times (tools/flow_trace/realpred.c) and splits every step of it.
Experiment tooling: uses oracle/ for the radius and the sample set; never part of the product path."""
import ctypes
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
import importlib
meshgen = importlib.import_module("gravomg_b200.meshgen")
from oracle.pyoracle import Oracle, available, build

path, name, scale = sys.argv[1], sys.argv[2], float(sys.argv[3])
so = os.path.join("/tmp", "flow_realpred.so")
subprocess.check_call(["gcc", "-O2", "-shared", "-fPIC", "-o", so, os.path.join(HERE, "realpred.c")])
if not available("orc"):
    build()
orc = Oracle("orc")
cfg = meshgen.make_config(name, scale)
V, A = cfg["V"], cfg["A"]
val = orc.to_edge_distance_matrix(A.indptr, A.indices, V)
edges, _ = orc.extract_edges(A.indptr, A.indices, val)
radius = 2.0 * orc.average_edge_length(V, edges)
n = V.shape[0]
sel = orc.fast_disc_sample(V, A.indptr, A.indices, radius)
inn = np.zeros(n, np.uint8); inn[sel] = 1
tr = np.fromfile(path, dtype=np.uint32).reshape(-1, 4).astype(np.int64)[:n]
push, pop, xpop, xdone = [np.ascontiguousarray(tr[:, i]) for i in range(4)]
lib = ctypes.CDLL(so); P = ctypes.c_void_p
pred = np.full(n, -1, np.int32)
lib.realpred(ctypes.c_int64(n), P(A.indptr.ctypes.data), P(A.indices.ctypes.data), P(val.ctypes.data), ctypes.c_double(radius),
             P(pop.ctypes.data), P(inn.ctypes.data), P(pred.ctypes.data))
v = int(np.argmax(pop)); chain = []
while v >= 0:
    chain.append(v); v = int(pred[v])
chain = chain[::-1]
rows = []
for u, v in zip(chain[:-1], chain[1:]):
    x = (u >> 12) != (v >> 12)
    rows.append((u, v, x, push[v] - pop[u], pop[v] - push[v], xpop[u] - pop[u] if x else 0, xdone[u] - xpop[u] if x else 0,
                 push[v] - xdone[u] if x else 0, pop[v] - pop[u]))
R = np.array(rows); S = R[R[:, 2] == 0]; X = R[R[:, 2] == 1]
pc = lambda a: np.percentile(a, [10, 50, 90, 99]).round(0)
print(f"critical chain: {len(R)} steps, {len(X)} cross a tile boundary; {1e-6 * (pop[chain[-1]] - pop[chain[0]]):.2f} ms")
print(f"same-tile steps: {len(S)}; mean pop->push {S[:, 3].mean():.0f} ns, push->pop (queue wait) {S[:, 4].mean():.0f} ns, "
      f"total {S[:, 8].mean():.0f} ns; sum {1e-6 * S[:, 8].sum():.2f} ms")
print("   total", pc(S[:, 8]), " pop->push", pc(S[:, 3]), " wait", pc(S[:, 4]))
print(f"cross-tile steps: {len(X)}; mean pop->xpop {X[:, 5].mean():.0f}, xpop->xdone {X[:, 6].mean():.0f}, xdone->push {X[:, 7].mean():.0f}, "
      f"push->pop {X[:, 4].mean():.0f}, total {X[:, 8].mean():.0f} ns; sum {1e-6 * X[:, 8].sum():.2f} ms")
for nm, c in (("pop->xpop", 5), ("xpop->xdone", 6), ("xdone->push", 7), ("push->pop", 4)):
    print("  ", nm, pc(X[:, c]))
# what a rank of the chain waits for between being queued and being popped: how many ranks were popped in between
order = np.argsort(pop, kind="stable")
rank_of = np.empty(n, np.int64); rank_of[order] = np.arange(n)
between = []
for v in chain[1:2000]:
    lo, hi = np.searchsorted(pop[order], [push[v], pop[v]])
    t = v >> 12
    between.append(int(np.count_nonzero((order[lo:hi] >> 12) == t)))
print("ranks of the SAME tile popped between a chain rank's push and its pop:", pc(np.array(between)))
for r in R[len(R) // 2:len(R) // 2 + 16]:
    print(r)
