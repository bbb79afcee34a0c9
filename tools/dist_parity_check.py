#!/usr/bin/env python
"""Multi-process check of the row-sharded build over NCCL (run under torchrun on a multi-GPU box):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        tools/dist_parity_check.py [C3 0.05]

Every rank builds the sharded hierarchy (its own range, NCCL all-gathers) AND the single-GPU hierarchy of
the same mesh, and compares: replicated arrays whole, P on its owned rows. Bit-exact or it fails.
Then the same with Galerkin operators (A_c pattern exact, values 1e-12) and a row-sharded V-cycle with
NCCL all-to-all halo exchanges against the single-GPU V-cycle (iterates to 1e-9)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist
from gravomg_b200 import capi, meshgen
from gravomg_b200.sharded import NativeComm, ShardedHierarchy, ShardedSolver, TorchComm, upload_sharded

name = sys.argv[1] if len(sys.argv) > 1 else "C3"
scale = float(sys.argv[2]) if len(sys.argv) > 2 else 0.05
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
stream = torch.cuda.Stream()
torch.cuda.set_stream(stream)                 # a real stream: the library never runs on the legacy stream 0
ctx = capi.Context(local, stream.cuda_stream)
cfg = meshgen.make_config(name, scale)
V, A = cfg["V"], cfg["A"]
inp = capi.Input(ctx, V, A.indptr, A.indices, None)
bad = 0
native = NativeComm(ctx)          # the library's own NCCL communicator: the C++ driver (gmg_shard_build_nccl)
for mode, nat in [(m, d) for m in ((capi.SAMPLER_HASHED, capi.SAMPLER_EXACT) if name != "C3" else (capi.SAMPLER_HASHED,))
                  for d in (None, native)]:
    single = capi.Hierarchy(ctx, inp, cfg["levels"], sampler_mode=mode, seed=3)
    sh = ShardedHierarchy(ctx, inp, cfg["levels"], world, [rank], comm=TorchComm(), sampler_mode=mode, seed=3, native=nat)
    assert sh.same_stream
    assert sh.num_levels == single.num_levels
    for lvl in range(single.num_levels):
        w = single.level_arrays(lvl)
        for key, which in (("sel", capi.H_SAMPLES), ("par", capi.H_PARENTS), ("ccp", capi.H_COARSE_COLPTR),
                           ("cri", capi.H_COARSE_ROWIDX), ("cval", capi.H_COARSE_VAL), ("cpts", capi.H_COARSE_POINTS)):
            if not np.array_equal(sh.fetch(lvl, which), w[key]):
                bad += 1
                print(f"rank {rank} mode {mode} level {lvl}: {key} differs", flush=True)
        i = sh.info(lvl)
        lo, hi = i["row_lo"], i["row_hi"]
        rp = sh.fetch(lvl, capi.H_P_ROWPTR)
        b, e = w["prp"][lo], w["prp"][hi]
        if not (np.array_equal(rp, w["prp"][lo:hi + 1] - b) and np.array_equal(sh.fetch(lvl, capi.H_P_COLIDX), w["pci"][b:e])
                and np.array_equal(sh.fetch(lvl, capi.H_P_VAL), w["pva"][b:e])):
            bad += 1
            print(f"rank {rank} mode {mode} level {lvl}: P rows [{lo},{hi}) differ", flush=True)
    print(f"rank {rank}/{world} mode {mode} driver {'c++/nccl' if nat else 'python/torch'}: {single.num_levels} levels, n={V.shape[0]}, exchanges {sh.exchanges}, "
          f"sharded {sh.build_ms:.2f} ms vs single {single.build_ms:.2f} ms, mismatches so far {bad}", flush=True)
    sh.close(); single.close()
# Galerkin operators and the V-cycle, on an input uploaded slice by slice and completed over NVLink
inp_a, h2d = upload_sharded(ctx, TorchComm(), V, A.indptr, A.indices, A.data)
pts_d, cp_d, ri_d = inp_a.fetch_arrays()
if not (np.array_equal(pts_d, V) and np.array_equal(cp_d, A.indptr) and np.array_equal(ri_d, A.indices)
        and np.array_equal(inp_a.fetch_operator(), A.data)):
    bad += 1
    print(f"rank {rank}: sharded upload differs from the host arrays", flush=True)
print(f"rank {rank}: sharded upload copied {h2d / 1e6:.1f} MB of {(V.nbytes + A.indptr.nbytes + A.indices.nbytes + A.data.nbytes) / 1e6:.1f} MB from the host", flush=True)
single = capi.Hierarchy(ctx, inp_a, cfg["levels"], sampler_mode=capi.SAMPLER_HASHED, seed=3, build_galerkin=True)
sh = ShardedHierarchy(ctx, inp_a, cfg["levels"], world, [rank], comm=TorchComm(), sampler_mode=capi.SAMPLER_HASHED, seed=3,
                      build_galerkin=True, native=native)
for lvl in range(single.num_levels):
    want, got = single.galerkin(lvl), sh.galerkin(lvl)
    if not (np.array_equal(got.indptr, want.indptr) and np.array_equal(got.indices, want.indices)
            and np.abs(got.data - want.data).max() <= 1e-12 * np.abs(want.data).max()):
        bad += 1
        print(f"rank {rank} level {lvl}: A_c differs", flush=True)
b = A @ (np.sin(3 * V[:, 0]) * np.cos(2 * V[:, 1]) + V[:, 2])
ref = capi.Solver(ctx, inp_a, single)
x_ref, res_ref = ref.vcycle(b, cycles=2, pre=2, post=2, coarse_sweeps=20)
single_ms = ctx.last_kernel_ms
sv = ShardedSolver(sh, inp_a)
x, res = sv.vcycle(b, cycles=2, pre=2, post=2, coarse_sweeps=20)
g = sv.geom[0][0]
own = slice(g["lo"], g["hi"])
if not (np.abs(res - res_ref).max() <= 1e-9 * res_ref[0] and np.abs(x[own] - x_ref[own]).max() <= 1e-9 * np.abs(x_ref).max()):
    bad += 1
    print(f"rank {rank}: V-cycle differs: {res} vs {res_ref}", flush=True)
print(f"rank {rank}/{world} galerkin+vcycle: build sharded {sh.build_ms:.2f} ms vs single {single.build_ms:.2f} ms; "
      f"V-cycle sharded {sv.last_ms / 2:.2f} ms ({sv.exchanges} halo exchanges, {sv.exchange_bytes / 1e6:.2f} MB sent by this rank) "
      f"vs single {single_ms / 2:.2f} ms; residuals {res[0]:.3e} -> {res[-1]:.3e}; mismatches so far {bad}", flush=True)
sv.close(); sh.close(); ref.close(); single.close()
t = torch.tensor([bad], device="cuda")
dist.all_reduce(t)
if rank == 0:
    print("DIST PARITY OK" if int(t[0]) == 0 else f"DIST PARITY FAILED ({int(t[0])} mismatches)", flush=True)
dist.destroy_process_group()
sys.exit(0 if int(t[0]) == 0 else 1)
