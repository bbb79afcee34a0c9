#!/usr/bin/env python
"""bench.py -- hierarchy-build throughput (Mverts/s) of the B200 path, with the HBM roofline of the
dominant stage, the end-to-end number through the C ABI with host buffers, and the reference's CPU
implementation timed beside it.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload C2|C3|C1] [--sampler exact|parallel]
    python bench.py --impl reference ...        # the reference's CPU code (oracle/_ref, else the C port)

A "step" is one full hierarchy build (all levels, all stages) over one synthetic mesh.
N = 1: config C2 (the configuration the metric is quoted on), single-GPU resident builder.
N > 1: one process per GPU (torchrun), ROW-SHARDED build (gravo-mg_b200/sharded.py) of ONE noisy torus
with 12.5 M vertices per GPU (N = 8 is the 100 M-vertex mesh of config C5), generated on the device;
exchanges are NCCL all-gathers over NVLink; weak scaling. `--workload T<millions>` builds a torus of that
many vertices at any N (e.g. T25 at N=1 and N=2 gives the strong-scaling pair).
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "hierarchy-build Mverts/s"
LEVELS = {"C1": 3, "C2": 4, "C3": 4, "C4": 4}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=None, help="C1|C2|C3|C4 (single GPU) or T<millions> (torus; sharded when N>1)")
    ap.add_argument("--per-gpu-mverts", type=float, default=12.5, help="default N>1 workload: torus vertices per GPU (millions)")
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--sampler", default=None, choices=["exact", "parallel", "hashed"],
                    help="default: exact on one GPU (the reference's sample set, the parity mode), hashed for the row-sharded "
                         "build (the exact greedy order is a cross-rank dependency chain; SURVEY.md 8(e))")
    ap.add_argument("--galerkin", type=int, default=None, help="1: also A_c = P^T A P on every level (default: 1 for C3 and the torus family)")
    ap.add_argument("--vcycle", type=int, default=None, help="1 (with --galerkin 1): one V-cycle Poisson solve and an SpMV, timed")
    ap.add_argument("--reorder", default="none", choices=["none", "morton"],
                    help="C4: renumber the cloud along the Morton curve on the device (gmg_morton_order) before the graph is built")
    ap.add_argument("--driver", default="native", choices=["native", "python"],
                    help="N > 1: the build's phase loop in C++ with the library's own NCCL communicator (gmg_shard_build_nccl), "
                         "or driven from Python through torch.distributed")
    ap.add_argument("--sharded", action="store_true", help="N = 1: run the workload through the row-sharded driver (world 1), "
                                                            "the 1-GPU point of an N > 1 curve (C4, torus family)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-scale", type=float, default=None, help="size of the in-line CPU baseline sample (default: the workload itself, which also gives the parity check)")
    a = ap.parse_args()
    world = max(a.gpus, int(os.environ.get("WORLD_SIZE", "1")))
    if a.workload is None:
        # N = 1: config C3, the largest single-GPU configuration of BASELINE.json and the north star's target
        # (10 M-vertex torus, P and A_c at every level, one V-cycle). N > 1: the same family row-sharded, 12.5 M
        # vertices per GPU (N = 8 is the 100 M-vertex mesh of config C5).
        a.workload = "C3" if world == 1 else f"T{a.per_gpu_mverts * world:g}"
    if world > 1 and a.workload not in ("C4",) and not a.workload.startswith("T"):
        raise SystemExit("bench.py: with N > 1 the workload is the torus family (T<millions>, the default) or C4; "
                         "C1-C3 are single-GPU configurations")
    if (world > 1 or a.sharded) and a.workload == "C4":
        a.reorder = "morton"        # the row-sharded build always renumbers a cloud (see run_b200_sharded)
    torus = a.workload == "C3" or a.workload.startswith("T")
    if a.galerkin is None:
        a.galerkin = 1 if torus else 0
    if a.vcycle is None:
        a.vcycle = 1 if (torus and a.galerkin) else 0
    if a.sampler is None:
        a.sampler = "hashed" if (world > 1 or a.sharded or a.workload.startswith("T")) else "exact"
    return a


def torus_dims(workload):
    """T<millions> -> (nu, nv) of a square noisy torus grid with about that many vertices."""
    total = float(workload[1:]) * 1e6
    side = max(8, int(round(np.sqrt(total))))
    return side, side


def workload_name(a):
    if a.workload.startswith("T"):
        nu, nv = torus_dims(a.workload)
        return f"noisy torus {nu}x{nv} ({nu * nv:,} verts, C5 family), 4 levels, ratio 8, generated on the device"
    base = {"C1": "C1: icosphere 6 subdivisions (40,962 verts), 3 levels",
            "C2": "C2: icosphere 9 subdivisions + vertex noise (2,621,442 verts), 4 levels, ratio 8",
            "C3": "C3: anisotropic noisy torus 10000x1000 (10,000,000 verts), 4 levels, ratio 8",
            "C4": "C4: kNN k=8 sphere cloud 20M points, 4 levels"}[a.workload]
    if a.galerkin:
        base += ", P + Galerkin A_c at every level"
    return base if a.scale == 1.0 else f"{base} [scaled x{a.scale}]"


def bench_config(a, world):
    """The `config` object; identical in the B200 arm and the reference arm."""
    cfg = {"workload": workload_name(a), "sampler": a.sampler, "galerkin": bool(a.galerkin), "vcycle": bool(a.vcycle),
           "n_gpus": world}
    if getattr(a, "reorder", "none") != "none":
        cfg["reorder"] = a.reorder
    return cfg


# ---- clocks sampler (B200_PROFILING.md recipe) ---------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc, self.nvml = index, [], None, None

    def start(self):
        # NVML in a thread when the binding is importable (no process spawn, no driver-wide stall while the
        # timed region runs); otherwise the nvidia-smi loop of the profiling recipe.
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nvml, self.handle = pynvml, pynvml.nvmlDeviceGetHandleByIndex(self.index)
            self.stop_flag = False
            self.thread = threading.Thread(target=self._poll_nvml, daemon=True)
            self.thread.start()
            return
        except Exception:
            self.nvml = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _poll_nvml(self):
        n = self.nvml
        bits = {"hw_slowdown": n.nvmlClocksEventReasonHwSlowdown if hasattr(n, "nvmlClocksEventReasonHwSlowdown") else 0x8,
                "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}
        while not self.stop_flag:
            try:
                sm = n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM)
                mx = n.nvmlDeviceGetMaxClockInfo(self.handle, n.NVML_CLOCK_SM)
                try:
                    reasons = n.nvmlDeviceGetCurrentClocksEventReasons(self.handle)
                except Exception:
                    reasons = n.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle)
                self.rows.append(["", str(sm), str(mx), "", ""] + ["Active" if reasons & b else "Not Active"
                                                                       for b in bits.values()])
            except Exception:
                pass
            time.sleep(0.05)

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if getattr(self, "nvml", None) is not None:
            self.stop_flag = True
            self.thread.join(timeout=1.0)
        elif self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        else:
            self.proc.terminate()
        sm = [float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) < 9:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ---- algorithmic bytes per stage (int32 indices, fp64 values, single touch) -------
def stage_bytes(i):
    n, nnz, m, nnz_c, T, nnzP, nnzAc = i["n"], i["nnz"], i["m"], i["nnz_c"], i["T"], i["nnz_p"], i["nnz_ac"]
    return {
        "edge_lengths": 28 * n + 12 * nnz,
        "sampling": 32 * n + 4 * nnz + 4 * m,
        "parents": 40 * n + 4 * nnz + 4 * m,
        "cluster_csr": 8 * n + 4 * m,
        "coarse_edges": 8 * n + 12 * nnz + 4 * m + 12 * nnz_c,
        "coarse_points": 28 * n + 28 * m,
        "triangle_fans": 32 * m + 12 * nnz_c + 48 * T,
        "prolongation_rows": 32 * n + 12 * nnzP + 32 * m + 4 * nnz_c + 48 * T,
        "galerkin": (12 * n + 4 * nnz + 8 * nnzP + 4 * m) + (12 * n + 12 * nnz + 24 * nnzP + 4 * m + 12 * nnzAc),
    }


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


# kernels of each stage (for roofline.traffic: DRAM bytes per launch from the committed ncu --set full capture)
STAGE_KERNELS = {"edge_lengths": ["edge_length_kernel"], "sampling": ["sampler_coherence_kernel<0>", "sampler_build_kernel<0>", "sampler_build_raw_kernel<0>", "sampler_flow_kernel<0, 0>"],
                 "parents": ["parents_sweep_kernel", "parents_pred_kernel"], "coarse_edges": ["coarse_rows_warp_kernel"],
                 "prolongation_rows": ["prolong_rows_kernel"], "galerkin_product2": ["product2_kernel", "compact_columns_kernel"],
                 "galerkin_product1": ["product1_kernel"],
                 "galerkin_transpose": ["count_columns_kernel", "transpose_fill_kernel", "sort_rows_kernel"]}


def ncu_traffic(workload, stage):
    """DRAM read+write bytes of the stage's main kernels at level 0, from profiles/traffic_<workload>_level0.json
    (written by tools/ncu_summary.py from an `ncu --set full` capture of this same command); None if absent."""
    p = os.path.join(ROOT, "profiles", f"traffic_{workload.lower()}_level0.json")
    if not os.path.exists(p) or stage not in STAGE_KERNELS:
        return None, None
    with open(p) as f:
        js = json.load(f)
    got = [js["kernels"][k]["dram_bytes"] for k in STAGE_KERNELS[stage] if k in js["kernels"]]
    return (int(sum(got)), js["capture"]) if got else (None, None)


# ---- CPU arm: the reference's own code on the host cores ---------------------------------------------
CPU_STAGES = ["edge_lengths", "sampling", "parents", "coarse_edges", "coarse_points", "triangle_fans", "prolongation_rows",
              "galerkin"]


def cpu_hierarchy(V, A, levels, galerkin=False, radii=None, kind=None):
    """One hierarchy build with the reference's CPU implementation (the UNMODIFIED sources on the Eigen shim when
    oracle/_ref exists, else the plain-C port), one thread: the reference has no threading. The call order is the
    demo's (test/main.cpp:68-121) looped over levels; A_c = P^T A P, which the reference does not contain, is our
    single-thread CSR product (oracle/solver_oracle.c). Returns (per-level results, per-stage seconds, total
    seconds, kind). radii: use these sampling radii (the GPU's, so that the two hierarchies can be compared bit
    for bit; averageEdgeLength is still computed and timed)."""
    from oracle.pyoracle import Oracle, available
    kind = kind or ("ref" if available("ref") else "orc")
    be = Oracle(kind)
    sec = {k: 0.0 for k in CPU_STAGES}

    def timed(stage, fn, *args):
        t0 = time.perf_counter()
        r = fn(*args)
        sec[stage] += time.perf_counter() - t0
        return r

    t_all = time.perf_counter()
    pts, cp, ri = V, A.indptr, A.indices
    val = timed("edge_lengths", be.to_edge_distance_matrix, cp, ri, pts)
    a_cur = (A.indptr, A.indices, A.data)
    res = []
    for lvl in range(levels - 1):
        r = {}
        edges, _ = timed("edge_lengths", be.extract_edges, cp, ri, val)
        r["radius_cpu"] = np.cbrt(8.0) * timed("edge_lengths", be.average_edge_length, pts, edges)
        del edges
        r["radius"] = r["radius_cpu"] if radii is None else radii[lvl]
        r["sel"] = timed("sampling", be.fast_disc_sample, pts, cp, ri, r["radius"])
        m = len(r["sel"])
        r["par"] = timed("parents", be.assign_parents, pts, cp, ri, r["sel"])
        r["ccp"], r["cri"], r["cval"] = timed("coarse_edges", be.extract_coarse_edges, cp, ri, val, m, r["par"])
        r["cpts"] = timed("coarse_points", be.coarse_from_mean_of_fine_children, pts, cp, ri, r["par"], m)
        # constructProlongation calls constructVoronoiTriangles itself (src/multigrid.cpp:272); the demo's own extra
        # call (test/main.cpp:107) only feeds the viewer and is timed as the triangle stage
        r["tri"], r["nrm"], r["fptr"], r["fan"] = timed("triangle_fans", be.construct_voronoi_triangles, r["cpts"],
                                                        r["ccp"], r["cri"], r["cval"])
        pr = timed("prolongation_rows", be.construct_prolongation, pts, r["cpts"], r["ccp"], r["cri"], r["cval"], r["par"], 0)
        r["prp"], r["pci"], r["pva"] = pr[0], pr[1], pr[2]
        if galerkin:
            r["ac"] = timed("galerkin", be.galerkin, *a_cur, r["prp"], r["pci"], r["pva"], m)
            a_cur = r["ac"]
        res.append(r)
        pts, cp, ri, val = r["cpts"], r["ccp"], r["cri"], r["cval"]
    return res, sec, time.perf_counter() - t_all, ("reference" if kind == "ref" else "port")


def compare_with_cpu(h, capi, cpu_levels, galerkin):
    """Bit-exact comparison of a GPU hierarchy (exact sampler) with the CPU run on the same input and radii."""
    keys = "sel par ccp cri cval cpts tri nrm fptr fan prp pci pva".split()
    out = {"levels": len(cpu_levels), "arrays": 0, "mismatches": []}
    ac_err = 0.0
    for lvl, w in enumerate(cpu_levels):
        g = h.level_arrays(lvl)
        for k in keys:
            x, y = np.asarray(g[k]), np.asarray(w[k])
            out["arrays"] += 1
            if x.shape != y.shape or not np.array_equal(x, y):
                out["mismatches"].append(f"L{lvl}.{k}")
        rel = abs(g["radius"] - w["radius_cpu"]) / w["radius_cpu"]
        if rel > 1e-12:
            out["mismatches"].append(f"L{lvl}.radius({rel:.1e})")
        if galerkin:
            gp, gi, gv = h.fetch(lvl, capi.H_AC_COLPTR), h.fetch(lvl, capi.H_AC_ROWIDX), h.fetch(lvl, capi.H_AC_VAL)
            wp, wi, wv = w["ac"]
            out["arrays"] += 3
            if not (np.array_equal(gp, wp) and np.array_equal(gi, wi)):
                out["mismatches"].append(f"L{lvl}.A_c pattern")
            else:
                ac_err = max(ac_err, float(np.abs(gv - wv).max(initial=0.0) / np.abs(wv).max()))
                if ac_err > 1e-12:
                    out["mismatches"].append(f"L{lvl}.A_c values({ac_err:.1e})")
                out.setdefault("ac_bit_identical", True)
                out["ac_bit_identical"] = out["ac_bit_identical"] and bool(np.array_equal(gv, wv))
    out["ac_max_rel_err"] = ac_err
    return out


def cpu_solver_seconds(A, levels_cpu):
    """Single-thread CSR SpMV and one Gauss-Seidel sweep on the level-0 operator (oracle/solver_oracle.c), timed."""
    from oracle.pyoracle import Oracle
    be = Oracle("orc")
    n = A.shape[0]
    x = np.linspace(0.0, 1.0, n)
    t0 = time.perf_counter(); be.spmv(A.indptr, A.indices, A.data, x); t_spmv = time.perf_counter() - t0
    rows = np.arange(n, dtype=np.int32)
    t0 = time.perf_counter(); be.gauss_seidel(A.indptr, A.indices, A.data, x, np.zeros(n), rows, 1); t_gs = time.perf_counter() - t0
    return {"spmv_s": t_spmv, "spmv_gbs": (20 * n + 12 * A.nnz) / t_spmv / 1e9, "gauss_seidel_sweep_s": t_gs,
            "what": "single-thread CSR kernels of ours (the reference has none), level-0 operator"}


def run_reference(a, rank, world):
    """The reference arm: same config, metric and unit as the B200 arm; every step is one whole hierarchy build by
    the reference's CPU code on a BOUNDED sample of the workload (the full C3 build takes about 100 s per step on one
    core; the reference is single-threaded), sized so that steps + warmup fit a few minutes."""
    if rank != 0:
        return
    from gravomg_b200 import meshgen
    config = bench_config(a, world)
    if a.workload.startswith("T"):      # the torus family: the CPU sample is the C3 generator at a bounded size
        nu, nv = torus_dims(a.workload)
        full_n, gen = nu * nv, "C3"
    else:
        full_n, gen = {"C1": 40962, "C2": 2621442, "C3": 10_000_000, "C4": 20_000_000}[a.workload] * a.scale, a.workload
    per_step_s = max(2.0, 240.0 / max(1, a.steps + a.warmup))
    rate = 0.10e6 if a.galerkin else 0.2e6          # verts per second of the CPU build, conservative
    target_n = min(full_n, per_step_s * rate)
    base_n = {"C1": 40962, "C2": 2621442, "C3": 10_000_000, "C4": 20_000_000}[gen]
    scale = min(1.0, target_n / base_n)
    if gen in ("C1", "C2"):             # icospheres come in subdivision steps: round the scale to a power of 4
        scale = min(1.0, 4.0 ** np.floor(np.log(max(scale, 1e-6)) / np.log(4.0)))
    cfg = meshgen.make_config(gen, scale)
    V, A = cfg["V"], cfg["A"]
    n = V.shape[0]
    times, kind, sec_acc = [], "port", None
    for s in range(a.warmup + a.steps):
        _, sec, dt, kind = cpu_hierarchy(V, A, LEVELS[gen], galerkin=bool(a.galerkin))
        if s >= a.warmup:
            times.append(dt)
            sec_acc = sec if sec_acc is None else {k: sec_acc[k] + v for k, v in sec.items()}
    ms = 1e3 * float(np.mean(times))
    value = n / (ms * 1e-3) / 1e6
    sample = (f"{gen} generator at scale {scale:g}: {n} verts, {LEVELS[gen]} levels, whole hierarchy"
              f"{' + A_c every level' if a.galerkin else ''} per step; reference sources on the Eigen shim, -O2, 1 thread")
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": "Mverts/s", "n_gpus": a.gpus,
            "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config,
            "cpu_baseline": {"value": value, "unit": "Mverts/s", "cores": 1, "kind": kind, "sample": sample,
                             "host_cores": os.cpu_count(), "sample_verts": n,
                             "stage_seconds": {k: round(v / max(1, a.steps), 4) for k, v in (sec_acc or {}).items()}},
            "e2e": {"value": value, "unit": "Mverts/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# ---- B200 arm ------------------------------------------------------------------------------------------
def run_b200(a, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from gravomg_b200 import meshgen
    from gravomg_b200 import capi
    from gravomg_b200.capi import Context, Input, Hierarchy

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)         # a real stream shared by the library and the timing events
    ctx = Context(local_rank, stream.cuda_stream)
    builder = None
    if a.workload == "C4":
        # the kNN graph is built on the device (gmg_input_from_knn; a host kd-tree needs minutes at 2e7 points) and
        # fetched back once, so that the timed steps, the end-to-end arm and the CPU arm all start from host arrays
        import scipy.sparse as sp
        V = meshgen.sphere_cloud(max(64, int(20_000_000 * a.scale)))
        morton_ms = None
        if a.reorder == "morton":
            t0 = time.perf_counter()
            perm = ctx.morton_order(V)
            morton_ms = 1e3 * (time.perf_counter() - t0)
            V = np.asfortranarray(V[perm])
        t0 = time.perf_counter()
        knn = Input.knn(ctx, V, 8)
        torch.cuda.synchronize()
        knn_ms = 1e3 * (time.perf_counter() - t0)
        _, cp0, ri0 = knn.fetch_arrays()
        knn.close()
        A = sp.csc_matrix((np.ones(ri0.shape[0]), ri0, cp0), shape=(V.shape[0], V.shape[0]))
        cfg = dict(name="C4", V=V, A=A, levels=LEVELS["C4"])
        builder = {"what": "exact kNN (k=8) by grid hashing + union-symmetrisation on the device, host points in, "
                           "wall clock of the call including the H2D copy of the points", "ms": knn_ms,
                   "points": int(V.shape[0]), "nnz": int(A.nnz), "Mpoints_per_s": V.shape[0] / (knn_ms * 1e-3) / 1e6,
                   "reorder": a.reorder, "morton_order_ms": morton_ms}
    else:
        cfg = meshgen.make_config(a.workload, a.scale)
    V, A = cfg["V"], cfg["A"]
    levels = LEVELS[a.workload]
    n = V.shape[0]
    MODES = {"exact": capi.SAMPLER_EXACT, "parallel": capi.SAMPLER_PARALLEL, "hashed": capi.SAMPLER_HASHED}
    mode = MODES[a.sampler]

    # pinned host copies of the inputs for the end-to-end arm
    def pinned(arr):
        t = torch.from_numpy(np.ascontiguousarray(arr)).pin_memory()
        return t, t.numpy()
    tV, hV = pinned(np.asfortranarray(V).T)             # (3, n) C-order == (n, 3) column-major
    hV = hV.T
    tcp, hcp = pinned(A.indptr.astype(np.int32))
    tri, hri = pinned(A.indices.astype(np.int32))
    tav, hav = pinned(A.data.astype(np.float64))

    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")   # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    inp = Input(ctx, hV, hcp, hri, hav if a.galerkin else None)

    def step_resident(m=mode):
        return Hierarchy(ctx, inp, levels, sampler_mode=m, build_galerkin=bool(a.galerkin))

    # pinned result buffers, allocated once and reused like a caller's own output arrays (rows of P hold <= 3 entries)
    pin_out = {}

    def pinned_out(key, count, dtype):
        if key not in pin_out or pin_out[key].shape[0] < count:
            tdt = {np.int32: torch.int32, np.float64: torch.float64}[dtype]
            pin_out[key] = torch.empty(max(count, 1), dtype=tdt).pin_memory().numpy()
        return pin_out[key]

    def step_e2e():
        # A's values are needed at the first Galerkin product only: they travel on the copy stream while the build runs
        i2 = Input(ctx, hV, hcp, hri, hav if a.galerkin else None, defer_values=True)
        # every level's P (and A_c) leaves the device on the copy stream as soon as the level is complete, while the
        # next level builds; the step ends when the last copy has landed in the pinned result buffers
        def stream_out(lvl, info):
            nl, nzp, m, nzac = info["n"], info["nnz_p"], info["m"], info["nnz_ac"]
            want = []
            # with A_c the library calls twice per level: P as soon as it is complete (it travels while A_c is formed),
            # A_c when the level is done
            if info["p_ready_only"] or not a.galerkin:
                want += [(capi.H_P_ROWPTR, nl + 1, np.int32), (capi.H_P_COLIDX, nzp, np.int32), (capi.H_P_VAL, nzp, np.float64)]
            if a.galerkin and not info["p_ready_only"]:
                want += [(capi.H_AC_COLPTR, m + 1, np.int32), (capi.H_AC_ROWIDX, nzac, np.int32), (capi.H_AC_VAL, nzac, np.float64)]
            return [(which, pinned_out((lvl, which), cnt, dt)[:cnt]) for which, cnt, dt in want]
        h = Hierarchy(ctx, i2, levels, sampler_mode=mode, build_galerkin=bool(a.galerkin), stream_out=stream_out)
        ctx.copy_sync()
        d2h = h.streamed_bytes
        nb = i2.h2d_bytes
        h.close(); i2.close()
        return nb, d2h

    def timed_builds(m, steps, warmup):
        for _ in range(warmup):
            step_resident(m).close()
        barrier()
        l0 = ctx.launch_count
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        infos, acc, gacc = None, None, None
        for s_ in range(steps):
            flush.fill_(s_ & 0xFF)                        # flush L2 between timed iterations
            torch.cuda.synchronize()
            ev[s_][0].record(stream)
            h = step_resident(m)
            ev[s_][1].record(stream)
            torch.cuda.synchronize()
            infos = h.info
            sm = np.array([i["stage_ms"] for i in h.info])
            gm = np.array([i["galerkin_ms"] for i in h.info])
            acc = sm if acc is None else acc + sm
            gacc = gm if gacc is None else gacc + gm
            h.close()
        barrier()
        ms_ = float(np.mean([e0.elapsed_time(e1) for e0, e1 in ev]))
        return ms_, infos, acc / steps, gacc / steps, ctx.launch_count - l0

    clocks = ClockSampler(local_rank)
    clocks.start()
    ms, infos, stage_ms, gal_ms, launches = timed_builds(mode, a.steps, a.warmup)
    clk = clocks.stop()
    # the other sampler order beside it (fewer steps): exact is the parity mode, parallel the throughput mode
    other = None
    other_name = "parallel" if a.sampler == "exact" else "exact"
    if a.scale == 1.0 or a.workload != "C4":
        o_ms, o_infos, _, _, _ = timed_builds(MODES[other_name], max(1, min(a.steps, 5)), 2)
        other = {"sampler": other_name, "ms_per_step": o_ms, "value": n / (o_ms * 1e-3) / 1e6, "unit": "Mverts/s",
                 "coarse_sizes": [i["m"] for i in o_infos]}

    # end-to-end through the C ABI with host buffers: H2D of the inputs + build + D2H of every P (and A_c)
    e2e_ms, h2d, d2h = [], 0, 0
    step_e2e()
    barrier()
    for s_ in range(max(1, min(a.steps, 5))):
        flush.fill_(s_)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        h2d, d2h = step_e2e()
        torch.cuda.synchronize()
        e2e_ms.append(1e3 * (time.perf_counter() - t0))
    e2e = float(np.mean(e2e_ms))

    if world > 1:
        t = torch.tensor([ms, e2e], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, e2e = float(t[0]), float(t[1])
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peak, peak_src = measured_peak()
    names = capi.STAGE_NAMES
    # the dominant kernel group at level 0 (the level that holds ~7/8 of the work); the Galerkin stage is split
    # into its transpose and its two products (T = P^T A, A_c = T P), each with the bytes it cannot avoid: every
    # input array read once, the result written once (T itself is an intermediate and is not counted)
    i0 = infos[0]
    sb = stage_bytes(i0)
    cand = {names[s_]: (float(stage_ms[0][s_]), sb[names[s_]]) for s_ in range(8)}
    if a.galerkin:
        n0, nnz0, m0, nzp, nzac = i0["n"], i0["nnz"], i0["m"], i0["nnz_p"], i0["nnz_ac"]
        cand["galerkin_transpose"] = (float(gal_ms[0][0]), 12 * nzp + 4 * m0 + 12 * nzp)
        cand["galerkin_product1"] = (float(gal_ms[0][1]), 8 * n0 + 12 * nnz0 + 12 * nzp)
        cand["galerkin_product2"] = (float(gal_ms[0][2]), 8 * m0 + 12 * nzp + 12 * nzac)
    dom = max(cand, key=lambda k: cand[k][0])
    dom_ms, bytes0 = cand[dom]
    achieved = bytes0 / (dom_ms * 1e-3) / 1e9
    traffic, traffic_src = ncu_traffic(a.workload, dom) if a.scale == 1.0 else (None, None)
    total_bytes = sum(sum(v for k, v in stage_bytes(i).items() if (k != "galerkin" or a.galerkin)) for i in infos)
    config = bench_config(a, world)
    config.update({"l2": "256 MB flush write between timed steps; inputs (>200 MB) exceed the 126 MB L2",
                   "multi_gpu": "replicas only (one independent build per rank)" if world > 1 else "single GPU",
                   "levels": [{k: i[k] for k in ("n", "nnz", "m", "nnz_c", "T", "nnz_p", "nnz_ac", "sampler_rounds",
                                                  "parent_rounds")} for i in infos]})
    line = {"metric": METRIC, "value": world * n / (ms * 1e-3) / 1e6, "unit": "Mverts/s", "n_gpus": world,
            "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config,
            "value_is": f"{a.sampler} sampler" + (" (the reference's own sample sets; the parity mode)" if a.sampler == "exact" else ""),
            "other_sampler": other,
            "clocks": clk, "gpu_launches": int(launches),
            "e2e": {"value": world * n / (e2e * 1e-3) / 1e6, "unit": "Mverts/s", "ms_per_step": e2e,
                    "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h)},
            "roofline": {"bound": "hbm",
                         "kernel": f"level-0 '{dom}' ({' + '.join(STAGE_KERNELS.get(dom, [dom]))})",
                         "achieved": achieved, "peak": peak, "peak_source": peak_src, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                         "algorithmic_bytes": int(bytes0), "stage_ms": dom_ms,
                         "per_stage_frac": {k: round(v[1] / (v[0] * 1e-3) / 1e9 / peak, 4) for k, v in cand.items() if v[0] > 0},
                         "whole_build": {"algorithmic_bytes": int(total_bytes),
                                         "achieved": total_bytes / (ms * 1e-3) / 1e9,
                                         "frac": total_bytes / (ms * 1e-3) / 1e9 / peak}},
            "stage_ms": {f"L{l}": {names[s_]: round(float(stage_ms[l][s_]), 4) for s_ in range(9)}
                         for l in range(len(infos))}}
    if builder:
        line["input_builder"] = builder
    if a.galerkin:
        # SpGEMM (P^T A P) of level 0 against the HBM roofline, from the build's own stage events
        gb = sb["galerkin"]
        gms = float(stage_ms[0][names.index("galerkin")])
        line["spgemm"] = {"ms": gms, "algorithmic_bytes": int(gb), "achieved": gb / (gms * 1e-3) / 1e9, "unit": "GB/s",
                          "frac": gb / (gms * 1e-3) / 1e9 / peak, "nnz_ac": int(i0["nnz_ac"]),
                          "transpose_ms": float(gal_ms[0][0]), "product1_ms": float(gal_ms[0][1]), "product2_ms": float(gal_ms[0][2]),
                          "what": "level-0 A_c = P^T A P: transpose R = P^T, T = P^T A (sizing pass + values), A_c = T P (sizing pass + values + "
                                  "compaction); bytes: the symbolic + numeric formulas of SURVEY.md 8(d)"}
    h_exact = None
    if a.galerkin and a.vcycle:
        from gravomg_b200.capi import Solver, spmv
        h = Hierarchy(ctx, inp, levels, sampler_mode=mode, build_galerkin=True)
        xs = np.sin(3.0 * V[:, 0]) * np.cos(2.0 * V[:, 1]) + V[:, 2]          # smooth function of position
        Acsr = A.tocsr()
        bvec = Acsr @ xs
        for _ in range(3):
            spmv(ctx, A.indptr, A.indices, A.data, xs)
        sp_ms = []
        for _ in range(5):
            spmv(ctx, A.indptr, A.indices, A.data, xs)
            sp_ms.append(ctx.last_kernel_ms)
        sbytes = 20 * n + 12 * A.nnz
        sms = float(np.median(sp_ms))
        line["spmv"] = {"ms": sms, "algorithmic_bytes": int(sbytes), "achieved": sbytes / (sms * 1e-3) / 1e9, "unit": "GB/s",
                        "frac": sbytes / (sms * 1e-3) / 1e9 / peak, "what": "y = A x on the level-0 operator, device time"}
        sol = Solver(ctx, inp, h)
        sol.vcycle(bvec, cycles=1)                                             # warm-up (colourings are built here)
        x1, res = sol.vcycle(bvec, cycles=1, pre=2, post=2)
        line["vcycle"] = {"ms": ctx.last_kernel_ms, "residual_before": float(res[0]), "residual_after": float(res[1]),
                          "reduction": float(res[1] / res[0]), "what": "one V(2,2) cycle on A x = b, multicolour Gauss-Seidel, "
                          "b = A x*, x* smooth; device time of the cycle"}
        sol.close()
        if a.sampler == "exact":
            h_exact = h
        else:
            h.close()
    if not a.no_cpu_baseline:
        # the reference's CPU code on the host cores, on the SAME workload at the same size by default: that run is
        # then also the parity check (GPU hierarchy, exact sampler, against the CPU one: every array, bit for bit)
        cscale = a.scale if a.cpu_scale is None else a.cpu_scale
        same = cscale == a.scale
        c2 = cfg if same else meshgen.make_config(a.workload, cscale)
        radii = None
        if same:
            if h_exact is None:
                h_exact = Hierarchy(ctx, inp, levels, sampler_mode=capi.SAMPLER_EXACT, build_galerkin=bool(a.galerkin))
            radii = [i["radius"] for i in h_exact.info]
        cpu_levels, sec, dt, kind = cpu_hierarchy(c2["V"], c2["A"], levels, galerkin=bool(a.galerkin), radii=radii)
        nn = c2["V"].shape[0]
        line["cpu_baseline"] = {"value": nn / dt / 1e6, "unit": "Mverts/s", "cores": 1, "kind": kind,
                                "host_cores": os.cpu_count(), "seconds": dt,
                                "stage_seconds": {k: round(v, 3) for k, v in sec.items()},
                                "sample": f"{a.workload} at scale {cscale:g}: {nn} verts, whole {levels}-level hierarchy"
                                          f"{' + A_c every level (our single-thread CSR product)' if a.galerkin else ''}, "
                                          f"one run, reference sources on the Eigen shim, -O2, 1 thread"}
        if a.galerkin:
            line["cpu_baseline"]["solver"] = cpu_solver_seconds(c2["A"], cpu_levels)
        if same:
            try:
                par = compare_with_cpu(h_exact, capi, cpu_levels, bool(a.galerkin))
                line["parity_checked"] = len(par["mismatches"]) == 0
                line["parity"] = par
            except Exception as e:                                  # a failed comparison must not hide the numbers
                line["parity_checked"] = False
                line["parity"] = {"error": repr(e)}
        else:
            line["parity_checked"] = False
            line["parity"] = {"skipped": "the CPU sample is a different size (--cpu-scale)"}
    if h_exact is not None:
        h_exact.close()
    if a.workload == "C3" and a.scale == 1.0 and world == 1:
        # the 1-GPU point of the N = 2, 4, 8 curve: the SAME workload family and code path as `bench.py --gpus N`
        # (12.5 M-vertex torus per GPU through the row-sharded driver, hashed sampler, A_c on every level), so that
        # scaling efficiency can be computed on one family: value(N) / (N * scale_anchor.value)
        try:
            from gravomg_b200.sharded import ShardedHierarchy
            nu, nv = torus_dims(f"T{a.per_gpu_mverts:g}")
            tin = Input.torus(ctx, nu, nv, mass_shift=1e-3)
            t_ms = []
            for s_ in range(5):
                flush.fill_(s_)
                torch.cuda.synchronize()
                sh = ShardedHierarchy(ctx, tin, 4, 1, [0], comm=None, sampler_mode=capi.SAMPLER_HASHED, build_galerkin=True)
                if s_ >= 2:
                    t_ms.append(sh.build_ms)
                sh.close()
            tin.close()
            t_mean = float(np.mean(t_ms))
            line["scale_anchor"] = {"workload": f"noisy torus {nu}x{nv} ({nu * nv:,} verts, C5 family) through the row-sharded driver, "
                                                f"world 1, hashed sampler, A_c on every level", "ms_per_step": t_mean,
                                    "value": nu * nv / (t_mean * 1e-3) / 1e6, "unit": "Mverts/s", "steps": len(t_ms)}
        except Exception as e:                                      # never lose the main line over the side measurement
            line["scale_anchor"] = {"error": repr(e)}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


# ---- B200 arm, N > 1 (and any torus workload): row-sharded build --------------------------------------
def run_b200_sharded(a, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from gravomg_b200 import capi
    from gravomg_b200.capi import Context, Input
    from gravomg_b200.sharded import NativeComm, ShardedHierarchy, ShardedSolver, TorchComm, upload_sharded

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)         # a real stream shared by the library, the NCCL exchanges and the timing events
    ctx = Context(local_rank, stream.cuda_stream)
    comm = TorchComm() if world > 1 else None
    native = NativeComm(ctx) if world > 1 and a.driver == "native" else None
    levels = 4
    mode = {"exact": capi.SAMPLER_EXACT, "parallel": capi.SAMPLER_PARALLEL, "hashed": capi.SAMPLER_HASHED}[a.sampler]
    galerkin = bool(a.galerkin or a.vcycle)
    builder = None
    if a.workload == "C4":
        # config C4 row-sharded: every rank builds the same kNN graph on its own device (deterministic: exact neighbours,
        # ties by index) from the cloud renumbered along the Morton curve -- contiguous row ranges of an unordered cloud
        # would have the whole cloud as their halo. A graph without values: P only, no A_c, no V-cycle.
        from gravomg_b200 import meshgen
        V = meshgen.sphere_cloud(max(64, int(20_000_000 * a.scale)))
        t0 = time.perf_counter()
        perm = ctx.morton_order(V)
        t1 = time.perf_counter()
        V = np.asfortranarray(V[perm])          # (host gather, not timed)
        t2 = time.perf_counter()
        inp = Input.knn(ctx, V, 8)
        torch.cuda.synchronize()
        builder = {"what": "Morton order + exact kNN (k=8) graph on the device, on every rank", "morton_order_ms": 1e3 * (t1 - t0),
                   "knn_ms": 1e3 * (time.perf_counter() - t2), "points": int(inp.n), "nnz": int(inp.nnz)}
        n = inp.n
        del V
    else:
        nu, nv = torus_dims(a.workload)
        n = nu * nv
        # the whole mesh, resident on every rank (with its system matrix when A_c is asked for)
        inp = Input.torus(ctx, nu, nv, mass_shift=1e-3 if galerkin else None)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def device_points(which):
        """x, y, z of an Input as device tensors (via pinned host memory: the C ABI hands out host copies)."""
        ctx.lib.gmg_input_fetch.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
        host = torch.empty(3 * which.n, dtype=torch.float64).pin_memory()
        ctx._check(ctx.lib.gmg_input_fetch(ctx.h, which.h, 0, ctypes.c_void_p(host.data_ptr())))
        return host.cuda().view(3, which.n)

    def build(which, profile=False):
        return ShardedHierarchy(ctx, which, levels, world, [rank], comm=comm, sampler_mode=mode, profile=profile,
                                build_galerkin=galerkin, native=None if profile else native)

    for _ in range(a.warmup):
        build(inp).close()
    barrier()
    clocks = ClockSampler(local_rank)
    clocks.start()
    launches0 = ctx.launch_count
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(a.steps)]
    infos, exchanges, xbytes = None, 0, 0
    barrier()
    t_all0 = time.perf_counter()
    for s in range(a.steps):
        flush.fill_(s & 0xFF)
        barrier()
        ev[s][0].record(stream)
        h = build(inp)
        ev[s][1].record(stream)
        torch.cuda.synchronize()
        infos = [h.info(l) for l in range(h.num_levels)]
        exchanges, xbytes = h.exchanges, h.exchange_bytes
        h.close()
    barrier()
    launches = ctx.launch_count - launches0
    clk = clocks.stop()
    ms = float(np.mean([e0.elapsed_time(e1) for e0, e1 in ev]))
    # one extra, untimed build with per-phase events: where the step goes (rank 0's view)
    hp = build(inp, profile=True)
    phase_ms = {k: round(v, 3) for k, v in sorted(hp.phase_ms().items(), key=lambda kv: -kv[1])}
    vc = None
    if a.vcycle:
        # one V(2,2) cycle on A x = b with b = A x*, x* smooth: row-sharded smoothing with halo exchanges
        sv = ShardedSolver(hp, inp)
        px = device_points(inp)                                               # x* from the device-resident positions
        sv.rhs_from_solution(torch.sin(3 * px[0]) * torch.cos(2 * px[1]) + px[2])
        del px
        sv.vcycle(None, cycles=1, pre=2, post=2, coarse_sweeps=20, fetch=False)   # warm-up (allocator, NCCL channels)
        ex0, xb0 = sv.exchanges, sv.exchange_bytes
        _, res = sv.vcycle(None, cycles=1, pre=2, post=2, coarse_sweeps=20, fetch=False)
        vms = torch.tensor([sv.last_ms], device="cuda", dtype=torch.float64)
        xb = torch.tensor([float(sv.exchange_bytes - xb0)], device="cuda", dtype=torch.float64)
        if world > 1:
            dist.all_reduce(vms, op=dist.ReduceOp.MAX)
            dist.all_reduce(xb, op=dist.ReduceOp.SUM)
        vc = {"ms": float(vms[0]), "residual_before": float(res[0]), "residual_after": float(res[1]),
              "sweeps": "2 pre + 2 post multicolour Gauss-Seidel per level, 20 on the coarsest",
              "halo_exchanges": int(sv.exchanges - ex0), "halo_bytes_all_ranks": int(xb[0]),
              "colours": [g[0]["colours"] for g in sv.geom]}
        sv.close()
    hp.close()

    # end to end: host buffers -> device (each rank its own slice, all-gathered over NVLink), build, owned rows of P -> host
    hV, hcp, hri = inp.fetch_arrays()
    tV = torch.from_numpy(hV.T).pin_memory(); hV = tV.numpy().T
    tcp = torch.from_numpy(hcp).pin_memory(); hcp = tcp.numpy()
    tri = torch.from_numpy(hri).pin_memory(); hri = tri.numpy()
    hA = None
    if galerkin:
        tA = torch.from_numpy(inp.fetch_operator()).pin_memory(); hA = tA.numpy()

    pin_out = {}

    def pinned_out(key, count, dtype):
        if key not in pin_out or pin_out[key].shape[0] < count:
            tdt = {np.int32: torch.int32, np.float64: torch.float64}[dtype]
            pin_out[key] = torch.empty(max(count, 1), dtype=tdt).pin_memory().numpy()
        return pin_out[key]

    def step_e2e():
        if world > 1:       # each rank uploads its own slice over PCIe; NVLink all-gathers complete the mesh
            i2, _ = upload_sharded(ctx, comm, hV, hcp, hri, hA)
        else:
            i2 = Input(ctx, hV, hcp, hri, hA)
        h2 = build(i2)
        d2h = 0
        for lvl in range(h2.num_levels):
            li = h2.info(lvl)
            own, nzp = li["row_hi"] - li["row_lo"], li["nnz_p"]
            for which, cnt, dt in ((capi.H_P_ROWPTR, own + 1, np.int32), (capi.H_P_COLIDX, nzp, np.int32),
                                   (capi.H_P_VAL, nzp, np.float64)):
                d2h += h2.fetch(lvl, which, out=pinned_out((lvl, which), cnt, dt)).nbytes
        nb = i2.h2d_bytes
        h2.close(); i2.close()
        return nb, d2h

    step_e2e()
    e2e_ms, h2d, d2h = [], 0, 0
    for s in range(max(1, min(a.steps, 2))):
        barrier()
        t0 = time.perf_counter()
        h2d, d2h = step_e2e()
        barrier()
        e2e_ms.append(1e3 * (time.perf_counter() - t0))
    e2e = float(np.mean(e2e_ms))

    if world > 1:
        t = torch.tensor([ms, e2e], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, e2e = float(t[0]), float(t[1])
        b = torch.tensor([float(h2d), float(d2h)], device="cuda", dtype=torch.float64)
        dist.all_reduce(b, op=dist.ReduceOp.SUM)
        h2d, d2h = int(b[0]), int(b[1])
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    peak, peak_src = measured_peak()
    for i in infos:
        i.setdefault("nnz_ac", 0)
    total_bytes = sum(sum(v for k, v in stage_bytes(i).items() if k != "galerkin" or galerkin) for i in infos)
    achieved = total_bytes / (ms * 1e-3) / 1e9
    line = {"metric": METRIC, "value": n / (ms * 1e-3) / 1e6, "unit": "Mverts/s", "n_gpus": world, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": ms, "higher_is_better": True,
            "scaling": "strong" if a.workload == "C4" else "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {**bench_config(a, world),
                       "l2": "256 MB flush write between timed steps; inputs exceed the 126 MB L2",
                       "multi_gpu": (f"row-sharded over {world} ranks: owned ranges of fine vertices, replicated level "
                                     f"graph, {exchanges} NCCL all-gathers ({xbytes / 1e6:.1f} MB gathered per rank) per build; phase loop "
                                     + ("in C++ with the library's own communicator (gmg_shard_build_nccl)" if native is not None
                                        else "driven from Python through torch.distributed"))
                       if world > 1 else "single GPU through the sharded driver (world 1)",
                       "verts_per_gpu": n // world,
                       "levels": [{k: i[k] for k in ("n", "nnz", "m", "nnz_c", "T", "nnz_ac", "sampler_rounds", "parent_rounds")}
                                  for i in infos]},
            "clocks": clk, "gpu_launches": int(launches), "phase_ms": phase_ms,
            "e2e": {"value": n / (e2e * 1e-3) / 1e6, "unit": "Mverts/s", "ms_per_step": e2e,
                    "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h)},
            "roofline": {"bound": "hbm", "kernel": "whole sharded build (all stages, all levels, exchanges included)",
                         "achieved": achieved / world, "peak": peak, "peak_source": peak_src, "unit": "GB/s per GPU",
                         "frac": achieved / world / peak, "traffic": None, "algorithmic_bytes": int(total_bytes)}}
    if vc is not None:
        line["vcycle"] = vc
    if builder is not None:
        line["input_builder"] = builder
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    a = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if a.impl == "reference":
        run_reference(a, rank, world)
    elif world > 1 or a.workload.startswith("T") or (a.sharded and a.workload == "C4"):
        run_b200_sharded(a, rank, world, local_rank)
    else:
        run_b200(a, rank, world, local_rank)


if __name__ == "__main__":
    main()
