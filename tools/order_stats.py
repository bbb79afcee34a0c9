"""CPU experiment behind DESIGN.md section 4 ("sampler orders"): sample count and coverage radius of the greedy rule
(reference src/sampling.cpp:22-50, run by the checker) under candidate precedence orders, against the index order.

    python tools/order_stats.py C3 0.1 [supers ...]

An order is given as a permutation; the graph is renumbered by it and the reference rule is run on the result.
TEST/EXPERIMENT INFRASTRUCTURE: uses oracle/, never part of the product path."""
import sys
import numpy as np
import scipy.sparse as sp

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
import importlib
meshgen = importlib.import_module("gravomg_b200.meshgen")
from oracle.pyoracle import Oracle, available, build

TILE = 4096


def interleaved_rank(n, supers):
    """GMG_SAMPLER_PARALLEL's order (k_sampler.cu order_rank) for `supers` super-tiles"""
    third = ((n + supers - 1) // supers + TILE - 1) // TILE * TILE
    v = np.arange(n, dtype=np.int64)
    sup = v // third
    inner = v - sup * third
    return ((inner // TILE) * supers + sup) * TILE + (inner % TILE)


def run_order(orc, V, A, radius, rank):
    n = V.shape[0]
    perm = np.argsort(rank, kind="stable")
    Ap = sp.csc_matrix((np.ones(A.nnz), A.indices, A.indptr), shape=(n, n))[perm][:, perm].tocsc()
    Ap.sort_indices()
    Vp = np.asfortranarray(V[perm])
    sel = orc.fast_disc_sample(Vp, Ap.indptr.astype(np.int32), Ap.indices.astype(np.int32), radius)
    return np.sort(perm[sel])


def coverage_radius(orc, V, cp, ri, sel):
    par = orc.assign_parents(V, cp, ri, sel)
    return float(np.linalg.norm(V - V[sel[par]], axis=1).max())


def main():
    name, scale = sys.argv[1], float(sys.argv[2])
    supers = [int(x) for x in sys.argv[3:]] or [2, 3]
    if not available("orc"):
        build()
    orc = Oracle("orc")
    cfg = meshgen.make_config(name, scale)
    V, A = cfg["V"], cfg["A"]
    n = V.shape[0]
    val = orc.to_edge_distance_matrix(A.indptr, A.indices, V)
    edges, _ = orc.extract_edges(A.indptr, A.indices, val)
    radius = 2.0 * orc.average_edge_length(V, edges)
    ref = orc.fast_disc_sample(V, A.indptr, A.indices, radius)
    r_ref = coverage_radius(orc, V, A.indptr, A.indices, ref)
    print(f"{name} x{scale}: n {n}, index order: {len(ref)} samples, coverage radius {r_ref:.6g}", flush=True)
    for s in supers:
        sel = run_order(orc, V, A, radius, interleaved_rank(n, s))
        r = coverage_radius(orc, V, A.indptr, A.indices, sel)
        print(f"  {s} interleaved super-tiles: {len(sel)} samples ({100 * (len(sel) - len(ref)) / len(ref):+.2f} %), "
              f"coverage radius {r:.6g} ({100 * (r - r_ref) / r_ref:+.2f} %)", flush=True)


if __name__ == "__main__":
    main()
