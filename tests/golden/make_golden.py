"""Generate the committed golden fixtures from the REFERENCE ITSELF (unmodified sources on the Eigen
shim, oracle/_ref/libgravomg_ref.so). Run in the build container, where /root/reference is mounted:

    python tests/golden/make_golden.py

Each .npz holds the inputs and every intermediate of one or two levels of the demo call order
(test/main.cpp:68-121), so the fixtures pin the oracle (and through it the CUDA path) on a box that
has no /root/reference."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from gravomg_b200 import meshgen  # noqa: E402
from oracle.pyoracle import Oracle  # noqa: E402
from pipeline import run_hierarchy  # noqa: E402


def cases():
    V, F = meshgen.icosphere(3)
    yield "icosphere3", V, meshgen.cotangent_laplacian(V, F), 3, 0
    V, F = meshgen.icosphere(4)
    V = meshgen.add_noise(V, meshgen.mean_edge_length(V, F))
    yield "icosphere4_noise", V, meshgen.cotangent_laplacian(V, F), 3, 0
    V, F = meshgen.torus_grid(48, 16)
    V = meshgen.add_noise(V, 2 * np.pi * 0.3 / 16)
    yield "torus48x16_noise", V, meshgen.cotangent_laplacian(V, F, 1e-3), 3, 0
    V = meshgen.sphere_cloud(1500)
    A = meshgen.knn_graph(V, 8)
    yield "knn1500_bary", V, A, 3, 0
    yield "knn1500_uniform", V, A, 2, 1
    yield "knn1500_invdist", V, A, 2, 2


def main():
    ref = Oracle("ref")
    for name, V, A, levels, weighting in cases():
        res = run_hierarchy(ref, V, A, levels, weighting=weighting)
        out = dict(V=np.asarray(V), colptr=A.indptr.astype(np.int32), rowidx=A.indices.astype(np.int32),
                   aval=A.data.astype(np.float64), levels=np.int64(levels), weighting=np.int64(weighting),
                   fine_val=ref.to_edge_distance_matrix(A.indptr, A.indices, V))
        for lvl, r in enumerate(res):
            for k, v in r.items():
                out[f"L{lvl}_{k}"] = np.asarray(v)
        path = os.path.join(HERE, name + ".npz")
        np.savez_compressed(path, **out)
        print(name, V.shape[0], "->", [len(r["sel"]) for r in res], os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    main()
