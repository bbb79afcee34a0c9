"""Headless dump of a hierarchy: what the reference's Polyscope viewer shows
(test/main.cpp:159-182 -- the point clouds of every level, the coarse triangles, the projected points)
written to files instead: .npy arrays plus an ASCII .ply per level that any mesh viewer opens. No graphics
stack is involved."""
import os

import numpy as np


def write_ply(path, points, triangles=None, scalars=None):
    """ASCII PLY: vertices (x y z [+ one float property per entry of `scalars`]) and optional triangle faces."""
    points = np.asarray(points, dtype=np.float64)
    tris = np.zeros((0, 3), dtype=np.int64) if triangles is None else np.asarray(triangles, dtype=np.int64).reshape(-1, 3)
    scalars = scalars or {}
    with open(path, "w") as f:
        f.write("ply\nformat ascii 1.0\ncomment gravomg-b200 headless dump\n")
        f.write(f"element vertex {points.shape[0]}\nproperty double x\nproperty double y\nproperty double z\n")
        for name in scalars:
            f.write(f"property double {name}\n")
        f.write(f"element face {tris.shape[0]}\nproperty list uchar int vertex_indices\nend_header\n")
        cols = [points] + [np.asarray(v, dtype=np.float64).reshape(-1, 1) for v in scalars.values()]
        np.savetxt(f, np.hstack(cols), fmt="%.17g")
        if tris.shape[0]:
            np.savetxt(f, np.hstack([np.full((tris.shape[0], 1), 3, dtype=np.int64), tris]), fmt="%d")


def dump_hierarchy(h, fine_points, outdir):
    """capi.Hierarchy -> outdir/level<k>_{coarse_points,parents,samples,P_rowptr,P_colidx,P_val,projected}.npy and
    outdir/level<k>_coarse.ply (coarse points + Voronoi triangles) / level<k>_fine.ply (fine points coloured by
    parent). Returns the list of files written."""
    os.makedirs(outdir, exist_ok=True)
    written = []
    pts = np.asarray(fine_points)
    for lvl in range(h.num_levels):
        a = h.level_arrays(lvl)
        proj = h.projected_points(lvl)                          # projectedPoints on the device (src/multigrid.cpp:500-510)
        for name, arr in (("coarse_points", a["cpts"]), ("parents", a["par"]), ("samples", a["sel"]),
                          ("P_rowptr", a["prp"]), ("P_colidx", a["pci"]), ("P_val", a["pva"]), ("projected", proj)):
            path = os.path.join(outdir, f"level{lvl}_{name}.npy")
            np.save(path, arr)
            written.append(path)
        path = os.path.join(outdir, f"level{lvl}_coarse.ply")
        write_ply(path, a["cpts"], a["tri"])
        written.append(path)
        path = os.path.join(outdir, f"level{lvl}_fine.ply")
        write_ply(path, pts, None, {"parent": a["par"], "displacement": np.linalg.norm(proj - pts, axis=1)})
        written.append(path)
        pts = a["cpts"]
    return written
