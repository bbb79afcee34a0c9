"""One level of the reference's demo call order (test/main.cpp:68-121) against any backend that
exposes the reference entry points (oracle.pyoracle.Oracle or gravomg_b200.capi.Context)."""
import numpy as np

STAGE_KEYS = ["sel", "par", "ccp", "cri", "cval", "cpts", "tri", "nrm", "fptr", "fan", "prp", "pci", "pva", "proj"]


def run_level(be, pts, cp, ri, val, radius=None, samples=None, weighting=0, ratio=8.0):
    out = {}
    if radius is None:
        edges, _ = be.extract_edges(cp, ri, val)
        radius = np.cbrt(ratio) * be.average_edge_length(pts, edges)
    out["radius"] = radius
    out["sel"] = be.fast_disc_sample(pts, cp, ri, radius) if samples is None else np.asarray(samples, dtype=np.int64)
    m = len(out["sel"])
    out["par"] = be.assign_parents(pts, cp, ri, out["sel"])
    out["ccp"], out["cri"], out["cval"] = be.extract_coarse_edges(cp, ri, val, m, out["par"])
    if weighting & 0x100:       # nested mode: the coarse points are the samples themselves
        out["cpts"] = np.asfortranarray(np.asarray(pts)[out["sel"]])
    else:
        out["cpts"] = be.coarse_from_mean_of_fine_children(pts, cp, ri, out["par"], m)
    out["tri"], out["nrm"], out["fptr"], out["fan"] = be.construct_voronoi_triangles(out["cpts"], out["ccp"], out["cri"], out["cval"])
    r = be.construct_prolongation(pts, out["cpts"], out["ccp"], out["cri"], out["cval"], out["par"], weighting)
    out["prp"], out["pci"], out["pva"] = r[0], r[1], r[2]
    out["proj"] = be.projected_points(out["prp"], out["pci"], out["pva"], out["cpts"])
    return out


def run_hierarchy(be, V, A, levels, weighting=0, radii=None):
    """levels-1 prolongations, feeding (coarse_points, coarse_edge_matrix) back in as the next fine level."""
    pts, cp, ri = V, A.indptr, A.indices
    val = be.to_edge_distance_matrix(cp, ri, pts)
    res = []
    for lvl in range(levels - 1):
        r = run_level(be, pts, cp, ri, val, radius=None if radii is None else radii[lvl], weighting=weighting)
        res.append(r)
        pts, cp, ri, val = r["cpts"], r["ccp"], r["cri"], r["cval"]
    return res


def assert_level_equal(a, b, keys=STAGE_KEYS, what=""):
    for k in keys:
        x, y = np.asarray(a[k]), np.asarray(b[k])
        assert x.shape == y.shape, f"{what}{k}: shape {x.shape} vs {y.shape}"
        if x.dtype.kind == "f":
            same = (x == y) | (np.isnan(x) & np.isnan(y))
            assert same.all(), f"{what}{k}: {np.count_nonzero(~same)} of {x.size} values differ, max abs {np.nanmax(np.abs(x - y))}"
        else:
            assert np.array_equal(x, y), f"{what}{k}: {np.count_nonzero(x != y)} of {x.size} entries differ"
