"""Device input builders (gravo-mg_b200/csrc/k_builders.cu) against their host statements in meshgen.py:
the exact kNN graph (scipy cKDTree), the cotangent Laplacian from a face list, the Morton order."""
import numpy as np
import pytest

from gravomg_b200 import capi, meshgen

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n,k,kind", [(5000, 8, "sphere"), (200_000, 8, "sphere"), (60_000, 5, "cube"), (30_000, 20, "sphere"),
                                      (4000, 8, "clustered")])
def test_knn_graph_equals_kdtree(ctx, n, k, kind):
    """Pattern equality with the host kd-tree graph: exact neighbours, union-symmetrised, diagonal, sorted."""
    rng = np.random.default_rng(5)
    if kind == "sphere":
        V = meshgen.sphere_cloud(n)
    elif kind == "cube":
        V = np.asfortranarray(rng.random((n, 3)))
    else:       # a few tight clusters far apart plus stragglers: occupancy far from uniform, rings must widen
        centres = rng.random((7, 3)) * 100.0
        V = np.asfortranarray(centres[rng.integers(0, 7, n)] + rng.normal(size=(n, 3)) * 0.01)
        V[:50] = rng.random((50, 3)) * 100.0
    want = meshgen.knn_graph(V, k)
    inp = capi.Input.knn(ctx, V, k)
    pts, cp, ri = inp.fetch_arrays()
    assert np.array_equal(pts, V)
    assert np.array_equal(cp, want.indptr), "column pointers differ"
    assert np.array_equal(ri, want.indices), "neighbour sets differ"
    # and it feeds the build
    h = capi.Hierarchy(ctx, inp, 3, sampler_mode=capi.SAMPLER_EXACT)
    assert h.num_levels == 2 and h.info[0]["m"] > 0
    h.close(); inp.close()


def test_knn_rejects_bad_arguments(ctx):
    V = meshgen.sphere_cloud(100)
    with pytest.raises(capi.GmgError):
        capi.Input.knn(ctx, V, 0)
    with pytest.raises(capi.GmgError):
        capi.Input.knn(ctx, V[:8], 8)
    W = V.copy(order="F"); W[3, 1] = np.nan
    with pytest.raises(capi.GmgError) as e:
        capi.Input.knn(ctx, W, 8)
    assert e.value.status == 2                           # GMG_ERR_INVALID, not a CUDA fault
    capi.Input.knn(ctx, V, 8).close()                    # and the context is still usable


@pytest.mark.parametrize("name", ["C1", "C2s", "C3s"])
def test_mesh_laplacian_equals_host(ctx, name):
    if name == "C1":
        V, F = meshgen.icosphere(6)
        shift = 0.0
    elif name == "C2s":
        V, F = meshgen.icosphere(5)
        V = meshgen.add_noise(V, meshgen.mean_edge_length(V, F))
        shift = 0.0
    else:
        V, F = meshgen.torus_grid(120, 40)
        V = meshgen.add_noise(V, min(2 * np.pi * 1.3 / 120, 2 * np.pi * 0.3 / 40))
        shift = 1e-3
    want = meshgen.cotangent_laplacian(V, F, mass_shift=shift)
    inp = capi.Input.mesh(ctx, V, F, mass_shift=shift)
    pts, cp, ri = inp.fetch_arrays()
    assert np.array_equal(cp, want.indptr) and np.array_equal(ri, want.indices)
    got = inp.fetch_operator()
    assert np.abs(got - want.data).max() <= 1e-11 * np.abs(want.data).max()
    # symmetric to rounding, rows sum to mass_shift * area >= 0
    import scipy.sparse as sp
    A = sp.csc_matrix((got, ri, cp), shape=want.shape)
    assert abs(A - A.T).max() <= 1e-12 * np.abs(got).max()
    inp.close()
    with pytest.raises(capi.GmgError):
        Fb = F.copy(); Fb[0, 1] = V.shape[0]
        capi.Input.mesh(ctx, V, Fb)


@pytest.mark.parametrize("n", [1000, 300_000])
def test_morton_order(ctx, n):
    rng = np.random.default_rng(9)
    V = np.asfortranarray(rng.random((n, 3)) * [3.0, 1.0, 0.5] - 1.0)
    V[n // 2] = V[n // 3]                                 # a duplicate: ties go by index
    perm = ctx.morton_order(V)
    lo = V.min(axis=0); ext = (V.max(axis=0) - lo).max()
    q = np.minimum(np.maximum((V - lo) * (2097151.0 / ext), 0.0), 2097151.0).astype(np.uint64)
    code = np.zeros(n, dtype=np.uint64)
    for bit in range(21):
        for d in range(3):
            code |= ((q[:, d] >> np.uint64(bit)) & np.uint64(1)) << np.uint64(3 * bit + d)
    want = np.lexsort((np.arange(n), code))
    assert np.array_equal(perm, want)
