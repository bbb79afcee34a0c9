"""Procedural inputs for the BASELINE.json configs: icosphere, noisy torus grid,
kNN point cloud; cotangent Laplacian; counter-based RNG.

All inputs are generated on the HOST, deterministically, so the same arrays go to the CPU checker
and to the GPU. The step before the hot path in the reference is an external library call
(test/main.cpp:68, buildPointCloudLaplacian); these generators stand in for it.

Conventions: points (n,3) float64 Fortran-ordered (Eigen::MatrixXd layout, include/gravomg/utility.h:13);
matrices scipy CSC with int32 indices and sorted row indices (Eigen::SparseMatrix<double>, utility.h:15).
"""
import numpy as np
import scipy.sparse as sp

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def splitmix64(x):
    """Vectorised splitmix64 finaliser on uint64 arrays."""
    with np.errstate(over="ignore"):
        x = (x + np.uint64(0x9E3779B97F4A7C15)) & _M64
        x = ((x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M64
        x = ((x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M64
        return x ^ (x >> np.uint64(31))


def uniform_pm1(seed, index, component):
    """Uniform in [-1, 1) keyed by (seed, vertex index, component); identical on every machine."""
    with np.errstate(over="ignore"):
        key = (np.asarray(index, dtype=np.uint64) * np.uint64(4) + np.uint64(component)
               + np.uint64(seed) * np.uint64(0x100000001B3))
    bits = splitmix64(splitmix64(key))
    return (bits >> np.uint64(11)).astype(np.float64) * (2.0 / (1 << 53)) - 1.0


def add_noise(V, h, amplitude=0.2, seed=42):
    """Displace vertex i by amplitude*h*(u0,u1,u2) (SURVEY.md 8(d) noise rule)."""
    idx = np.arange(V.shape[0], dtype=np.uint64)
    out = np.array(V, dtype=np.float64, order="F")
    for c in range(3):
        out[:, c] += amplitude * h * uniform_pm1(seed, idx, c)
    return out


def icosphere(subdivisions):
    """Unit icosphere. Vertices of level k keep their indices at level k+1; midpoints are appended
    in ascending (min,max) edge order. Returns V (n,3) F-order, F (nf,3) int64."""
    t = (1.0 + np.sqrt(5.0)) / 2.0
    V = np.array([[-1, t, 0], [1, t, 0], [-1, -t, 0], [1, -t, 0], [0, -1, t], [0, 1, t],
                  [0, -1, -t], [0, 1, -t], [t, 0, -1], [t, 0, 1], [-t, 0, -1], [-t, 0, 1]], dtype=np.float64)
    V /= np.linalg.norm(V, axis=1, keepdims=True)
    F = np.array([[0, 11, 5], [0, 5, 1], [0, 1, 7], [0, 7, 10], [0, 10, 11], [1, 5, 9], [5, 11, 4], [11, 10, 2],
                  [10, 7, 6], [7, 1, 8], [3, 9, 4], [3, 4, 2], [3, 2, 6], [3, 6, 8], [3, 8, 9], [4, 9, 5],
                  [2, 4, 11], [6, 2, 10], [8, 6, 7], [9, 8, 1]], dtype=np.int64)
    for _ in range(subdivisions):
        n = V.shape[0]
        e = np.concatenate([F[:, [0, 1]], F[:, [1, 2]], F[:, [2, 0]]], axis=0)
        key = np.minimum(e[:, 0], e[:, 1]) * n + np.maximum(e[:, 0], e[:, 1])
        ukey, inv = np.unique(key, return_inverse=True)
        a, b = ukey // n, ukey % n
        mid = V[a] + V[b]
        mid /= np.linalg.norm(mid, axis=1, keepdims=True)
        V = np.concatenate([V, mid], axis=0)
        nf = F.shape[0]
        m01, m12, m20 = n + inv[:nf], n + inv[nf:2 * nf], n + inv[2 * nf:]
        F = np.concatenate([np.stack([F[:, 0], m01, m20], 1), np.stack([F[:, 1], m12, m01], 1),
                            np.stack([F[:, 2], m20, m12], 1), np.stack([m01, m12, m20], 1)], axis=0)
    return np.asfortranarray(V), F


def torus_grid(nu, nv, R=1.0, r=0.3):
    """Closed torus, nu steps around the big circle and nv around the tube; vertex (i,j) has index
    i*nv + j; every quad is split on the same diagonal. Returns V, F."""
    i = np.arange(nu, dtype=np.int64)[:, None]
    j = np.arange(nv, dtype=np.int64)[None, :]
    u = 2.0 * np.pi * i / nu
    v = 2.0 * np.pi * j / nv
    x = (R + r * np.cos(v)) * np.cos(u)
    y = (R + r * np.cos(v)) * np.sin(u)
    z = r * np.sin(v) * np.ones_like(u)
    V = np.asfortranarray(np.stack([x.ravel(), y.ravel(), z.ravel()], axis=1))
    i1 = (i + 1) % nu
    j1 = (j + 1) % nv
    a = (i * nv + j).ravel(); b = (i1 * nv + j).ravel(); c = (i1 * nv + j1).ravel(); d = (i * nv + j1).ravel()
    F = np.concatenate([np.stack([a, b, c], 1), np.stack([a, c, d], 1)], axis=0)
    return V, F


def mean_edge_length(V, F):
    e = np.concatenate([F[:, [0, 1]], F[:, [1, 2]], F[:, [2, 0]]], axis=0)
    return float(np.mean(np.linalg.norm(V[e[:, 0]] - V[e[:, 1]], axis=1)))


def _finish_csc(M):
    M = sp.csc_matrix(M)
    M.sum_duplicates()
    M.sort_indices()
    M.indptr = M.indptr.astype(np.int32)
    M.indices = M.indices.astype(np.int32)
    return M


def cotangent_laplacian(V, F, mass_shift=0.0):
    """Symmetric positive semi-definite cotangent Laplacian (+ mass_shift * lumped mass), CSC with
    int32 indices. Its PATTERN (adjacency + diagonal) is the graph the hot path works on."""
    V = np.asarray(V); n = V.shape[0]
    I, J, W = [], [], []
    area = np.zeros(n)
    for k in range(3):
        a, b, c = F[:, k], F[:, (k + 1) % 3], F[:, (k + 2) % 3]
        u, w = V[b] - V[a], V[c] - V[a]
        cr = np.linalg.norm(np.cross(u, w), axis=1)
        cot = np.einsum("ij,ij->i", u, w) / np.maximum(cr, 1e-300)   # cot of angle at a, opposite edge (b,c)
        I += [b, c]; J += [c, b]; W += [-0.5 * cot, -0.5 * cot]
        if k == 0:
            for q in (a, b, c):
                np.add.at(area, q, cr / 6.0)
    I = np.concatenate(I); J = np.concatenate(J); W = np.concatenate(W)
    L = sp.coo_matrix((W, (I, J)), shape=(n, n)).tocsc()
    d = -np.asarray(L.sum(axis=1)).ravel() + mass_shift * area
    return _finish_csc(L + sp.diags(d, format="csc"))


def adjacency_with_diagonal(n, rows, cols):
    """Symmetrised pattern (rows<->cols) plus the diagonal, all values 1.0, CSC int32."""
    r = np.concatenate([rows, cols, np.arange(n)])
    c = np.concatenate([cols, rows, np.arange(n)])
    M = sp.coo_matrix((np.ones(r.shape[0]), (r, c)), shape=(n, n)).tocsc()
    M.data[:] = 1.0
    return _finish_csc(M)


def sphere_cloud(n, seed=42):
    """n points uniform on the unit sphere from the counter RNG (z uniform, azimuth uniform)."""
    idx = np.arange(n, dtype=np.uint64)
    z = uniform_pm1(seed, idx, 0)
    phi = np.pi * (uniform_pm1(seed, idx, 1) + 1.0)
    s = np.sqrt(np.maximum(0.0, 1.0 - z * z))
    return np.asfortranarray(np.stack([s * np.cos(phi), s * np.sin(phi), z], axis=1))


def knn_graph(V, k=8):
    """Exact k nearest neighbours, union-symmetrised, plus the diagonal (config C4)."""
    from scipy.spatial import cKDTree
    n = V.shape[0]
    _, nb = cKDTree(np.ascontiguousarray(V)).query(np.ascontiguousarray(V), k=k + 1, workers=-1)
    rows = np.repeat(np.arange(n), k)
    cols = nb[:, 1:].ravel()
    return adjacency_with_diagonal(n, rows, cols)


def make_config(name, scale=1.0):
    """BASELINE.json configs by short name. `scale` < 1 shrinks a config for tests.
    Returns dict(V, A, levels, name); A is the system matrix whose pattern is the level-0 graph."""
    if name == "C1":                                   # icosphere 6 subdivisions, no noise, 3 levels
        V, F = icosphere(6 if scale >= 1 else max(1, int(round(6 + np.log2(scale) / 2))))
        return dict(name=name, V=V, A=cotangent_laplacian(V, F), levels=3)
    if name == "C2":                                   # icosphere 9 subdivisions + noise, 4 levels
        V, F = icosphere(9 if scale >= 1 else max(1, int(round(9 + np.log2(scale) / 2))))
        V = add_noise(V, mean_edge_length(V, F))
        return dict(name=name, V=V, A=cotangent_laplacian(V, F), levels=4)
    if name == "C3":                                   # torus 10000 x 1000 + noise, SPD operator
        nu, nv = max(8, int(10000 * np.sqrt(scale))), max(8, int(1000 * np.sqrt(scale)))
        V, F = torus_grid(nu, nv)
        h = min(2 * np.pi * 1.3 / nu, 2 * np.pi * 0.3 / nv)
        V = add_noise(V, h)
        return dict(name=name, V=V, A=cotangent_laplacian(V, F, mass_shift=1e-3), levels=4)
    if name == "C4":                                   # kNN cloud
        n = max(64, int(20_000_000 * scale))
        V = sphere_cloud(n)
        return dict(name=name, V=V, A=knn_graph(V, 8), levels=4)
    raise ValueError(name)
