"""SpMV, multicolour Gauss-Seidel and the V-cycle [not in reference; BASELINE.json north star].
Checker: scipy for A x, a numpy restatement of Gauss-Seidel in (colour, row) order, tolerance 1e-12."""
import numpy as np
import pytest
import scipy.sparse as sp

from gravomg_b200 import capi, meshgen

pytestmark = pytest.mark.gpu


def numpy_colour_gs(A, b, x, colours, nc, sweeps):
    A = sp.csr_matrix(A)
    d = A.diagonal()
    x = x.copy()
    for _ in range(sweeps):
        for c in range(nc):
            rows = np.nonzero(colours == c)[0]
            Ax = A[rows] @ x
            x[rows] = (b[rows] - (Ax - d[rows] * x[rows])) / d[rows]
    return x


@pytest.mark.parametrize("name,scale", [("C1", 1.0), ("C3", 0.004)])
def test_spmv_and_gauss_seidel(ctx, name, scale):
    from gravomg_b200 import capi
    cfg = meshgen.make_config(name, scale)
    A = cfg["A"]
    n = A.shape[0]
    rng = np.random.default_rng(0)
    x = rng.standard_normal(n)
    y = capi.spmv(ctx, A.indptr, A.indices, A.data, x)
    want = A.T @ x
    assert np.abs(y - want).max() <= 1e-12 * np.abs(want).max()
    As = (A + 1e-3 * sp.identity(n)).tocsc(); As.sort_indices()
    b = rng.standard_normal(n)
    x1, colours, nc = capi.gauss_seidel(ctx, As.indptr, As.indices, As.data, b, np.zeros(n), sweeps=3)
    # proper colouring
    coo = As.tocoo(); off = coo.row != coo.col
    assert np.all(colours[coo.row[off]] != colours[coo.col[off]]) and colours.min() >= 0 and nc <= 16
    want = numpy_colour_gs(As, b, np.zeros(n), colours, nc, 3)
    assert np.abs(x1 - want).max() <= 1e-12 * np.abs(want).max()


def test_vcycle_poisson(ctx):
    """One V-cycle (2 pre + 2 post sweeps) on the SPD torus operator of config C3 (scaled): the
    residual must drop, and repeated cycles must converge."""
    from gravomg_b200.capi import Input, Hierarchy, Solver
    cfg = meshgen.make_config("C3", 0.01)
    V, A = cfg["V"], cfg["A"]
    n = V.shape[0]
    xstar = np.sin(3 * V[:, 0]) * np.cos(2 * V[:, 1]) + V[:, 2]
    b = A @ xstar
    inp = Input(ctx, V, A.indptr, A.indices, A.data)
    h = Hierarchy(ctx, inp, cfg["levels"], build_galerkin=True)
    s = Solver(ctx, inp, h)
    x, res = s.vcycle(b, cycles=8, pre=2, post=2, coarse_sweeps=50)
    assert res[1] < 0.5 * res[0]
    assert res[-1] < 1e-3 * res[0], res
    assert np.abs(A @ x - b).max() <= 1.01 * res[-1]
    factors = res[1:] / res[:-1]
    assert factors.max() < 0.9
    s.close(); h.close(); inp.close()


def test_pcg_with_vcycle_preconditioner(ctx):
    """Conjugate gradients preconditioned by one symmetric V(2,2) cycle on the SPD torus operator: converges to
    1e-10 in a few tens of iterations (plain CG needs hundreds), to the solution scipy's direct solve gives."""
    import scipy.sparse.linalg as spla
    from gravomg_b200.capi import Input, Hierarchy, Solver
    cfg = meshgen.make_config("C3", 0.01)
    V, A = cfg["V"], cfg["A"]
    xstar = np.sin(3 * V[:, 0]) * np.cos(2 * V[:, 1]) + V[:, 2]
    b = A @ xstar
    inp = Input(ctx, V, A.indptr, A.indices, A.data)
    h = Hierarchy(ctx, inp, cfg["levels"], build_galerkin=True)
    s = Solver(ctx, inp, h)
    x, res = s.pcg(b, rel_tol=1e-10, max_iter=60)
    assert len(res) - 1 <= 40, len(res)
    assert res[-1] <= 1e-10 * np.linalg.norm(b)
    assert np.abs(A @ x - b).max() <= 10 * res[-1]
    ref = spla.spsolve(A.tocsc(), b)
    assert np.abs(x - ref).max() <= 1e-6 * np.abs(ref).max()
    # the preconditioner matters: unpreconditioned CG is nowhere near after the same number of iterations
    _, info = spla.cg(A, b, rtol=1e-10, maxiter=len(res) - 1)
    assert info != 0
    s.close(); h.close(); inp.close()


def test_colouring_is_proper_beyond_64_colours(ctx):
    """Rows with far more than 64 neighbours (Galerkin operators of kNN-32 clouds, deep coarse levels): no stored
    entry may join two rows of the same colour, or the 'multicolour' sweep is no longer a Gauss-Seidel sweep.
    Checked on A_c of a kNN cloud and on a banded matrix whose rows need ~100 colours."""
    import scipy.sparse as sp
    from oracle.pyoracle import Oracle
    cfg = meshgen.make_config("C4", 1500 / 20_000_000)
    V, A0 = cfg["V"], cfg["A"]
    inp = capi.Input(ctx, V, A0.indptr, A0.indices, A0.data + 0.0)
    h = capi.Hierarchy(ctx, inp, 3, build_galerkin=True)
    mats = [h.galerkin(0).tocsr()]
    h.close(); inp.close()
    n = 600
    band = sp.diags([np.full(n - abs(k), -1.0 if k else 250.0) for k in range(-100, 101)], list(range(-100, 101)), format="csr")
    mats.append(band)
    for A in mats:
        A = sp.csr_matrix(A); A.sort_indices()
        n = A.shape[0]
        A = A + sp.diags(np.asarray(np.abs(A).sum(axis=1)).ravel() + 1.0)           # diagonally dominant: the sweep is well defined
        A = sp.csr_matrix(A); A.sort_indices()
        b = np.ones(n)
        x1, colours, nc = capi.gauss_seidel(ctx, A.indptr, A.indices, A.data, b, np.zeros(n), sweeps=2)
        coo = A.tocoo()
        off = coo.row != coo.col
        assert colours.min() >= 0 and colours.max() == nc - 1
        assert np.all(colours[coo.row[off]] != colours[coo.col[off]]), "two coupled rows share a colour"
        order = np.lexsort((np.arange(n), colours)).astype(np.int32)
        want = Oracle("orc").gauss_seidel(A.indptr, A.indices, A.data, b, np.zeros(n), order, sweeps=2)
        assert np.abs(x1 - want).max() <= 1e-12 * np.abs(want).max()
    assert nc > 64        # the banded matrix needs more colours than one 64-bit mask holds
