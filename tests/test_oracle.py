"""The oracle (plain-C restatement) is pinned two ways: against the committed golden fixtures, which
are outputs of the reference itself (tests/golden/make_golden.py), and, where oracle/_ref has been
built, against the unmodified reference sources directly on fresh inputs. CPU only."""
import numpy as np
import pytest

from gravomg_b200 import meshgen
from golden_util import GOLDEN, load_golden
from pipeline import run_hierarchy, assert_level_equal, STAGE_KEYS


@pytest.mark.parametrize("name", GOLDEN)
def test_oracle_matches_golden(orc, name):
    V, A, levels, weighting, fine_val, want = load_golden(name)
    assert np.array_equal(orc.to_edge_distance_matrix(A.indptr, A.indices, V), fine_val)
    got = run_hierarchy(orc, V, A, levels, weighting=weighting)
    assert len(got) == len(want)
    for lvl, (g, w) in enumerate(zip(got, want)):
        assert g["radius"] == w["radius"], f"{name} level {lvl}: radius"
        assert_level_equal(g, w, what=f"{name} level {lvl} ")


@pytest.mark.parametrize("name,scale", [("C1", 1.0), ("C2", 1 / 256), ("C3", 0.002), ("C4", 0.0005)])
def test_oracle_matches_reference_on_shim(orc, ref, name, scale):
    cfg = meshgen.make_config(name, scale)
    V, A = cfg["V"], cfg["A"]
    got = run_hierarchy(orc, V, A, cfg["levels"])
    want = run_hierarchy(ref, V, A, cfg["levels"])
    for lvl, (g, w) in enumerate(zip(got, want)):
        assert g["radius"] == w["radius"]
        assert_level_equal(g, w, what=f"{name} level {lvl} ")


def test_invariants(orc):
    """Properties that follow from the reference code, no second implementation needed."""
    cfg = meshgen.make_config("C4", 0.0005)
    V, A = cfg["V"], cfg["A"]
    res = run_hierarchy(orc, V, A, 3)
    for r in res:
        n, m = len(r["par"]), len(r["sel"])
        assert np.all(np.diff(r["sel"]) > 0)                                   # ascending selection
        assert np.array_equal(r["par"][r["sel"]], np.arange(m))                # parents[samples[c]] == c
        rows = np.diff(r["prp"])
        assert rows.min() >= 1 and rows.max() <= 3
        sums = np.add.reduceat(r["pva"], r["prp"][:-1])
        assert np.allclose(sums, 1.0, atol=1e-12)                              # rows of P sum to 1
        import scipy.sparse as sp
        C = sp.csc_matrix((r["cval"], r["cri"], r["ccp"]), shape=(m, m))
        pat = C.copy(); pat.data[:] = 1
        assert (pat != pat.T).nnz == 0 and C.diagonal().sum() == 0             # symmetric pattern, empty diagonal
        assert np.all((r["tri"][:, 0] < r["tri"][:, 1]) & (r["tri"][:, 1] < r["tri"][:, 2]))
        key = r["tri"][:, 0] * m * m + r["tri"][:, 1] * m + r["tri"][:, 2]
        assert np.all(np.diff(key) > 0)                                        # lexicographic triangle ids
        for v in range(0, m, max(1, m // 50)):
            assert np.all(np.diff(r["fan"][r["fptr"][v]:r["fptr"][v + 1]]) > 0)


def test_edge_cases(orc):
    # empty graph, isolated vertices (degree-0 parents), a single edge (degree-1 parents)
    V = np.asfortranarray(np.array([[0., 0, 0], [1, 0, 0], [5, 5, 5], [1.5, 0.2, 0]]))
    import scipy.sparse as sp
    A = sp.csc_matrix(np.array([[1., 1, 0, 0], [1, 1, 0, 1], [0, 0, 1, 0], [0, 1, 0, 1]]))
    A.sort_indices()
    val = orc.to_edge_distance_matrix(A.indptr, A.indices, V)
    sel = orc.fast_disc_sample(V, A.indptr, A.indices, 0.6)
    assert list(sel) == [0, 1, 2]
    par = orc.assign_parents(V, A.indptr, A.indices, sel)
    assert list(par) == [0, 1, 2, 1]
    ccp, cri, cval = orc.extract_coarse_edges(A.indptr, A.indices, val, 3, par)
    assert list(ccp) == [0, 1, 2, 2] and list(cri) == [1, 0]
    cpts = orc.coarse_from_mean_of_fine_children(V, A.indptr, A.indices, par, 3)
    prp, pci, pva, rc = orc.construct_prolongation(V, cpts, ccp, cri, cval, par, 0)
    assert rc == 0 and list(np.diff(prp)) == [2, 2, 1, 2]
    assert np.allclose(np.add.reduceat(pva, prp[:-1]), 1.0)


def test_headless_ply_dump(tmp_path):
    """gravo-mg_b200/dump.py write_ply: the header counts and the body must round-trip (the viewer replacement)."""
    import numpy as np
    from gravomg_b200.dump import write_ply
    pts = np.array([[0., 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1.5]])
    tris = np.array([[0, 1, 2], [0, 2, 3]])
    p = tmp_path / "t.ply"
    write_ply(str(p), pts, tris, {"parent": np.array([0, 0, 1, 1])})
    lines = p.read_text().splitlines()
    assert lines[0] == "ply" and "element vertex 4" in lines and "element face 2" in lines
    body = lines[lines.index("end_header") + 1:]
    assert len(body) == 6
    assert np.allclose([float(v) for v in body[3].split()], [0, 0, 1.5, 1])
    assert body[5].split() == ["3", "0", "2", "3"]


@pytest.mark.parametrize("name,scale", [("C1", 1.0), ("C3", 0.004), ("C4", 0.001)])
def test_solver_oracle_matches_scipy(orc, name, scale):
    """The parts the reference lacks (oracle/solver_oracle.c): P^T A P with the structural pattern of an un-pruned
    product and sorted indices, SpMV, Gauss-Seidel -- pinned against scipy / numpy, on every level of a hierarchy."""
    import scipy.sparse as sp
    cfg = meshgen.make_config(name, scale)
    V, A = cfg["V"], cfg["A"]
    res = run_hierarchy(orc, V, A, cfg["levels"])
    Acur = A.tocsc()
    for r in res:
        n, m = len(r["par"]), len(r["sel"])
        P = sp.csr_matrix((r["pva"], r["pci"], r["prp"]), shape=(n, m))
        ptr, idx, val = orc.galerkin(Acur.indptr, Acur.indices, Acur.data, P.indptr, P.indices, P.data, m)
        got = sp.csc_matrix((val, idx, ptr), shape=(m, m))
        assert np.all(np.diff(ptr) >= 0) and got.has_sorted_indices
        # structural pattern: every (j, i) reachable through stored entries of P, A, P -- with 0/1 matrices
        ones = lambda M: sp.csr_matrix((np.ones(M.nnz), M.indices, M.indptr), shape=M.shape) if M.format == "csr" else \
            sp.csc_matrix((np.ones(M.nnz), M.indices, M.indptr), shape=M.shape)
        pat = (ones(P).T @ ones(Acur) @ ones(P)).tocsc()
        pat.sort_indices()
        assert np.array_equal(pat.indptr, ptr) and np.array_equal(pat.indices, idx)
        want = (P.T @ Acur @ P).tocsc()
        assert np.abs((got - want).tocoo().data).max(initial=0.0) <= 1e-13 * np.abs(want.data).max()
        Acur = got
    n = A.shape[0]
    x = np.random.default_rng(1).standard_normal(n)
    assert np.abs(orc.spmv(A.indptr, A.indices, A.data, x) - A.T @ x).max() <= 1e-12 * np.abs(A.T @ x).max()
    As = (A + sp.identity(n)).tocsr(); As.sort_indices()
    b = np.ones(n)
    got = orc.gauss_seidel(As.indptr, As.indices, As.data, b, np.zeros(n), np.arange(n, dtype=np.int32), sweeps=2)
    Lo = sp.tril(As, format="csr")
    want = np.zeros(n)
    for _ in range(2):
        want = sp.linalg.spsolve_triangular(Lo, b - (As - Lo) @ want, lower=True)
    assert np.abs(got - want).max() <= 1e-10 * np.abs(want).max()


def test_oracle_options_semantics(orc):
    """The restatement of the two options the reference leaves as todo (selection bounds, nested mode): the
    properties their definitions promise, on C1."""
    cfg = meshgen.make_config("C1")
    V, A = cfg["V"], cfg["A"]
    val = orc.to_edge_distance_matrix(A.indptr, A.indices, V)
    edges, _ = orc.extract_edges(A.indptr, A.indices, val)
    radius = 2.0 * orc.average_edge_length(V, edges)
    plain = orc.fast_disc_sample(V, A.indptr, A.indices, radius)
    sel, used = orc.fast_disc_sample_bounded(V, A.indptr, A.indices, radius, 0, 50)
    assert np.array_equal(sel, plain[:50]) and used == radius
    sel, used = orc.fast_disc_sample_bounded(V, A.indptr, A.indices, radius, 2 * len(plain), 2 ** 62)
    assert len(sel) >= 2 * len(plain) and used in (radius / 2, radius / 4, radius / 8)
    assert np.array_equal(sel, orc.fast_disc_sample(V, A.indptr, A.indices, used))
    nested = run_hierarchy(orc, V, A, cfg["levels"], weighting=0x100)[0]
    s = nested["sel"]
    assert np.array_equal(nested["cpts"], V[s])
    assert np.all(np.diff(nested["prp"])[s] == 1) and np.all(nested["pva"][nested["prp"][s]] == 1.0)
    assert np.array_equal(nested["pci"][nested["prp"][s]], np.arange(len(s)))
    rows = np.add.reduceat(nested["pva"], nested["prp"][:-1])
    assert np.abs(rows - 1.0).max() < 1e-12
