"""Parity of the CUDA path (through the C ABI) with the CPU oracle on identical inputs.
Integer outputs bit-exact; floating-point outputs are compared bit-exact too (the kernels follow the
oracle's association order with FMA contraction off), which is stricter than the 1e-12 the north star asks."""
import numpy as np
import pytest

from gravomg_b200 import meshgen
from pipeline import run_level, run_hierarchy, assert_level_equal

pytestmark = pytest.mark.gpu

CASES = [("C1", 1.0), ("C2", 1.0 / 64), ("C3", 0.004), ("C4", 0.002)]


@pytest.mark.parametrize("name,scale", CASES)
def test_hierarchy_bit_exact(ctx, orc, name, scale):
    cfg = meshgen.make_config(name, scale)
    V, A = cfg["V"], cfg["A"]
    want = run_hierarchy(orc, V, A, cfg["levels"])
    # radius is an input (the reference's summation order is sequential); everything else is computed
    got = run_hierarchy(ctx, V, A, cfg["levels"], radii=[w["radius"] for w in want])
    for lvl, (g, w) in enumerate(zip(got, want)):
        assert_level_equal(g, w, what=f"{name} level {lvl} ")


@pytest.mark.parametrize("weighting", [1, 2])
def test_weighting_schemes(ctx, orc, weighting):
    cfg = meshgen.make_config("C4", 0.001)
    V, A = cfg["V"], cfg["A"]
    val = orc.to_edge_distance_matrix(A.indptr, A.indices, V)
    w = run_level(orc, V, A.indptr, A.indices, val, weighting=weighting)
    g = run_level(ctx, V, A.indptr, A.indices, val, radius=w["radius"], weighting=weighting)
    assert_level_equal(g, w, what=f"weighting {weighting} ")


def test_average_edge_length_close(ctx, orc):
    cfg = meshgen.make_config("C1", 1.0)
    V, A = cfg["V"], cfg["A"]
    val = orc.to_edge_distance_matrix(A.indptr, A.indices, V)
    e_o, v_o = orc.extract_edges(A.indptr, A.indices, val)
    e_g, v_g = ctx.extract_edges(A.indptr, A.indices, val)
    assert np.array_equal(e_o, e_g) and np.array_equal(v_o, v_g)
    a, b = orc.average_edge_length(V, e_o), ctx.average_edge_length(V, e_g)
    assert abs(a - b) <= 1e-12 * abs(a)


@pytest.mark.parametrize("name,scale", [("C1", 1.0), ("C3", 0.004), ("C4", 0.002)])
def test_resident_hierarchy_matches_oracle(ctx, orc, name, scale):
    """gmg_hierarchy_build (everything resident in HBM) against the oracle looped over levels, with the
    radii the GPU used fed to the oracle (the mean edge length is a different summation order)."""
    from gravomg_b200.capi import Input, Hierarchy
    cfg = meshgen.make_config(name, scale)
    V, A = cfg["V"], cfg["A"]
    inp = Input(ctx, V, A.indptr, A.indices, A.data)
    h = Hierarchy(ctx, inp, cfg["levels"], build_galerkin=True)
    assert h.num_levels == cfg["levels"] - 1
    radii = [i["radius"] for i in h.info]
    want = run_hierarchy(orc, V, A, cfg["levels"], radii=radii)
    # the GPU's own radius agrees with the reference's to rounding
    ref_radii = [w["radius"] for w in run_hierarchy(orc, V, A, cfg["levels"])] if name == "C1" else None
    keys = [k for k in ("sel par ccp cri cval cpts tri nrm fptr fan prp pci pva").split()]
    Acur = A.tocsc()
    for lvl, w in enumerate(want):
        g = h.level_arrays(lvl)
        assert_level_equal(g, w, keys=keys, what=f"{name} level {lvl} ")
        # projectedPoints from the resident arrays (GMG_H_PROJECTED), bit-exact like the per-function call
        assert np.array_equal(h.projected_points(lvl), orc.projected_points(w["prp"], w["pci"], w["pva"], w["cpts"]))
        # Galerkin operator: the structural (un-pruned) pattern must be EQUAL to the CPU product's, values within
        # 1e-12; scipy (which sums duplicates and keeps explicit zeros too) cross-checks the CPU product itself
        P = h.prolongation(lvl)
        ac_want = orc.galerkin(Acur.indptr, Acur.indices, Acur.data, P.indptr, P.indices, P.data, P.shape[1])
        Ac = h.galerkin(lvl)
        assert np.array_equal(Ac.indptr, ac_want[0]) and np.array_equal(Ac.indices, ac_want[1]), f"{name} level {lvl}: A_c pattern"
        scale_v = np.abs(ac_want[2]).max()
        assert np.abs(Ac.data - ac_want[2]).max(initial=0.0) <= 1e-12 * scale_v
        Ac_ref = (P.T @ Acur @ P).tocsc()
        assert np.abs((Ac - Ac_ref).tocoo().data).max(initial=0.0) <= 1e-12 * scale_v
        Acur = Ac
    if ref_radii is not None:
        for a, b in zip(radii, ref_radii):
            assert abs(a - b) <= 1e-12 * abs(b)


from golden_util import GOLDEN, load_golden  # noqa: E402


@pytest.mark.parametrize("name", GOLDEN)
def test_cuda_matches_golden_fixtures(ctx, name):
    """Against outputs of the reference itself, committed under tests/golden/ (no oracle in the loop)."""
    V, A, levels, weighting, fine_val, want = load_golden(name)
    assert np.array_equal(ctx.to_edge_distance_matrix(A.indptr, A.indices, V), fine_val)
    got = run_hierarchy(ctx, V, A, levels, weighting=weighting, radii=[w["radius"] for w in want])
    for lvl, (g, w) in enumerate(zip(got, want)):
        assert_level_equal(g, w, what=f"{name} level {lvl} ")


def test_edge_cases_and_errors(ctx, orc):
    import scipy.sparse as sp
    from gravomg_b200 import capi
    from gravomg_b200.capi import GmgError
    V = np.asfortranarray(np.array([[0., 0, 0], [1, 0, 0], [5, 5, 5], [1.5, 0.2, 0]]))
    A = sp.csc_matrix(np.array([[1., 1, 0, 0], [1, 1, 0, 1], [0, 0, 1, 0], [0, 1, 0, 1]]))
    A.sort_indices()
    val = orc.to_edge_distance_matrix(A.indptr, A.indices, V)
    w = run_level(orc, V, A.indptr, A.indices, val, radius=0.6)     # degree-0 and degree-1 parents
    g = run_level(ctx, V, A.indptr, A.indices, val, radius=0.6)
    assert_level_equal(g, w, what="tiny ")
    # empty problem
    e_cp = np.zeros(1, dtype=np.int32); e_ri = np.zeros(0, dtype=np.int32)
    assert len(ctx.fast_disc_sample(np.zeros((0, 3)), e_cp, e_ri, 1.0)) == 0
    # asymmetric pattern is rejected, not silently mis-handled
    B = sp.csc_matrix(np.array([[1., 1, 0], [0, 1, 0], [0, 0, 1]])); B.sort_indices()
    with pytest.raises(GmgError) as e:
        ctx.fast_disc_sample(V[:3], B.indptr, B.indices, 1.0)
    assert e.value.status == 5
    # duplicate points: the reference's heap order becomes dynamic; reported, not guessed
    Vd = V.copy(); Vd[3] = Vd[1]
    with pytest.raises(GmgError) as e:
        ctx.assign_parents(Vd, A.indptr, A.indices, np.array([0, 2]))
    assert e.value.status == 4
    # out-of-range parent index
    with pytest.raises(GmgError) as e:
        ctx.extract_coarse_edges(A.indptr, A.indices, val, 2, np.array([0, 1, 2, 1]))
    assert e.value.status == 2
    # a column pointer that runs backwards is caught before any kernel indexes through it, and the context survives
    bad_cp = A.indptr.copy(); bad_cp[2] = bad_cp[1] - 1
    with pytest.raises(GmgError) as e:
        ctx.fast_disc_sample(V, bad_cp, A.indices, 1.0)
    assert e.value.status == 2
    with pytest.raises(GmgError) as e:
        capi.Input(ctx, V, bad_cp, A.indices, None)
    assert e.value.status == 2
    assert len(ctx.fast_disc_sample(V, A.indptr, A.indices, 0.6)) == len(w["sel"])
    # index arrays the kernels use as addresses are range-checked at the boundary
    with pytest.raises(GmgError) as e:
        ctx.average_edge_length(V, np.array([[0, 1], [1, 7]], dtype=np.int32))
    assert e.value.status == 2
    with pytest.raises(GmgError) as e:
        ctx.projected_points(w["prp"], np.where(w["pci"] == 0, 99, w["pci"]).astype(np.int32), w["pva"], w["cpts"])
    assert e.value.status == 2
    bad_inner = A.indices.copy(); bad_inner[3] = 9
    with pytest.raises(GmgError) as e:
        capi.spmv(ctx, A.indptr, bad_inner, A.data, np.ones(4))
    assert e.value.status == 2


def test_levels_streamed_out_during_the_build(ctx):
    """on_level / gmg_hierarchy_fetch_async: P of every level copied out on the copy stream while the next level
    builds must equal the ordinary fetch after the build."""
    import torch
    from gravomg_b200 import capi
    cfg = meshgen.make_config("C3", 0.01)
    V, A = cfg["V"], cfg["A"]
    inp = capi.Input(ctx, V, A.indptr, A.indices, None)
    got = {}

    def stream_out(level, info):
        bufs = [(capi.H_P_ROWPTR, torch.empty(info["n"] + 1, dtype=torch.int32).pin_memory().numpy()),
                (capi.H_P_COLIDX, torch.empty(info["nnz_p"], dtype=torch.int32).pin_memory().numpy()),
                (capi.H_P_VAL, torch.empty(info["nnz_p"], dtype=torch.float64).pin_memory().numpy())]
        got[level] = bufs
        return bufs

    h = capi.Hierarchy(ctx, inp, cfg["levels"], sampler_mode=capi.SAMPLER_PARALLEL, stream_out=stream_out)
    ctx.copy_sync()
    assert sorted(got) == list(range(h.num_levels)) and h.streamed_bytes > 0
    for lvl, bufs in got.items():
        for which, arr in bufs:
            assert np.array_equal(arr, h.fetch(lvl, which)), (lvl, which)
    # int64 outputs are converted on the way out and have no asynchronous path
    h.close(); inp.close()
    assert ctx.trim() >= 0


def test_streaming_with_galerkin_and_deferred_values(ctx):
    """With build_galerkin the level callback comes twice (P ready, then A_c ready), and A's values may be uploaded
    on the copy stream after the input exists (gmg_input_set_values_async): same hierarchy as the plain calls."""
    import torch
    from gravomg_b200 import capi
    cfg = meshgen.make_config("C3", 0.01)
    V, A = cfg["V"], cfg["A"]
    ref_inp = capi.Input(ctx, V, A.indptr, A.indices, A.data)
    ref = capi.Hierarchy(ctx, ref_inp, cfg["levels"], build_galerkin=True)
    pinned_a = torch.from_numpy(np.ascontiguousarray(A.data)).pin_memory().numpy()
    inp = capi.Input(ctx, V, A.indptr, A.indices, pinned_a, defer_values=True)
    got, calls = {}, []

    def stream_out(level, info):
        calls.append((level, info["p_ready_only"]))
        if info["p_ready_only"]:
            ids = [(capi.H_P_ROWPTR, info["n"] + 1, torch.int32), (capi.H_P_COLIDX, info["nnz_p"], torch.int32),
                   (capi.H_P_VAL, info["nnz_p"], torch.float64)]
        else:
            assert info["nnz_ac"] > 0
            ids = [(capi.H_AC_COLPTR, info["m"] + 1, torch.int32), (capi.H_AC_ROWIDX, info["nnz_ac"], torch.int32),
                   (capi.H_AC_VAL, info["nnz_ac"], torch.float64)]
        bufs = [(w, torch.empty(cnt, dtype=dt).pin_memory().numpy()) for w, cnt, dt in ids]
        got.setdefault(level, []).extend(bufs)
        return bufs

    h = capi.Hierarchy(ctx, inp, cfg["levels"], build_galerkin=True, stream_out=stream_out)
    ctx.copy_sync()
    assert calls == [(l, e) for l in range(h.num_levels) for e in (True, False)]
    for lvl, bufs in got.items():
        assert len(bufs) == 6
        for which, arr in bufs:
            assert np.array_equal(arr, ref.fetch(lvl, which)), (lvl, which)
    assert np.array_equal(inp.fetch_operator(), A.data)
    # a solver on the deferred input
    b = A @ (np.sin(3 * V[:, 0]) * np.cos(2 * V[:, 1]) + V[:, 2])
    x, res = capi.Solver(ctx, inp, h).vcycle(b, cycles=2)
    x_ref, res_ref = capi.Solver(ctx, ref_inp, ref).vcycle(b, cycles=2)
    assert res[-1] < 0.2 * res[0] and np.array_equal(res, res_ref) and np.array_equal(x, x_ref)
    h.close(); inp.close(); ref.close(); ref_inp.close()
