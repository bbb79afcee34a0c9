"""The two options the reference marks as todo: selection bounds for the sampler
(src/sampling.cpp:11-12) and nested mode with exact-match weight 1.0 (src/multigrid.cpp:291-292). The reference has
no code for them, so parity is against the plain-C restatement only (oracle/gravomg_oracle.c states the semantics)."""
import numpy as np
import pytest

from gravomg_b200 import capi, meshgen
from pipeline import assert_level_equal, run_hierarchy, run_level

pytestmark = pytest.mark.gpu
KEYS = "sel par ccp cri cval cpts tri nrm fptr fan prp pci pva".split()


def _radius(orc, V, A):
    val = orc.to_edge_distance_matrix(A.indptr, A.indices, V)
    edges, _ = orc.extract_edges(A.indptr, A.indices, val)
    return 2.0 * orc.average_edge_length(V, edges)


@pytest.mark.parametrize("name,scale", [("C1", 1.0), ("C4", 0.002), ("C3", 0.01)])
def test_sampler_bounds(ctx, orc, name, scale):
    cfg = meshgen.make_config(name, scale)
    V, A = cfg["V"], cfg["A"]
    n = V.shape[0]
    radius = _radius(orc, V, A)
    plain = orc.fast_disc_sample(V, A.indptr, A.indices, radius)
    m0 = len(plain)
    # no bounds: the plain sampler
    sel, used = ctx.fast_disc_sample_bounded(V, A.indptr, A.indices, radius)
    assert np.array_equal(sel, plain) and used == radius
    # max: the head of the ascending selection
    sel, used = ctx.fast_disc_sample_bounded(V, A.indptr, A.indices, radius, max_selection=m0 // 3)
    want, wused = orc.fast_disc_sample_bounded(V, A.indptr, A.indices, radius, 0, m0 // 3)
    assert np.array_equal(sel, want) and np.array_equal(sel, plain[:m0 // 3]) and used == wused == radius
    # min: the radius is halved until there are enough samples
    for need in (m0 + 1, 3 * m0, n, n + 5):
        sel, used = ctx.fast_disc_sample_bounded(V, A.indptr, A.indices, radius, min_selection=need)
        want, wused = orc.fast_disc_sample_bounded(V, A.indptr, A.indices, radius, need, 2 ** 62)
        assert np.array_equal(sel, want), f"min_selection {need}"
        assert used == wused and used < radius and (len(sel) >= need or len(sel) == n)
    # both at once, and the parallel order goes through the same loop
    sel, used = ctx.fast_disc_sample_bounded(V, A.indptr, A.indices, radius, min_selection=2 * m0, max_selection=2 * m0 + 7)
    want, wused = orc.fast_disc_sample_bounded(V, A.indptr, A.indices, radius, 2 * m0, 2 * m0 + 7)
    assert np.array_equal(sel, want) and used == wused and len(sel) <= 2 * m0 + 7
    sel, _ = ctx.fast_disc_sample_bounded(V, A.indptr, A.indices, radius, max_selection=10, mode=capi.SAMPLER_HASHED, seed=3)
    assert len(sel) == 10 and np.all(np.diff(sel) > 0)
    with pytest.raises(capi.GmgError):
        ctx.fast_disc_sample_bounded(V, A.indptr, A.indices, radius, min_selection=-1)


@pytest.mark.parametrize("name,scale,weighting", [("C1", 1.0, capi.BARYCENTRIC), ("C4", 0.002, capi.INVDIST),
                                                  ("C3", 0.01, capi.UNIFORM), ("C2", 0.02, capi.BARYCENTRIC)])
def test_nested_mode(ctx, orc, name, scale, weighting):
    """Coarse points = sample positions, every sample's row is (its own coarse index, 1.0), everything else as usual:
    per function against the restatement, and through the resident builder."""
    cfg = meshgen.make_config(name, scale)
    V, A = cfg["V"], cfg["A"]
    w = weighting | capi.NESTED
    want = run_hierarchy(orc, V, A, cfg["levels"], weighting=w)
    # per function
    lvl0 = want[0]
    cpts = ctx.coarse_from_samples(V, lvl0["sel"])
    assert np.array_equal(cpts, V[lvl0["sel"]])
    prp, pci, pva = ctx.construct_prolongation(V, cpts, lvl0["ccp"], lvl0["cri"], lvl0["cval"], lvl0["par"], w)[:3]
    assert np.array_equal(prp, lvl0["prp"]) and np.array_equal(pci, lvl0["pci"]) and np.array_equal(pva, lvl0["pva"])
    sel = lvl0["sel"]
    assert np.all(np.diff(prp)[sel] == 1) and np.array_equal(pci[prp[sel]], np.arange(len(sel))) and np.all(pva[prp[sel]] == 1.0)
    # not the same operator as the plain mode
    plain = run_level(orc, V, A.indptr, A.indices, orc.to_edge_distance_matrix(A.indptr, A.indices, V), weighting=weighting)
    assert not np.array_equal(plain["cpts"], lvl0["cpts"])
    # the resident builder
    inp = capi.Input(ctx, V, A.indptr, A.indices, None)
    h = capi.Hierarchy(ctx, inp, cfg["levels"], weighting=w, radius_override=[x["radius"] for x in want])
    assert h.num_levels == len(want)
    for lvl, x in enumerate(want):
        assert_level_equal(h.level_arrays(lvl), x, keys=KEYS, what=f"{name} nested level {lvl} ")
    h.close(); inp.close()


def test_nested_mode_sharded(ctx):
    from gravomg_b200.sharded import ShardedHierarchy
    cfg = meshgen.make_config("C3", 0.004)
    V, A = cfg["V"], cfg["A"]
    w = capi.BARYCENTRIC | capi.NESTED
    inp = capi.Input(ctx, V, A.indptr, A.indices, None)
    h = capi.Hierarchy(ctx, inp, cfg["levels"], weighting=w, sampler_mode=capi.SAMPLER_HASHED, seed=7)
    sh = ShardedHierarchy(ctx, inp, cfg["levels"], 3, [0, 1, 2], comm=None, weighting=w, sampler_mode=capi.SAMPLER_HASHED, seed=7)
    keys = "sel par ccp cri cval cpts prp pci pva".split()
    for lvl in range(h.num_levels):
        assert_level_equal(sh.level_arrays(lvl), h.level_arrays(lvl), keys=keys, what=f"sharded nested level {lvl} ")
    sh.close(); h.close(); inp.close()
