"""Row-sharded build (gmg_shard_*, gravo-mg_b200/sharded.py) against the single-GPU build.

Several VIRTUAL ranks run in lock-step on the one GPU with the loop-back exchange; the phases, the
owned ranges and the buffers are exactly those of the multi-process NCCL run. The sharded build must
reproduce the single-GPU hierarchy bit for bit (indices AND values), for every world size, because
the greedy sample set, the shortest-path distances and the tie rules are order-independent facts.
"""
import numpy as np
import pytest

from gravomg_b200 import capi, meshgen
from pipeline import assert_level_equal, run_hierarchy

pytestmark = pytest.mark.gpu
# triangles and fans are rank-local in a sharded build (each rank enumerates what its rows of P read);
# the rows of P, which are computed FROM those fans, are compared instead
KEYS = "sel par ccp cri cval cpts prp pci pva".split()


def _single(ctx, V, A, levels, mode, radii=None):
    inp = capi.Input(ctx, V, A.indptr, A.indices, None)
    h = capi.Hierarchy(ctx, inp, levels, sampler_mode=mode, seed=7, radius_override=radii)
    out = [h.level_arrays(l) for l in range(h.num_levels)]
    h.close(); inp.close()
    return out


@pytest.mark.parametrize("name,scale,world,mode", [
    ("C1", 1.0, 2, capi.SAMPLER_EXACT), ("C1", 1.0, 3, capi.SAMPLER_HASHED), ("C3", 0.004, 4, capi.SAMPLER_HASHED),
    ("C4", 0.002, 2, capi.SAMPLER_HASHED), ("C3", 0.004, 8, capi.SAMPLER_EXACT)])
def test_sharded_equals_single_gpu(ctx, name, scale, world, mode):
    from gravomg_b200.sharded import ShardedHierarchy
    cfg = meshgen.make_config(name, scale)
    V, A = cfg["V"], cfg["A"]
    want = _single(ctx, V, A, cfg["levels"], mode)
    inp = capi.Input(ctx, V, A.indptr, A.indices, None)
    sh = ShardedHierarchy(ctx, inp, cfg["levels"], world, list(range(world)), comm=None, sampler_mode=mode, seed=7)
    assert sh.num_levels == len(want)
    for lvl, w in enumerate(want):
        g = sh.level_arrays(lvl)
        assert g["radius"] == w["radius"]
        assert_level_equal(g, w, keys=KEYS, what=f"{name} world {world} level {lvl} ")
        # every rank holds the same replicated coarse data
        for k in range(1, world):
            assert np.array_equal(sh.fetch(lvl, capi.H_PARENTS, k), w["par"])
            assert np.array_equal(sh.fetch(lvl, capi.H_COARSE_ROWIDX, k), w["cri"])
        # the owned row ranges tile [0, n)
        los = [sh.info(lvl, k)["row_lo"] for k in range(world)]
        his = [sh.info(lvl, k)["row_hi"] for k in range(world)]
        assert los[0] == 0 and his[-1] == sh.info(lvl)["n"] and all(a == b for a, b in zip(his[:-1], los[1:]))
    assert sh.exchanges > 0
    sh.close(); inp.close()


@pytest.mark.parametrize("name,scale,world", [("C1", 1.0, 2), ("C3", 0.004, 4), ("C3", 0.004, 7)])
def test_sharded_galerkin_equals_single_gpu(ctx, name, scale, world):
    """A_c = P^T A P on every level of the sharded build (rows of P gathered, owned outer vectors formed per
    rank, A_c gathered): same pattern as the single-GPU product, values to 1e-12 (the accumulation order inside
    a row's hash table is not fixed), and equal to scipy's P^T A P."""
    from gravomg_b200.sharded import ShardedHierarchy
    cfg = meshgen.make_config(name, scale)
    V, A = cfg["V"], cfg["A"]
    inp = capi.Input(ctx, V, A.indptr, A.indices, A.data)
    h = capi.Hierarchy(ctx, inp, cfg["levels"], sampler_mode=capi.SAMPLER_HASHED, seed=3, build_galerkin=True)
    sh = ShardedHierarchy(ctx, inp, cfg["levels"], world, list(range(world)), comm=None, sampler_mode=capi.SAMPLER_HASHED,
                          seed=3, build_galerkin=True)
    assert sh.num_levels == h.num_levels
    Acur = A.tocsc()
    for lvl in range(h.num_levels):
        want = h.galerkin(lvl)
        assert sh.info(lvl)["nnz_ac"] == want.nnz
        for k in (0, world - 1):
            got = sh.galerkin(lvl, k)
            assert np.array_equal(got.indptr, want.indptr) and np.array_equal(got.indices, want.indices)
            assert np.abs(got.data - want.data).max() <= 1e-12 * np.abs(want.data).max()
        P = h.prolongation(lvl)
        ref = (P.T @ Acur @ P).tocsc()
        D = (got - ref).tocoo()
        assert np.abs(D.data).max(initial=0.0) <= 1e-12 * np.abs(ref.data).max()
        # P itself is still the single-GPU P
        assert_level_equal(sh.level_arrays(lvl), h.level_arrays(lvl), keys=KEYS, what=f"{name} galerkin level {lvl} ")
        Acur = want
    sh.close(); h.close(); inp.close()


@pytest.mark.parametrize("world,replicate_below", [(2, 0), (5, 0), (3, 2_000_000), (4, 5_000)])
def test_sharded_vcycle_equals_single_gpu(ctx, world, replicate_below):
    """Row-sharded multicolour Gauss-Seidel / V-cycle with halo exchanges of ghost entries (packed all-to-all;
    loop-back here) against the single-GPU V-cycle on the same hierarchy: same colouring, same (colour, row)
    sweep order, so the iterates agree to rounding (A_c's accumulation order is the only difference)."""
    from gravomg_b200.sharded import ShardedHierarchy, ShardedSolver
    cfg = meshgen.make_config("C3", 0.01)
    V, A = cfg["V"], cfg["A"]
    n = V.shape[0]
    b = A @ (np.sin(3 * V[:, 0]) * np.cos(2 * V[:, 1]) + V[:, 2])
    inp = capi.Input(ctx, V, A.indptr, A.indices, A.data)
    h = capi.Hierarchy(ctx, inp, cfg["levels"], sampler_mode=capi.SAMPLER_HASHED, seed=5, build_galerkin=True)
    ref = capi.Solver(ctx, inp, h)
    x_ref, res_ref = ref.vcycle(b, cycles=3, pre=2, post=2, coarse_sweeps=20)
    sh = ShardedHierarchy(ctx, inp, cfg["levels"], world, list(range(world)), comm=None, sampler_mode=capi.SAMPLER_HASHED,
                          seed=5, build_galerkin=True)
    sv = ShardedSolver(sh, inp, replicate_below=replicate_below)
    assert sv.num_levels == h.num_levels + 1
    flags = [g[0]["replicated"] for g in sv.geom]
    assert flags[0] is False and flags[-1] is True and flags == sorted(flags)      # a suffix of the levels
    assert sum(flags) == (1 if replicate_below == 0 else sum(1 for g in sv.geom[1:] if g[0]["n"] <= replicate_below) or 1)
    x, res = sv.vcycle(b, cycles=3, pre=2, post=2, coarse_sweeps=20)
    assert not np.isnan(x).any()
    assert np.abs(res - res_ref).max() <= 1e-9 * res_ref[0], (res, res_ref)
    assert np.abs(x - x_ref).max() <= 1e-9 * np.abs(x_ref).max()
    assert res[-1] < 0.05 * res[0]
    assert sv.exchanges > 0 and sv.exchange_bytes > 0
    # the halo is a boundary layer, not the vector: far fewer ghosts than vertices
    ghosts = sum(int(g.numel()) for g in sv.plan_a[0].ghost)
    assert 0 < ghosts < 0.5 * n * world
    sv.close(); sh.close(); ref.close(); h.close(); inp.close()


def test_sharded_exact_matches_oracle(ctx, orc):
    """The sharded build in exact-sampler mode is the reference's hierarchy: checked against the oracle."""
    from gravomg_b200.sharded import ShardedHierarchy
    cfg = meshgen.make_config("C1")
    V, A = cfg["V"], cfg["A"]
    inp = capi.Input(ctx, V, A.indptr, A.indices, None)
    sh = ShardedHierarchy(ctx, inp, cfg["levels"], 4, [0, 1, 2, 3], comm=None, sampler_mode=capi.SAMPLER_EXACT)
    radii = [sh.info(l)["radius"] for l in range(sh.num_levels)]
    want = run_hierarchy(orc, V, A, cfg["levels"], radii=radii)
    for lvl, w in enumerate(want):
        assert_level_equal(sh.level_arrays(lvl), w, keys=KEYS, what=f"sharded C1 level {lvl} ")
    sh.close(); inp.close()


def test_device_torus_generator(ctx):
    """gmg_generate_torus against the host generator: identical pattern; positions equal up to the last
    ulp of sin/cos (the noise is bit-identical by construction)."""
    nu, nv = 50, 20
    V, F = meshgen.torus_grid(nu, nv)
    V = meshgen.add_noise(V, min(2 * np.pi * 1.3 / nu, 2 * np.pi * 0.3 / nv))
    A = meshgen.cotangent_laplacian(V, F)
    inp = capi.Input.torus(ctx, nu, nv)
    pts, cp, ri = inp.fetch_arrays()
    assert np.array_equal(cp, A.indptr) and np.array_equal(ri, A.indices)
    assert np.abs(pts - V).max() < 1e-15
    # the system matrix of C3/C5 on the same pattern (cotangent Laplacian + 1e-3 lumped mass)
    inp_a = capi.Input.torus(ctx, nu, nv, mass_shift=1e-3)
    want = meshgen.cotangent_laplacian(pts, F, mass_shift=1e-3)
    got = inp_a.fetch_operator()
    assert np.array_equal(want.indices, ri)
    assert np.abs(got - want.data).max() <= 1e-11 * np.abs(want.data).max()
    inp_a.close()
    # and it feeds the build
    h = capi.Hierarchy(ctx, inp, 3, sampler_mode=capi.SAMPLER_HASHED)
    assert h.num_levels == 2 and h.info[0]["m"] > 0
    h.close(); inp.close()


def test_sharded_errors_are_raised_by_every_rank(ctx):
    """A duplicate point in ONE rank's slice: the zero-length-edge flag is local to that rank's pull kernel, but the
    level-end checks run on the statuses of all ranks (LEVEL_STATUS / G(STATUS) / LEVEL_END), so every rank --
    here the first virtual rank, whose own slice is clean -- raises the same error instead of walking into the
    next level's all-gather alone."""
    from gravomg_b200.sharded import ShardedHierarchy
    cfg = meshgen.make_config("C3", 0.004)
    V, A = np.array(cfg["V"], order="F"), cfg["A"]
    n = V.shape[0]
    world = 4
    v = int(n * 0.6)                                   # a vertex of rank 2's slice
    nb = A.indices[A.indptr[v]:A.indptr[v + 1]]
    u = int(nb[nb != v][0])
    assert n // 2 < min(u, v) and max(u, v) < 3 * n // 4
    V[u] = V[v]
    inp = capi.Input(ctx, V, A.indptr, A.indices, None)
    with pytest.raises(capi.GmgError) as e:
        ShardedHierarchy(ctx, inp, cfg["levels"], world, list(range(world)), comm=None, sampler_mode=capi.SAMPLER_HASHED, seed=7)
    assert e.value.status == 4
    with pytest.raises(capi.GmgError) as e:            # the single-GPU build reports the same thing
        capi.Hierarchy(ctx, inp, cfg["levels"], sampler_mode=capi.SAMPLER_HASHED, seed=7)
    assert e.value.status == 4
    inp.close()
