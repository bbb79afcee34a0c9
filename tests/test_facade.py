"""The C++ facade (namespace GravoMG, the reference's signatures, Eigen-typed) driven by a program
written the way a caller of the reference writes it (gravo-mg_b200/facade/facade_test.cpp), compared
with the golden outputs of the reference itself."""
import os
import struct
import subprocess

import numpy as np
import pytest

from golden_util import load_golden

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "gravo-mg_b200", "lib", "facade_test")


def test_facade_headers_keep_the_reference_include_paths():
    inc = os.path.join(ROOT, "gravo-mg_b200", "facade", "include", "gravomg")
    for h in ("multigrid.h", "sampling.h", "utility.h"):
        assert os.path.exists(os.path.join(inc, h))
    with open(os.path.join(inc, "b200_facade.h")) as f:
        text = f.read()
    for name in ("fastDiscSample", "assignParents", "averageEdgeLength", "extractCoarseEdges",
                 "coarseFromMeanOfFineChildren", "constructVoronoiTriangles", "constructProlongation",
                 "projectedPoints", "toEdgeDistanceMatrix", "extractEdges", "scaleMesh", "inTriangle",
                 "uniformWeights", "inverseDistanceWeights"):
        assert name in text


def _write(f, arr):
    arr = np.ascontiguousarray(arr)
    f.write(struct.pack("qq", arr.size, arr.dtype.itemsize))
    f.write(arr.tobytes())


def _read(f, dtype):
    count, size = struct.unpack("qq", f.read(16))
    assert size == np.dtype(dtype).itemsize
    return np.frombuffer(f.read(count * size), dtype=dtype)


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["icosphere4_noise", "knn1500_bary", "knn1500_invdist"])
def test_facade_program_matches_reference(tmp_path, ctx, name):
    assert os.path.exists(EXE), "facade_test not built: run python __graft_entry__.py"
    V, A, levels, weighting, fine_val, want = load_golden(name)
    fin, fout = tmp_path / "in.bin", tmp_path / "out.bin"
    with open(fin, "wb") as f:
        _write(f, np.array([V.shape[0], levels, weighting], dtype=np.int64))
        _write(f, np.array([w["radius"] for w in want], dtype=np.float64))
        _write(f, np.asfortranarray(V).T.copy())          # column-major n x 3
        _write(f, A.indptr.astype(np.int32)); _write(f, A.indices.astype(np.int32)); _write(f, A.data.astype(np.float64))
    r = subprocess.run([EXE, str(fin), str(fout)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    with open(fout, "rb") as f:
        for lvl, w in enumerate(want):
            own_radius = _read(f, np.float64)[0]
            assert abs(own_radius - w["radius"]) <= 1e-12 * w["radius"]
            got = dict(sel=_read(f, np.int64), par=_read(f, np.int64), ccp=_read(f, np.int32), cri=_read(f, np.int32),
                       cval=_read(f, np.float64))
            m = len(got["sel"])
            got["cpts"] = _read(f, np.float64).reshape((m, 3), order="F")
            got["tri"] = _read(f, np.int64).reshape(-1, 3); got["nrm"] = _read(f, np.float64).reshape(-1, 3)
            got["fptr"] = _read(f, np.int64); got["fan"] = _read(f, np.int64)
            got["prp"] = _read(f, np.int32); got["pci"] = _read(f, np.int32); got["pva"] = _read(f, np.float64)
            got["proj"] = _read(f, np.float64).reshape((len(got["par"]), 3), order="F")
            for k, v in got.items():
                assert np.array_equal(v, np.asarray(w[k])), f"{name} level {lvl} {k}"
        # resident builder through the facade == resident builder through the Python mirror
        from gravomg_b200.capi import Input, Hierarchy
        h = Hierarchy(ctx, Input(ctx, V, A.indptr, A.indices, A.data), levels, weighting=weighting, build_galerkin=True)
        nl = _read(f, np.int64)[0]
        assert nl == h.num_levels
        for lvl in range(nl):
            assert _read(f, np.float64)[0] == h.info[lvl]["radius"]
            P = h.prolongation(lvl)
            assert np.array_equal(_read(f, np.int32), P.indptr) and np.array_equal(_read(f, np.int32), P.indices)
            assert np.array_equal(_read(f, np.float64), P.data)
            Ac = h.galerkin(lvl)
            assert np.array_equal(_read(f, np.int32), Ac.indptr) and np.array_equal(_read(f, np.int32), Ac.indices)
            assert np.allclose(_read(f, np.float64), Ac.data, rtol=1e-12, atol=1e-12 * np.abs(Ac.data).max())
        # GravoMG::knnGraph (gmg_input_from_knn through the facade) == the host kd-tree graph
        from gravomg_b200 import meshgen
        G = meshgen.knn_graph(np.asfortranarray(V), 8)
        assert np.array_equal(_read(f, np.int32), G.indptr) and np.array_equal(_read(f, np.int32), G.indices)
        assert np.array_equal(_read(f, np.float64), np.ones(G.nnz))
