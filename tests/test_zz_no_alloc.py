"""gmg_average_edge_length_noalloc (csrc/k_noalloc.cu): the entry point that allocates nothing -- device pointers in and
out, scratch in a fixed __device__ array -- against the allocating gmg_average_edge_length, through the C ABI, on two
calls in a row (the second smaller than the first: the array keeps the first call's partial sums and they must not
matter). Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

from gravomg_b200 import capi, meshgen

pytestmark = pytest.mark.gpu


def _edge_list(orc, cfg):
    V, A = cfg["V"], cfg["A"]
    val = orc.to_edge_distance_matrix(A.indptr, A.indices, V)
    edges, _ = orc.extract_edges(A.indptr, A.indices, val)
    return np.asfortranarray(V), np.asfortranarray(edges, dtype=np.int32)


def test_no_alloc_entry_point_matches_the_allocating_one(orc):
    import torch
    stream = torch.cuda.Stream()
    with torch.cuda.stream(stream):
        ctx = capi.Context(0, stream.cuda_stream)          # the library's work and torch's tensors on ONE stream
        lib = ctx.lib
        lib.gmg_average_edge_length_noalloc.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
        lib.gmg_average_edge_length_noalloc.restype = C.c_int
        d_out = torch.zeros(1, dtype=torch.float64, device="cuda")
        for name, scale in (("C3", 0.02), ("C1", 1.0)):   # 1.4 M edge rows (684 partial sums), then 0.29 M (140)
            V, edges = _edge_list(orc, meshgen.make_config(name, scale))
            n, ne = V.shape[0], edges.shape[0]
            want = ctx.average_edge_length(V, edges)
            # device copies made by the CALLER: column-major n x 3 packed, column-major n_edges x 2
            d_pts = torch.from_numpy(np.ascontiguousarray(V.T)).cuda()
            d_edges = torch.from_numpy(np.ascontiguousarray(edges.T)).cuda()
            d_out.fill_(-1.0)
            before = torch.cuda.memory_allocated()
            rc = lib.gmg_average_edge_length_noalloc(ctx.h, n, C.c_void_p(d_pts.data_ptr()), ne, C.c_void_p(d_edges.data_ptr()),
                                                     C.c_void_p(d_out.data_ptr()))
            assert rc == 0, lib.gmg_last_error(ctx.h).decode()
            assert torch.cuda.memory_allocated() == before
            stream.synchronize()
            got = float(d_out.cpu()[0])
            assert abs(got - want) <= 1e-12 * abs(want), (name, got, want)
            assert got == want, "same summation tree: the two entry points should agree bit for bit"
        # beyond the fixed scratch: refused with a clear message, nothing launched
        rc = lib.gmg_average_edge_length_noalloc(ctx.h, 10, C.c_void_p(d_pts.data_ptr()), 134_217_729, C.c_void_p(d_edges.data_ptr()),
                                                 C.c_void_p(d_out.data_ptr()))
        assert rc == 2 and b"134,217,728" in lib.gmg_last_error(ctx.h)
        rc = lib.gmg_average_edge_length_noalloc(ctx.h, 10, None, 5, C.c_void_p(d_edges.data_ptr()), C.c_void_p(d_out.data_ptr()))
        assert rc == 2
        stream.synchronize()
        ctx.close()
