"""ctypes binding of the C ABI (include/gravomg_b200.h) -- the host-side mirror of the reference's
entry points, one method per reference function with the same argument meaning:

    reference (namespace GravoMG)            here (class Context)
    toEdgeDistanceMatrix(matrix, points)     to_edge_distance_matrix(colptr, rowidx, points)
    extractEdges(matrix)                     extract_edges(colptr, rowidx, val)
    averageEdgeLength(positions, neighbors)  average_edge_length(points, edges)
    fastDiscSample(pos, edge_matrix, radius) fast_disc_sample(points, colptr, rowidx, radius)
    assignParents(...)                       assign_parents(points, colptr, rowidx, samples)
    extractCoarseEdges(...)                  extract_coarse_edges(colptr, rowidx, val, m, parents)
    coarseFromMeanOfFineChildren(...)        coarse_from_mean_of_fine_children(points, colptr, rowidx, parents, m)
    constructVoronoiTriangles(...)           construct_voronoi_triangles(points, colptr, rowidx, val)
    constructProlongation(...)               construct_prolongation(fine, coarse, c_colptr, c_rowidx, c_val, parents, weighting)
    projectedPoints(weights, coarse_points)  projected_points(rowptr, colidx, val, coarse_points)

There is no CPU fallback: if the CUDA library is missing or no GPU is usable, construction raises.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libgravomg_b200.so")

BARYCENTRIC, UNIFORM, INVDIST = 0, 1, 2
NESTED = 0x100          # OR-ed into a weighting (include/gravomg_b200.h GMG_NESTED)
LEVEL_P_READY = 0x40000000
SAMPLER_EXACT, SAMPLER_PARALLEL, SAMPLER_HASHED = 0, 1, 2

STATUS_NAMES = {0: "GMG_OK", 1: "GMG_ERR_CUDA", 2: "GMG_ERR_INVALID", 3: "GMG_ERR_FALLBACK_RATIO",
                4: "GMG_ERR_ZERO_LENGTH_EDGE", 5: "GMG_ERR_ASYMMETRIC", 6: "GMG_ERR_STATE", 7: "GMG_ERR_NCCL"}

_i32p, _i64p, _f64p = C.POINTER(C.c_int32), C.POINTER(C.c_int64), C.POINTER(C.c_double)


class GmgError(RuntimeError):
    def __init__(self, status, message):
        super().__init__(f"{STATUS_NAMES.get(status, status)}: {message}")
        self.status = status


class ProlongationStats(C.Structure):
    _fields_ = [("no_triangle_found", C.c_int64), ("edges_found", C.c_int64), ("fallback_count", C.c_int64)]


_lib = None


def load_library():
    """Load the CUDA library; raises if it has not been built (no fallback exists)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise FileNotFoundError(f"{LIB_PATH} is missing: run `python -m gravomg_b200.build` "
                                    "(the product path has no CPU fallback)")
        _lib = C.CDLL(LIB_PATH)
        _lib.gmg_last_error.restype = C.c_char_p
        _lib.gmg_last_error.argtypes = [C.c_void_p]
        _lib.gmg_last_kernel_ms.restype = C.c_double
        _lib.gmg_last_kernel_ms.argtypes = [C.c_void_p]
        _lib.gmg_launch_count.restype = C.c_int64
        _lib.gmg_launch_count.argtypes = [C.c_void_p]
        _lib.gmg_destroy.argtypes = [C.c_void_p]
        _lib.gmg_destroy.restype = None
    return _lib


def _p(a, ty):
    return a.ctypes.data_as(ty)


def _pts(a):
    a = np.asarray(a, dtype=np.float64)
    if a.ndim != 2 or a.shape[1] != 3:
        raise ValueError("points must have shape (n, 3)")
    return np.asfortranarray(a)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _i64(a):
    return np.ascontiguousarray(a, dtype=np.int64)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


class Context:
    """One CUDA stream worth of state; one calling thread at a time."""

    def __init__(self, device=0, stream=None):
        self.lib = load_library()
        # stream: a cudaStream_t handle to run on (e.g. torch.cuda.Stream().cuda_stream); None or 0 makes
        # the library create its own stream (0 is the legacy default stream, which it never uses)
        self.stream_handle = int(stream or 0)
        h = C.c_void_p()
        rc = self.lib.gmg_create(C.c_int(device), C.c_void_p(stream or 0), C.byref(h))
        if rc != 0:
            raise GmgError(rc, self.lib.gmg_last_error(None).decode())
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            self.lib.gmg_destroy(self.h)
            self.h = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def _check(self, rc):
        if rc != 0:
            raise GmgError(rc, self.lib.gmg_last_error(self.h).decode())

    def copy_sync(self):
        """Wait for the copies issued on the copy stream (Hierarchy(stream_out=...))."""
        self.lib.gmg_copy_sync.argtypes = [C.c_void_p]
        self._check(self.lib.gmg_copy_sync(self.h))

    def trim(self):
        """Return the device buffers this context has parked for reuse; returns how many bytes that was."""
        n = C.c_int64()
        self.lib.gmg_trim.argtypes = [C.c_void_p, C.POINTER(C.c_int64)]
        self._check(self.lib.gmg_trim(self.h, C.byref(n)))
        return int(n.value)

    @property
    def last_kernel_ms(self):
        return float(self.lib.gmg_last_kernel_ms(self.h))

    @property
    def launch_count(self):
        return int(self.lib.gmg_launch_count(self.h))

    # ---- reference entry points ---------------------------------------------------------------
    def to_edge_distance_matrix(self, colptr, rowidx, points):
        pts, colptr, rowidx = _pts(points), _i32(colptr), _i32(rowidx)
        n = pts.shape[0]
        out = np.empty(rowidx.shape[0], dtype=np.float64)
        self._check(self.lib.gmg_to_edge_distance_matrix(self.h, C.c_int64(n), _p(colptr, _i32p), _p(rowidx, _i32p),
                                                         _p(pts, _f64p), C.c_int64(n), _p(out, _f64p)))
        return out

    def extract_edges(self, colptr, rowidx, val):
        colptr, rowidx, val = _i32(colptr), _i32(rowidx), _f64(val)
        n, nnz = colptr.shape[0] - 1, rowidx.shape[0]
        edges = np.empty((nnz, 2), dtype=np.int32, order="F")
        values = np.empty(nnz, dtype=np.float64)
        self._check(self.lib.gmg_extract_edges(self.h, C.c_int64(n), _p(colptr, _i32p), _p(rowidx, _i32p),
                                               _p(val, _f64p), _p(edges, _i32p), _p(values, _f64p)))
        return edges, values

    def average_edge_length(self, points, edges):
        pts = _pts(points)
        edges = np.asfortranarray(edges, dtype=np.int32)
        out = C.c_double()
        self._check(self.lib.gmg_average_edge_length(self.h, C.c_int64(pts.shape[0]), _p(pts, _f64p),
                                                     C.c_int64(pts.shape[0]), C.c_int64(edges.shape[0]),
                                                     _p(edges, _i32p), C.byref(out)))
        return out.value

    def fast_disc_sample_bounded(self, points, colptr, rowidx, radius, min_selection=0, max_selection=2 ** 62,
                                 mode=SAMPLER_EXACT, seed=42):
        """(selection, radius used): fastDiscSample with the min/max bounds of src/sampling.cpp:11-12."""
        pts, colptr, rowidx = _pts(points), _i32(colptr), _i32(rowidx)
        n = pts.shape[0]
        sel = np.empty(max(n, 1), dtype=np.int64)
        m, used = C.c_int64(), C.c_double()
        self.lib.gmg_fast_disc_sample_bounded.argtypes = [C.c_void_p, C.c_int64, _f64p, C.c_int64, _i32p, _i32p, C.c_double,
                                                          C.c_int64, C.c_int64, C.c_int, C.c_uint64, _i64p,
                                                          C.POINTER(C.c_int64), C.POINTER(C.c_double)]
        self._check(self.lib.gmg_fast_disc_sample_bounded(self.h, n, _p(pts, _f64p), n, _p(colptr, _i32p), _p(rowidx, _i32p),
                                                          float(radius), int(min_selection), int(max_selection), int(mode),
                                                          int(seed), _p(sel, _i64p), C.byref(m), C.byref(used)))
        return sel[:m.value].copy(), used.value

    def coarse_from_samples(self, points, samples):
        """Nested mode: the coarse points are the samples' positions (gmg_coarse_from_samples)."""
        pts = _pts(points)
        s = np.ascontiguousarray(samples, dtype=np.int64)
        out = np.empty((s.shape[0], 3), dtype=np.float64, order="F")
        self.lib.gmg_coarse_from_samples.argtypes = [C.c_void_p, C.c_int64, _f64p, C.c_int64, C.c_int64, _i64p, _f64p, C.c_int64]
        self._check(self.lib.gmg_coarse_from_samples(self.h, pts.shape[0], _p(pts, _f64p), pts.shape[0], s.shape[0],
                                                     _p(s, _i64p), _p(out, _f64p), max(s.shape[0], 1)))
        return out

    def morton_order(self, points):
        """Permutation sorting the points along the Morton curve of their bounding cube (gmg_morton_order)."""
        pts = _pts(points)
        perm = np.empty(pts.shape[0], dtype=np.int32)
        self.lib.gmg_morton_order.argtypes = [C.c_void_p, C.c_int64, _f64p, C.c_int64, _i32p]
        self._check(self.lib.gmg_morton_order(self.h, pts.shape[0], _p(pts, _f64p), pts.shape[0], _p(perm, _i32p)))
        return perm

    def fast_disc_sample(self, points, colptr, rowidx, radius, mode=SAMPLER_EXACT, seed=42):
        pts, colptr, rowidx = _pts(points), _i32(colptr), _i32(rowidx)
        n = pts.shape[0]
        sel = np.empty(max(n, 1), dtype=np.int64)
        m = C.c_int64()
        self._check(self.lib.gmg_fast_disc_sample(self.h, C.c_int64(n), _p(pts, _f64p), C.c_int64(n),
                                                  _p(colptr, _i32p), _p(rowidx, _i32p), C.c_double(radius),
                                                  C.c_int(mode), C.c_uint64(seed), _p(sel, _i64p), C.byref(m)))
        return sel[:m.value].copy()

    def assign_parents(self, points, colptr, rowidx, samples):
        pts, colptr, rowidx, samples = _pts(points), _i32(colptr), _i32(rowidx), _i64(samples)
        n = pts.shape[0]
        par = np.empty(n, dtype=np.int64)
        self._check(self.lib.gmg_assign_parents(self.h, C.c_int64(n), _p(pts, _f64p), C.c_int64(n), _p(colptr, _i32p),
                                                _p(rowidx, _i32p), C.c_int64(samples.shape[0]), _p(samples, _i64p),
                                                _p(par, _i64p)))
        return par

    def _fetch_sparse(self, nnz):
        inner = np.empty(nnz, dtype=np.int32)
        val = np.empty(nnz, dtype=np.float64)
        self._check(self.lib.gmg_fetch_sparse(self.h, _p(inner, _i32p), _p(val, _f64p)))
        return inner, val

    def extract_coarse_edges(self, colptr, rowidx, val, m, parents):
        colptr, rowidx, val, parents = _i32(colptr), _i32(rowidx), _f64(val), _i64(parents)
        n = colptr.shape[0] - 1
        ocp = np.empty(m + 1, dtype=np.int32)
        nnz = C.c_int64()
        self._check(self.lib.gmg_extract_coarse_edges(self.h, C.c_int64(n), _p(colptr, _i32p), _p(rowidx, _i32p),
                                                      _p(val, _f64p), C.c_int64(m), _p(parents, _i64p),
                                                      _p(ocp, _i32p), C.byref(nnz)))
        ri, va = self._fetch_sparse(nnz.value)
        return ocp, ri, va

    def coarse_from_mean_of_fine_children(self, points, colptr, rowidx, parents, m):
        pts, colptr, rowidx, parents = _pts(points), _i32(colptr), _i32(rowidx), _i64(parents)
        n = pts.shape[0]
        out = np.empty((m, 3), dtype=np.float64, order="F")
        self._check(self.lib.gmg_coarse_from_mean_of_fine_children(
            self.h, C.c_int64(n), _p(pts, _f64p), C.c_int64(n), _p(colptr, _i32p), _p(rowidx, _i32p),
            _p(parents, _i64p), C.c_int64(m), _p(out, _f64p), C.c_int64(max(m, 1))))
        return out

    def construct_voronoi_triangles(self, points, colptr, rowidx, val):
        pts, colptr, rowidx, val = _pts(points), _i32(colptr), _i32(rowidx), _f64(val)
        m = pts.shape[0]
        T = C.c_int64()
        fan_ptr = np.empty(m + 1, dtype=np.int64)
        self._check(self.lib.gmg_construct_voronoi_triangles(self.h, C.c_int64(m), _p(pts, _f64p), C.c_int64(m),
                                                             _p(colptr, _i32p), _p(rowidx, _i32p), _p(val, _f64p),
                                                             C.byref(T), _p(fan_ptr, _i64p)))
        T = T.value
        tri = np.empty((T, 3), dtype=np.int64)
        nrm = np.empty((T, 3), dtype=np.float64)
        fan = np.empty(3 * T, dtype=np.int64)
        self._check(self.lib.gmg_fetch_triangles(self.h, _p(tri, _i64p), _p(nrm, _f64p), _p(fan, _i64p)))
        return tri, nrm, fan_ptr, fan

    def construct_prolongation(self, fine_points, coarse_points, c_colptr, c_rowidx, c_val, parents,
                               weighting=BARYCENTRIC, return_stats=False):
        fp, cp = _pts(fine_points), _pts(coarse_points)
        c_colptr, c_rowidx, c_val, parents = _i32(c_colptr), _i32(c_rowidx), _f64(c_val), _i64(parents)
        n, m = fp.shape[0], cp.shape[0]
        rowptr = np.empty(n + 1, dtype=np.int32)
        nnz = C.c_int64()
        st = ProlongationStats()
        self._check(self.lib.gmg_construct_prolongation(
            self.h, C.c_int64(n), _p(fp, _f64p), C.c_int64(n), C.c_int64(m), _p(cp, _f64p), C.c_int64(max(m, 1)),
            _p(c_colptr, _i32p), _p(c_rowidx, _i32p), _p(c_val, _f64p), _p(parents, _i64p), C.c_int(weighting),
            _p(rowptr, _i32p), C.byref(nnz), C.byref(st)))
        ci, va = self._fetch_sparse(nnz.value)
        if return_stats:
            return rowptr, ci, va, dict(no_triangle_found=st.no_triangle_found, edges_found=st.edges_found,
                                        fallback_count=st.fallback_count)
        return rowptr, ci, va

    def projected_points(self, rowptr, colidx, val, coarse_points):
        cp, rowptr, colidx, val = _pts(coarse_points), _i32(rowptr), _i32(colidx), _f64(val)
        n, m = rowptr.shape[0] - 1, cp.shape[0]
        out = np.empty((n, 3), dtype=np.float64, order="F")
        self._check(self.lib.gmg_projected_points(self.h, C.c_int64(n), C.c_int64(m), _p(rowptr, _i32p),
                                                  _p(colidx, _i32p), _p(val, _f64p), _p(cp, _f64p),
                                                  C.c_int64(max(m, 1)), _p(out, _f64p), C.c_int64(max(n, 1))))
        return out


# ---- resident hierarchy builder (gmg_hierarchy_*) ------------------------------------------------
LEVEL_CALLBACK = C.CFUNCTYPE(None, C.c_void_p, C.c_void_p, C.c_int)


class HierarchyOptions(C.Structure):
    _fields_ = [("levels", C.c_int32), ("ratio", C.c_double), ("weighting", C.c_int32), ("sampler_mode", C.c_int32),
                ("seed", C.c_uint64), ("radius_policy", C.c_int32), ("radius_override", _f64p),
                ("build_galerkin", C.c_int32), ("on_level", LEVEL_CALLBACK), ("user", C.c_void_p)]


class LevelInfo(C.Structure):
    _fields_ = [("n", C.c_int64), ("nnz", C.c_int64), ("m", C.c_int64), ("nnz_c", C.c_int64), ("T", C.c_int64),
                ("nnz_p", C.c_int64), ("nnz_ac", C.c_int64), ("radius", C.c_double), ("sampler_rounds", C.c_int32),
                ("parent_rounds", C.c_int32), ("stats", ProlongationStats), ("stage_ms", C.c_double * 9),
                ("galerkin_ms", C.c_double * 3)]

    def as_dict(self):
        d = {k: getattr(self, k) for k in ("n", "nnz", "m", "nnz_c", "T", "nnz_p", "nnz_ac", "radius",
                                           "sampler_rounds", "parent_rounds")}
        d["no_triangle_found"] = self.stats.no_triangle_found
        d["edges_found"] = self.stats.edges_found
        d["fallback_count"] = self.stats.fallback_count
        d["stage_ms"] = list(self.stage_ms)
        d["galerkin_ms"] = list(self.galerkin_ms)
        return d


STAGE_NAMES = ["edge_lengths", "sampling", "parents", "cluster_csr", "coarse_edges", "coarse_points",
               "triangle_fans", "prolongation_rows", "galerkin"]

(H_SAMPLES, H_PARENTS, H_COARSE_COLPTR, H_COARSE_ROWIDX, H_COARSE_VAL, H_COARSE_POINTS, H_TRIANGLES, H_NORMALS,
 H_FAN_PTR, H_FAN, H_P_ROWPTR, H_P_COLIDX, H_P_VAL, H_AC_COLPTR, H_AC_ROWIDX, H_AC_VAL, H_FINE_VAL,
 H_PROJECTED) = range(18)


class Input:
    """A level-0 problem resident in HBM (points + the pattern of A, optionally A's values)."""

    def __init__(self, ctx, points, colptr, rowidx, a_val=None, defer_values=False):
        """defer_values: copy A's values on the copy stream after the call has returned (gmg_input_set_values_async);
        a_val must then stay alive (and should be pinned) until a call that uses the values has returned."""
        self.ctx = ctx
        pts, colptr, rowidx = _pts(points), _i32(colptr), _i32(rowidx)
        self.n, self.nnz = pts.shape[0], rowidx.shape[0]
        a = None if a_val is None else _f64(a_val)
        h = C.c_void_p()
        ctx._check(ctx.lib.gmg_input_create(ctx.h, C.c_int64(self.n), _p(pts, _f64p), C.c_int64(self.n),
                                            _p(colptr, _i32p), _p(rowidx, _i32p),
                                            _p(a, _f64p) if a is not None and not defer_values else None, C.byref(h)))
        self.h = h
        if a is not None and defer_values:
            self._values = a                      # keeps the host array alive
            ctx.lib.gmg_input_set_values_async.argtypes = [C.c_void_p, C.c_void_p, _f64p]
            ctx._check(ctx.lib.gmg_input_set_values_async(ctx.h, h, _p(a, _f64p)))
        self.h2d_bytes = pts.nbytes + colptr.nbytes + rowidx.nbytes + (a.nbytes if a is not None else 0)

    @classmethod
    def device_empty(cls, ctx, n, nnz, with_a=False):
        """Arrays allocated on the device, to be filled by the caller (sharded.upload_sharded) and committed."""
        self = cls.__new__(cls)
        self.ctx = ctx
        h = C.c_void_p()
        ctx.lib.gmg_input_create_device.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_int, C.POINTER(C.c_void_p)]
        ctx._check(ctx.lib.gmg_input_create_device(ctx.h, n, nnz, 1 if with_a else 0, C.byref(h)))
        self.h = h
        self.n, self.nnz = int(n), int(nnz)
        self.h2d_bytes = 0
        return self

    def device_ptr(self, which):
        ptr = C.c_void_p()
        self.ctx.lib.gmg_input_device_ptr.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p)]
        rc = self.ctx.lib.gmg_input_device_ptr(self.h, which, C.byref(ptr))
        if rc != 0:
            raise GmgError(rc, "gmg_input_device_ptr")
        return ptr.value

    def commit(self):
        self.ctx.lib.gmg_input_commit.argtypes = [C.c_void_p, C.c_void_p]
        self.ctx._check(self.ctx.lib.gmg_input_commit(self.ctx.h, self.h))

    @classmethod
    def torus(cls, ctx, nu, nv, R=1.0, r=0.3, amplitude=0.2, seed=42, mass_shift=None):
        """The noisy torus grid of configs C3/C5 generated on the device (gmg_generate_torus); with mass_shift
        also its system matrix, cotangent Laplacian + mass_shift * lumped mass (gmg_input_torus_operator)."""
        self = cls.__new__(cls)
        self.ctx = ctx
        h = C.c_void_p()
        ctx.lib.gmg_generate_torus.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_double, C.c_double, C.c_double,
                                               C.c_uint64, C.POINTER(C.c_void_p)]
        ctx._check(ctx.lib.gmg_generate_torus(ctx.h, nu, nv, R, r, amplitude, seed, C.byref(h)))
        self.h = h
        self.n, self.nnz = nu * nv, 7 * nu * nv
        self.h2d_bytes = 0
        if mass_shift is not None:
            ctx.lib.gmg_input_torus_operator.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_double]
            ctx._check(ctx.lib.gmg_input_torus_operator(ctx.h, h, nu, nv, float(mass_shift)))
        return self

    @classmethod
    def knn(cls, ctx, points, k=8):
        """Exact kNN graph of a point cloud built on the device (gmg_input_from_knn): union-symmetrised + diagonal."""
        self = cls.__new__(cls)
        self.ctx = ctx
        pts = _pts(points)
        h = C.c_void_p()
        ctx.lib.gmg_input_from_knn.argtypes = [C.c_void_p, C.c_int64, _f64p, C.c_int64, C.c_int32, C.POINTER(C.c_void_p)]
        ctx._check(ctx.lib.gmg_input_from_knn(ctx.h, pts.shape[0], _p(pts, _f64p), pts.shape[0], int(k), C.byref(h)))
        self.h = h
        self._sizes()
        self.h2d_bytes = pts.nbytes
        return self

    @classmethod
    def mesh(cls, ctx, points, faces, mass_shift=0.0):
        """Triangle mesh -> adjacency + diagonal pattern and the cotangent Laplacian (+ mass_shift * lumped mass),
        assembled on the device (gmg_input_from_mesh)."""
        self = cls.__new__(cls)
        self.ctx = ctx
        pts = _pts(points)
        F = np.ascontiguousarray(faces, dtype=np.int32)
        h = C.c_void_p()
        ctx.lib.gmg_input_from_mesh.argtypes = [C.c_void_p, C.c_int64, _f64p, C.c_int64, C.c_int64, _i32p, C.c_double,
                                                C.POINTER(C.c_void_p)]
        ctx._check(ctx.lib.gmg_input_from_mesh(ctx.h, pts.shape[0], _p(pts, _f64p), pts.shape[0], F.shape[0], _p(F, _i32p),
                                               float(mass_shift), C.byref(h)))
        self.h = h
        self._sizes()
        self.h2d_bytes = pts.nbytes + F.nbytes
        return self

    def _sizes(self):
        n, nnz = C.c_int64(), C.c_int64()
        self.ctx.lib.gmg_input_size.argtypes = [C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
        self.ctx.lib.gmg_input_size(self.h, C.byref(n), C.byref(nnz))
        self.n, self.nnz = int(n.value), int(nnz.value)

    def fetch_operator(self):
        """A values (nnz) of an input that has them, copied back to the host."""
        lib = self.ctx.lib
        lib.gmg_input_fetch.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        val = np.empty(self.nnz, dtype=np.float64)
        self.ctx._check(lib.gmg_input_fetch(self.ctx.h, self.h, 3, val.ctypes.data_as(C.c_void_p)))
        return val

    def fetch_arrays(self):
        """(points (n,3) F-order, colptr, rowidx) copied back to the host."""
        lib = self.ctx.lib
        lib.gmg_input_fetch.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        pts = np.empty((self.n, 3), dtype=np.float64, order="F")
        cp = np.empty(self.n + 1, dtype=np.int32)
        ri = np.empty(self.nnz, dtype=np.int32)
        for which, arr in ((0, pts), (1, cp), (2, ri)):
            self.ctx._check(lib.gmg_input_fetch(self.ctx.h, self.h, which, arr.ctypes.data_as(C.c_void_p)))
        return pts, cp, ri

    def close(self):
        if C is None:       # interpreter shutdown: module globals are already gone
            return
        if getattr(self, "h", None) and getattr(self.ctx, "h", None):
            self.ctx.lib.gmg_input_destroy.argtypes = [C.c_void_p]
            self.ctx.lib.gmg_input_destroy(self.h)
        self.h = None

    __del__ = close


class Hierarchy:
    """levels-1 prolongations (and optionally Galerkin operators) built and kept on the device."""

    def __init__(self, ctx, inp, levels, ratio=8.0, weighting=BARYCENTRIC, sampler_mode=SAMPLER_EXACT, seed=42,
                 radius_policy=0, radius_override=None, build_galerkin=False, stream_out=None):
        """stream_out(level, info) -> [(array id, pinned 1-D numpy array), ...]: called as soon as a level is
        complete on the device; the listed arrays are copied out on the context's copy stream while the next
        level builds (gmg_hierarchy_fetch_async). Wait with ctx.copy_sync() before reading them."""
        self.ctx = ctx
        ro = None
        if radius_override is not None:
            ro = np.zeros(max(levels - 1, 1), dtype=np.float64)
            ro[:len(radius_override)] = radius_override
        self._ro = ro
        opt = HierarchyOptions(levels, ratio, weighting, sampler_mode, seed, radius_policy,
                               _p(ro, _f64p) if ro is not None else None, 1 if build_galerkin else 0)
        self.streamed_bytes = 0
        if stream_out is not None:
            lib = ctx.lib
            lib.gmg_hierarchy_level_info.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
            lib.gmg_hierarchy_fetch_async.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]

            def on_level(_user, hptr, level):
                # with build_galerkin the library calls twice per level: (level | LEVEL_P_READY) when P is complete,
                # then the plain level when A_c is; stream_out sees info["p_ready_only"] = True on the first call
                early = bool(level & LEVEL_P_READY)
                level &= ~LEVEL_P_READY
                li = LevelInfo()
                if lib.gmg_hierarchy_level_info(hptr, level, C.byref(li)) != 0:
                    return
                info = li.as_dict()
                info["p_ready_only"] = early
                for which, out in stream_out(level, info):
                    if lib.gmg_hierarchy_fetch_async(ctx.h, hptr, level, which, out.ctypes.data_as(C.c_void_p)) == 0:
                        self.streamed_bytes += out.nbytes
            self._cb = LEVEL_CALLBACK(on_level)          # keep the trampoline alive for the duration of the build
            opt.on_level = self._cb
        h = C.c_void_p()
        ctx._check(ctx.lib.gmg_hierarchy_build(ctx.h, inp.h, C.byref(opt), C.byref(h)))
        self.h = h
        self.build_ms = ctx.last_kernel_ms
        ctx.lib.gmg_hierarchy_num_levels.argtypes = [C.c_void_p]
        self.num_levels = int(ctx.lib.gmg_hierarchy_num_levels(h))
        self.info = []
        for lvl in range(self.num_levels):
            li = LevelInfo()
            ctx.lib.gmg_hierarchy_level_info.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
            ctx._check(ctx.lib.gmg_hierarchy_level_info(h, lvl, C.byref(li)))
            self.info.append(li.as_dict())

    def close(self):
        if C is None:
            return
        if getattr(self, "h", None) and getattr(self.ctx, "h", None):
            self.ctx.lib.gmg_hierarchy_destroy.argtypes = [C.c_void_p]
            self.ctx.lib.gmg_hierarchy_destroy(self.h)
        self.h = None

    __del__ = close

    def fetch(self, level, which, out=None):
        """Copy one array of a level to the host. `out`: optional preallocated (e.g. pinned) 1-D numpy array of
        the right dtype and at least the right length -- device-to-pinned copies run at PCIe speed, copies
        into fresh pageable memory are staged and several times slower."""
        i = self.info[level]
        n, m, T = i["n"], i["m"], i["T"]
        shape_dtype = {
            H_SAMPLES: (m, np.int64), H_PARENTS: (n, np.int64), H_COARSE_COLPTR: (m + 1, np.int32),
            H_COARSE_ROWIDX: (i["nnz_c"], np.int32), H_COARSE_VAL: (i["nnz_c"], np.float64),
            H_COARSE_POINTS: (3 * m, np.float64), H_TRIANGLES: (3 * T, np.int64), H_NORMALS: (3 * T, np.float64),
            H_FAN_PTR: (m + 1, np.int64), H_FAN: (3 * T, np.int64), H_P_ROWPTR: (n + 1, np.int32),
            H_P_COLIDX: (i["nnz_p"], np.int32), H_P_VAL: (i["nnz_p"], np.float64), H_AC_COLPTR: (m + 1, np.int32),
            H_AC_ROWIDX: (i["nnz_ac"], np.int32), H_AC_VAL: (i["nnz_ac"], np.float64), H_FINE_VAL: (i["nnz"], np.float64),
            H_PROJECTED: (3 * n, np.float64),
        }[which]
        if out is None:
            out = np.empty(max(shape_dtype[0], 1), dtype=shape_dtype[1])
        elif out.dtype != shape_dtype[1] or out.ndim != 1 or out.shape[0] < shape_dtype[0] or not out.flags.c_contiguous:
            raise ValueError("fetch: `out` must be a contiguous 1-D array of the array's dtype and length")
        self.ctx._check(self.ctx.lib.gmg_hierarchy_fetch(self.ctx.h, self.h, C.c_int(level), C.c_int(which),
                                                         out.ctypes.data_as(C.c_void_p)))
        out = out[:shape_dtype[0]]
        if which == H_COARSE_POINTS:
            return out.reshape((m, 3), order="F")
        if which == H_PROJECTED:
            return out.reshape((n, 3), order="F")
        if which in (H_TRIANGLES, H_NORMALS):
            return out.reshape(T, 3)
        return out

    def level_arrays(self, level):
        """Everything the reference's per-level calls would have returned, keyed like tests/pipeline.py."""
        f = self.fetch
        return dict(radius=self.info[level]["radius"], sel=f(level, H_SAMPLES), par=f(level, H_PARENTS),
                    ccp=f(level, H_COARSE_COLPTR), cri=f(level, H_COARSE_ROWIDX), cval=f(level, H_COARSE_VAL),
                    cpts=f(level, H_COARSE_POINTS), tri=f(level, H_TRIANGLES), nrm=f(level, H_NORMALS),
                    fptr=f(level, H_FAN_PTR), fan=f(level, H_FAN), prp=f(level, H_P_ROWPTR),
                    pci=f(level, H_P_COLIDX), pva=f(level, H_P_VAL))

    def projected_points(self, level):
        """projectedPoints(P, coarse_points) of a level, computed on the device (src/multigrid.cpp:500-510)."""
        return self.fetch(level, H_PROJECTED)

    def prolongation(self, level):
        import scipy.sparse as sp
        i = self.info[level]
        return sp.csr_matrix((self.fetch(level, H_P_VAL), self.fetch(level, H_P_COLIDX), self.fetch(level, H_P_ROWPTR)),
                             shape=(i["n"], i["m"]))

    def galerkin(self, level):
        import scipy.sparse as sp
        i = self.info[level]
        return sp.csc_matrix((self.fetch(level, H_AC_VAL), self.fetch(level, H_AC_ROWIDX),
                              self.fetch(level, H_AC_COLPTR)), shape=(i["m"], i["m"]))


# ---- SpMV, multicolour Gauss-Seidel, V-cycle --------------------------------------------------------
def spmv(ctx, outer, inner, val, x, n_inner=None):
    """y[i] = sum_k val[k] * x[inner[k]] over outer vector i (A x for CSR arrays, A^T x for CSC arrays)."""
    outer, inner, val, x = _i32(outer), _i32(inner), _f64(val), _f64(x)
    n = outer.shape[0] - 1
    y = np.empty(n, dtype=np.float64)
    ctx._check(ctx.lib.gmg_spmv(ctx.h, C.c_int64(n), C.c_int64(x.shape[0] if n_inner is None else n_inner),
                                _p(outer, _i32p), _p(inner, _i32p), _p(val, _f64p), _p(x, _f64p), _p(y, _f64p)))
    return y


def gauss_seidel(ctx, outer, inner, val, b, x, sweeps=1):
    """Multicolour Gauss-Seidel sweeps; returns (x_new, colours, num_colours)."""
    outer, inner, val, b = _i32(outer), _i32(inner), _f64(val), _f64(b)
    x = np.array(x, dtype=np.float64)
    n = outer.shape[0] - 1
    colours = np.empty(n, dtype=np.int32)
    nc = C.c_int32()
    ctx._check(ctx.lib.gmg_gauss_seidel(ctx.h, C.c_int64(n), _p(outer, _i32p), _p(inner, _i32p), _p(val, _f64p),
                                        _p(b, _f64p), _p(x, _f64p), C.c_int(sweeps), _p(colours, _i32p), C.byref(nc)))
    return x, colours, nc.value


class Solver:
    """V-cycle on A x = b over a hierarchy built with build_galerkin=True."""

    def __init__(self, ctx, inp, hierarchy):
        self.ctx, self._inp, self._h = ctx, inp, hierarchy
        h = C.c_void_p()
        ctx._check(ctx.lib.gmg_solver_create(ctx.h, inp.h, hierarchy.h, C.byref(h)))
        self.h = h

    def vcycle(self, b, x0=None, cycles=1, pre=2, post=2, coarse_sweeps=50):
        b = _f64(b)
        x = np.zeros_like(b) if x0 is None else np.array(x0, dtype=np.float64)
        res = np.empty(cycles + 1, dtype=np.float64)
        self.ctx._check(self.ctx.lib.gmg_solver_vcycle(self.ctx.h, self.h, _p(b, _f64p), _p(x, _f64p), C.c_int(cycles),
                                                       C.c_int(pre), C.c_int(post), C.c_int(coarse_sweeps),
                                                       _p(res, _f64p)))
        return x, res

    def pcg(self, b, x0=None, rel_tol=1e-8, max_iter=100, sweeps=2, coarse_sweeps=20):
        """Conjugate gradients preconditioned by one symmetric V(sweeps, sweeps) cycle. Returns (x, residual
        norms [iterations + 1])."""
        b = _f64(b)
        x = np.zeros_like(b) if x0 is None else np.array(x0, dtype=np.float64)
        res = np.zeros(max_iter + 1, dtype=np.float64)
        it = C.c_int(0)
        lib = self.ctx.lib
        lib.gmg_solver_pcg.argtypes = [C.c_void_p, C.c_void_p, _f64p, _f64p, C.c_double, C.c_int, C.c_int, C.c_int,
                                       C.POINTER(C.c_int), _f64p]
        self.ctx._check(lib.gmg_solver_pcg(self.ctx.h, self.h, _p(b, _f64p), _p(x, _f64p), rel_tol, max_iter, sweeps,
                                           coarse_sweeps, C.byref(it), _p(res, _f64p)))
        return x, res[:it.value + 1]

    def close(self):
        if C is None:
            return
        if getattr(self, "h", None) and getattr(self.ctx, "h", None):
            self.ctx.lib.gmg_solver_destroy.argtypes = [C.c_void_p]
            self.ctx.lib.gmg_solver_destroy(self.h)
        self.h = None

    __del__ = close
