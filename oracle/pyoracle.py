"""TEST INFRASTRUCTURE ONLY -- ctypes binding of the two CPU checkers (oracle/oracle_api.h).

May be imported only by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference
arm. The product package never imports it.

    Oracle("orc")  -> oracle/liboracle.so            plain-C restatement
    Oracle("ref")  -> oracle/_ref/libgravomg_ref.so  unmodified reference sources on the Eigen shim

Array conventions: points are float64 arrays of shape (n, 3) in Fortran order (column-major, like
Eigen::MatrixXd); edge matrices are scipy-style CSC triples (colptr int32, rowidx int32, val f64).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_PATHS = {"orc": os.path.join(_HERE, "liboracle.so"), "ref": os.path.join(_HERE, "_ref", "libgravomg_ref.so")}

_i32p = C.POINTER(C.c_int32)
_i64p = C.POINTER(C.c_int64)
_f64p = C.POINTER(C.c_double)


def build(quiet=True):
    """Compile liboracle.so and, where /root/reference is mounted, oracle/_ref."""
    subprocess.run(["make", "-C", _HERE, "all"], check=True,
                   stdout=subprocess.DEVNULL if quiet else None)


def available(kind):
    return os.path.exists(_PATHS[kind])


def _p(a, ty):
    return a.ctypes.data_as(ty)


def fpoints(a):
    """(n,3) float64 column-major copy/view."""
    a = np.asarray(a, dtype=np.float64)
    assert a.ndim == 2 and a.shape[1] == 3
    return np.asfortranarray(a)


class Oracle:
    def __init__(self, kind="orc"):
        if not os.path.exists(_PATHS[kind]) and kind == "orc":
            build()
        self.kind = kind
        self.lib = C.CDLL(_PATHS[kind])
        self._f = lambda name: getattr(self.lib, f"{kind}_{name}")
        self._f("average_edge_length").restype = C.c_double
        self._f("fast_disc_sample").restype = C.c_int64
        self._f("construct_voronoi_triangles").restype = C.c_int64
        self._f("construct_prolongation").restype = C.c_int
        self._f("free").argtypes = [C.c_void_p]

    def _take(self, ptr, count, dtype):
        ty = {np.int32: C.c_int32, np.int64: C.c_int64, np.float64: C.c_double}[dtype]
        out = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(ty)), shape=(max(count, 1),))[:count].copy()
        self._f("free")(C.cast(ptr, C.c_void_p))
        return out

    @staticmethod
    def _csc(colptr, rowidx, val=None):
        colptr = np.ascontiguousarray(colptr, dtype=np.int32)
        rowidx = np.ascontiguousarray(rowidx, dtype=np.int32)
        if val is not None:
            val = np.ascontiguousarray(val, dtype=np.float64)
        return colptr, rowidx, val

    def to_edge_distance_matrix(self, colptr, rowidx, points):
        pts = fpoints(points); n = pts.shape[0]
        colptr, rowidx, _ = self._csc(colptr, rowidx)
        out = np.empty(rowidx.shape[0], dtype=np.float64)
        self._f("to_edge_distance_matrix")(C.c_int64(n), _p(colptr, _i32p), _p(rowidx, _i32p), _p(pts, _f64p),
                                           _p(out, _f64p))
        return out

    def extract_edges(self, colptr, rowidx, val):
        colptr, rowidx, val = self._csc(colptr, rowidx, val)
        n = colptr.shape[0] - 1; nnz = rowidx.shape[0]
        edges = np.empty((nnz, 2), dtype=np.int32, order="F")
        values = np.empty(nnz, dtype=np.float64)
        self._f("extract_edges")(C.c_int64(n), _p(colptr, _i32p), _p(rowidx, _i32p), _p(val, _f64p),
                                 _p(edges, _i32p), _p(values, _f64p))
        return edges, values

    def average_edge_length(self, points, edges):
        pts = fpoints(points)
        edges = np.asfortranarray(edges, dtype=np.int32)
        return float(self._f("average_edge_length")(C.c_int64(pts.shape[0]), _p(pts, _f64p),
                                                    C.c_int64(edges.shape[0]), _p(edges, _i32p)))

    def fast_disc_sample(self, points, colptr, rowidx, radius):
        pts = fpoints(points); n = pts.shape[0]
        colptr, rowidx, _ = self._csc(colptr, rowidx)
        sel = np.empty(max(n, 1), dtype=np.int64)
        m = self._f("fast_disc_sample")(C.c_int64(n), _p(pts, _f64p), _p(colptr, _i32p), _p(rowidx, _i32p),
                                        C.c_double(radius), _p(sel, _i64p))
        return sel[:m].copy()

    def fast_disc_sample_bounded(self, points, colptr, rowidx, radius, min_selection, max_selection):
        """(selection, radius used): the sampler with the min/max bounds the reference marks as a todo (orc only)."""
        pts = fpoints(points); n = pts.shape[0]
        colptr, rowidx, _ = self._csc(colptr, rowidx)
        sel = np.empty(max(n, 1), dtype=np.int64)
        used = C.c_double()
        f = self.lib.orc_fast_disc_sample_bounded
        f.restype = C.c_int64
        m = f(C.c_int64(n), _p(pts, _f64p), _p(colptr, _i32p), _p(rowidx, _i32p), C.c_double(radius),
              C.c_int64(min_selection), C.c_int64(max_selection), _p(sel, _i64p), C.byref(used))
        return sel[:m].copy(), used.value

    def assign_parents(self, points, colptr, rowidx, samples):
        pts = fpoints(points); n = pts.shape[0]
        colptr, rowidx, _ = self._csc(colptr, rowidx)
        samples = np.ascontiguousarray(samples, dtype=np.int64)
        par = np.empty(n, dtype=np.int64)
        self._f("assign_parents")(C.c_int64(n), _p(pts, _f64p), _p(colptr, _i32p), _p(rowidx, _i32p),
                                  C.c_int64(samples.shape[0]), _p(samples, _i64p), _p(par, _i64p))
        return par

    def extract_coarse_edges(self, colptr, rowidx, val, m, parents):
        colptr, rowidx, val = self._csc(colptr, rowidx, val)
        n = colptr.shape[0] - 1
        parents = np.ascontiguousarray(parents, dtype=np.int64)
        ocp = np.empty(m + 1, dtype=np.int32)
        ori, ova = _i32p(), _f64p()
        self._f("extract_coarse_edges")(C.c_int64(n), _p(colptr, _i32p), _p(rowidx, _i32p), _p(val, _f64p),
                                        C.c_int64(m), _p(parents, _i64p), _p(ocp, _i32p), C.byref(ori), C.byref(ova))
        nnz = int(ocp[m])
        return ocp, self._take(ori, nnz, np.int32), self._take(ova, nnz, np.float64)

    def coarse_from_mean_of_fine_children(self, points, colptr, rowidx, parents, m):
        pts = fpoints(points); n = pts.shape[0]
        colptr, rowidx, _ = self._csc(colptr, rowidx)
        parents = np.ascontiguousarray(parents, dtype=np.int64)
        out = np.empty((m, 3), dtype=np.float64, order="F")
        self._f("coarse_from_mean_of_fine_children")(C.c_int64(n), _p(pts, _f64p), _p(colptr, _i32p),
                                                     _p(rowidx, _i32p), _p(parents, _i64p), C.c_int64(m),
                                                     _p(out, _f64p))
        return out

    def construct_voronoi_triangles(self, points, colptr, rowidx, val):
        pts = fpoints(points); m = pts.shape[0]
        colptr, rowidx, val = self._csc(colptr, rowidx, val)
        tri, nrm, fan = _i64p(), _f64p(), _i64p()
        fan_ptr = np.empty(m + 1, dtype=np.int64)
        T = self._f("construct_voronoi_triangles")(C.c_int64(m), _p(pts, _f64p), _p(colptr, _i32p), _p(rowidx, _i32p),
                                                   _p(val, _f64p), C.byref(tri), C.byref(nrm), _p(fan_ptr, _i64p),
                                                   C.byref(fan))
        return (self._take(tri, 3 * T, np.int64).reshape(T, 3), self._take(nrm, 3 * T, np.float64).reshape(T, 3),
                fan_ptr, self._take(fan, 3 * T, np.int64))

    def construct_prolongation(self, fine_points, coarse_points, c_colptr, c_rowidx, c_val, parents, weighting=0):
        fp = fpoints(fine_points); cp = fpoints(coarse_points)
        n, m = fp.shape[0], cp.shape[0]
        c_colptr, c_rowidx, c_val = self._csc(c_colptr, c_rowidx, c_val)
        parents = np.ascontiguousarray(parents, dtype=np.int64)
        rowptr = np.empty(n + 1, dtype=np.int32)
        col, val = _i32p(), _f64p()
        rc = self._f("construct_prolongation")(C.c_int64(n), _p(fp, _f64p), C.c_int64(m), _p(cp, _f64p),
                                               _p(c_colptr, _i32p), _p(c_rowidx, _i32p), _p(c_val, _f64p),
                                               _p(parents, _i64p), C.c_int(weighting), _p(rowptr, _i32p),
                                               C.byref(col), C.byref(val))
        nnz = int(rowptr[n])
        return rowptr, self._take(col, nnz, np.int32), self._take(val, nnz, np.float64), rc

    # ---- parts the reference lacks (oracle/solver_oracle.c; only in the plain-C library) ---------------
    def _solver_lib(self):
        if self.kind == "orc":
            return self.lib
        if not hasattr(self, "_orc_lib"):
            if not os.path.exists(_PATHS["orc"]):
                build()
            self._orc_lib = C.CDLL(_PATHS["orc"])
        return self._orc_lib

    def galerkin(self, a_ptr, a_idx, a_val, p_rowptr, p_colidx, p_val, m):
        """A_c = P^T A P by outer vector (structural pattern, sorted inner indices): (ptr, idx, val)."""
        lib = self._solver_lib()
        a_ptr, a_idx, a_val = self._csc(a_ptr, a_idx, a_val)
        p_rowptr, p_colidx, p_val = self._csc(p_rowptr, p_colidx, p_val)
        n = a_ptr.shape[0] - 1
        optr = np.empty(m + 1, dtype=np.int32)
        oi, ov = _i32p(), _f64p()
        lib.orc_galerkin(C.c_int64(n), _p(a_ptr, _i32p), _p(a_idx, _i32p), _p(a_val, _f64p), C.c_int64(m),
                         _p(p_rowptr, _i32p), _p(p_colidx, _i32p), _p(p_val, _f64p), _p(optr, _i32p),
                         C.byref(oi), C.byref(ov))
        nnz = int(optr[m])
        lib.orc_solver_free.argtypes = [C.c_void_p]

        def take(ptr, ty, dt):
            out = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(ty)), shape=(max(nnz, 1),))[:nnz].copy()
            lib.orc_solver_free(C.cast(ptr, C.c_void_p))
            return out.astype(dt, copy=False)
        return optr, take(oi, C.c_int32, np.int32), take(ov, C.c_double, np.float64)

    def spmv(self, outer, inner, val, x):
        lib = self._solver_lib()
        outer, inner, val = self._csc(outer, inner, val)
        x = np.ascontiguousarray(x, dtype=np.float64)
        n = outer.shape[0] - 1
        y = np.empty(n, dtype=np.float64)
        lib.orc_spmv(C.c_int64(n), _p(outer, _i32p), _p(inner, _i32p), _p(val, _f64p), _p(x, _f64p), _p(y, _f64p))
        return y

    def gauss_seidel(self, outer, inner, val, b, x, rows, sweeps=1):
        """Gauss-Seidel sweeps in the row order `rows` (the multicolour kernel's (colour, row) order)."""
        lib = self._solver_lib()
        outer, inner, val = self._csc(outer, inner, val)
        b = np.ascontiguousarray(b, dtype=np.float64)
        x = np.array(x, dtype=np.float64)
        rows = np.ascontiguousarray(rows, dtype=np.int32)
        n = outer.shape[0] - 1
        lib.orc_gauss_seidel(C.c_int64(n), _p(outer, _i32p), _p(inner, _i32p), _p(val, _f64p), _p(b, _f64p),
                             _p(rows, _i32p), C.c_int(sweeps), _p(x, _f64p))
        return x

    def projected_points(self, rowptr, colidx, val, coarse_points):
        cp = fpoints(coarse_points); m = cp.shape[0]
        rowptr = np.ascontiguousarray(rowptr, dtype=np.int32); n = rowptr.shape[0] - 1
        colidx = np.ascontiguousarray(colidx, dtype=np.int32)
        val = np.ascontiguousarray(val, dtype=np.float64)
        out = np.empty((n, 3), dtype=np.float64, order="F")
        self._f("projected_points")(C.c_int64(n), C.c_int64(m), _p(rowptr, _i32p), _p(colidx, _i32p), _p(val, _f64p),
                                    _p(cp, _f64p), _p(out, _f64p))
        return out
