// TEST INFRASTRUCTURE ONLY -- never linked into the product path.
//
// extern "C" raw-array driver over the UNMODIFIED reference entry points (namespace GravoMG,
// /root/reference/include/gravomg/{multigrid,sampling,utility}.h), compiled together with
// /root/reference/src/*.cpp against oracle/eigen_shim. No reference source is copied: this file
// only marshals arrays into the reference's Eigen-typed signatures and back.
#define ORACLE_PFX(name) ref_##name
#include "oracle_api.h"

#include <cstdlib>
#include <cstring>

#include <gravomg/multigrid.h>
#include <gravomg/sampling.h>
#include <gravomg/utility.h>

using Eigen::Index;
using GravoMG::EdgeMatrix;
using GravoMG::PointMatrix;

namespace {

PointMatrix points_in(int64_t n, const double* p) {
    PointMatrix m(n, 3);
    std::memcpy(m.data(), p, sizeof(double) * size_t(3 * n));
    return m;
}
EdgeMatrix edges_in(int64_t n, const int32_t* colptr, const int32_t* rowidx, const double* val) {
    if (val) return EdgeMatrix::fromCompressed(n, n, colptr, rowidx, val);
    std::vector<double> ones(size_t(colptr[n]), 1.0);  // pattern-only callers never read values
    return EdgeMatrix::fromCompressed(n, n, colptr, rowidx, ones.data());
}
template <typename T> T* dup(const T* src, size_t count) {
    T* p = static_cast<T*>(std::malloc(sizeof(T) * (count ? count : 1)));
    if (count) std::memcpy(p, src, sizeof(T) * count);
    return p;
}
template <typename Mat> void sparse_out(Mat m, int32_t* out_outer, int32_t** out_inner, double** out_val) {
    m.makeCompressed();
    std::memcpy(out_outer, m.outerIndexPtr(), sizeof(int32_t) * size_t(m.outerSize() + 1));
    *out_inner = dup(m.innerIndexPtr(), size_t(m.nonZeros()));
    *out_val = dup(m.valuePtr(), size_t(m.nonZeros()));
}

}  // namespace

extern "C" {

void ref_free(void* p) { std::free(p); }

void ref_to_edge_distance_matrix(int64_t n, const int32_t* colptr, const int32_t* rowidx, const double* points,
                                 double* out_val) {
    auto d = GravoMG::toEdgeDistanceMatrix(edges_in(n, colptr, rowidx, nullptr), points_in(n, points));
    d.makeCompressed();
    std::memcpy(out_val, d.valuePtr(), sizeof(double) * size_t(d.nonZeros()));
}

void ref_extract_edges(int64_t n, const int32_t* colptr, const int32_t* rowidx, const double* val,
                       int32_t* out_edges, double* out_values) {
    auto [e, v] = GravoMG::extractEdges(edges_in(n, colptr, rowidx, val));
    std::memcpy(out_edges, e.data(), sizeof(int32_t) * size_t(e.size()));
    std::memcpy(out_values, v.data(), sizeof(double) * size_t(v.size()));
}

double ref_average_edge_length(int64_t n, const double* points, int64_t n_edges, const int32_t* edges) {
    GravoMG::EdgeList e(n_edges, 2);
    std::memcpy(e.data(), edges, sizeof(int32_t) * size_t(2 * n_edges));
    return GravoMG::averageEdgeLength(points_in(n, points), e);
}

int64_t ref_fast_disc_sample(int64_t n, const double* points, const int32_t* colptr, const int32_t* rowidx,
                             double radius, int64_t* selection_out) {
    auto s = GravoMG::fastDiscSample(points_in(n, points), edges_in(n, colptr, rowidx, nullptr), radius);
    for (size_t i = 0; i < s.size(); ++i) selection_out[i] = s[i];
    return int64_t(s.size());
}

void ref_assign_parents(int64_t n, const double* points, const int32_t* colptr, const int32_t* rowidx, int64_t m,
                        const int64_t* samples, int64_t* parents_out) {
    std::vector<Index> s(samples, samples + m);
    auto p = GravoMG::assignParents(points_in(n, points), edges_in(n, colptr, rowidx, nullptr), s);
    for (size_t i = 0; i < p.size(); ++i) parents_out[i] = p[i];
}

void ref_extract_coarse_edges(int64_t n, const int32_t* colptr, const int32_t* rowidx, const double* val, int64_t m,
                              const int64_t* parents, int32_t* out_colptr, int32_t** out_rowidx, double** out_val) {
    std::vector<Index> samples(size_t(m), 0);  // used for its size only (src/multigrid.cpp:141)
    std::vector<Index> par(parents, parents + n);
    PointMatrix dummy(n, 3);  // used for rows() only (src/multigrid.cpp:142)
    sparse_out(GravoMG::extractCoarseEdges(dummy, edges_in(n, colptr, rowidx, val), samples, par), out_colptr,
               out_rowidx, out_val);
}

void ref_coarse_from_mean_of_fine_children(int64_t n, const double* points, const int32_t* colptr,
                                           const int32_t* rowidx, const int64_t* parents, int64_t m,
                                           double* out_points) {
    std::vector<Index> par(parents, parents + n);
    auto c = GravoMG::coarseFromMeanOfFineChildren(points_in(n, points), edges_in(n, colptr, rowidx, nullptr), par,
                                                   size_t(m));
    std::memcpy(out_points, c.data(), sizeof(double) * size_t(3 * m));
}

int64_t ref_construct_voronoi_triangles(int64_t m, const double* points, const int32_t* colptr, const int32_t* rowidx,
                                        const double* val, int64_t** out_tri, double** out_normal,
                                        int64_t* out_fan_ptr, int64_t** out_fan) {
    auto [tris, fans] = GravoMG::constructVoronoiTriangles(points_in(m, points), edges_in(m, colptr, rowidx, val));
    const size_t T = tris.size();
    *out_tri = static_cast<int64_t*>(std::malloc(sizeof(int64_t) * (3 * T + 1)));
    *out_normal = static_cast<double*>(std::malloc(sizeof(double) * (3 * T + 1)));
    *out_fan = static_cast<int64_t*>(std::malloc(sizeof(int64_t) * (3 * T + 1)));
    for (size_t t = 0; t < T; ++t)
        for (int k = 0; k < 3; ++k) {
            (*out_tri)[3 * t + k] = tris[t].first[k];
            (*out_normal)[3 * t + k] = tris[t].second[k];
        }
    int64_t pos = 0;
    for (int64_t v = 0; v < m; ++v) {
        out_fan_ptr[v] = pos;
        for (size_t t : fans[size_t(v)]) (*out_fan)[pos++] = int64_t(t);
    }
    out_fan_ptr[m] = pos;
    return int64_t(T);
}

int ref_construct_prolongation(int64_t n, const double* fine_points, int64_t m, const double* coarse_points,
                               const int32_t* c_colptr, const int32_t* c_rowidx, const double* c_val,
                               const int64_t* parents, int weighting, int32_t* out_rowptr, int32_t** out_colidx,
                               double** out_val) {
    std::vector<Index> par(parents, parents + n);
    sparse_out(GravoMG::constructProlongation(points_in(n, fine_points), points_in(m, coarse_points),
                                              edges_in(m, c_colptr, c_rowidx, c_val), par,
                                              static_cast<GravoMG::Weighting>(weighting)),
               out_rowptr, out_colidx, out_val);
    return 0;
}

void ref_projected_points(int64_t n, int64_t m, const int32_t* rowptr, const int32_t* colidx, const double* val,
                          const double* coarse_points, double* out_points) {
    auto P = GravoMG::ProlongationOperator::fromCompressed(n, m, rowptr, colidx, val);
    auto r = GravoMG::projectedPoints(P, points_in(m, coarse_points));
    std::memcpy(out_points, r.data(), sizeof(double) * size_t(3 * n));
}

}  // extern "C"
