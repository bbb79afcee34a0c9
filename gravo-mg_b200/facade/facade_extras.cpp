// Facade, part 2: the three host-side helpers the reference exposes in multigrid.h, and the
// resident multi-level build (buildHierarchy -> gmg_hierarchy_*).
#include "gravomg/b200_facade.h"

#include <algorithm>
#include <cmath>

#include "gravomg_b200.h"

namespace GravoMG {

// reference src/multigrid.cpp:18-55. Projects p onto the triangle's plane, fills the barycentric
// coordinates, maintains the caller's edge map; >= 0 iff the projection is inside.
double inTriangle(const Point& p, std::span<Index, 3> tri, const Normal& n, const PointMatrix& pos,
                  Eigen::RowVector3d& bary, std::map<Index, float>& insideEdge) {
    const Point a = pos.row(tri[0]), b = pos.row(tri[1]), c = pos.row(tri[2]);
    const Point ap = p - a, ab = b - a, ac = c - a;
    const double height = ap.dot(n);
    const Point q = p - height * n;
    const double area2 = ab.cross(ac).dot(n);
    bary(0) = (c - b).cross(q - b).dot(n) / area2;
    bary(1) = (a - c).cross(q - c).dot(n) / area2;
    bary(2) = 1. - bary(0) - bary(1);
    insideEdge.try_emplace(tri[1], float((ap - ap.dot(ab) * ab).norm()));
    insideEdge.try_emplace(tri[2], float((ap - ap.dot(ac) * ac).norm()));
    if (bary(0) < 0. || bary(1) < 0.) insideEdge[tri[1]] = -1.f;
    if (bary(0) < 0. || bary(2) < 0.) insideEdge[tri[2]] = -1.f;
    const bool inside = bary(0) >= 0. && bary(1) >= 0. && bary(2) >= 0.;
    return inside ? std::fabs(height) : -1.;
}

// reference src/multigrid.cpp:57-61
std::vector<double> uniformWeights(const int& n_points) { return std::vector<double>(size_t(n_points), 1. / n_points); }

// reference src/multigrid.cpp:63-75
std::vector<double> inverseDistanceWeights(const Eigen::MatrixXd& pos, const Eigen::RowVector3d& source,
                                           const std::span<Index>& edges) {
    std::vector<double> w(edges.size());
    double total = 0.;
    for (size_t j = 0; j < edges.size(); ++j) {
        const Point d = source - pos.row(edges[j]);
        w[j] = 1. / std::max(1e-8, d.norm());
        total += w[j];
    }
    for (double& x : w) x = x / total;
    return w;
}

namespace {
struct InputHandle { gmg_input* h = nullptr; ~InputHandle() { gmg_input_destroy(h); } };
struct HierarchyHandle { gmg_hierarchy* h = nullptr; ~HierarchyHandle() { gmg_hierarchy_destroy(h); } };

void must(int rc, gmg_ctx* c, const char* where) {
    if (rc != GMG_OK) throw B200Error(rc, std::string(where) + ": " + gmg_last_error(c));
}
template <typename Mat> Mat sparse_from(Index rows, Index cols, const std::vector<int>& outer,
                                        const std::vector<int>& inner, const std::vector<double>& values) {
#ifdef GRAVOMG_EIGEN_SHIM_SPARSE
    return Mat::fromCompressed(rows, cols, outer.data(), inner.data(), values.data());
#else
    return Mat(Eigen::Map<const Mat>(rows, cols, Index(inner.size()), outer.data(), inner.data(), values.data()));
#endif
}
}  // namespace

std::vector<HierarchyLevel> buildHierarchy(const PointMatrix& points, const EdgeMatrix& A_in,
                                           const HierarchyOptions& o) {
    EdgeMatrix A = A_in;
    A.makeCompressed();
    gmg_ctx* c = nullptr;
    must(gmg_create(0, nullptr, &c), nullptr, "gmg_create");
    struct Ctx { gmg_ctx* c; ~Ctx() { gmg_destroy(c); } } guard{c};
    InputHandle in;
    must(gmg_input_create(c, points.rows(), points.data(), points.rows(), A.outerIndexPtr(), A.innerIndexPtr(),
                          o.galerkin ? A.valuePtr() : nullptr, &in.h), c, "gmg_input_create");
    gmg_hierarchy_options opt{};
    opt.levels = o.levels; opt.ratio = o.ratio; opt.weighting = int(o.weighting) | (o.nested ? GMG_NESTED : 0);
    opt.sampler_mode = o.parallel_sampler ? GMG_SAMPLER_PARALLEL : GMG_SAMPLER_EXACT;
    opt.seed = 42; opt.radius_policy = 0; opt.radius_override = nullptr; opt.build_galerkin = o.galerkin ? 1 : 0;
    HierarchyHandle h;
    must(gmg_hierarchy_build(c, in.h, &opt, &h.h), c, "gmg_hierarchy_build");
    std::vector<HierarchyLevel> out(size_t(gmg_hierarchy_num_levels(h.h)));
    for (int l = 0; l < int(out.size()); ++l) {
        gmg_level_info info{};
        must(gmg_hierarchy_level_info(h.h, l, &info), c, "gmg_hierarchy_level_info");
        HierarchyLevel& L = out[size_t(l)];
        L.radius = info.radius;
        L.samples.resize(size_t(info.m)); L.parents.resize(size_t(info.n));
        must(gmg_hierarchy_fetch(c, h.h, l, GMG_H_SAMPLES, L.samples.data()), c, "fetch samples");
        must(gmg_hierarchy_fetch(c, h.h, l, GMG_H_PARENTS, L.parents.data()), c, "fetch parents");
        L.coarse_points = PointMatrix(info.m, Index(3));
        must(gmg_hierarchy_fetch(c, h.h, l, GMG_H_COARSE_POINTS, L.coarse_points.data()), c, "fetch coarse points");
        std::vector<int> outer(static_cast<size_t>(info.m) + 1), inner(static_cast<size_t>(info.nnz_c));
        std::vector<double> val(static_cast<size_t>(info.nnz_c));
        must(gmg_hierarchy_fetch(c, h.h, l, GMG_H_COARSE_COLPTR, outer.data()), c, "fetch");
        must(gmg_hierarchy_fetch(c, h.h, l, GMG_H_COARSE_ROWIDX, inner.data()), c, "fetch");
        must(gmg_hierarchy_fetch(c, h.h, l, GMG_H_COARSE_VAL, val.data()), c, "fetch");
        L.coarse_edge_matrix = sparse_from<EdgeMatrix>(info.m, info.m, outer, inner, val);
        outer.assign(size_t(info.n) + 1, 0); inner.assign(size_t(info.nnz_p), 0); val.assign(size_t(info.nnz_p), 0.);
        must(gmg_hierarchy_fetch(c, h.h, l, GMG_H_P_ROWPTR, outer.data()), c, "fetch");
        must(gmg_hierarchy_fetch(c, h.h, l, GMG_H_P_COLIDX, inner.data()), c, "fetch");
        must(gmg_hierarchy_fetch(c, h.h, l, GMG_H_P_VAL, val.data()), c, "fetch");
        L.P = sparse_from<ProlongationOperator>(info.n, info.m, outer, inner, val);
        if (o.galerkin) {
            outer.assign(size_t(info.m) + 1, 0); inner.assign(size_t(info.nnz_ac), 0); val.assign(size_t(info.nnz_ac), 0.);
            must(gmg_hierarchy_fetch(c, h.h, l, GMG_H_AC_COLPTR, outer.data()), c, "fetch");
            must(gmg_hierarchy_fetch(c, h.h, l, GMG_H_AC_ROWIDX, inner.data()), c, "fetch");
            must(gmg_hierarchy_fetch(c, h.h, l, GMG_H_AC_VAL, val.data()), c, "fetch");
            L.A_coarse = sparse_from<EdgeMatrix>(info.m, info.m, outer, inner, val);
        }
    }
    return out;
}

namespace {
EdgeMatrix graph_from_input(gmg_ctx* c, gmg_input* in, bool with_values) {
    int64_t n = 0, nnz = 0;
    must(gmg_input_size(in, &n, &nnz), c, "gmg_input_size");
    std::vector<int> outer(size_t(n) + 1), inner(static_cast<size_t>(nnz));
    std::vector<double> val(static_cast<size_t>(nnz), 1.0);
    must(gmg_input_fetch(c, in, 1, outer.data()), c, "fetch colptr");
    must(gmg_input_fetch(c, in, 2, inner.data()), c, "fetch rowidx");
    if (with_values) must(gmg_input_fetch(c, in, 3, val.data()), c, "fetch values");
    return sparse_from<EdgeMatrix>(Index(n), Index(n), outer, inner, val);
}
}  // namespace

EdgeMatrix knnGraph(const PointMatrix& points, int k) {
    gmg_ctx* c = nullptr;
    must(gmg_create(0, nullptr, &c), nullptr, "gmg_create");
    struct Ctx { gmg_ctx* c; ~Ctx() { gmg_destroy(c); } } guard{c};
    InputHandle in;
    must(gmg_input_from_knn(c, points.rows(), points.data(), points.rows(), k, &in.h), c, "gmg_input_from_knn");
    return graph_from_input(c, in.h, false);
}

EdgeMatrix cotangentLaplacian(const PointMatrix& points, const std::vector<int>& faces, double mass_shift) {
    gmg_ctx* c = nullptr;
    must(gmg_create(0, nullptr, &c), nullptr, "gmg_create");
    struct Ctx { gmg_ctx* c; ~Ctx() { gmg_destroy(c); } } guard{c};
    InputHandle in;
    must(gmg_input_from_mesh(c, points.rows(), points.data(), points.rows(), int64_t(faces.size() / 3), faces.data(), mass_shift,
                             &in.h), c, "gmg_input_from_mesh");
    return graph_from_input(c, in.h, true);
}

}  // namespace GravoMG
