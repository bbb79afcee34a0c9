// Implementation of the C++ facade (see include/gravomg/b200_facade.h): flattens Eigen objects to
// the raw arrays of include/gravomg_b200.h, calls the CUDA library, rebuilds Eigen results.
// No algorithm runs on the host here except the three tiny helpers the reference also exposes as
// host functions (inTriangle, uniformWeights, inverseDistanceWeights) and scaleMesh.
#include "gravomg/b200_facade.h"

#include <algorithm>
#include <cmath>
#include <atomic>

#include "gravomg_b200.h"

namespace GravoMG {
namespace {

// One context (= one CUDA stream and its retained results) PER CALLING THREAD: the reference's functions are stateless
// and may be called from several threads at once; here a call such as extractCoarseEdges is two C-ABI calls (compute,
// then fetch of the retained result) on one context, so threads must not share one. Created on first use, destroyed
// when the thread ends.
std::atomic<int> g_device{0};
struct ThreadContext {
    gmg_ctx* c = nullptr;
    ~ThreadContext() { if (c) gmg_destroy(c); }
};
thread_local ThreadContext t_ctx;

gmg_ctx* context() {
    if (!t_ctx.c) {
        const int rc = gmg_create(g_device.load(), nullptr, &t_ctx.c);
        if (rc != GMG_OK) throw B200Error(rc, std::string("gmg_create: ") + gmg_last_error(nullptr));
    }
    return t_ctx.c;
}

void check(int rc, const char* where) {
    if (rc != GMG_OK) throw B200Error(rc, std::string(where) + ": " + gmg_last_error(t_ctx.c));
}

// compressed view of a column- or row-major Eigen sparse matrix (copy only when it is uncompressed)
template <typename Mat> struct Compressed {
    Mat copy;
    const Mat* m;
    explicit Compressed(const Mat& in) : m(&in) {
        if (!in.isCompressed()) { copy = in; copy.makeCompressed(); m = &copy; }
    }
    const int* outer() const { return m->outerIndexPtr(); }
    const int* inner() const { return m->innerIndexPtr(); }
    const double* values() const { return m->valuePtr(); }
};

template <typename Mat>
Mat make_sparse(Index rows, Index cols, const std::vector<int>& outer, const std::vector<int>& inner,
                const std::vector<double>& values) {
#ifdef GRAVOMG_EIGEN_SHIM_SPARSE
    return Mat::fromCompressed(rows, cols, outer.data(), inner.data(), values.data());
#else
    return Mat(Eigen::Map<const Mat>(rows, cols, Index(inner.size()), outer.data(), inner.data(), values.data()));
#endif
}

}  // namespace

void setDevice(int device) { g_device.store(device); }     // for contexts created from now on

// reference src/utility.cpp:8-48: translate to the origin, scale by the longest axis, centre
void scaleMesh(Eigen::MatrixXd& V, const Eigen::MatrixXi&, double scaleRatio) {
    const Index n = V.rows(), d = V.cols();
    std::vector<double> lo(static_cast<size_t>(d)), extent(static_cast<size_t>(d));
    double longest = 0.0;
    for (Index c = 0; c < d; ++c) {
        lo[size_t(c)] = V.col(c).minCoeff();
        extent[size_t(c)] = V.col(c).maxCoeff() - lo[size_t(c)];
        longest = c == 0 ? extent[0] : std::max(longest, extent[size_t(c)]);
    }
    const double s = scaleRatio / longest;
    for (Index c = 0; c < d; ++c) {
        for (Index i = 0; i < n; ++i) V(i, c) = (V(i, c) - lo[size_t(c)]) * s;
        const double half = 0.5 * V.col(c).maxCoeff();
        for (Index i = 0; i < n; ++i) V(i, c) = V(i, c) - half;
    }
}

EdgeMatrix toEdgeDistanceMatrix(const EdgeMatrix& matrix, const PointMatrix& points) {
    Compressed<EdgeMatrix> m(matrix);
    const Index n = matrix.outerSize(), nnz = matrix.nonZeros();
    std::vector<double> out(static_cast<size_t>(nnz));
    check(gmg_to_edge_distance_matrix(context(), n, m.outer(), m.inner(), points.data(), points.rows(), out.data()),
          "toEdgeDistanceMatrix");
    std::vector<int> outer(m.outer(), m.outer() + n + 1), inner(m.inner(), m.inner() + nnz);
    return make_sparse<EdgeMatrix>(matrix.rows(), matrix.cols(), outer, inner, out);
}

std::pair<EdgeList, Eigen::VectorXd> extractEdges(const EdgeMatrix& matrix) {
    Compressed<EdgeMatrix> m(matrix);
    const Index nnz = matrix.nonZeros();
    EdgeList edges(nnz, 2);
    Eigen::VectorXd values(nnz);
    check(gmg_extract_edges(context(), matrix.outerSize(), m.outer(), m.inner(), m.values(), edges.data(), values.data()),
          "extractEdges");
    return {edges, values};
}

std::vector<Index> fastDiscSample(const Eigen::MatrixXd& pos, const Eigen::SparseMatrix<double>& edge_matrix,
                                  const double& radius) {
    Compressed<EdgeMatrix> m(edge_matrix);
    std::vector<Index> sel(size_t(std::max<Index>(pos.rows(), 1)));
    int64_t count = 0;
    static_assert(sizeof(Index) == sizeof(int64_t), "Eigen::Index must be 64-bit");
    check(gmg_fast_disc_sample(context(), pos.rows(), pos.data(), pos.rows(), m.outer(), m.inner(), radius,
                               GMG_SAMPLER_EXACT, 0, reinterpret_cast<int64_t*>(sel.data()), &count),
          "fastDiscSample");
    sel.resize(size_t(count));
    return sel;
}

std::vector<Index> fastDiscSample(const Eigen::MatrixXd& pos, const Eigen::SparseMatrix<double>& edge_matrix,
                                  const double& radius, std::size_t min_selection, std::size_t max_selection) {
    Compressed<EdgeMatrix> m(edge_matrix);
    std::vector<Index> sel(size_t(std::max<Index>(pos.rows(), 1)));
    int64_t count = 0;
    const auto clamp = [](std::size_t v) { return int64_t(std::min<std::size_t>(v, std::size_t(1) << 62)); };
    check(gmg_fast_disc_sample_bounded(context(), pos.rows(), pos.data(), pos.rows(), m.outer(), m.inner(), radius,
                                       clamp(min_selection), clamp(max_selection), GMG_SAMPLER_EXACT, 0,
                                       reinterpret_cast<int64_t*>(sel.data()), &count, nullptr),
          "fastDiscSample");
    sel.resize(size_t(count));
    return sel;
}

std::vector<Index> assignParents(const Eigen::MatrixXd& fine_points, const EdgeMatrix& fine_edge_matrix,
                                 const std::vector<Index>& coarse_samples) {
    Compressed<EdgeMatrix> m(fine_edge_matrix);
    std::vector<Index> parents(size_t(fine_points.rows()));
    check(gmg_assign_parents(context(), fine_points.rows(), fine_points.data(), fine_points.rows(), m.outer(), m.inner(),
                             int64_t(coarse_samples.size()), reinterpret_cast<const int64_t*>(coarse_samples.data()),
                             reinterpret_cast<int64_t*>(parents.data())),
          "assignParents");
    return parents;
}

double averageEdgeLength(const PointMatrix& positions, const EdgeList& neighbors) {
    double mean = 0.0;
    check(gmg_average_edge_length(context(), positions.rows(), positions.data(), positions.rows(), neighbors.rows(),
                                  neighbors.data(), &mean),
          "averageEdgeLength");
    return mean;
}

Eigen::SparseMatrix<double> extractCoarseEdges(const PointMatrix& fine_points, const EdgeMatrix& fine_edge_matrix,
                                               const std::vector<Index>& coarse_samples,
                                               const std::vector<Index>& fine_to_nearest_coarse) {
    Compressed<EdgeMatrix> m(fine_edge_matrix);
    const Index mc = Index(coarse_samples.size());   // used for its size only (src/multigrid.cpp:141)
    std::vector<int> colptr(size_t(mc) + 1);
    int64_t nnz = 0;
    check(gmg_extract_coarse_edges(context(), fine_points.rows(), m.outer(), m.inner(), m.values(), mc,
                                   reinterpret_cast<const int64_t*>(fine_to_nearest_coarse.data()), colptr.data(), &nnz),
          "extractCoarseEdges");
    std::vector<int> rowidx(static_cast<size_t>(nnz));
    std::vector<double> val(static_cast<size_t>(nnz));
    check(gmg_fetch_sparse(context(), rowidx.data(), val.data()), "extractCoarseEdges");
    return make_sparse<EdgeMatrix>(mc, mc, colptr, rowidx, val);
}

PointMatrix coarseFromMeanOfFineChildren(const PointMatrix& fine_points, const EdgeMatrix& fine_edge_matrix,
                                         const std::vector<Index>& fine_to_nearest_coarse,
                                         std::size_t num_coarse_points) {
    Compressed<EdgeMatrix> m(fine_edge_matrix);
    PointMatrix out(Index(num_coarse_points), Index(3));
    check(gmg_coarse_from_mean_of_fine_children(context(), fine_points.rows(), fine_points.data(), fine_points.rows(),
                                                m.outer(), m.inner(),
                                                reinterpret_cast<const int64_t*>(fine_to_nearest_coarse.data()),
                                                int64_t(num_coarse_points), out.data(), std::max<Index>(out.rows(), 1)),
          "coarseFromMeanOfFineChildren");
    return out;
}

std::pair<std::vector<TriangleWithNormal>, std::vector<std::vector<size_t>>> constructVoronoiTriangles(
    const PointMatrix& points, const EdgeMatrix& edge_matrix) {
    Compressed<EdgeMatrix> m(edge_matrix);
    const Index mc = points.rows();
    int64_t T = 0;
    std::vector<int64_t> fan_ptr(size_t(mc) + 1);
    check(gmg_construct_voronoi_triangles(context(), mc, points.data(), points.rows(), m.outer(), m.inner(), m.values(),
                                          &T, fan_ptr.data()),
          "constructVoronoiTriangles");
    std::vector<int64_t> tri(static_cast<size_t>(3 * T)), fan(static_cast<size_t>(3 * T));
    std::vector<double> nrm(static_cast<size_t>(3 * T));
    check(gmg_fetch_triangles(context(), tri.data(), nrm.data(), fan.data()), "constructVoronoiTriangles");
    std::pair<std::vector<TriangleWithNormal>, std::vector<std::vector<size_t>>> out;
    out.first.reserve(size_t(T));
    for (int64_t t = 0; t < T; ++t)
        out.first.emplace_back(Triangle{tri[3 * t], tri[3 * t + 1], tri[3 * t + 2]},
                               Normal{nrm[3 * t], nrm[3 * t + 1], nrm[3 * t + 2]});
    out.second.resize(size_t(mc));
    for (Index v = 0; v < mc; ++v) out.second[size_t(v)].assign(fan.begin() + fan_ptr[v], fan.begin() + fan_ptr[v + 1]);
    return out;
}

ProlongationOperator constructProlongation(const PointMatrix& fine_points, const PointMatrix& coarse_points,
                                           const EdgeMatrix& coarse_edge_matrix,
                                           const std::vector<Index>& fine_to_nearest_coarse,
                                           Weighting weighting_scheme) {
    return constructProlongation(fine_points, coarse_points, coarse_edge_matrix, fine_to_nearest_coarse, weighting_scheme, false);
}

PointMatrix coarseFromSamples(const PointMatrix& fine_points, const std::vector<Index>& coarse_samples) {
    PointMatrix out(Index(coarse_samples.size()), Index(3));
    check(gmg_coarse_from_samples(context(), fine_points.rows(), fine_points.data(), fine_points.rows(),
                                  int64_t(coarse_samples.size()), reinterpret_cast<const int64_t*>(coarse_samples.data()),
                                  out.data(), std::max<Index>(out.rows(), 1)),
          "coarseFromSamples");
    return out;
}

ProlongationOperator constructProlongation(const PointMatrix& fine_points, const PointMatrix& coarse_points,
                                           const EdgeMatrix& coarse_edge_matrix,
                                           const std::vector<Index>& fine_to_nearest_coarse,
                                           Weighting weighting_scheme, bool nested) {
    Compressed<EdgeMatrix> m(coarse_edge_matrix);
    const Index n = fine_points.rows(), mc = coarse_points.rows();
    std::vector<int> rowptr(size_t(n) + 1);
    int64_t nnz = 0;
    gmg_prolongation_stats stats{};
    check(gmg_construct_prolongation(context(), n, fine_points.data(), fine_points.rows(), mc, coarse_points.data(),
                                     std::max<Index>(coarse_points.rows(), 1), m.outer(), m.inner(), m.values(),
                                     reinterpret_cast<const int64_t*>(fine_to_nearest_coarse.data()),
                                     int(weighting_scheme) | (nested ? GMG_NESTED : 0), rowptr.data(), &nnz, &stats),
          "constructProlongation");
    std::vector<int> colidx(static_cast<size_t>(nnz));
    std::vector<double> val(static_cast<size_t>(nnz));
    check(gmg_fetch_sparse(context(), colidx.data(), val.data()), "constructProlongation");
    return make_sparse<ProlongationOperator>(n, mc, rowptr, colidx, val);
}

PointMatrix projectedPoints(const ProlongationOperator& weights, const PointMatrix& coarse_points) {
    Compressed<ProlongationOperator> w(weights);
    PointMatrix out(weights.rows(), Index(3));
    check(gmg_projected_points(context(), weights.rows(), coarse_points.rows(), w.outer(), w.inner(), w.values(),
                               coarse_points.data(), std::max<Index>(coarse_points.rows(), 1), out.data(),
                               std::max<Index>(out.rows(), 1)),
          "projectedPoints");
    return out;
}

}  // namespace GravoMG
