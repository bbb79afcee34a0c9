// C++ facade of gravomg-b200: the reference's public API (namespace GravoMG, Eigen types) backed by
// the B200 CUDA library through the C ABI of include/gravomg_b200.h.
//
// A caller of the reference relinks against libgravomg_b200.so + gravomg_facade.cpp and changes
// nothing else. Signatures mirror, entry by entry:
//     reference include/gravomg/utility.h:10-24    types, scaleMesh, toEdgeDistanceMatrix, extractEdges
//     reference include/gravomg/sampling.h:14-18   fastDiscSample
//     reference include/gravomg/multigrid.h:12-65  Weighting, inTriangle, uniformWeights,
//                                                  inverseDistanceWeights, assignParents, averageEdgeLength,
//                                                  extractCoarseEdges, coarseFromMeanOfFineChildren,
//                                                  constructVoronoiTriangles, constructProlongation,
//                                                  projectedPoints
// Differences, all deliberate:
//   * failures surface as GravoMG::B200Error (a std::runtime_error carrying the gmg_status) instead
//     of undefined behaviour or an assert;
//   * sparse results come back compressed (the reference leaves extractCoarseEdges' result in
//     Eigen's uncompressed mode; coefficient access and iteration give identical values);
//   * buildHierarchy (new) keeps all levels on the device.
// Works with real Eigen >= 3.4 or with the repository's test shim (oracle/eigen_shim).
#pragma once

#include <array>
#include <cstddef>
#include <map>
#include <span>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include <Eigen/Eigen>

namespace GravoMG {

using Eigen::Index;
using Point = Eigen::RowVector3d;
using Normal = Eigen::RowVector3d;
using PointMatrix = Eigen::MatrixXd;
using EdgeList = Eigen::MatrixXi;
using EdgeMatrix = Eigen::SparseMatrix<double>;
using Triangle = std::array<Index, 3>;
using TriangleWithNormal = std::pair<Triangle, Normal>;
using ProlongationOperator = Eigen::SparseMatrix<double, Eigen::RowMajor>;

enum class Weighting { BARYCENTRIC = 0, UNIFORM = 1, INVDIST = 2 };

// Thrown when the CUDA library reports a non-zero gmg_status.
class B200Error : public std::runtime_error {
public:
    B200Error(int status, const std::string& what) : std::runtime_error(what), status_(status) {}
    int status() const { return status_; }
private:
    int status_;
};

// Device the facade's process-wide context lives on (default 0). Call before the first API call.
void setDevice(int device);

// ---- utility.h ---------------------------------------------------------------------------------
void scaleMesh(Eigen::MatrixXd& V, const Eigen::MatrixXi& F, double scaleRatio = 1.0);
EdgeMatrix toEdgeDistanceMatrix(const EdgeMatrix& matrix, const PointMatrix& points);
std::pair<EdgeList, Eigen::VectorXd> extractEdges(const EdgeMatrix& matrix);

// ---- sampling.h --------------------------------------------------------------------------------
std::vector<Index> fastDiscSample(const Eigen::MatrixXd& pos, const Eigen::SparseMatrix<double>& edge_matrix,
                                  const double& radius);
// the signature the reference leaves commented out (src/sampling.cpp:11-12): fewer than min_selection samples ->
// the radius is halved and the sampling repeated; more than max_selection -> the loop stops selecting there
std::vector<Index> fastDiscSample(const Eigen::MatrixXd& pos, const Eigen::SparseMatrix<double>& edge_matrix,
                                  const double& radius, std::size_t min_selection, std::size_t max_selection);

// ---- multigrid.h -------------------------------------------------------------------------------
// Host-side helpers, kept for API completeness (the device path has its own copies).
double inTriangle(const Point& p, std::span<Index, 3> tri, const Normal& triNormal, const PointMatrix& pos,
                  Eigen::RowVector3d& bary, std::map<Index, float>& insideEdge);
std::vector<double> uniformWeights(const int& n_points);
std::vector<double> inverseDistanceWeights(const Eigen::MatrixXd& pos, const Eigen::RowVector3d& source,
                                           const std::span<Index>& edges);

std::vector<Index> assignParents(const Eigen::MatrixXd& fine_points, const EdgeMatrix& fine_edge_matrix,
                                 const std::vector<Index>& coarse_samples);

double averageEdgeLength(const PointMatrix& positions, const EdgeList& neighbors);

Eigen::SparseMatrix<double> extractCoarseEdges(const PointMatrix& fine_points, const EdgeMatrix& fine_edge_matrix,
                                               const std::vector<Index>& coarse_samples,
                                               const std::vector<Index>& fine_to_nearest_coarse);

PointMatrix coarseFromMeanOfFineChildren(const PointMatrix& fine_points, const EdgeMatrix& fine_edge_matrix,
                                         const std::vector<Index>& fine_to_nearest_coarse,
                                         std::size_t num_coarse_points);

std::pair<std::vector<TriangleWithNormal>, std::vector<std::vector<size_t>>> constructVoronoiTriangles(
    const PointMatrix& points, const EdgeMatrix& edge_matrix);

ProlongationOperator constructProlongation(const PointMatrix& fine_points, const PointMatrix& coarse_points,
                                           const EdgeMatrix& coarse_edge_matrix,
                                           const std::vector<Index>& fine_to_nearest_coarse,
                                           Weighting weighting_scheme);
// nested mode (the reference's todo at src/multigrid.cpp:291-292): a fine point that coincides with its parent's
// coarse point gets weight 1.0 on it; pair it with coarseFromSamples
ProlongationOperator constructProlongation(const PointMatrix& fine_points, const PointMatrix& coarse_points,
                                           const EdgeMatrix& coarse_edge_matrix,
                                           const std::vector<Index>& fine_to_nearest_coarse,
                                           Weighting weighting_scheme, bool nested);
PointMatrix coarseFromSamples(const PointMatrix& fine_points, const std::vector<Index>& coarse_samples);

PointMatrix projectedPoints(const ProlongationOperator& weights, const PointMatrix& coarse_points);

// ---- new: resident multi-level build [not in reference] --------------------------------------------
struct HierarchyOptions {
    int levels = 4;                 // including the finest
    double ratio = 8.0;             // radius = cbrt(ratio) * averageEdgeLength (test/main.cpp:74)
    Weighting weighting = Weighting::BARYCENTRIC;
    bool parallel_sampler = false;  // false: the reference's index-order greedy set, bit-exact
    bool galerkin = false;          // also A_c = P^T A P per level
    bool nested = false;            // coarse points = sample positions, exact matches get weight 1.0
};
struct HierarchyLevel {
    std::vector<Index> samples, parents;
    PointMatrix coarse_points;
    EdgeMatrix coarse_edge_matrix;
    ProlongationOperator P;
    EdgeMatrix A_coarse;            // empty unless options.galerkin
    double radius = 0.0;
};
// A's pattern is the level-0 graph (test/main.cpp:68-69); its values are used only for the Galerkin operator.
std::vector<HierarchyLevel> buildHierarchy(const PointMatrix& points, const EdgeMatrix& A,
                                           const HierarchyOptions& options = {});

// Level-0 graphs built on the device from raw geometry (gmg_input_from_knn / gmg_input_from_mesh; what the
// reference's driver gets from an external buildPointCloudLaplacian or a mesh loader, test/main.cpp:68).
// knnGraph: exact k nearest neighbours, union-symmetrised, plus the diagonal, all stored values 1.
// cotangentLaplacian: faces is an nf x 3 row-major index array; adjacency + diagonal pattern, cotangent weights
// + mass_shift * lumped mass.
EdgeMatrix knnGraph(const PointMatrix& points, int k);
EdgeMatrix cotangentLaplacian(const PointMatrix& points, const std::vector<int>& faces, double mass_shift = 0.0);

}  // namespace GravoMG
