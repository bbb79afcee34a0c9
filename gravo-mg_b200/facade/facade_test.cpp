// Self-check driver for the C++ facade: runs the reference's demo call order (test/main.cpp:68-121)
// through namespace GravoMG exactly as a caller of the reference would write it, on arrays read
// from a file, and dumps every intermediate for tests/test_facade.py to compare with the oracle.
//   facade_test <in.bin> <out.bin>
// File format: a sequence of records [int64 count][int64 element size][payload].
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <thread>

#include <gravomg/multigrid.h>
#include <gravomg/sampling.h>

using Eigen::Index;

template <typename T> std::vector<T> read_rec(FILE* f) {
    int64_t hdr[2];
    if (fread(hdr, sizeof(int64_t), 2, f) != 2 || hdr[1] != int64_t(sizeof(T))) { std::fprintf(stderr, "bad record\n"); std::exit(2); }
    std::vector<T> v(static_cast<size_t>(hdr[0]));
    if (hdr[0] && fread(v.data(), sizeof(T), v.size(), f) != v.size()) { std::fprintf(stderr, "short record\n"); std::exit(2); }
    return v;
}
template <typename T> void write_rec(FILE* f, const T* p, size_t count) {
    int64_t hdr[2] = {int64_t(count), int64_t(sizeof(T))};
    fwrite(hdr, sizeof(int64_t), 2, f);
    if (count) fwrite(p, sizeof(T), count, f);
}
template <typename Mat> void write_sparse(FILE* f, Mat m) {
    m.makeCompressed();
    write_rec(f, m.outerIndexPtr(), size_t(m.outerSize() + 1));
    write_rec(f, m.innerIndexPtr(), size_t(m.nonZeros()));
    write_rec(f, m.valuePtr(), size_t(m.nonZeros()));
}

int main(int argc, char** argv) {
    if (argc < 3) return 2;
    FILE* in = fopen(argv[1], "rb");
    FILE* out = fopen(argv[2], "wb");
    if (!in || !out) return 2;
    auto meta = read_rec<int64_t>(in);           // n, levels, weighting
    auto radii = read_rec<double>(in);
    auto pts = read_rec<double>(in);
    auto colptr = read_rec<int32_t>(in);
    auto rowidx = read_rec<int32_t>(in);
    auto aval = read_rec<double>(in);
    const Index n = meta[0];
    try {
        GravoMG::PointMatrix fine_points(n, 3);
        std::memcpy(fine_points.data(), pts.data(), sizeof(double) * pts.size());
#ifdef GRAVOMG_EIGEN_SHIM_SPARSE
        GravoMG::EdgeMatrix L = GravoMG::EdgeMatrix::fromCompressed(n, n, colptr.data(), rowidx.data(), aval.data());
#else
        GravoMG::EdgeMatrix L = Eigen::Map<const GravoMG::EdgeMatrix>(n, n, Index(rowidx.size()), colptr.data(), rowidx.data(), aval.data());
#endif
        auto fine_edge_matrix = GravoMG::toEdgeDistanceMatrix(L, fine_points);
        {   // the reference's functions are stateless: two threads at once must get what one thread gets
            const double r0 = radii[0] > 0 ? radii[0] : 0.5;
            auto run = [&](std::vector<Index>& par, GravoMG::EdgeMatrix& ce) {
                auto s = GravoMG::fastDiscSample(fine_points, fine_edge_matrix, r0);
                par = GravoMG::assignParents(fine_points, fine_edge_matrix, s);
                ce = GravoMG::extractCoarseEdges(fine_points, fine_edge_matrix, s, par);
            };
            std::vector<Index> p0, p1, p2;
            GravoMG::EdgeMatrix c0, c1, c2;
            run(p0, c0);
            std::thread ta([&] { run(p1, c1); }), tb([&] { run(p2, c2); });
            ta.join(); tb.join();
            auto same = [](const GravoMG::EdgeMatrix& x, const GravoMG::EdgeMatrix& y) {
                if (x.rows() != y.rows() || x.nonZeros() != y.nonZeros()) return false;
                for (Index k = 0; k < x.nonZeros(); ++k)
                    if (x.innerIndexPtr()[k] != y.innerIndexPtr()[k] || x.valuePtr()[k] != y.valuePtr()[k]) return false;
                return true;
            };
            if (p1 != p0 || p2 != p0 || !same(c1, c0) || !same(c2, c0)) {
                std::fprintf(stderr, "facade: concurrent callers disagree with a single caller\n");
                return 2;
            }
        }
        for (int lvl = 0; lvl + 1 < meta[1]; ++lvl) {
            auto [edge_pairs, edge_distances] = GravoMG::extractEdges(fine_edge_matrix);
            const double own_radius = std::cbrt(8.0) * GravoMG::averageEdgeLength(fine_points, edge_pairs);
            const double radius = radii[size_t(lvl)] > 0 ? radii[size_t(lvl)] : own_radius;
            auto samples = GravoMG::fastDiscSample(fine_points, fine_edge_matrix, radius);
            auto parents = GravoMG::assignParents(fine_points, fine_edge_matrix, samples);
            auto coarse_edge_matrix = GravoMG::extractCoarseEdges(fine_points, fine_edge_matrix, samples, parents);
            auto coarse_points = GravoMG::coarseFromMeanOfFineChildren(fine_points, fine_edge_matrix, parents, samples.size());
            auto [triangles, fans] = GravoMG::constructVoronoiTriangles(coarse_points, coarse_edge_matrix);
            auto U = GravoMG::constructProlongation(fine_points, coarse_points, coarse_edge_matrix, parents,
                                                    static_cast<GravoMG::Weighting>(meta[2]));
            auto projected = GravoMG::projectedPoints(U, coarse_points);
            write_rec(out, &own_radius, 1);
            write_rec(out, samples.data(), samples.size());
            write_rec(out, parents.data(), parents.size());
            write_sparse(out, coarse_edge_matrix);
            write_rec(out, coarse_points.data(), size_t(coarse_points.size()));
            std::vector<int64_t> tri; std::vector<double> nrm; std::vector<int64_t> fan_flat, fan_ptr{0};
            for (auto& [t, nn] : triangles) { for (int k = 0; k < 3; ++k) { tri.push_back(t[k]); nrm.push_back(nn[k]); } }
            for (auto& l : fans) { for (size_t t : l) fan_flat.push_back(int64_t(t)); fan_ptr.push_back(int64_t(fan_flat.size())); }
            write_rec(out, tri.data(), tri.size());
            write_rec(out, nrm.data(), nrm.size());
            write_rec(out, fan_ptr.data(), fan_ptr.size());
            write_rec(out, fan_flat.data(), fan_flat.size());
            write_sparse(out, U);
            write_rec(out, projected.data(), size_t(projected.size()));
            fine_points = coarse_points;
            fine_edge_matrix = coarse_edge_matrix;
        }
        // the resident builder through the facade must give the same P on level 0
        GravoMG::HierarchyOptions ho;
        ho.levels = int(meta[1]); ho.weighting = static_cast<GravoMG::Weighting>(meta[2]); ho.galerkin = true;
        GravoMG::PointMatrix p0(n, 3);
        std::memcpy(p0.data(), pts.data(), sizeof(double) * pts.size());
        auto H = GravoMG::buildHierarchy(p0, L, ho);
        int64_t nl = int64_t(H.size());
        write_rec(out, &nl, 1);
        for (auto& lv : H) { write_rec(out, &lv.radius, 1); write_sparse(out, lv.P); write_sparse(out, lv.A_coarse); }
        // the device kNN builder through the facade: written out for the test to compare with the host kd-tree graph
        auto G = GravoMG::knnGraph(p0, 8);
        write_sparse(out, G);
    } catch (const GravoMG::B200Error& e) {
        std::fprintf(stderr, "B200Error %d: %s\n", e.status(), e.what());
        return 1;
    }
    fclose(in); fclose(out);
    return 0;
}
