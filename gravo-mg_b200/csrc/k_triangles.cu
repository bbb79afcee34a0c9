// Voronoi-neighbour triangle-fan discovery (GravoMG::constructVoronoiTriangles, reference
// src/multigrid.cpp:209-263).
//
// The reference emits, for v0 ascending, v1 over column v0 with v1 >= v0 (:225), v2 over the rest
// of the column with v2 >= v0 (:228-232), a triangle when coeff(v1, v2) != 0 (:237). Emission
// order is the triangle id. Here every stored entry (v0, slot a) counts its triangles (binary
// search of row v1 in column v2, value test), an exclusive scan over the entries gives the ids in
// exactly the reference's lexicographic order, and a second pass writes triangles and normals.
// Per-vertex incident lists ("fans") are built by count/scan/fill and sorted by id, and are stored
// pre-rotated: for owner c the fan entry holds the two other vertices in the cyclic order that
// std::rotate(:360) would leave them in.
// Algorithmic bytes: 32m + 12nnz_c + 48T.
#include "stages.h"

namespace gmg {

constexpr int kTLanes = 8;

__device__ __forceinline__ bool has_nonzero(const int32_t* __restrict__ colptr, const int32_t* __restrict__ rowidx,
                                            const double* __restrict__ val, int32_t row, int32_t col) {
    const int32_t b = __ldg(colptr + col), e = __ldg(colptr + col + 1);
    const int32_t pos = lower_bound_i32(rowidx, b, e, row);
    return pos < e && __ldg(rowidx + pos) == row && __ldg(val + pos) != 0;
}

template <bool WRITE>
__global__ void __launch_bounds__(kBlock) triangle_kernel(int32_t m, const double* __restrict__ pts,
                                                          const int32_t* __restrict__ colptr,
                                                          const int32_t* __restrict__ rowidx,
                                                          const double* __restrict__ val,
                                                          int32_t* __restrict__ entry_count,   // !WRITE: out
                                                          const int32_t* __restrict__ entry_off,  // WRITE: in
                                                          int32_t* __restrict__ tri, double* __restrict__ normal,
                                                          int32_t* __restrict__ fan_count,
                                                          const uint8_t* __restrict__ v0_mask) {
    const int64_t gid = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    const int32_t v0 = int32_t(gid / kTLanes);
    const int sub = int(gid % kTLanes);
    if (v0 >= m) return;
    const int32_t b = __ldg(colptr + v0), e = __ldg(colptr + v0 + 1);
    if (v0_mask && !v0_mask[v0]) {                 // row-sharded build: this rank does not need v0's triangles
        if (!WRITE) for (int32_t a = b + sub; a < e; a += kTLanes) entry_count[a] = 0;
        return;
    }
    for (int32_t a = b + sub; a < e; a += kTLanes) {
        const int32_t v1 = __ldg(rowidx + a);
        int32_t count = 0;
        if (v1 >= v0) {                                                          // :225
            int32_t t = WRITE ? entry_off[a] : 0;
            for (int32_t k = a + 1; k < e; ++k) {                                // :228
                const int32_t v2 = __ldg(rowidx + k);
                if (v2 < v0) continue;                                           // :232
                if (!has_nonzero(colptr, rowidx, val, v1, v2)) continue;         // :237
                if (WRITE) {
                    const V3 p0 = ld3(pts, m, v0);
                    const V3 nn = normalized(cross(ld3(pts, m, v1) - p0, ld3(pts, m, v2) - p0));   // :240-242
                    tri[3 * int64_t(t)] = v0; tri[3 * int64_t(t) + 1] = v1; tri[3 * int64_t(t) + 2] = v2;
                    normal[3 * int64_t(t)] = nn.x; normal[3 * int64_t(t) + 1] = nn.y; normal[3 * int64_t(t) + 2] = nn.z;
                    atomicAdd(fan_count + v0, 1); atomicAdd(fan_count + v1, 1); atomicAdd(fan_count + v2, 1);
                    ++t;
                }
                ++count;
            }
        }
        if (!WRITE) entry_count[a] = count;
    }
}

__global__ void __launch_bounds__(kBlock) fan_fill_kernel(int64_t T, const int32_t* __restrict__ tri,
                                                          const int32_t* __restrict__ fan_ptr,
                                                          int32_t* __restrict__ cursor, int32_t* __restrict__ fan_tri) {
    const int64_t t = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (t >= T) return;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const int32_t v = tri[3 * t + k];
        fan_tri[fan_ptr[v] + atomicAdd(cursor + v, 1)] = int32_t(t);
    }
}

// One thread per vertex: sort its triangle ids ascending, then store the rotated other two vertices.
__global__ void __launch_bounds__(kBlock) fan_finish_kernel(int32_t m, const int32_t* __restrict__ tri,
                                                            const int32_t* __restrict__ fan_ptr,
                                                            int32_t* __restrict__ fan_tri, int32_t* __restrict__ o1,
                                                            int32_t* __restrict__ o2) {
    const int64_t v = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (v >= m) return;
    const int32_t b = fan_ptr[v], e = fan_ptr[v + 1];
    for (int32_t i = b + 1; i < e; ++i) {
        const int32_t x = fan_tri[i];
        int32_t j = i - 1;
        while (j >= b && fan_tri[j] > x) { fan_tri[j + 1] = fan_tri[j]; --j; }
        fan_tri[j + 1] = x;
    }
    for (int32_t i = b; i < e; ++i) {
        const int64_t t = fan_tri[i];
        const int32_t a0 = tri[3 * t], a1 = tri[3 * t + 1], a2 = tri[3 * t + 2];
        // rotate left until the owner is first (:360); a degenerate triangle that lists the owner
        // twice stops at the first match, like the reference's while loop
        if (a0 == v) { o1[i] = a1; o2[i] = a2; }
        else if (a1 == v) { o1[i] = a2; o2[i] = a0; }
        else { o1[i] = a0; o2[i] = a1; }
    }
}

void voronoi_triangles(gmg_ctx* c, int32_t m, int64_t nnz, const double* pts, const int32_t* colptr,
                       const int32_t* rowidx, const double* val, Triangles& out, const uint8_t* v0_mask) {
    out.fan_ptr.alloc(c, int64_t(m) + 1);
    out.T = 0;
    if (m == 0 || nnz == 0) {
        out.fan_ptr.zero();
        out.tri.alloc(c, 0); out.normal.alloc(c, 0); out.fan_tri.alloc(c, 0); out.fan_o1.alloc(c, 0); out.fan_o2.alloc(c, 0);
        return;
    }
    const int grid = div_up(int64_t(m) * kTLanes, kBlock);
    DevBuf<int32_t> entry(c, nnz), fan_count(c, int64_t(m) + 1);
    DevBuf<int64_t> total(c, 1);
    GMG_LAUNCH(c, triangle_kernel<false>, grid, kBlock, 0, m, pts, colptr, rowidx, val, entry.p, nullptr, nullptr,
               nullptr, nullptr, v0_mask);
    exclusive_scan_i32(c, entry.p, entry.p, nnz, total.p);
    total.download(&out.T, 1);
    sync(c);
    GMG_REQUIRE(out.T < (int64_t(1) << 31) / 3, GMG_ERR_INVALID, "constructVoronoiTriangles: too many triangles for 32-bit ids");
    out.tri.alloc(c, 3 * out.T); out.normal.alloc(c, 3 * out.T);
    out.fan_tri.alloc(c, 3 * out.T); out.fan_o1.alloc(c, 3 * out.T); out.fan_o2.alloc(c, 3 * out.T);
    fan_count.zero();
    GMG_LAUNCH(c, triangle_kernel<true>, grid, kBlock, 0, m, pts, colptr, rowidx, val, nullptr, entry.p, out.tri.p,
               out.normal.p, fan_count.p, v0_mask);
    exclusive_scan_i32(c, fan_count.p, out.fan_ptr.p, int64_t(m) + 1, nullptr);
    fan_count.zero();
    if (out.T > 0) GMG_LAUNCH(c, fan_fill_kernel, div_up(out.T, kBlock), kBlock, 0, out.T, out.tri.p, out.fan_ptr.p, fan_count.p, out.fan_tri.p);
    GMG_LAUNCH(c, fan_finish_kernel, div_up(m, kBlock), kBlock, 0, m, out.tri.p, out.fan_ptr.p, out.fan_tri.p,
               out.fan_o1.p, out.fan_o2.p);
}

}  // namespace gmg
