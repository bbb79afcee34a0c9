// Streaming utility kernels: exclusive scan, edge lengths (GravoMG::toEdgeDistanceMatrix,
// reference src/utility.cpp:50-56), edge-list flattening (extractEdges, src/utility.cpp:58-71),
// mean edge length (averageEdgeLength, src/multigrid.cpp:127-133) and the pattern-symmetry check.
#include "stages.h"

namespace gmg {

// ---------------------------------------------------------------------------------------------
// Exclusive scan: reduce-then-scan over tiles of 2048 (256 threads x 8 items).
// ---------------------------------------------------------------------------------------------
constexpr int kScanItems = 8;
constexpr int kScanTile = kBlock * kScanItems;

__device__ __forceinline__ int32_t block_exclusive_scan(int32_t v, int32_t& block_total) {
    __shared__ int32_t warp_sums[kBlock / 32];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int32_t incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        int32_t t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += t;
    }
    if (lane == 31) warp_sums[wid] = incl;
    __syncthreads();
    if (wid == 0) {
        int32_t w = lane < kBlock / 32 ? warp_sums[lane] : 0;
#pragma unroll
        for (int d = 1; d < kBlock / 32; d <<= 1) {
            int32_t t = __shfl_up_sync(0xffffffffu, w, d);
            if (lane >= d) w += t;
        }
        if (lane < kBlock / 32) warp_sums[lane] = w;
    }
    __syncthreads();
    const int32_t warp_off = wid ? warp_sums[wid - 1] : 0;
    block_total = warp_sums[kBlock / 32 - 1];
    __syncthreads();
    return warp_off + incl - v;
}

__global__ void __launch_bounds__(kBlock) scan_tile_sums(const int32_t* __restrict__ in, int64_t n,
                                                         int32_t* __restrict__ tile_sums) {
    const int64_t base = int64_t(blockIdx.x) * kScanTile;
    int32_t s = 0;
#pragma unroll
    for (int k = 0; k < kScanItems; ++k) {
        int64_t i = base + k * kBlock + threadIdx.x;
        if (i < n) s += in[i];
    }
    int32_t total;
    block_exclusive_scan(s, total);
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}

// One block scans up to any number of tile sums sequentially in chunks of kBlock.
__global__ void __launch_bounds__(kBlock) scan_small(int32_t* __restrict__ a, int64_t n, int64_t* __restrict__ total) {
    int64_t carry = 0;
    for (int64_t base = 0; base < n; base += kBlock) {
        int64_t i = base + threadIdx.x;
        int32_t v = i < n ? a[i] : 0;
        int32_t bt;
        int32_t ex = block_exclusive_scan(v, bt);
        if (i < n) a[i] = int32_t(carry + ex);
        carry += bt;
    }
    if (threadIdx.x == 0 && total) *total = carry;
}

__global__ void __launch_bounds__(kBlock) scan_apply(const int32_t* __restrict__ in, int32_t* __restrict__ out,
                                                     int64_t n, const int32_t* __restrict__ tile_offsets) {
    // blocked arrangement: thread t owns items [t*8, t*8+8) of the tile, so the scan is in order
    const int64_t base = int64_t(blockIdx.x) * kScanTile + int64_t(threadIdx.x) * kScanItems;
    int32_t v[kScanItems];
    int32_t s = 0;
#pragma unroll
    for (int k = 0; k < kScanItems; ++k) {
        v[k] = (base + k < n) ? in[base + k] : 0;
        s += v[k];
    }
    int32_t bt;
    int32_t ex = block_exclusive_scan(s, bt) + tile_offsets[blockIdx.x];
#pragma unroll
    for (int k = 0; k < kScanItems; ++k) {
        if (base + k < n) out[base + k] = ex;
        ex += v[k];
    }
}

void exclusive_scan_i32(gmg_ctx* c, const int32_t* in, int32_t* out, int64_t n, int64_t* d_total) {
    if (n <= 0) {
        if (d_total) GMG_CUDA(cudaMemsetAsync(d_total, 0, sizeof(int64_t), c->stream));
        return;
    }
    const int tiles = div_up(n, kScanTile);
    DevBuf<int32_t> sums(c, tiles);
    GMG_LAUNCH(c, scan_tile_sums, tiles, kBlock, 0, in, n, sums.p);
    GMG_LAUNCH(c, scan_small, 1, kBlock, 0, sums.p, int64_t(tiles), d_total);
    GMG_LAUNCH(c, scan_apply, tiles, kBlock, 0, in, out, n, sums.p);
}

// ---------------------------------------------------------------------------------------------
// K0 edge lengths: 8 lanes per column walk its stored entries; a warp covers 4 adjacent columns,
// so index reads and length writes are contiguous. Position gathers are the irregular part.
// Algorithmic bytes: 28n + 12nnz.
// ---------------------------------------------------------------------------------------------
constexpr int kGroup = 8;

__global__ void __launch_bounds__(kBlock, 8) edge_length_kernel(int32_t n, const int32_t* __restrict__ colptr,
                                                             const int32_t* __restrict__ rowidx,
                                                             const double* __restrict__ pts,
                                                             double* __restrict__ out_len,
                                                             const uint8_t* __restrict__ col_mask) {
    const int64_t gid = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    const int32_t col = int32_t(gid / kGroup);
    const int sub = int(gid % kGroup);
    if (col >= n) return;
    if (col_mask && !col_mask[col]) return;        // row-sharded build: a column this rank never reads
    const int32_t b = __ldg(colptr + col), e = __ldg(colptr + col + 1);
    const V3 pc = ld3(pts, n, col);
    for (int32_t k = b + sub; k < e; k += kGroup) {
        const int32_t r = __ldg(rowidx + k);
        out_len[k] = norm(ld3(pts, n, r) - pc);   // (points.row(row) - points.row(col)).norm()
    }
}

void edge_lengths(gmg_ctx* c, int32_t n, const int32_t* colptr, const int32_t* rowidx, const double* pts,
                  double* out_len, const uint8_t* col_mask) {
    if (n == 0) return;
    GMG_LAUNCH(c, edge_length_kernel, div_up(int64_t(n) * kGroup, kBlock), kBlock, 0, n, colptr, rowidx, pts, out_len,
               col_mask);
}

// extractEdges: edge e of column c, row r -> edges(e,0) = c, edges(e,1) = r; values copied.
__global__ void __launch_bounds__(kBlock) extract_edges_kernel(int32_t n, const int32_t* __restrict__ colptr,
                                                               const int32_t* __restrict__ rowidx,
                                                               const double* __restrict__ val, int64_t nnz,
                                                               int32_t* __restrict__ out_edges,
                                                               double* __restrict__ out_values) {
    const int64_t gid = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    const int32_t col = int32_t(gid / kGroup);
    const int sub = int(gid % kGroup);
    if (col >= n) return;
    const int32_t b = __ldg(colptr + col), e = __ldg(colptr + col + 1);
    for (int32_t k = b + sub; k < e; k += kGroup) {
        out_edges[k] = col;
        out_edges[nnz + k] = __ldg(rowidx + k);
        out_values[k] = __ldg(val + k);
    }
}

void extract_edges_dev(gmg_ctx* c, int32_t n, const int32_t* colptr, const int32_t* rowidx, const double* val,
                       int64_t nnz, int32_t* out_edges, double* out_values) {
    if (n == 0) return;
    GMG_LAUNCH(c, extract_edges_kernel, div_up(int64_t(n) * kGroup, kBlock), kBlock, 0, n, colptr, rowidx, val, nnz,
               out_edges, out_values);
}

// Mean edge length: fixed-shape deterministic reduction. Each block sums a contiguous slab with a
// fixed tree; block partials are then summed by one block with the same tree.
__device__ __forceinline__ double block_sum(double v) {
    __shared__ double ws[kBlock / 32];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v += __shfl_down_sync(0xffffffffu, v, d);
    if (lane == 0) ws[wid] = v;
    __syncthreads();
    double r = 0.0;
    if (wid == 0) {
        r = lane < kBlock / 32 ? ws[lane] : 0.0;
#pragma unroll
        for (int d = 4; d > 0; d >>= 1) r += __shfl_down_sync(0xffffffffu, r, d);
    }
    __syncthreads();
    return r;  // valid in thread 0
}

constexpr int kSumPerThread = 8;

__global__ void __launch_bounds__(kBlock) edge_list_length_partial(int32_t n, const double* __restrict__ pts,
                                                                   int64_t n_edges, const int32_t* __restrict__ edges,
                                                                   double* __restrict__ partial) {
    const int64_t base = int64_t(blockIdx.x) * kBlock * kSumPerThread;
    double s = 0.0;
#pragma unroll
    for (int k = 0; k < kSumPerThread; ++k) {
        const int64_t e = base + int64_t(k) * kBlock + threadIdx.x;
        if (e < n_edges) s += norm(ld3(pts, n, __ldg(edges + n_edges + e)) - ld3(pts, n, __ldg(edges + e)));
    }
    s = block_sum(s);
    if (threadIdx.x == 0) partial[blockIdx.x] = s;
}

__global__ void __launch_bounds__(kBlock) sum_partials(const double* __restrict__ partial, int64_t count,
                                                       double denom, double* __restrict__ out) {
    double s = 0.0;
    for (int64_t i = threadIdx.x; i < count; i += kBlock) s += partial[i];
    s = block_sum(s);
    if (threadIdx.x == 0) *out = s / denom;
}

void average_edge_length_dev(gmg_ctx* c, int32_t n, const double* pts, int64_t n_edges, const int32_t* edges,
                             double* d_out) {
    const int blocks = n_edges > 0 ? div_up(n_edges, int64_t(kBlock) * kSumPerThread) : 1;
    DevBuf<double> partial(c, blocks);
    if (n_edges > 0) GMG_LAUNCH(c, edge_list_length_partial, blocks, kBlock, 0, n, pts, n_edges, edges, partial.p);
    else partial.zero();
    GMG_LAUNCH(c, sum_partials, 1, kBlock, 0, partial.p, int64_t(blocks), double(n_edges - n), d_out);
}

// Mean stored length inside the hierarchy builders. The summation tree is defined on COLUMN GROUPS so that
// a row-sharded build reproduces it bit for bit: partial[g] = sum of the entries of columns [32g, 32g+32)
// (one warp, lane-strided, fixed shuffle tree); the partials are then folded 2048 at a time by fixed block
// trees until one value is left. Rank boundaries are multiples of 32 columns, so every group has one owner.
__global__ void __launch_bounds__(kBlock) length_group_partial(int32_t n, int32_t g_lo, int32_t g_hi,
                                                               const int32_t* __restrict__ colptr,
                                                               const double* __restrict__ len, double* __restrict__ partial) {
    const int lane = threadIdx.x & 31;
    const int64_t g = int64_t(g_lo) + ((int64_t(blockIdx.x) * kBlock + threadIdx.x) >> 5);
    if (g >= g_hi) return;
    const int64_t c0 = g * 32, c1 = c0 + 32 < n ? c0 + 32 : n;
    const int32_t b = colptr[c0], e = colptr[c1];
    double s = 0.0;
    for (int32_t k = b + lane; k < e; k += 32) s += __ldg(len + k);
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) s += __shfl_down_sync(0xffffffffu, s, d);
    if (lane == 0) partial[g] = s;
}

__global__ void __launch_bounds__(kBlock) fold_partials(const double* __restrict__ in, int64_t count, double* __restrict__ out) {
    const int64_t base = int64_t(blockIdx.x) * kBlock * kSumPerThread;
    double s = 0.0;
#pragma unroll
    for (int k = 0; k < kSumPerThread; ++k) {
        const int64_t e = base + int64_t(k) * kBlock + threadIdx.x;
        if (e < count) s += in[e];
    }
    s = block_sum(s);
    if (threadIdx.x == 0) out[blockIdx.x] = s;
}

void length_group_partials(gmg_ctx* c, int32_t n, int32_t g_lo, int32_t g_hi, const int32_t* colptr, const double* len,
                           double* partial) {
    if (g_hi > g_lo) GMG_LAUNCH(c, length_group_partial, div_up(int64_t(g_hi - g_lo) * 32, kBlock), kBlock, 0, n, g_lo, g_hi,
                                colptr, len, partial);
}

void mean_from_group_partials(gmg_ctx* c, int64_t groups, const double* partial, double denom, double* d_out) {
    const int64_t tile = int64_t(kBlock) * kSumPerThread;
    DevBuf<double> a, b;
    const double* cur = partial;
    int64_t count = groups;
    bool use_a = true;
    while (count > tile) {
        const int64_t blocks = (count + tile - 1) / tile;
        DevBuf<double>& dst = use_a ? a : b;
        dst.alloc(c, blocks);
        GMG_LAUNCH(c, fold_partials, int(blocks), kBlock, 0, cur, count, dst.p);
        cur = dst.p; count = blocks; use_a = !use_a;
    }
    GMG_LAUNCH(c, sum_partials, 1, kBlock, 0, cur, count, denom, d_out);
}

void mean_length_dev(gmg_ctx* c, int32_t n, int64_t nnz, const int32_t* colptr, const double* len, double denom,
                     double* d_out) {
    (void)nnz;
    const int32_t groups = (n + 31) / 32;
    DevBuf<double> partial(c, groups > 0 ? groups : 1);
    length_group_partials(c, n, 0, groups, colptr, len, partial.p);
    mean_from_group_partials(c, groups, partial.p, denom, d_out);
}

// Pattern checks: row indices strictly ascending and in range; (r,c) stored => (c,r) stored.
__global__ void __launch_bounds__(kBlock) check_pattern_kernel(int32_t n, const int32_t* __restrict__ colptr,
                                                               const int32_t* __restrict__ rowidx,
                                                               int32_t* __restrict__ flags) {
    const int64_t gid = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    const int32_t col = int32_t(gid / kGroup);
    const int sub = int(gid % kGroup);
    if (col >= n) return;
    const int32_t b = __ldg(colptr + col), e = __ldg(colptr + col + 1);
    if (e < b) { flags[0] = 1; return; }
    for (int32_t k = b + sub; k < e; k += kGroup) {
        const int32_t r = __ldg(rowidx + k);
        if (r < 0 || r >= n || (k > b && __ldg(rowidx + k - 1) >= r)) { flags[0] = 1; continue; }
        const int32_t rb = __ldg(colptr + r), re = __ldg(colptr + r + 1);
        const int32_t pos = lower_bound_i32(rowidx, rb, re, col);
        if (pos >= re || __ldg(rowidx + pos) != col) flags[1] = 1;
    }
}

// colptr alone, before anything indexes rowidx with it: non-decreasing, starting at 0
__global__ void __launch_bounds__(kBlock) check_colptr_kernel(int32_t n, const int32_t* __restrict__ colptr, int32_t* __restrict__ flag) {
    const int64_t i = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (i < n && (colptr[i] > colptr[i + 1] || colptr[i] < 0)) *flag = 1;
}

// every idx[k] in [0, upper) -- for index arrays that arrive from the caller and are used as addresses
__global__ void __launch_bounds__(kBlock) check_range_kernel(int64_t count, const int32_t* __restrict__ idx, int32_t upper,
                                                             int32_t* __restrict__ flag) {
    const int64_t k = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (k < count && (idx[k] < 0 || idx[k] >= upper)) *flag = 1;
}
void check_index_range(gmg_ctx* c, int64_t count, const int32_t* idx, int64_t upper, const char* what) {
    if (count <= 0) return;
    DevBuf<int32_t> flag(c, 1);
    flag.zero();
    GMG_LAUNCH(c, check_range_kernel, div_up(count, kBlock), kBlock, 0, count, idx, int32_t(upper), flag.p);
    int32_t bad = 0;
    flag.download(&bad, 1);
    sync(c);
    if (bad) throw Error{GMG_ERR_INVALID, std::string(what) + ": index out of range"};
}
// an outer-pointer array on the device: starts at 0 (checked on the host), non-decreasing
void check_outer_pointers(gmg_ctx* c, int64_t n, const int32_t* outer, const char* what) {
    if (n <= 0) return;
    DevBuf<int32_t> flag(c, 1);
    flag.zero();
    GMG_LAUNCH(c, check_colptr_kernel, div_up(n, kBlock), kBlock, 0, int32_t(n), outer, flag.p);
    int32_t bad = 0;
    flag.download(&bad, 1);
    sync(c);
    if (bad) throw Error{GMG_ERR_INVALID, std::string(what) + ": outer pointers must be non-decreasing"};
}

void check_pattern(gmg_ctx* c, int32_t n, const int32_t* colptr, const int32_t* rowidx, bool need_symmetric) {
    if (n == 0) return;
    DevBuf<int32_t> flags(c, 2);
    flags.zero();
    {
        GMG_LAUNCH(c, check_colptr_kernel, div_up(n, kBlock), kBlock, 0, n, colptr, flags.p);
        int32_t bad = 0;
        flags.download(&bad, 1);
        sync(c);
        GMG_REQUIRE(bad == 0, GMG_ERR_INVALID, "edge matrix: colptr must be non-decreasing");
    }
    GMG_LAUNCH(c, check_pattern_kernel, div_up(int64_t(n) * kGroup, kBlock), kBlock, 0, n, colptr, rowidx, flags.p);
    int32_t h[2];
    flags.download(h, 2);
    sync(c);
    GMG_REQUIRE(h[0] == 0, GMG_ERR_INVALID, "edge matrix: row indices must be in range and strictly ascending per column");
    GMG_REQUIRE(!need_symmetric || h[1] == 0, GMG_ERR_ASYMMETRIC, "edge matrix pattern is not symmetric");
}

}  // namespace gmg
