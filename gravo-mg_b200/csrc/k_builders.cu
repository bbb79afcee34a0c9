// Level-0 input builders on the device [not in reference; SURVEY.md 8(f) rank 2: the step BEFORE the hot path.
// The reference's driver gets its graph from an external buildPointCloudLaplacian / a mesh loader
// (test/main.cpp:68); at 2 x 10^7 points a host kd-tree takes minutes while the hierarchy takes milliseconds].
//
//   gmg_input_from_knn      exact k nearest neighbours of a point cloud by grid hashing, union-symmetrised, plus the
//                           diagonal (config C4's graph: gravo-mg_b200/meshgen.py knn_graph is the host statement)
//   gmg_morton_order        permutation that sorts points along the Morton curve (ties by index)
//   gmg_input_from_mesh     cotangent Laplacian (+ mass_shift * lumped mass) of a triangle mesh from its face list
//                           (meshgen.py cotangent_laplacian is the host statement)
//
// kNN. Points are binned into cubic cells of edge h, cells into a table of B >= 2n buckets by a hash of the cell
// coordinates (a dense grid over the bounding box of a surface cloud is almost all empty cells); a counting sort
// lays the points out bucket by bucket (positions packed x,y,z,index per point so that a bucket is one contiguous
// read). A query scans the (2r+1)^3 cells around its own, r = 1 first, and keeps the k best (distance^2, index)
// pairs in registers; everything within r*h of the query lies in the scanned cells, so the answer is exact as soon
// as the k-th distance is <= r*h, else r grows. Two cells of a scan that share a bucket, or foreign cells that
// hash into a scanned bucket, only add candidates (a candidate already in the list is skipped). h is chosen by
// measurement, not by assuming a dimension: bin, look at the mean occupancy of occupied buckets, rescale, twice.
// Ties in distance are broken by index, so the result does not depend on the order inside a bucket.
#include "hierarchy_types.h"

#include <cmath>
#include <cstring>
#include <memory>
#include <utility>

using namespace gmg;

namespace {

constexpr int kMaxK = 32;

struct Grid {
    double ox, oy, oz, inv_h, h;
    uint32_t bucket_mask;
};
__device__ __forceinline__ int32_t cell_of(double x, double o, double inv_h) { return int32_t(floor((x - o) * inv_h)); }
__device__ __forceinline__ uint32_t bucket_of(const Grid& g, int32_t cx, int32_t cy, int32_t cz) {
    uint32_t h = uint32_t(cx) * 0x8DA6B343u ^ uint32_t(cy) * 0xD8163841u ^ uint32_t(cz) * 0xCB1AB31Fu;
    h ^= h >> 15; h *= 0x2C1B3C6Du; h ^= h >> 12;
    return h & g.bucket_mask;
}

// doubles as order-preserving 64-bit keys (for atomicMin / atomicMax)
__device__ __forceinline__ unsigned long long ordered(double v) {
    const unsigned long long b = (unsigned long long)__double_as_longlong(v);
    return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}
inline double unordered(unsigned long long k) {
    const unsigned long long b = (k >> 63) ? (k & 0x7FFFFFFFFFFFFFFFull) : ~k;
    double v;
    memcpy(&v, &b, sizeof v);
    return v;
}

// box[0..2] = min, box[3..5] = max (as ordered keys), box[6] = a coordinate is not finite
__global__ void __launch_bounds__(kBlock) bounds_kernel(int32_t n, const double* __restrict__ pts, unsigned long long* box) {
    const int64_t i = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    unsigned long long lo[3], hi[3];
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        const double x = i < n ? pts[int64_t(d) * n + i] : 0.0;
        if (!isfinite(x)) box[6] = 1ull;                  // NaN compares false with everything: it would slip through the extent
        const unsigned long long key = i < n ? ordered(x) : 0ull;
        lo[d] = i < n ? key : ~0ull;
        hi[d] = key;
#pragma unroll
        for (int s = 16; s > 0; s >>= 1) {
            const unsigned long long a = __shfl_down_sync(0xffffffffu, lo[d], s), b = __shfl_down_sync(0xffffffffu, hi[d], s);
            lo[d] = a < lo[d] ? a : lo[d];
            hi[d] = b > hi[d] ? b : hi[d];
        }
    }
    if ((threadIdx.x & 31) == 0)
#pragma unroll
        for (int d = 0; d < 3; ++d) { atomicMin(box + d, lo[d]); atomicMax(box + 3 + d, hi[d]); }
}

__global__ void __launch_bounds__(kBlock) bin_count_kernel(int32_t n, const double* __restrict__ pts, Grid g, int32_t* __restrict__ count) {
    const int64_t i = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (i >= n) return;
    const uint32_t b = bucket_of(g, cell_of(pts[i], g.ox, g.inv_h), cell_of(pts[n + i], g.oy, g.inv_h), cell_of(pts[2 * int64_t(n) + i], g.oz, g.inv_h));
    atomicAdd(count + b, 1);
}
// stats[0] = occupied buckets, stats[1] = largest bucket
__global__ void __launch_bounds__(kBlock) bin_stats_kernel(int64_t buckets, const int32_t* __restrict__ count, unsigned long long* __restrict__ stats) {
    const int64_t b = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    const int32_t c = b < buckets ? count[b] : 0;
    const unsigned occ = __ballot_sync(0xffffffffu, c > 0);
    const int32_t mx = __reduce_max_sync(0xffffffffu, c);
    if ((threadIdx.x & 31) == 0 && occ) { atomicAdd(stats, (unsigned long long)__popc(occ)); atomicMax(stats + 1, (unsigned long long)mx); }
}
// packed[pos] = (x, y, z, index as the bits of a double)
__global__ void __launch_bounds__(kBlock) bin_fill_kernel(int32_t n, const double* __restrict__ pts, Grid g, const int32_t* __restrict__ start,
                                                          int32_t* __restrict__ cursor, double4* __restrict__ packed) {
    const int64_t i = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (i >= n) return;
    const double x = pts[i], y = pts[n + i], z = pts[2 * int64_t(n) + i];
    const uint32_t b = bucket_of(g, cell_of(x, g.ox, g.inv_h), cell_of(y, g.oy, g.inv_h), cell_of(z, g.oz, g.inv_h));
    const int32_t pos = start[b] + atomicAdd(cursor + b, 1);
    packed[pos] = make_double4(x, y, z, __longlong_as_double((long long)i));
}

// One thread per query, in bucket order (the threads of a warp then scan the same few buckets). out: k neighbour
// indices per point, nearest first, the point itself excluded.
template <int K>
__global__ void __launch_bounds__(128) knn_query_kernel(int32_t n, int32_t k, Grid g, const int32_t* __restrict__ start,
                                                        const double4* __restrict__ packed, int32_t* __restrict__ out,
                                                        int32_t* __restrict__ stats) {
    const int64_t t = int64_t(blockIdx.x) * 128 + threadIdx.x;
    if (t >= n) return;
    const double4 q = packed[t];
    const int32_t self = int32_t(__double_as_longlong(q.w));
    const int32_t cx = cell_of(q.x, g.ox, g.inv_h), cy = cell_of(q.y, g.oy, g.inv_h), cz = cell_of(q.z, g.oz, g.inv_h);
    double bd[K];
    int32_t bi[K];
    for (int ring = 1;; ++ring) {
#pragma unroll
        for (int s = 0; s < K; ++s) { bd[s] = 1e300; bi[s] = 0x7fffffff; }
        const bool everything = ring > 6;            // give up on the grid: every point is a candidate
        const int32_t b_lo = everything ? 0 : -ring, b_hi = everything ? 0 : ring;
        for (int32_t dz = b_lo; dz <= b_hi; ++dz)
            for (int32_t dy = b_lo; dy <= b_hi; ++dy)
                for (int32_t dx = b_lo; dx <= b_hi; ++dx) {
                    int32_t lo, hi;
                    if (everything) { lo = 0; hi = n; }
                    else { const uint32_t b = bucket_of(g, cx + dx, cy + dy, cz + dz); lo = __ldg(start + b); hi = __ldg(start + b + 1); }
                    for (int32_t e = lo; e < hi; ++e) {
                        const double2 pxy = __ldg(reinterpret_cast<const double2*>(packed + e));
                        const double2 pzw = __ldg(reinterpret_cast<const double2*>(packed + e) + 1);
                        const int32_t id = int32_t(__double_as_longlong(pzw.y));
                        const double ex = pxy.x - q.x, ey = pxy.y - q.y, ez = pzw.x - q.z;
                        const double d = (ex * ex + ey * ey) + ez * ez;
                        if (id == self || d > bd[K - 1] || (d == bd[K - 1] && id >= bi[K - 1])) continue;
                        // insert (d, id) into the sorted list; an equal pair is the same point seen through a second bucket
                        bool dup = false;
#pragma unroll
                        for (int s = 0; s < K; ++s) dup = dup || (bi[s] == id);
                        if (dup) continue;
                        double cd = d;
                        int32_t ci = id;
#pragma unroll
                        for (int s = 0; s < K; ++s) {
                            const bool before = cd < bd[s] || (cd == bd[s] && ci < bi[s]);
                            const double td = bd[s]; const int32_t ti = bi[s];
                            if (before) { bd[s] = cd; bi[s] = ci; cd = td; ci = ti; }
                        }
                    }
                }
        const double reach = double(ring) * g.h;
        double kth = 0.0;                              // (the list holds the K best; k <= K of them are asked for)
#pragma unroll
        for (int s = 0; s < K; ++s) kth = s == k - 1 ? bd[s] : kth;
        if (everything || kth <= reach * reach) {
            if (ring > 1) atomicAdd(stats + (everything ? 1 : 0), 1);
            break;
        }
    }
#pragma unroll
    for (int s = 0; s < K; ++s)
        if (s < k) out[int64_t(self) * k + s] = bi[s];
}

// ---- union-symmetrised pattern with the diagonal ------------------------------------------------------------
__global__ void __launch_bounds__(kBlock) reverse_count_kernel(int64_t edges, const int32_t* __restrict__ nb, int32_t* __restrict__ rev) {
    const int64_t e = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (e < edges) atomicAdd(rev + nb[e], 1);
}
// slot of column c: room for the diagonal, its k neighbours and everybody that names it
__global__ void __launch_bounds__(kBlock) slot_size_kernel(int32_t n, int32_t k, const int32_t* __restrict__ rev, int32_t* __restrict__ ub) {
    const int64_t c = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (c <= n) ub[c] = c < n ? 1 + k + rev[c] : 0;
}
__global__ void __launch_bounds__(kBlock) slot_fill_kernel(int32_t n, int32_t k, const int32_t* __restrict__ nb, const int32_t* __restrict__ sptr,
                                                           int32_t* __restrict__ cursor, int32_t* __restrict__ slot) {
    const int64_t c = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (c >= n) return;
    int32_t* mine = slot + sptr[c];
    mine[0] = int32_t(c);
    for (int32_t s = 0; s < k; ++s) {
        const int32_t j = nb[c * k + s];
        mine[1 + s] = j;
        slot[sptr[j] + 1 + k + atomicAdd(cursor + j, 1)] = int32_t(c);
    }
}
// sort the slot, drop duplicates (i names j and j names i), count
__global__ void __launch_bounds__(kBlock) slot_unique_kernel(int32_t n, const int32_t* __restrict__ sptr, int32_t* __restrict__ slot,
                                                             int32_t* __restrict__ count) {
    const int64_t c = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (c > n) return;
    if (c == n) { count[c] = 0; return; }
    int32_t* a = slot + sptr[c];
    const int32_t len = sptr[c + 1] - sptr[c];
    for (int32_t i = 1; i < len; ++i) {
        const int32_t x = a[i];
        int32_t j = i - 1;
        while (j >= 0 && a[j] > x) { a[j + 1] = a[j]; --j; }
        a[j + 1] = x;
    }
    int32_t u = 0;
    for (int32_t i = 0; i < len; ++i)
        if (i == 0 || a[i] != a[u - 1]) a[u++] = a[i];
    count[c] = u;
}
__global__ void __launch_bounds__(kBlock) slot_compact_kernel(int32_t n, const int32_t* __restrict__ sptr, const int32_t* __restrict__ slot,
                                                              const int32_t* __restrict__ colptr, int32_t* __restrict__ rowidx) {
    const int64_t c = (int64_t(blockIdx.x) * kBlock + threadIdx.x) >> 3;
    const int sub = threadIdx.x & 7;
    if (c >= n) return;
    const int32_t src = sptr[c], dst = colptr[c], cnt = colptr[c + 1] - dst;
    for (int32_t s = sub; s < cnt; s += 8) rowidx[dst + s] = slot[src + s];
}

Grid make_grid(const double lo[3], double h, uint32_t bucket_mask) {
    return Grid{lo[0], lo[1], lo[2], 1.0 / h, h, bucket_mask};
}

// pts: packed x|y|z on the device. Fills in->colptr / rowidx / nnz.
void knn_pattern(gmg_ctx* c, gmg_input* in, int32_t k) {
    const int32_t n = in->n;
    GMG_REQUIRE(k >= 1 && k <= kMaxK, GMG_ERR_INVALID, "kNN: k must be in 1..32");
    GMG_REQUIRE(n > k, GMG_ERR_INVALID, "kNN: fewer points than k + 1");
    GMG_REQUIRE((int64_t(n) * (2 * k + 1)) < (int64_t(1) << 31), GMG_ERR_INVALID, "kNN: too many points for 32-bit entry offsets");
    Trace tr(c, "knn");
    // bounding box
    DevBuf<unsigned long long> box(c, 8);
    {
        unsigned long long init[8] = {~0ull, ~0ull, ~0ull, 0ull, 0ull, 0ull, 0ull, 0ull};
        box.upload(init, 8);
        GMG_LAUNCH(c, bounds_kernel, div_up(n, kBlock), kBlock, 0, n, in->pts.p, box.p);
    }
    unsigned long long hbox[8];
    GMG_CUDA(cudaMemcpyAsync(hbox, box.p, sizeof hbox, cudaMemcpyDeviceToHost, c->stream));
    sync(c);
    double lo[3], ext = 0.0;
    for (int d = 0; d < 3; ++d) { lo[d] = unordered(hbox[d]); const double e = unordered(hbox[3 + d]) - lo[d]; ext = e > ext ? e : ext; }
    GMG_REQUIRE(std::isfinite(ext) && hbox[6] == 0, GMG_ERR_INVALID, "kNN: non-finite coordinates");
    if (!(ext > 0.0)) ext = 1.0;
    // buckets and the cell size: start from the volume guess, then let the measured occupancy correct it
    int64_t buckets = 1;
    while (buckets < 2 * int64_t(n)) buckets <<= 1;
    DevBuf<int32_t> count(c, buckets + 1), start(c, buckets + 1);
    DevBuf<unsigned long long> stats(c, 2);
    const double target = k > 2 ? 0.5 * k : 1.0;          // points per occupied cell
    double h = ext * std::cbrt(target / double(n));
    Grid g{};
    for (int pass = 0; pass < 4; ++pass) {
        g = make_grid(lo, h, uint32_t(buckets - 1));
        count.zero(); stats.zero();
        GMG_LAUNCH(c, bin_count_kernel, div_up(n, kBlock), kBlock, 0, n, in->pts.p, g, count.p);
        GMG_LAUNCH(c, bin_stats_kernel, div_up(buckets, kBlock), kBlock, 0, buckets, count.p, stats.p);
        unsigned long long hs[2];
        GMG_CUDA(cudaMemcpyAsync(hs, stats.p, sizeof hs, cudaMemcpyDeviceToHost, c->stream));
        sync(c);
        const double mean = double(n) / double(hs[0] ? hs[0] : 1);
        if (getenv("GMG_TRACE")) fprintf(stderr, "[gmg] knn: h %.6g, %llu occupied buckets, mean %.2f, largest %llu\n", h, hs[0], mean, hs[1]);
        if (pass == 3 || (mean > 0.6 * target && mean < 1.6 * target)) break;
        // occupancy grows like h^d with d the (unknown) dimension of the cloud: assume a surface on the way up
        // (too few per cell) and a volume on the way down, which errs on the side of larger cells
        h *= mean < target ? std::sqrt(target / mean) : std::cbrt(target / mean);
    }
    exclusive_scan_i32(c, count.p, start.p, buckets + 1, nullptr);
    DevBuf<double4> packed(c, n);
    count.zero();
    GMG_LAUNCH(c, bin_fill_kernel, div_up(n, kBlock), kBlock, 0, n, in->pts.p, g, start.p, count.p, packed.p);
    tr.mark("binning");
    DevBuf<int32_t> nb(c, int64_t(n) * k), qstats(c, 2);
    qstats.zero();
    const int grid = div_up(n, 128);
    if (k <= 8) GMG_LAUNCH(c, knn_query_kernel<8>, grid, 128, 0, n, k, g, start.p, packed.p, nb.p, qstats.p);
    else if (k <= 16) GMG_LAUNCH(c, knn_query_kernel<16>, grid, 128, 0, n, k, g, start.p, packed.p, nb.p, qstats.p);
    else GMG_LAUNCH(c, knn_query_kernel<32>, grid, 128, 0, n, k, g, start.p, packed.p, nb.p, qstats.p);
    tr.mark("queries");
    if (getenv("GMG_TRACE")) {
        int32_t hq[2];
        qstats.download(hq, 2);
        sync(c);
        fprintf(stderr, "[gmg] knn: %d queries needed a wider scan, %d fell back to all points\n", hq[0], hq[1]);
    }
    // symmetrise
    DevBuf<int32_t> rev(c, n), ub(c, int64_t(n) + 1), sptr(c, int64_t(n) + 1), ucount(c, int64_t(n) + 1);
    DevBuf<int64_t> total(c, 2);
    rev.zero();
    GMG_LAUNCH(c, reverse_count_kernel, div_up(int64_t(n) * k, kBlock), kBlock, 0, int64_t(n) * k, nb.p, rev.p);
    GMG_LAUNCH(c, slot_size_kernel, div_up(int64_t(n) + 1, kBlock), kBlock, 0, n, k, rev.p, ub.p);
    exclusive_scan_i32(c, ub.p, sptr.p, int64_t(n) + 1, total.p);
    DevBuf<int32_t> slot(c, int64_t(n) * (2 * k + 1));
    rev.zero();
    GMG_LAUNCH(c, slot_fill_kernel, div_up(n, kBlock), kBlock, 0, n, k, nb.p, sptr.p, rev.p, slot.p);
    GMG_LAUNCH(c, slot_unique_kernel, div_up(int64_t(n) + 1, kBlock), kBlock, 0, n, sptr.p, slot.p, ucount.p);
    in->colptr.alloc(c, int64_t(n) + 1);
    exclusive_scan_i32(c, ucount.p, in->colptr.p, int64_t(n) + 1, total.p + 1);
    int64_t ht[2];
    total.download(ht, 2);
    sync(c);
    in->nnz = ht[1];
    in->rowidx.alloc(c, in->nnz);
    GMG_LAUNCH(c, slot_compact_kernel, div_up(int64_t(n) * 8, kBlock), kBlock, 0, n, sptr.p, slot.p, in->colptr.p, in->rowidx.p);
    tr.mark("symmetrise");
}

// ---- Morton order -------------------------------------------------------------------------------------------
// 21 bits per axis, interleaved x (lowest), y, z
__device__ __forceinline__ unsigned long long spread21(unsigned long long v) {
    v &= 0x1FFFFFull;
    v = (v | (v << 32)) & 0x001F00000000FFFFull;
    v = (v | (v << 16)) & 0x001F0000FF0000FFull;
    v = (v | (v << 8)) & 0x100F00F00F00F00Full;
    v = (v | (v << 4)) & 0x10C30C30C30C30C3ull;
    v = (v | (v << 2)) & 0x1249249249249249ull;
    return v;
}
__global__ void __launch_bounds__(kBlock) morton_key_kernel(int32_t n, const double* __restrict__ pts, double ox, double oy, double oz,
                                                            double scale, unsigned long long* __restrict__ key, int32_t* __restrict__ idx) {
    const int64_t i = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (i >= n) return;
    const double q[3] = {(pts[i] - ox) * scale, (pts[n + i] - oy) * scale, (pts[2 * int64_t(n) + i] - oz) * scale};
    unsigned long long code = 0;
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        const double c = fmin(fmax(q[d], 0.0), 2097151.0);
        code |= spread21((unsigned long long)c) << d;
    }
    key[i] = code;
    idx[i] = int32_t(i);
}

// Stable least-significant-digit radix sort of (key, index) pairs, 8 bits per pass. One WARP owns a tile of
// kSortTile consecutive elements in both kernels of a pass: it histograms its tile, a scan over (digit, tile)
// turns the histograms into start offsets, and the same warp then walks its tile 32 elements at a time, in order,
// giving every element the next free place of its digit (lanes with equal digits are ranked by lane with
// __match_any_sync) -- which is what makes the pass stable.
constexpr int kSortTile = 2048;
__global__ void __launch_bounds__(kBlock) radix_hist_kernel(int64_t n, const unsigned long long* __restrict__ key, int shift, int64_t tiles,
                                                            int32_t* __restrict__ hist) {
    __shared__ int32_t s_hist[kBlock / 32][256];
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t tile = int64_t(blockIdx.x) * (kBlock / 32) + w;
    for (int d = lane; d < 256; d += 32) s_hist[w][d] = 0;
    __syncwarp();
    if (tile < tiles) {
        const int64_t lo = tile * kSortTile, hi = lo + kSortTile < n ? lo + kSortTile : n;
        for (int64_t e = lo + lane; e < hi; e += 32) atomicAdd(&s_hist[w][(key[e] >> shift) & 255u], 1);
        __syncwarp();
        for (int d = lane; d < 256; d += 32) hist[int64_t(d) * tiles + tile] = s_hist[w][d];
    }
}
__global__ void __launch_bounds__(kBlock) radix_scatter_kernel(int64_t n, const unsigned long long* __restrict__ key, const int32_t* __restrict__ idx,
                                                               int shift, int64_t tiles, const int32_t* __restrict__ offset,
                                                               unsigned long long* __restrict__ key_out, int32_t* __restrict__ idx_out) {
    __shared__ int32_t s_next[kBlock / 32][256];
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t tile = int64_t(blockIdx.x) * (kBlock / 32) + w;
    if (tile >= tiles) return;
    for (int d = lane; d < 256; d += 32) s_next[w][d] = offset[int64_t(d) * tiles + tile];
    __syncwarp();
    const int64_t lo = tile * kSortTile, hi = lo + kSortTile < n ? lo + kSortTile : n;
    for (int64_t e0 = lo; e0 < hi; e0 += 32) {
        const int64_t e = e0 + lane;
        const bool have = e < hi;
        const unsigned long long kk = have ? key[e] : 0ull;
        const int d = have ? int((kk >> shift) & 255u) : 256 + lane;       // idle lanes match nobody
        const unsigned same = __match_any_sync(0xffffffffu, d);
        if (have) {
            const int32_t pos = s_next[w][d] + __popc(same & ((1u << lane) - 1u));
            key_out[pos] = kk;
            idx_out[pos] = idx[e];
        }
        __syncwarp();
        if (have && (__ffs(same) - 1) == lane) s_next[w][d] += __popc(same);
        __syncwarp();
    }
}
// keys/indices end up sorted in (key, idx); the scratch pair must hold n elements too
void radix_sort_pairs(gmg_ctx* c, int64_t n, DevBuf<unsigned long long>& key, DevBuf<int32_t>& idx, int bits) {
    if (n <= 1) return;
    const int64_t tiles = (n + kSortTile - 1) / kSortTile;
    GMG_REQUIRE(256 * tiles < (int64_t(1) << 31), GMG_ERR_INVALID, "radix sort: too many elements");
    DevBuf<unsigned long long> key2(c, n);
    DevBuf<int32_t> idx2(c, n), hist(c, 256 * tiles), offset(c, 256 * tiles);
    const int grid = div_up(tiles, kBlock / 32);
    for (int shift = 0; shift < bits; shift += 8) {
        GMG_LAUNCH(c, radix_hist_kernel, grid, kBlock, 0, n, key.p, shift, tiles, hist.p);
        exclusive_scan_i32(c, hist.p, offset.p, 256 * tiles, nullptr);
        GMG_LAUNCH(c, radix_scatter_kernel, grid, kBlock, 0, n, key.p, idx.p, shift, tiles, offset.p, key2.p, idx2.p);
        std::swap(key, key2);
        std::swap(idx, idx2);
    }
}

// ---- cotangent Laplacian of a triangle mesh -----------------------------------------------------------------
// faces: nf x 3 vertex indices, row-major
__global__ void __launch_bounds__(kBlock) face_check_count_kernel(int32_t n, int64_t nf, const int32_t* __restrict__ faces,
                                                                  int32_t* __restrict__ deg, int32_t* __restrict__ flags) {
    const int64_t f = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (f >= nf) return;
    const int32_t a = faces[3 * f], b = faces[3 * f + 1], c = faces[3 * f + 2];
    if (a < 0 || b < 0 || c < 0 || a >= n || b >= n || c >= n || a == b || b == c || a == c) { flags[0] = 1; return; }
    atomicAdd(deg + a, 1); atomicAdd(deg + b, 1); atomicAdd(deg + c, 1);
}
__global__ void __launch_bounds__(kBlock) face_fill_kernel(int64_t nf, const int32_t* __restrict__ faces, const int32_t* __restrict__ fptr,
                                                           int32_t* __restrict__ cursor, int32_t* __restrict__ flist) {
    const int64_t f = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (f >= nf) return;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const int32_t v = faces[3 * f + k];
        flist[fptr[v] + atomicAdd(cursor + v, 1)] = int32_t(f);
    }
}
// Row v of L in its slot (2 * faces(v) + 1 entries at most): faces ascending; in face (v, b, c) the angle at c is
// opposite edge (v, b) and the angle at b opposite (v, c); w(v,x) += cot / 2, area(v) += |cross| / 6; then
// L(v,x) = -w(v,x), L(v,v) = sum of w + mass_shift * area (meshgen.py cotangent_laplacian). Sorted by column.
__global__ void __launch_bounds__(kBlock) laplacian_rows_kernel(int32_t n, const double* __restrict__ pts, const int32_t* __restrict__ faces,
                                                                const int32_t* __restrict__ fptr, int32_t* __restrict__ flist,
                                                                double mass_shift, const int32_t* __restrict__ sptr,
                                                                int32_t* __restrict__ sidx, double* __restrict__ sval, int32_t* __restrict__ count) {
    const int64_t v = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (v > n) return;
    if (v == n) { count[v] = 0; return; }
    int32_t* fl = flist + fptr[v];
    const int32_t nfv = fptr[v + 1] - fptr[v];
    for (int32_t i = 1; i < nfv; ++i) {           // the fill's arrival order is not reproducible: sort
        const int32_t x = fl[i];
        int32_t j = i - 1;
        while (j >= 0 && fl[j] > x) { fl[j + 1] = fl[j]; --j; }
        fl[j + 1] = x;
    }
    int32_t* idx = sidx + sptr[v];
    double* val = sval + sptr[v];
    int32_t cnt = 0;
    double area = 0.0;
    const V3 pv = ld3(pts, n, int32_t(v));
    for (int32_t i = 0; i < nfv; ++i) {
        const int32_t f = fl[i];
        const int32_t t[3] = {faces[3 * int64_t(f)], faces[3 * int64_t(f) + 1], faces[3 * int64_t(f) + 2]};
        const int at = t[0] == int32_t(v) ? 0 : (t[1] == int32_t(v) ? 1 : 2);
        const int32_t b = t[(at + 1) % 3], cc = t[(at + 2) % 3];
        const V3 pb = ld3(pts, n, b), pc = ld3(pts, n, cc);
        const V3 ub = pv - pb, wb = pc - pb;      // at b: opposite (v, c)
        const V3 uc = pv - pc, wc = pb - pc;      // at c: opposite (v, b)
        const double cot_b = dot(ub, wb) / fmax(norm(cross(ub, wb)), 1e-300);
        const double cot_c = dot(uc, wc) / fmax(norm(cross(uc, wc)), 1e-300);
        area += norm(cross(pb - pv, pc - pv)) / 6.0;
        const int32_t key[2] = {b, cc};
        const double add[2] = {0.5 * cot_c, 0.5 * cot_b};
#pragma unroll
        for (int e = 0; e < 2; ++e) {
            int32_t s = 0;
            while (s < cnt && idx[s] != key[e]) ++s;
            if (s == cnt) { idx[cnt] = key[e]; val[cnt] = add[e]; ++cnt; }
            else val[s] = val[s] + add[e];
        }
    }
    double diag = mass_shift * area;
    for (int32_t s = 0; s < cnt; ++s) { diag += val[s]; val[s] = -val[s]; }
    idx[cnt] = int32_t(v); val[cnt] = diag; ++cnt;
    for (int32_t i = 1; i < cnt; ++i) {
        const int32_t x = idx[i]; const double y = val[i];
        int32_t j = i - 1;
        while (j >= 0 && idx[j] > x) { idx[j + 1] = idx[j]; val[j + 1] = val[j]; --j; }
        idx[j + 1] = x; val[j + 1] = y;
    }
    count[v] = cnt;
}
__global__ void __launch_bounds__(kBlock) slot_size_faces_kernel(int32_t n, const int32_t* __restrict__ deg, int32_t* __restrict__ ub) {
    const int64_t v = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (v <= n) ub[v] = v < n ? 2 * deg[v] + 1 : 0;
}
__global__ void __launch_bounds__(kBlock) slot_compact_pairs_kernel(int32_t n, const int32_t* __restrict__ sptr, const int32_t* __restrict__ sidx,
                                                                    const double* __restrict__ sval, const int32_t* __restrict__ colptr,
                                                                    int32_t* __restrict__ rowidx, double* __restrict__ val) {
    const int64_t c = (int64_t(blockIdx.x) * kBlock + threadIdx.x) >> 3;
    const int sub = threadIdx.x & 7;
    if (c >= n) return;
    const int32_t src = sptr[c], dst = colptr[c], cnt = colptr[c + 1] - dst;
    for (int32_t s = sub; s < cnt; s += 8) { rowidx[dst + s] = sidx[src + s]; val[dst + s] = sval[src + s]; }
}

template <typename F> int guarded_b(gmg_ctx* c, F&& body) {
    if (!c) return GMG_ERR_STATE;
    try {
        GMG_CUDA(cudaSetDevice(c->device));
        body();
        c->last_error.clear();
        return GMG_OK;
    } catch (const Error& e) {
        c->last_error = e.what;
        abort_sync(c);
        return e.status;
    } catch (const std::exception& e) {
        c->last_error = e.what();
        abort_sync(c);
        return GMG_ERR_INVALID;
    }
}

}  // namespace

extern "C" {

int gmg_input_from_knn(gmg_ctx* c, int64_t n, const double* points, int64_t ld, int32_t k, gmg_input** out) {
    if (!out) return GMG_ERR_INVALID;
    *out = nullptr;
    return guarded_b(c, [&] {
        GMG_REQUIRE(points && n > 0 && n < (int64_t(1) << 31) && ld >= n, GMG_ERR_INVALID, "kNN: bad point array");
        std::unique_ptr<gmg_input> in(new gmg_input());
        in->ctx = c; in->n = int32_t(n);
        in->pts = points_to_device(c, points, n, ld);
        knn_pattern(c, in.get(), k);
        sync(c);
        *out = in.release();
    });
}

int gmg_morton_order(gmg_ctx* c, int64_t n, const double* points, int64_t ld, int32_t* perm_out) {
    return guarded_b(c, [&] {
        GMG_REQUIRE(points && perm_out && n > 0 && n < (int64_t(1) << 31) && ld >= n, GMG_ERR_INVALID, "Morton order: bad arguments");
        DevBuf<double> pts = points_to_device(c, points, n, ld);
        DevBuf<unsigned long long> box(c, 8);
        unsigned long long init[8] = {~0ull, ~0ull, ~0ull, 0ull, 0ull, 0ull, 0ull, 0ull}, hbox[8];
        box.upload(init, 8);
        GMG_LAUNCH(c, bounds_kernel, div_up(n, kBlock), kBlock, 0, int32_t(n), pts.p, box.p);
        GMG_CUDA(cudaMemcpyAsync(hbox, box.p, sizeof hbox, cudaMemcpyDeviceToHost, c->stream));
        sync(c);
        double lo[3], ext = 0.0;
        for (int d = 0; d < 3; ++d) { lo[d] = unordered(hbox[d]); const double e = unordered(hbox[3 + d]) - lo[d]; ext = e > ext ? e : ext; }
        GMG_REQUIRE(std::isfinite(ext) && hbox[6] == 0, GMG_ERR_INVALID, "Morton order: non-finite coordinates");
        const double scale = ext > 0.0 ? 2097151.0 / ext : 0.0;
        DevBuf<unsigned long long> key(c, n);
        DevBuf<int32_t> idx(c, n);
        GMG_LAUNCH(c, morton_key_kernel, div_up(n, kBlock), kBlock, 0, int32_t(n), pts.p, lo[0], lo[1], lo[2], scale, key.p, idx.p);
        radix_sort_pairs(c, n, key, idx, 64);
        GMG_CUDA(cudaMemcpyAsync(perm_out, idx.p, sizeof(int32_t) * size_t(n), cudaMemcpyDeviceToHost, c->stream));
        sync(c);
    });
}

int gmg_input_from_mesh(gmg_ctx* c, int64_t n, const double* points, int64_t ld, int64_t nf, const int32_t* faces, double mass_shift,
                        gmg_input** out) {
    if (!out) return GMG_ERR_INVALID;
    *out = nullptr;
    return guarded_b(c, [&] {
        GMG_REQUIRE(points && faces && n > 0 && n < (int64_t(1) << 31) && ld >= n && nf > 0, GMG_ERR_INVALID, "mesh input: bad arguments");
        GMG_REQUIRE(6 * nf + n < (int64_t(1) << 31), GMG_ERR_INVALID, "mesh input: too many faces for 32-bit entry offsets");
        std::unique_ptr<gmg_input> in(new gmg_input());
        in->ctx = c; in->n = int32_t(n);
        in->pts = points_to_device(c, points, n, ld);
        DevBuf<int32_t> F = to_device(c, faces, 3 * nf);
        DevBuf<int32_t> deg(c, n + 1), fptr(c, n + 1), flags(c, 1), ub(c, n + 1), sptr(c, n + 1), count(c, n + 1);
        DevBuf<int64_t> total(c, 2);
        deg.zero(); flags.zero();
        GMG_LAUNCH(c, face_check_count_kernel, div_up(nf, kBlock), kBlock, 0, int32_t(n), nf, F.p, deg.p, flags.p);
        int32_t bad = 0;
        flags.download(&bad, 1);
        sync(c);
        GMG_REQUIRE(bad == 0, GMG_ERR_INVALID, "mesh input: a face has an index out of range or a repeated vertex");
        exclusive_scan_i32(c, deg.p, fptr.p, n + 1, nullptr);
        GMG_LAUNCH(c, slot_size_faces_kernel, div_up(n + 1, kBlock), kBlock, 0, int32_t(n), deg.p, ub.p);
        exclusive_scan_i32(c, ub.p, sptr.p, n + 1, total.p);
        DevBuf<int32_t> flist(c, 3 * nf), sidx(c, 6 * nf + n);
        DevBuf<double> sval(c, 6 * nf + n);
        deg.zero();
        GMG_LAUNCH(c, face_fill_kernel, div_up(nf, kBlock), kBlock, 0, nf, F.p, fptr.p, deg.p, flist.p);
        GMG_LAUNCH(c, laplacian_rows_kernel, div_up(n + 1, kBlock), kBlock, 0, int32_t(n), in->pts.p, F.p, fptr.p, flist.p, mass_shift,
                   sptr.p, sidx.p, sval.p, count.p);
        in->colptr.alloc(c, n + 1);
        exclusive_scan_i32(c, count.p, in->colptr.p, n + 1, total.p + 1);
        int64_t ht[2];
        total.download(ht, 2);
        sync(c);
        in->nnz = ht[1];
        in->rowidx.alloc(c, in->nnz);
        in->a_val.alloc(c, in->nnz);
        GMG_LAUNCH(c, slot_compact_pairs_kernel, div_up(n * 8, kBlock), kBlock, 0, int32_t(n), sptr.p, sidx.p, sval.p, in->colptr.p,
                   in->rowidx.p, in->a_val.p);
        in->has_a = true;
        sync(c);
        *out = in.release();
    });
}

}  // extern "C"
