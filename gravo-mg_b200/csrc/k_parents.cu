// Nearest-source (parent) assignment as a frontier-parallel sweep (GravoMG::assignParents,
// reference src/multigrid.cpp:77-125, a multi-source Dijkstra with a (distance, index) min-heap).
//
// What the sequential code computes, and how it is reproduced bit-for-bit in parallel:
//   1. dist[v] = min over source->v paths of the left-to-right sum fl(edge + path) (:107-108).
//      Rounded addition is monotone, so a label-correcting sweep relaxing with the same
//      fl(len + dist[u]) converges to the same bits. Non-negative doubles order like their bit
//      patterns, so the relaxation is one atomicMin on a 64-bit word.
//   2. parents[v] is inherited from the neighbour whose offer FIRST reached dist[v] (strict <, :111).
//      When every edge strictly increases the path length, heap pops happen in lexicographic
//      (dist, index) order, hence pred[v] = argmin_{u : fl(len_uv + dist[u]) == dist[v]} (dist[u], u).
//      parents[v] = parents[pred[v]]; sources are roots. An edge with fl(len + d) == d (duplicate
//      points) makes the pop order dynamic: detected and reported as GMG_ERR_ZERO_LENGTH_EDGE.
//   3. Unreached vertices keep parent 0 (:82).
// "Neighbours of v" are the stored rows of column v; with a symmetric pattern these are also the
// vertices that relax INTO v.
//
// Algorithmic bytes: 40n + 4nnz + 4m (+ 8nnz for the cached lengths).
#include "stages.h"

namespace gmg {

constexpr int kPLanes = 8;
constexpr unsigned long long kInfBits = 0x7FEFFFFFFFFFFFFFull;  // DBL_MAX, src/multigrid.cpp:84

struct ParentArgs {
    int32_t n;
    const int32_t* colptr;
    const int32_t* rowidx;
    const double* len;
    unsigned long long* dist;  // bit patterns of non-negative doubles
    int32_t* stamp;            // round in which a vertex's distance last improved (-1: never, 0: source)
    int32_t* counters;         // [0..2] rotating "something improved" flags; [3] rounds
};

__global__ void __launch_bounds__(kBlock) parents_init_kernel(int32_t n, unsigned long long* __restrict__ dist,
                                                              int32_t* __restrict__ stamp, int32_t* __restrict__ srcid) {
    const int64_t i = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (i < n) { dist[i] = kInfBits; stamp[i] = -1; srcid[i] = -1; }
}

// Later sources overwrite earlier ones exactly like the loop at :89-93 (duplicates: last wins).
__global__ void __launch_bounds__(kBlock) parents_seed_kernel(int32_t m, const int32_t* __restrict__ samples,
                                                              unsigned long long* __restrict__ dist,
                                                              int32_t* __restrict__ srcid, int32_t* __restrict__ stamp) {
    const int64_t cidx = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (cidx >= m) return;
    const int32_t v = samples[cidx];
    dist[v] = 0ull;
    stamp[v] = 0;
    atomicMax(srcid + v, int32_t(cidx));
}

// Round r relaxes out of every vertex whose distance improved in round r-1. The frontier is not a
// queue: every warp reads 32 stamps (coalesced), ballots the active ones and hands them to its four
// 8-lane groups. A queue would put one atomic per improved vertex on a single tail word, which
// serialises in the L2 atomic unit; the stamp scan costs 4 bytes per vertex and round instead.
__global__ void __launch_bounds__(kBlock) parents_sweep_kernel(ParentArgs a) {
    cg::grid_group grid = cg::this_grid();
    const int lane = threadIdx.x & 31;
    const int grp = lane / kPLanes;
    const int sub = lane % kPLanes;
    const int64_t warps = (int64_t(gridDim.x) * kBlock) >> 5;
    const int64_t warp = (int64_t(blockIdx.x) * kBlock + threadIdx.x) >> 5;
    int round = 1;
    for (;; ++round) {
        bool improved = false;
        for (int64_t base = warp * 32; base < a.n; base += warps * 32) {
            const int64_t mine = base + lane;
            const bool active = mine < a.n && *(volatile int32_t*)(a.stamp + mine) == round - 1;
            unsigned mask = __ballot_sync(0xffffffffu, active);
            while (mask) {
                unsigned m = mask;
                for (int i = 0; i < grp; ++i) m &= m - 1;               // the grp-th set bit is ours
                if (m) {
                    const int32_t u = int32_t(base) + (__ffs(m) - 1);
                    const double du = __longlong_as_double((long long)*(volatile unsigned long long*)(a.dist + u));
                    const int32_t b = __ldg(a.colptr + u), e = __ldg(a.colptr + u + 1);
                    for (int32_t k = b + sub; k < e; k += kPLanes) {
                        const int32_t v = __ldg(a.rowidx + k);
                        const double cand = __ldg(a.len + k) + du;                  // edge + path, :108
                        if (!(cand >= 0.0)) continue;                               // NaN never wins a strict <
                        const unsigned long long cb = (unsigned long long)__double_as_longlong(cand);
                        if (cb < *(volatile unsigned long long*)(a.dist + v) && cb < atomicMin(a.dist + v, cb)) {
                            *(volatile int32_t*)(a.stamp + v) = round;              // every writer stores the same value
                            improved = true;
                        }
                    }
                }
#pragma unroll
                for (int i = 0; i < 32 / kPLanes; ++i) mask &= mask - 1;
            }
        }
        if (improved) *(volatile int32_t*)(a.counters + round % 3) = 1;
        grid.sync();
        const bool more = *(volatile int32_t*)(a.counters + round % 3) != 0;
        if (blockIdx.x == 0 && threadIdx.x == 0) a.counters[(round + 2) % 3] = 0;   // flag of round+2; round+1's is live
        if (!more) break;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) a.counters[3] = round;
}

// pred[v] by the (dist, index) rule; also detects edges that do not increase the path length.
__global__ void __launch_bounds__(kBlock) parents_pred_kernel(int32_t lo, int32_t n, const int32_t* __restrict__ colptr,
                                                              const int32_t* __restrict__ rowidx,
                                                              const double* __restrict__ len,
                                                              const unsigned long long* __restrict__ dist,
                                                              const int32_t* __restrict__ srcid,
                                                              int32_t* __restrict__ pred, int32_t* __restrict__ flags) {
    const int64_t gid = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    const int32_t v = lo + int32_t(gid / kPLanes);                  // vertices [lo, n)
    const int sub = int(gid % kPLanes);
    if (v >= n) return;
    const unsigned long long dvb = dist[v];
    const bool is_source = srcid[v] >= 0;
    unsigned long long best_d = ~0ull;
    int32_t best_u = 0x7fffffff;
    const int32_t b = __ldg(colptr + v), e = __ldg(colptr + v + 1);
    for (int32_t k = b + sub; k < e; k += kPLanes) {
        const int32_t u = __ldg(rowidx + k);
        if (u == v) continue;
        const unsigned long long dub = dist[u];
        if (dub == kInfBits) continue;
        const double du = __longlong_as_double((long long)dub);
        const double cand = __ldg(len + k) + du;
        if (!(cand > du)) flags[0] = 1;          // zero-length (or NaN) edge next to a reached vertex
        const unsigned long long cb = (unsigned long long)__double_as_longlong(cand);
        if (cb == dvb && (dub < best_d || (dub == best_d && u < best_u))) { best_d = dub; best_u = u; }
    }
#pragma unroll
    for (int d = kPLanes / 2; d > 0; d >>= 1) {
        const unsigned long long od = __shfl_xor_sync(0xffffffffu, best_d, d, kPLanes);
        const int32_t ou = __shfl_xor_sync(0xffffffffu, best_u, d, kPLanes);
        if (od < best_d || (od == best_d && ou < best_u)) { best_d = od; best_u = ou; }
    }
    if (sub == 0) {
        // sources and unreached vertices are roots; a reached non-source always has an offerer
        if (is_source || dvb == kInfBits) pred[v] = v;
        else if (best_u == 0x7fffffff) { pred[v] = v; flags[1] = 1; }
        else pred[v] = best_u;
    }
}

__global__ void __launch_bounds__(kBlock) parents_resolve_kernel(int32_t lo, int32_t n, int32_t cap, const int32_t* __restrict__ pred,
                                                                 const int32_t* __restrict__ srcid,
                                                                 int32_t* __restrict__ parents) {
    const int64_t i = int64_t(lo) + int64_t(blockIdx.x) * kBlock + threadIdx.x;   // vertices [lo, n)
    if (i >= n) return;
    int32_t u = int32_t(i);
    // dist strictly decreases along pred, so this walk ends at a root; the step cap only matters when a
    // zero-length edge (reported separately) has closed a cycle
    for (int32_t steps = 0; steps < cap; ++steps) {
        const int32_t p = pred[u];
        if (p == u) break;
        u = p;
    }
    const int32_t s = srcid[u];
    parents[i] = s >= 0 ? s : 0;     // unreached: value-initialised 0 (:82)
}

void assign_parents_dev(gmg_ctx* c, int32_t n, const int32_t* colptr, const int32_t* rowidx, const double* len,
                        int32_t m, const int32_t* samples, int32_t* parents, ParentStats* stats,
                        StageStatus* status) {
    if (stats) stats->rounds = 0;
    if (n == 0) return;
    StageStatus local;
    StageStatus* st = status ? status : &local;
    if (!status) local.begin(c);
    int32_t* flags = st->dev.p + StageStatus::kZeroEdge;     // [0] zero-length edge, [1] no offerer
    DevBuf<unsigned long long> dist(c, n);
    DevBuf<int32_t> stamp(c, n), srcid(c, n), counters(c, 4), pred(c, n);
    counters.zero();
    GMG_LAUNCH(c, parents_init_kernel, div_up(n, kBlock), kBlock, 0, n, dist.p, stamp.p, srcid.p);
    if (m > 0) {
        GMG_LAUNCH(c, parents_seed_kernel, div_up(m, kBlock), kBlock, 0, m, samples, dist.p, srcid.p, stamp.p);
        ParentArgs a{n, colptr, rowidx, len, dist.p, stamp.p, counters.p};
        int per_sm = 0;
        GMG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, parents_sweep_kernel, kBlock, 0));
        GMG_REQUIRE(per_sm > 0, GMG_ERR_CUDA, "parent sweep kernel does not fit on an SM");
        int grid = per_sm * c->sm_count;
        const int needed = div_up(n, kBlock);
        if (grid > needed) grid = needed;
        void* args[] = {&a};
        GMG_CUDA(cudaLaunchCooperativeKernel((void*)parents_sweep_kernel, dim3(grid), dim3(kBlock), args, 0, c->stream));
        ++c->launches;
    }
    GMG_LAUNCH(c, parents_pred_kernel, div_up(int64_t(n) * kPLanes, kBlock), kBlock, 0, 0, n, colptr, rowidx, len, dist.p,
               srcid.p, pred.p, flags);
    GMG_LAUNCH(c, parents_resolve_kernel, div_up(n, kBlock), kBlock, 0, 0, n, n, pred.p, srcid.p, parents);
    GMG_CUDA(cudaMemcpyAsync(st->dev.p + StageStatus::kParentRounds, counters.p + 3, sizeof(int32_t),
                             cudaMemcpyDeviceToDevice, c->stream));
    if (status) return;                       // the caller fetches and checks once per level
    local.fetch(c);
    if (stats) stats->rounds = local.host[StageStatus::kParentRounds];
    local.raise_parents();
}

void StageStatus::raise_parents() const {
    GMG_REQUIRE(host[kZeroEdge] == 0, GMG_ERR_ZERO_LENGTH_EDGE,
                "assignParents: an edge does not increase the path length (duplicate points?); "
                "the reference's heap order is then dynamic and is not reproduced");
    GMG_REQUIRE(host[kNoOffer] == 0, GMG_ERR_ASYMMETRIC, "assignParents: a reached vertex has no offering neighbour (asymmetric pattern)");
}

// ---------------------------------------------------------------------------------------------
// Row-sharded variant. A rank may only write the vertices it owns, so the relaxation is a PULL:
// an owned vertex whose neighbour improved in the previous round recomputes
// min(dist[v], min_u fl(len_uv + dist[u])) over the improved neighbours -- the same rounded sums,
// hence the same fixed point, as the push sweep above. dist (uint64 bit patterns) and the
// "improved" bytes are replicated; the caller all-gathers the owned slices after every round.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlock) shard_parents_init_kernel(int32_t n, unsigned long long* __restrict__ dist,
                                                                    uint8_t* __restrict__ changed, int32_t* __restrict__ srcid) {
    const int64_t i = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (i < n) { dist[i] = kInfBits; changed[i] = 0; srcid[i] = -1; }
}
__global__ void __launch_bounds__(kBlock) shard_parents_seed_kernel(int32_t m, const int32_t* __restrict__ samples,
                                                                    unsigned long long* __restrict__ dist,
                                                                    uint8_t* __restrict__ changed, int32_t* __restrict__ srcid) {
    const int64_t cidx = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (cidx >= m) return;
    const int32_t v = samples[cidx];
    dist[v] = 0ull;
    changed[v] = 1;
    atomicMax(srcid + v, int32_t(cidx));
}
__global__ void __launch_bounds__(kBlock) shard_parents_pull_kernel(int32_t lo, int32_t hi, const int32_t* __restrict__ colptr,
                                                                    const int32_t* __restrict__ rowidx,
                                                                    const double* __restrict__ len,
                                                                    unsigned long long* dist,
                                                                    const uint8_t* __restrict__ changed_in,
                                                                    uint8_t* __restrict__ changed_out) {
    const int64_t gid = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    const int32_t v = lo + int32_t(gid / kPLanes);
    const int sub = int(gid % kPLanes);
    if (v >= hi) return;
    const unsigned long long dv = *(volatile unsigned long long*)(dist + v);
    unsigned long long best = dv;
    const int32_t b = __ldg(colptr + v), e = __ldg(colptr + v + 1);
    for (int32_t k = b + sub; k < e; k += kPLanes) {
        const int32_t u = __ldg(rowidx + k);
        if (!changed_in[u]) continue;                               // only an improved neighbour can offer less
        const double cand = __ldg(len + k) + __longlong_as_double((long long)*(volatile unsigned long long*)(dist + u));
        if (!(cand >= 0.0)) continue;
        const unsigned long long cb = (unsigned long long)__double_as_longlong(cand);
        best = cb < best ? cb : best;
    }
#pragma unroll
    for (int d = kPLanes / 2; d > 0; d >>= 1) {
        const unsigned long long o = __shfl_xor_sync(0xffffffffu, best, d, kPLanes);
        best = o < best ? o : best;
    }
    if (sub == 0) {
        if (best < dv) { *(volatile unsigned long long*)(dist + v) = best; changed_out[v] = 1; }
        else changed_out[v] = 0;
    }
}

void shard_parents_init(gmg_ctx* c, int32_t n, int32_t m, const int32_t* samples, unsigned long long* dist,
                        uint8_t* changed, int32_t* srcid) {
    if (n > 0) GMG_LAUNCH(c, shard_parents_init_kernel, div_up(n, kBlock), kBlock, 0, n, dist, changed, srcid);
    if (m > 0) GMG_LAUNCH(c, shard_parents_seed_kernel, div_up(m, kBlock), kBlock, 0, m, samples, dist, changed, srcid);
}
void shard_parents_pull(gmg_ctx* c, int32_t lo, int32_t hi, const int32_t* colptr, const int32_t* rowidx,
                        const double* len, unsigned long long* dist, const uint8_t* changed_in, uint8_t* changed_out) {
    if (hi > lo) GMG_LAUNCH(c, shard_parents_pull_kernel, div_up(int64_t(hi - lo) * kPLanes, kBlock), kBlock, 0, lo, hi,
                            colptr, rowidx, len, dist, changed_in, changed_out);
}
void shard_parents_pred(gmg_ctx* c, int32_t n, int32_t lo, int32_t hi, const int32_t* colptr, const int32_t* rowidx,
                        const double* len, const unsigned long long* dist, const int32_t* srcid, int32_t* pred,
                        int32_t* d_flags) {
    (void)n;
    if (hi > lo) GMG_LAUNCH(c, parents_pred_kernel, div_up(int64_t(hi - lo) * kPLanes, kBlock), kBlock, 0, lo, hi, colptr,
                            rowidx, len, dist, srcid, pred, d_flags);
}
void shard_parents_resolve(gmg_ctx* c, int32_t n, int32_t lo, int32_t hi, const int32_t* pred, const int32_t* srcid,
                           int32_t* parents) {
    if (hi > lo) GMG_LAUNCH(c, parents_resolve_kernel, div_up(hi - lo, kBlock), kBlock, 0, lo, hi, n, pred, srcid, parents);
}

}  // namespace gmg
