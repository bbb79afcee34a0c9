// Coarse-point sampling over the vertex adjacency graph (GravoMG::fastDiscSample, reference
// src/sampling.cpp:7-53).
//
// The reference walks vertices in index order; an unvisited vertex is selected and marks as visited
// every vertex it "covers": one hop with length < radius (:35-38) or two hops whose lengths sum to
// < radius (:43-45). On a symmetric pattern "covers" is a symmetric relation (|a-b| and fl(x+y) are
// symmetric bit-for-bit), so the selected set is the lexicographically-first maximal independent
// set of that relation. This file computes exactly that set with a dependency-driven sweep:
//
//   cnt[i]   = number of covering PATHS into i from vertices that precede i     (count kernel)
//   i is IN  as soon as cnt[i] drops to 0 (every preceding coverer is OUT)
//   i is OUT as soon as a preceding coverer is IN
//
// IN vertices mark their later coverees OUT; OUT vertices decrement the counters of their later
// coverees. Both fronts live in append-only queues processed by one persistent cooperative kernel
// (grid.sync between the two half-rounds), so a whole level is one launch. Every vertex's two-hop
// neighbourhood is walked exactly twice. "Precedes" is index order (GMG_SAMPLER_EXACT, the
// reference's set bit-for-bit) or a hashed priority (GMG_SAMPLER_PARALLEL: same rule, O(log n) depth).
//
// Path counting is consistent because paths are symmetric too: i->k->j is enumerated from i when
// counting and as j->k->i when j retires, with the same predicate bits.
//
// Algorithmic bytes: 32n + 4nnz + 4m.
#include "stages.h"

namespace gmg {

constexpr int kLanes = 8;  // lanes cooperating on one vertex (mesh columns hold ~7 entries)

__device__ __forceinline__ uint32_t mix32(uint32_t x, uint32_t seed) {
    x ^= seed;
    x *= 0x9E3779B1u; x ^= x >> 16;
    x *= 0x85EBCA6Bu; x ^= x >> 13;
    x *= 0xC2B2AE35u; x ^= x >> 16;
    return x;
}

template <int MODE> __device__ __forceinline__ bool precedes(int32_t a, int32_t b, uint32_t seed) {
    if (MODE == GMG_SAMPLER_EXACT) return a < b;
    const uint32_t ha = mix32(uint32_t(a), seed), hb = mix32(uint32_t(b), seed);
    return ha < hb || (ha == hb && a < b);
}

enum : int32_t { kUndecided = 0, kIn = 1, kOut = 2 };

struct SamplerArgs {
    int32_t n;
    const int32_t* colptr;
    const int32_t* rowidx;
    const double* len;
    double radius;
    uint32_t seed;
    int32_t* state;
    int32_t* cnt;
    int32_t* inq;       // vertices that became IN, in discovery order
    int32_t* outq;      // vertices that became OUT
    int32_t* counters;  // [0] in_tail, [1] out_tail, [2] rounds
};

// Walk every covering path out of `self` and call f(target) for each path whose target differs
// from self. Lanes of the group split the first hop.
template <typename F>
__device__ __forceinline__ void for_each_covered(const SamplerArgs& a, int32_t self, int sub, F f) {
    const int32_t b = __ldg(a.colptr + self), e = __ldg(a.colptr + self + 1);
    for (int32_t k = b + sub; k < e; k += kLanes) {
        const int32_t j = __ldg(a.rowidx + k);
        const double d = __ldg(a.len + k);
        if (d < a.radius) {                                            // src/sampling.cpp:35
            if (j != self) f(j);
            const int32_t b2 = __ldg(a.colptr + j), e2 = __ldg(a.colptr + j + 1);
            for (int32_t k2 = b2; k2 < e2; ++k2) {                     // :43
                const int32_t t = __ldg(a.rowidx + k2);
                if (d + __ldg(a.len + k2) < a.radius && t != self) f(t);  // :44
            }
        }
    }
}

template <int MODE>
__global__ void __launch_bounds__(kBlock) sampler_count_kernel(SamplerArgs a) {
    const int64_t gid = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    const int32_t i = int32_t(gid / kLanes);
    const int sub = int(gid % kLanes);
    if (i >= a.n) return;
    int32_t c = 0;
    for_each_covered(a, i, sub, [&](int32_t t) { c += precedes<MODE>(t, i, a.seed) ? 1 : 0; });
#pragma unroll
    for (int d = kLanes / 2; d > 0; d >>= 1) c += __shfl_xor_sync(0xffffffffu, c, d, kLanes);
    if (sub == 0) {
        a.cnt[i] = c;
        if (c == 0) {
            a.state[i] = kIn;
            a.inq[warp_append(a.counters + 0)] = i;
        } else {
            a.state[i] = kUndecided;
        }
    }
}

template <int MODE>
__global__ void __launch_bounds__(kBlock) sampler_sweep_kernel(SamplerArgs a) {
    cg::grid_group grid = cg::this_grid();
    const int64_t groups = int64_t(gridDim.x) * (kBlock / kLanes);
    const int64_t group = (int64_t(blockIdx.x) * kBlock + threadIdx.x) / kLanes;
    const int sub = threadIdx.x % kLanes;
    int32_t in_head = 0, out_head = 0;
    int rounds = 0;
    for (;;) {
        // half-round 1: newly IN vertices knock out everything they cover
        const int32_t in_tail = *(volatile int32_t*)(a.counters + 0);
        for (int64_t q = in_head + group; q < in_tail; q += groups) {
            const int32_t j = a.inq[q];
            for_each_covered(a, j, sub, [&](int32_t t) {
                if (atomicCAS(a.state + t, kUndecided, kOut) == kUndecided) a.outq[warp_append(a.counters + 1)] = t;
            });
        }
        in_head = in_tail;
        grid.sync();
        // half-round 2: newly OUT vertices release the vertices that were waiting on them
        const int32_t out_tail = *(volatile int32_t*)(a.counters + 1);
        for (int64_t q = out_head + group; q < out_tail; q += groups) {
            const int32_t j = a.outq[q];
            for_each_covered(a, j, sub, [&](int32_t t) {
                if (precedes<MODE>(j, t, a.seed) && atomicSub(a.cnt + t, 1) == 1) {
                    a.state[t] = kIn;  // every preceding coverer is OUT, so nobody can mark t OUT
                    a.inq[warp_append(a.counters + 0)] = t;
                }
            });
        }
        out_head = out_tail;
        ++rounds;
        grid.sync();
        if (*(volatile int32_t*)(a.counters + 0) == in_head) break;  // no new IN => fixed point
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) a.counters[2] = rounds;
}

__global__ void __launch_bounds__(kBlock) flag_in_kernel(int32_t n, const int32_t* __restrict__ state,
                                                         int32_t* __restrict__ flag) {
    const int64_t i = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (i < n) flag[i] = state[i] == kIn ? 1 : 0;
}
__global__ void __launch_bounds__(kBlock) compact_kernel(int32_t n, const int32_t* __restrict__ state,
                                                         const int32_t* __restrict__ pos, int32_t* __restrict__ out) {
    const int64_t i = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (i < n && state[i] == kIn) out[pos[i]] = int32_t(i);
}

template <int MODE> static void run_sampler(gmg_ctx* c, SamplerArgs a) {
    GMG_LAUNCH(c, sampler_count_kernel<MODE>, div_up(int64_t(a.n) * kLanes, kBlock), kBlock, 0, a);
    int per_sm = 0;
    GMG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, sampler_sweep_kernel<MODE>, kBlock, 0));
    GMG_REQUIRE(per_sm > 0, GMG_ERR_CUDA, "sampler sweep kernel does not fit on an SM");
    int grid = per_sm * c->sm_count;
    const int needed = div_up(int64_t(a.n) * kLanes, kBlock);
    if (grid > needed) grid = needed;
    void* args[] = {&a};
    GMG_CUDA(cudaLaunchCooperativeKernel((void*)sampler_sweep_kernel<MODE>, dim3(grid), dim3(kBlock), args, 0, c->stream));
    ++c->launches;
}

int32_t disc_sample(gmg_ctx* c, int32_t n, const int32_t* colptr, const int32_t* rowidx, const double* len,
                    double radius, int mode, uint64_t seed, DevBuf<int32_t>& selection, SamplerStats* stats) {
    if (n == 0) { selection.alloc(c, 0); if (stats) stats->rounds = 0; return 0; }
    GMG_REQUIRE(mode == GMG_SAMPLER_EXACT || mode == GMG_SAMPLER_PARALLEL, GMG_ERR_INVALID, "unknown sampler mode");
    DevBuf<int32_t> state(c, n), cnt(c, n), inq(c, n), outq(c, n), counters(c, 4);
    counters.zero();
    SamplerArgs a{n, colptr, rowidx, len, radius, uint32_t(seed ^ (seed >> 32)) | 1u,
                  state.p, cnt.p, inq.p, outq.p, counters.p};
    if (mode == GMG_SAMPLER_EXACT) run_sampler<GMG_SAMPLER_EXACT>(c, a);
    else run_sampler<GMG_SAMPLER_PARALLEL>(c, a);
    // ascending compaction of the IN flags (cnt is reused as the flag/offset array)
    DevBuf<int64_t> total(c, 1);
    GMG_LAUNCH(c, flag_in_kernel, div_up(n, kBlock), kBlock, 0, n, state.p, cnt.p);
    exclusive_scan_i32(c, cnt.p, cnt.p, n, total.p);
    int64_t m = 0;
    int32_t h_counters[4];
    total.download(&m, 1);
    counters.download(h_counters, 4);
    sync(c);
    selection.alloc(c, m);
    GMG_LAUNCH(c, compact_kernel, div_up(n, kBlock), kBlock, 0, n, state.p, cnt.p, selection.p);
    if (stats) stats->rounds = h_counters[2];
    return int32_t(m);
}

}  // namespace gmg
