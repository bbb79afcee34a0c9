// Coarse-point sampling over the vertex adjacency graph (GravoMG::fastDiscSample, reference
// src/sampling.cpp:7-53).
//
// The reference walks vertices in index order; an unvisited vertex is selected and marks as visited
// every vertex it "covers": one hop with length < radius (:35-38) or two hops whose lengths sum to
// < radius (:43-45). On a symmetric pattern "covers" is a symmetric relation (|a-b| and fl(x+y) are
// symmetric bit-for-bit), so the selected set is the lexicographically-first maximal independent
// set of that relation. This file computes exactly that set as a dependency-driven DATAFLOW, the
// way a synchronisation-free sparse triangular solve works:
//
//   cnt[i]   = number of covering PATHS into i from vertices that precede i
//   i is IN  as soon as cnt[i] drops to 0 (every preceding coverer is OUT)
//   i is OUT as soon as a preceding coverer is IN
//
// Kernel 1 (build) walks every vertex's two-hop neighbourhood ONCE (the only irregular pass),
// counts the preceding paths and writes the targets of the LATER paths into a contiguous per-vertex
// list (staged in shared memory, block-aggregated bump allocation). Only later coverees are ever
// needed: an IN vertex knocks its later coverees OUT, an OUT vertex decrements their counters.
// Kernel 2 (flow) has no grid-wide barrier and no queue: lanes take TICKETS in precedence order
// (one atomicAdd per warp), ticket r owns vertex unrank(r), waits until that vertex is decided
// (poll with back-off on its own state word), then applies its list. A vertex is decided by the
// processing of vertices that precede it, i.e. of EARLIER tickets, whose owners are running, so
// the wait always ends and blocks that are not resident yet hold no tickets. The work per vertex
// is O(list length) whatever the dependency depth (the index order of a row-major 10000 x 1000
// grid has chains of 24,000 steps).
// "Precedes" is index order (GMG_SAMPLER_EXACT, the reference's set bit-for-bit) or the order of a
// seeded BIJECTIVE hash of the index on k = ceil(log2 n) bits (GMG_SAMPLER_PARALLEL: same rule,
// O(log n) depth); the hash is invertible so that tickets can be handed out in its order.
//
// Path counting is consistent because paths are symmetric too: i->k->j is enumerated from i when
// counting and as j->k->i when j's list is written, with the same predicate bits.
//
// Algorithmic bytes: 32n + 4nnz + 4m.
#include "stages.h"

namespace gmg {

constexpr int kLanes = 8;    // lanes cooperating on one vertex in the build kernel (mesh columns hold ~7 entries)
constexpr int kStage = 16;   // later-path targets a lane can stage in shared memory

// Invertible mixing of k-bit integers: xor-seed, then (odd multiply, xorshift by ceil(k/2)) twice and a
// last odd multiply, all modulo 2^k. Every step is a bijection on k bits; the xorshift is its own
// inverse because twice the shift is >= k.
struct Order {
    uint32_t mask, shift, seed;
    uint32_t c1, c2, c3, i1, i2, i3;   // odd multipliers and their inverses modulo 2^32
};
__host__ __device__ __forceinline__ uint32_t order_rank(const Order& o, uint32_t v) {
    uint32_t x = (v ^ o.seed) & o.mask;
    x = (x * o.c1) & o.mask; x ^= x >> o.shift;
    x = (x * o.c2) & o.mask; x ^= x >> o.shift;
    return (x * o.c3) & o.mask;
}
__host__ __device__ __forceinline__ uint32_t order_unrank(const Order& o, uint32_t r) {
    uint32_t x = (r * o.i3) & o.mask;
    x ^= x >> o.shift; x = (x * o.i2) & o.mask;
    x ^= x >> o.shift; x = (x * o.i1) & o.mask;
    return (x ^ o.seed) & o.mask;
}
static uint32_t inverse_odd(uint32_t c) {       // Newton iteration for c^-1 modulo 2^32
    uint32_t x = c;
    for (int i = 0; i < 5; ++i) x *= 2u - c * x;
    return x;
}
static Order make_order(int32_t n, uint64_t seed) {
    Order o{};
    int k = 1;
    while ((int64_t(1) << k) < int64_t(n)) ++k;
    o.mask = k >= 32 ? 0xFFFFFFFFu : ((1u << k) - 1u);
    o.shift = uint32_t((k + 1) / 2);
    const uint64_t z = (seed + 0x9E3779B97F4A7C15ull) * 0xBF58476D1CE4E5B9ull;
    o.seed = uint32_t(z >> 32) & o.mask;
    o.c1 = 0x9E3779B1u; o.c2 = 0x85EBCA6Bu | 1u; o.c3 = 0xC2B2AE35u | 1u;
    o.i1 = inverse_odd(o.c1); o.i2 = inverse_odd(o.c2); o.i3 = inverse_odd(o.c3);
    return o;
}

enum : int32_t { kUndecided = 0, kIn = 1, kOut = 2 };

struct SamplerArgs {
    int32_t n;
    const int32_t* colptr;
    const int32_t* rowidx;
    const double* len;
    double radius;
    Order order;
    int32_t* state;
    int32_t* cnt;
    long long* lptr;             // start of the later-coveree list of every vertex
    int32_t* lcnt;               // its length (paths, duplicates kept)
    int32_t* lst;                // the lists
    unsigned long long* cursor;  // bump allocator over lst
    unsigned int* ticket;        // next ticket of the flow kernel
    unsigned int tickets;        // n (exact) or 2^k (parallel)
};

template <int MODE> __device__ __forceinline__ uint32_t rank_of(const SamplerArgs& a, int32_t v) {
    return MODE == GMG_SAMPLER_EXACT ? uint32_t(v) : order_rank(a.order, uint32_t(v));
}

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
            // four entries per step, loads issued before any is consumed: the kernel waits on memory
            // latency (long-scoreboard stalls), so what counts is how many loads a lane has in flight
            constexpr int kUnroll = 4;
            for (int32_t k2 = b2; k2 < e2; k2 += kUnroll) {            // :43
                int32_t t[kUnroll];
                double l[kUnroll];
#pragma unroll
                for (int u = 0; u < kUnroll; ++u) {
                    const bool in = k2 + u < e2;
                    t[u] = in ? __ldg(a.rowidx + k2 + u) : self;
                    l[u] = in ? __ldg(a.len + k2 + u) : 0.0;
                }
#pragma unroll
                for (int u = 0; u < kUnroll; ++u)
                    if (d + l[u] < a.radius && t[u] != self) f(t[u]);  // :44
            }
        }
    }
}

template <int MODE>
__global__ void __launch_bounds__(kBlock) sampler_build_kernel(SamplerArgs a) {
    static_assert(kBlock / kLanes == 32, "one warp scans the group totals");
    __shared__ int32_t s_group[kBlock / kLanes];
    __shared__ unsigned long long s_base;
    __shared__ int32_t s_stage[kBlock][kStage + 1];   // +1: conflict-free lane-private rows
    const int64_t gid = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    const int32_t i = int32_t(gid / kLanes);
    const int sub = threadIdx.x % kLanes;
    const int g = threadIdx.x / kLanes;
    const bool active = i < a.n;
    int32_t prev = 0, later = 0;
    if (active) {
        const uint32_t ri = rank_of<MODE>(a, i);
        int32_t* stage = s_stage[threadIdx.x];
        for_each_covered(a, i, sub, [&](int32_t t) {
            if (rank_of<MODE>(a, t) < ri) ++prev;
            else { if (later < kStage) stage[later] = t; ++later; }
        });
    }
    const bool overflow = __any_sync(0xffffffffu, later > kStage);   // per warp; re-walk is rare
    const unsigned group_over = __ballot_sync(0xffffffffu, later > kStage) & (0xFFu << ((threadIdx.x & 31) & ~(kLanes - 1)));
#pragma unroll
    for (int d = kLanes / 2; d > 0; d >>= 1) prev += __shfl_xor_sync(0xffffffffu, prev, d, kLanes);
    int32_t incl = later;
#pragma unroll
    for (int d = 1; d < kLanes; d <<= 1) {
        const int32_t t = __shfl_up_sync(0xffffffffu, incl, d, kLanes);
        if (sub >= d) incl += t;
    }
    const int32_t group_total = __shfl_sync(0xffffffffu, incl, kLanes - 1, kLanes);
    if (sub == 0) s_group[g] = group_total;
    __syncthreads();
    if (threadIdx.x < 32) {
        const int32_t v = s_group[threadIdx.x];
        int32_t s = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int32_t t = __shfl_up_sync(0xffffffffu, s, d);
            if (int(threadIdx.x) >= d) s += t;
        }
        s_group[threadIdx.x] = s - v;
        if (threadIdx.x == 31) s_base = atomicAdd(a.cursor, (unsigned long long)s);
    }
    __syncthreads();
    if (!active) return;
    const long long start = (long long)s_base + s_group[g];
    int32_t* dst = a.lst + start + (incl - later);
    if (!(overflow && group_over)) {
        const int32_t* stage = s_stage[threadIdx.x];
        for (int32_t w = 0; w < later; ++w) dst[w] = stage[w];
    } else {                                          // some lane of this vertex outgrew its stage: walk again
        const uint32_t ri = rank_of<MODE>(a, i);
        int32_t w = 0;
        for_each_covered(a, i, sub, [&](int32_t t) { if (!(rank_of<MODE>(a, t) < ri)) dst[w++] = t; });
    }
    if (sub == 0) {
        a.lptr[i] = start;
        a.lcnt[i] = group_total;
        a.cnt[i] = prev;
        a.state[i] = prev == 0 ? kIn : kUndecided;
    }
}

template <int MODE>
__global__ void __launch_bounds__(kBlock) sampler_flow_kernel(SamplerArgs a) {
    const int lane = threadIdx.x & 31;
    for (;;) {
        unsigned int base = 0;
        if (lane == 0) base = atomicAdd(a.ticket, 32u);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base >= a.tickets) break;
        // The warp polls its 32 state words TOGETHER (one coalesced request per poll in index order):
        // with per-lane polling, 300k waiting lanes saturate the L2 request rate on deep dependency chains
        // and slow the very atomics they are waiting for.
        const unsigned int r = base + lane;
        uint32_t v = 0;
        bool pending = r < a.tickets;
        if (pending) { v = MODE == GMG_SAMPLER_EXACT ? r : order_unrank(a.order, r); pending = v < uint32_t(a.n); }
        long long p = 0;
        int32_t c = 0;
        if (pending) { p = a.lptr[v]; c = a.lcnt[v]; }
        unsigned ns = 0;
        while (__any_sync(0xffffffffu, pending)) {
            const int32_t st = pending ? *(volatile int32_t*)(a.state + v) : kUndecided;
            if (pending && st != kUndecided) {
                pending = false;
                if (st == kIn) {
                    // a selected vertex knocks out everything it covers (src/sampling.cpp:38,45); later
                    // coverees of an IN vertex are undecided or already OUT, never IN
                    for (int32_t k = 0; k < c; k += 4) {
                        int32_t t[4];
#pragma unroll
                        for (int u = 0; u < 4; ++u) t[u] = k + u < c ? a.lst[p + k + u] : -1;
#pragma unroll
                        for (int u = 0; u < 4; ++u) if (t[u] >= 0) *(volatile int32_t*)(a.state + t[u]) = kOut;
                    }
                } else {
                    for (int32_t k = 0; k < c; k += 4) {
                        int32_t t[4], old[4];
#pragma unroll
                        for (int u = 0; u < 4; ++u) t[u] = k + u < c ? a.lst[p + k + u] : -1;
#pragma unroll
                        for (int u = 0; u < 4; ++u) old[u] = t[u] >= 0 ? atomicSub(a.cnt + t[u], 1) : 0;
                        // every preceding coverer of t is OUT, so nobody can mark t OUT: t is selected (:25-28)
#pragma unroll
                        for (int u = 0; u < 4; ++u) if (old[u] == 1) *(volatile int32_t*)(a.state + t[u]) = kIn;
                    }
                }
                ns = 0;
            } else if (ns) {
                __nanosleep(ns);
            }
            ns = ns == 0 ? 32 : (ns < 512 ? ns << 1 : ns);
        }
    }
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

// sum over columns of deg^2: with nnz it bounds the number of covering paths
__global__ void __launch_bounds__(kBlock) degree_square_kernel(int32_t n, const int32_t* __restrict__ colptr,
                                                               unsigned long long* __restrict__ out) {
    const int64_t i = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    unsigned long long v = 0;
    if (i < n) { const unsigned long long d = (unsigned long long)(colptr[i + 1] - colptr[i]); v = d * d; }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v += __shfl_down_sync(0xffffffffu, v, d);
    if ((threadIdx.x & 31) == 0 && v) atomicAdd(out, v);
}

void degree_square_sum(gmg_ctx* c, int32_t n, const int32_t* colptr, unsigned long long* d_out) {
    if (n > 0) GMG_LAUNCH(c, degree_square_kernel, div_up(n, kBlock), kBlock, 0, n, colptr, d_out);
}

template <int MODE> static void run_sampler(gmg_ctx* c, SamplerArgs a) {
    GMG_LAUNCH(c, sampler_build_kernel<MODE>, div_up(int64_t(a.n) * kLanes, kBlock), kBlock, 0, a);
    int per_sm = 0;
    GMG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, sampler_flow_kernel<MODE>, kBlock, 0));
    GMG_REQUIRE(per_sm > 0, GMG_ERR_CUDA, "sampler flow kernel does not fit on an SM");
    int grid = per_sm * c->sm_count;
    const int needed = div_up(int64_t(a.tickets), kBlock);
    if (grid > needed) grid = needed;
    GMG_LAUNCH(c, sampler_flow_kernel<MODE>, grid < 1 ? 1 : grid, kBlock, 0, a);
}

int32_t disc_sample(gmg_ctx* c, int32_t n, const int32_t* colptr, const int32_t* rowidx, const double* len,
                    double radius, int mode, uint64_t seed, DevBuf<int32_t>& selection, SamplerStats* stats,
                    int64_t path_bound) {
    if (stats) stats->rounds = 0;
    if (n == 0) { selection.alloc(c, 0); return 0; }
    GMG_REQUIRE(mode == GMG_SAMPLER_EXACT || mode == GMG_SAMPLER_PARALLEL, GMG_ERR_INVALID, "unknown sampler mode");
    // bound on the later-path lists: every path has a mirror image, exactly one of the two is "later"
    DevBuf<unsigned long long> cursor(c, 3);          // [0] list cursor, [1] sum deg^2, [2] ticket counter
    cursor.zero();
    if (path_bound < 0) {
        degree_square_sum(c, n, colptr, cursor.p + 1);
        unsigned long long hsq[2];
        int32_t nnz = 0;
        cursor.download(hsq, 2);
        GMG_CUDA(cudaMemcpyAsync(&nnz, colptr + n, sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
        sync(c);
        path_bound = int64_t(hsq[1]) + nnz;
    }
    const int64_t list_cap = path_bound / 2 + 1;
    DevBuf<int32_t> state(c, n), cnt(c, n), lcnt(c, n), lst(c, list_cap);
    DevBuf<long long> lptr(c, n);
    const Order order = make_order(n, seed);
    const unsigned int tickets = mode == GMG_SAMPLER_EXACT ? unsigned(n) : (order.mask == 0xFFFFFFFFu ? 0xFFFFFFFFu : order.mask + 1u);
    SamplerArgs a{n, colptr, rowidx, len, radius, order, state.p, cnt.p, lptr.p, lcnt.p, lst.p, cursor.p,
                  reinterpret_cast<unsigned int*>(cursor.p + 2), tickets};
    if (mode == GMG_SAMPLER_EXACT) run_sampler<GMG_SAMPLER_EXACT>(c, a);
    else run_sampler<GMG_SAMPLER_PARALLEL>(c, a);
    // ascending compaction of the IN flags (cnt is reused as the flag/offset array)
    DevBuf<int64_t> total(c, 1);
    GMG_LAUNCH(c, flag_in_kernel, div_up(n, kBlock), kBlock, 0, n, state.p, cnt.p);
    exclusive_scan_i32(c, cnt.p, cnt.p, n, total.p);
    int64_t m = 0;
    total.download(&m, 1);
    sync(c);
    selection.alloc(c, m);
    GMG_LAUNCH(c, compact_kernel, div_up(n, kBlock), kBlock, 0, n, state.p, cnt.p, selection.p);
    return int32_t(m);
}

// ---------------------------------------------------------------------------------------------
// Row-sharded variant (one rank owns the vertices [lo, hi); the state array is replicated and
// all-gathered between rounds by the caller). The same greedy set is computed -- the
// lexicographically-first maximal independent set under the chosen order is unique -- but by
// PULL rounds, because a rank can only write the vertices it owns: an undecided owned vertex looks
// at the states of its PRECEDING coverers; one IN makes it OUT, all OUT makes it IN. Rounds needed
// = dependency depth (O(log n) in the hashed order, ~10 for an index-ordered subdivision mesh).
// ---------------------------------------------------------------------------------------------
template <int MODE>
__global__ void __launch_bounds__(kBlock) shard_path_bound_kernel(SamplerArgs a, int32_t lo, int32_t hi,
                                                                  unsigned long long* __restrict__ out) {
    const int64_t i = int64_t(lo) + int64_t(blockIdx.x) * kBlock + threadIdx.x;
    unsigned long long v = 0;
    if (i < hi) {
        for (int32_t k = a.colptr[i]; k < a.colptr[i + 1]; ++k) {
            const int32_t j = a.rowidx[k];
            v += 1ull + (unsigned long long)(a.colptr[j + 1] - a.colptr[j]);
        }
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v += __shfl_down_sync(0xffffffffu, v, d);
    if ((threadIdx.x & 31) == 0 && v) atomicAdd(out, v);
}

// Lists of PRECEDING coverers of the owned vertices; owned vertices without any are IN at once.
template <int MODE>
__global__ void __launch_bounds__(kBlock) shard_build_kernel(SamplerArgs a, int32_t lo, int32_t hi,
                                                             uint8_t* __restrict__ state8) {
    __shared__ int32_t s_group[kBlock / kLanes];
    __shared__ unsigned long long s_base;
    const int64_t gid = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    const int32_t i = lo + int32_t(gid / kLanes);
    const int sub = threadIdx.x % kLanes;
    const int g = threadIdx.x / kLanes;
    const bool active = i < hi;
    int32_t prev = 0;
    uint32_t ri = 0;
    if (active) {
        ri = rank_of<MODE>(a, i);
        for_each_covered(a, i, sub, [&](int32_t t) { if (rank_of<MODE>(a, t) < ri) ++prev; });
    }
    int32_t incl = prev;
#pragma unroll
    for (int d = 1; d < kLanes; d <<= 1) {
        const int32_t t = __shfl_up_sync(0xffffffffu, incl, d, kLanes);
        if (sub >= d) incl += t;
    }
    const int32_t group_total = __shfl_sync(0xffffffffu, incl, kLanes - 1, kLanes);
    if (sub == 0) s_group[g] = group_total;
    __syncthreads();
    if (threadIdx.x < 32) {
        const int32_t v = s_group[threadIdx.x];
        int32_t s = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int32_t t = __shfl_up_sync(0xffffffffu, s, d);
            if (int(threadIdx.x) >= d) s += t;
        }
        s_group[threadIdx.x] = s - v;
        if (threadIdx.x == 31) s_base = atomicAdd(a.cursor, (unsigned long long)s);
    }
    __syncthreads();
    if (!active) return;
    const long long start = (long long)s_base + s_group[g];
    int32_t* dst = a.lst + start + (incl - prev);
    int32_t w = 0;
    for_each_covered(a, i, sub, [&](int32_t t) { if (rank_of<MODE>(a, t) < ri) dst[w++] = t; });
    if (sub == 0) {
        a.lptr[i - lo] = start;
        a.lcnt[i - lo] = group_total;
        state8[i] = group_total == 0 ? uint8_t(kIn) : uint8_t(kUndecided);
    }
}

__global__ void __launch_bounds__(kBlock) shard_pull_kernel(int32_t lo, int32_t hi, const long long* __restrict__ lptr,
                                                            const int32_t* __restrict__ lcnt,
                                                            const int32_t* __restrict__ lst, uint8_t* state8) {
    const int64_t v = int64_t(lo) + int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (v >= hi || state8[v] != kUndecided) return;
    const long long p = lptr[v - lo];
    const int32_t c = lcnt[v - lo];
    bool all_out = true;
    for (int32_t k = 0; k < c; ++k) {
        const uint8_t st = *(volatile uint8_t*)(state8 + lst[p + k]);
        if (st == kIn) { state8[v] = uint8_t(kOut); return; }          // a preceding coverer is selected (:38,45)
        all_out = all_out && st == kOut;
    }
    if (all_out) state8[v] = uint8_t(kIn);                               // nobody before it covers it (:25-28)
}

// how many of the n bytes equal `what`: 16 bytes per thread and load, one atomic per block
__global__ void __launch_bounds__(kBlock) count_state_kernel(int64_t n, const uint8_t* __restrict__ state8, uint8_t what,
                                                             unsigned long long* __restrict__ out) {
    __shared__ unsigned int s_warp[kBlock / 32];
    const int64_t chunks = n / 16;
    unsigned int c = 0;
    for (int64_t q = int64_t(blockIdx.x) * kBlock + threadIdx.x; q < chunks; q += int64_t(gridDim.x) * kBlock) {
        const uint4 w = __ldg(reinterpret_cast<const uint4*>(state8) + q);
        const unsigned int words[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
        for (int k = 0; k < 4; ++k)
#pragma unroll
            for (int b = 0; b < 4; ++b) c += ((words[k] >> (8 * b)) & 0xFFu) == what ? 1u : 0u;
    }
    if (blockIdx.x == 0 && threadIdx.x < n - chunks * 16) c += state8[chunks * 16 + threadIdx.x] == what ? 1u : 0u;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) c += __shfl_down_sync(0xffffffffu, c, d);
    if ((threadIdx.x & 31) == 0) s_warp[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned int t = 0;
        for (int w = 0; w < kBlock / 32; ++w) t += s_warp[w];
        if (t) atomicAdd(out, (unsigned long long)t);
    }
}
__global__ void __launch_bounds__(kBlock) flag_in8_kernel(int32_t n, const uint8_t* __restrict__ state8,
                                                          int32_t* __restrict__ flag) {
    const int64_t i = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (i < n) flag[i] = state8[i] == kIn ? 1 : 0;
}
__global__ void __launch_bounds__(kBlock) compact8_kernel(int32_t n, const uint8_t* __restrict__ state8,
                                                          const int32_t* __restrict__ pos, int32_t* __restrict__ out) {
    const int64_t i = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (i < n && state8[i] == kIn) out[pos[i]] = int32_t(i);
}

void shard_sampler_build(gmg_ctx* c, int32_t n, int32_t lo, int32_t hi, const int32_t* colptr, const int32_t* rowidx,
                         const double* len, double radius, int mode, uint64_t seed, uint8_t* state8, ShardSampler& out) {
    GMG_REQUIRE(mode == GMG_SAMPLER_EXACT || mode == GMG_SAMPLER_PARALLEL, GMG_ERR_INVALID, "unknown sampler mode");
    out.lo = lo; out.hi = hi;
    const int32_t own = hi - lo;
    out.lptr.alloc(c, own); out.lcnt.alloc(c, own);
    if (own <= 0) { out.lst.alloc(c, 0); return; }
    DevBuf<unsigned long long> cursor(c, 2);
    cursor.zero();
    SamplerArgs a{n, colptr, rowidx, len, radius, make_order(n, seed), nullptr, nullptr, out.lptr.p, out.lcnt.p, nullptr,
                  cursor.p, nullptr, 0};
    if (mode == GMG_SAMPLER_EXACT) GMG_LAUNCH(c, shard_path_bound_kernel<GMG_SAMPLER_EXACT>, div_up(own, kBlock), kBlock, 0, a, lo, hi, cursor.p + 1);
    else GMG_LAUNCH(c, shard_path_bound_kernel<GMG_SAMPLER_PARALLEL>, div_up(own, kBlock), kBlock, 0, a, lo, hi, cursor.p + 1);
    unsigned long long hb[2];
    cursor.download(hb, 2);
    sync(c);
    out.lst.alloc(c, int64_t(hb[1]) + 1);
    a.lst = out.lst.p;
    const int grid = div_up(int64_t(own) * kLanes, kBlock);
    if (mode == GMG_SAMPLER_EXACT) GMG_LAUNCH(c, shard_build_kernel<GMG_SAMPLER_EXACT>, grid, kBlock, 0, a, lo, hi, state8);
    else GMG_LAUNCH(c, shard_build_kernel<GMG_SAMPLER_PARALLEL>, grid, kBlock, 0, a, lo, hi, state8);
}

void shard_sampler_round(gmg_ctx* c, const ShardSampler& s, uint8_t* state8) {
    const int32_t own = s.hi - s.lo;
    if (own > 0) GMG_LAUNCH(c, shard_pull_kernel, div_up(own, kBlock), kBlock, 0, s.lo, s.hi, s.lptr.p, s.lcnt.p, s.lst.p, state8);
}

int64_t count_state(gmg_ctx* c, int64_t n, const uint8_t* state8, int what) {
    DevBuf<unsigned long long> out(c, 1);
    out.zero();
    if (n > 0) {
        const int64_t want = div_up(n / 16 + 1, kBlock);
        const int grid = int(want < 4 * int64_t(c->sm_count) ? want : 4 * int64_t(c->sm_count));
        GMG_LAUNCH(c, count_state_kernel, grid, kBlock, 0, n, state8, uint8_t(what), out.p);
    }
    unsigned long long h = 0;
    out.download(&h, 1);
    sync(c);
    return int64_t(h);
}

int32_t compact_selected(gmg_ctx* c, int32_t n, const uint8_t* state8, DevBuf<int32_t>& selection) {
    if (n == 0) { selection.alloc(c, 0); return 0; }
    DevBuf<int32_t> pos(c, n);
    DevBuf<int64_t> total(c, 1);
    GMG_LAUNCH(c, flag_in8_kernel, div_up(n, kBlock), kBlock, 0, n, state8, pos.p);
    exclusive_scan_i32(c, pos.p, pos.p, n, total.p);
    int64_t m = 0;
    total.download(&m, 1);
    sync(c);
    selection.alloc(c, m);
    GMG_LAUNCH(c, compact8_kernel, div_up(n, kBlock), kBlock, 0, n, state8, pos.p, selection.p);
    return int32_t(m);
}

}  // namespace gmg
