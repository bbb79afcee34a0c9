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

// Precedence order.
//   GMG_SAMPLER_EXACT     rank(v) = v: the reference's order.
//   GMG_SAMPLER_HASHED    every vertex index goes through a seeded invertible mixing of k-bit integers: xor-seed,
//                         then (odd multiply, xorshift by ceil(k/2)) twice and a last odd multiply, all modulo 2^k;
//                         every step is a bijection on k bits and the xorshift is its own inverse because twice the
//                         shift is >= k. Dependency depth O(log n).
//   GMG_SAMPLER_PARALLEL  the reference's index-order sweep restarted in `supers` (2) contiguous SUPER-TILES of
//                         `ssize` vertices (a multiple of the flow tile) that are swept CONCURRENTLY: the flow tiles
//                         of the super-tiles are interleaved in the order, rank = ((flow tile in its super-tile) *
//                         supers + super-tile) * kTile + offset. Every super-tile starts at once (its first rows
//                         precede the last rows of its neighbour), so the dependency depth is half the index
//                         order's; the packing is the reference's except along the seams (why: DESIGN.md section 4,
//                         "sampler orders"). The seed is not used.
struct Order {
    uint32_t mask, shift, seed;
    uint32_t c1, c2, c3, i1, i2, i3;   // odd multipliers and their inverses modulo 2^32
    uint32_t supers, ssize;            // PARALLEL: number and size of the super-tiles (supers == 0: HASHED)
    uint32_t kbits;                    // bits of a rank; ranks live in [0, 2^kbits)
};
constexpr int kOrderTileLog = 12;      // = kTileLog of the tile flow kernel
constexpr int32_t kParallelMinVertices = 2000000;   // GMG_SAMPLER_PARALLEL: two concurrent sweeps from here on
__host__ __device__ __forceinline__ uint32_t order_rank(const Order& o, uint32_t v) {
    if (o.supers) {
        const uint32_t sup = (v >= o.ssize ? 1u : 0u) + (v >= 2u * o.ssize ? 1u : 0u);
        const uint32_t inner = v - sup * o.ssize;
        return (((inner >> kOrderTileLog) * o.supers + sup) << kOrderTileLog) | (inner & ((1u << kOrderTileLog) - 1u));
    }
    uint32_t x = (v ^ o.seed) & o.mask;
    x = (x * o.c1) & o.mask; x ^= x >> o.shift;
    x = (x * o.c2) & o.mask; x ^= x >> o.shift;
    x = (x * o.c3) & o.mask;
    return x;
}
// (a rank that stands for no vertex comes back as a value >= n)
__host__ __device__ __forceinline__ uint32_t order_unrank(const Order& o, uint32_t r) {
    if (o.supers) {
        const uint32_t q = r >> kOrderTileLog;
        const uint32_t ft = q / o.supers, sup = q - ft * o.supers;
        const uint32_t inner = (ft << kOrderTileLog) | (r & ((1u << kOrderTileLog) - 1u));
        return inner < o.ssize ? sup * o.ssize + inner : 0xFFFFFFFFu;
    }
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
static int ceil_log2(int64_t v) { int k = 0; while ((int64_t(1) << k) < v) ++k; return k; }
static Order make_order(int32_t n, uint64_t seed, int mode) {
    Order o{};
    if (mode == GMG_SAMPLER_EXACT) {            // identity; only kbits is read
        o.kbits = uint32_t(ceil_log2(n > 1 ? n : 2));
        return o;
    }
    if (mode == GMG_SAMPLER_PARALLEL) {
        // every additional sweep start costs samples (the packing of a sweep tightens over its first rows): measured
        // on the noisy torus with tools/order_stats.py, two sweeps lose 2.3 % at 1 M vertices, 1.4 % at 2.5 M and
        // 1.1 % at 10 M, three lose 3.2 % / 2.3 % at 1 M / 10 M. The north star allows 2 %: two sweeps from 2 M
        // vertices on, the plain index order below that.
        constexpr int64_t tile = int64_t(1) << kOrderTileLog;
        const int64_t want = n >= kParallelMinVertices ? 2 : 1;
        const int64_t part = ((int64_t(n) + want - 1) / want + tile - 1) / tile * tile;
        o.ssize = uint32_t(part);
        o.supers = uint32_t((int64_t(n) + part - 1) / part);
        o.kbits = uint32_t(ceil_log2(part * int64_t(o.supers)));
        if (o.kbits == 0) o.kbits = 1;
        return o;
    }
    const int kt = ceil_log2(n);
    o.mask = kt >= 32 ? 0xFFFFFFFFu : ((1u << kt) - 1u);
    o.shift = uint32_t((kt + 1) / 2);
    if (o.shift == 0) o.shift = 1;
    const uint64_t z = (seed + 0x9E3779B97F4A7C15ull) * 0xBF58476D1CE4E5B9ull;
    o.seed = uint32_t(z >> 32) & o.mask;
    o.c1 = 0x9E3779B1u; o.c2 = 0x85EBCA6Bu | 1u; o.c3 = 0xC2B2AE35u | 1u;
    o.i1 = inverse_odd(o.c1); o.i2 = inverse_odd(o.c2); o.i3 = inverse_odd(o.c3);
    o.kbits = uint32_t(kt);
    if (o.kbits == 0) o.kbits = 1;
    return o;
}

// ---- the flow's state, all indexed by RANK (so that a tile of consecutive tickets is contiguous) -----------
// wext[r]   preceding covering paths into r from vertices of OTHER flow tiles; bit 31 = knocked OUT by a selected
//           vertex of another tile. Only ever changed by fire-and-forget reductions (atomicSub / atomicOr).
// cint[r]   preceding covering paths from vertices of r's OWN flow tile (decremented in shared memory)
// lst       later coverees of every vertex as rank | (multiplicity - 1) << kbits: duplicates are merged (a mesh
//           vertex has ~24 later paths but ~8 distinct later coverees), same-tile targets first
// lcnt[r]   entries of the list | same-tile entries << 16 (0xFFFF: list not partitioned nor merged, see the build kernel)
constexpr int kTileLog = kOrderTileLog;
constexpr int kTile = 1 << kTileLog;          // ranks per flow tile
constexpr int kFlowThreads = 1024;
constexpr int kOwn = kTile / kFlowThreads;    // ranks per thread
constexpr int kListCap = 40960;               // same-tile list entries (uint16) a block can stage
constexpr int kExtCap = 16384;                // other-tile list entries (uint32) a block can stage
constexpr uint32_t kOutFlag = 0x80000000u;
constexpr uint32_t kUnpartitioned = 0xFFFFu;

struct SamplerArgs {
    int32_t n;
    const int32_t* colptr;
    const int32_t* rowidx;
    const double* len;
    double radius;
    Order order;
    uint32_t* wext;
    uint32_t* cint;
    long long* lptr;             // start of the later-coveree list of every rank
    uint32_t* lcnt;
    uint32_t* lst;               // the lists
    unsigned long long* cursor;  // bump allocator over lst
    unsigned int* ticket;        // next tile of the flow kernel
    unsigned int tiles;          // flow tiles: ceil(ranks / kTile)
    uint8_t* res;                // [ranks] 1 = selected
    int32_t* flags;              // [0] a vertex has more than 65535 list entries
    unsigned long long* stats;   // [0] same-tile list entries, [1] all list entries (decides which flow kernel runs)
    unsigned int* epoch;         // [tiles] tile flow: entries reserved in the mailbox of tile t
    int force_lane;              // GMG_SAMPLER_HASHED: no tile ever holds two coverers, always the per-lane flow
    unsigned long long* dbg;     // GMG_SAMPLER_DEBUG: 8 words per tile (timeline of the tile flow), else null
    uint32_t* trace;             // GMG_SAMPLER_TRACE: 4 words per rank (ns since t0: queued, popped, ext popped, ext done), else null
    const unsigned long long* trace_t0;
    uint16_t* mbox;              // [tiles * kTile] tile flow: mailbox of every tile (rank offset | knocked-out << kTileLog), 0xFFFF = empty
};

__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

// Which flow kernel runs is decided ON THE DEVICE from the build kernel's counts (no host round trip): the
// tile-cooperative kernel pays off when at least a quarter of the list entries stay inside a flow tile, i.e.
// when the precedence order is spatially coherent (grids, Morton orders, the coarse levels of such meshes).
__device__ __forceinline__ bool use_lane_flow(const SamplerArgs& a) {
    if (a.force_lane) return true;
    const unsigned long long same = *(volatile unsigned long long*)(a.stats + 0);
    const unsigned long long all = *(volatile unsigned long long*)(a.stats + 1);
    return same * 4ull < all;
}

// Which BUILD kernel runs is decided on the device too, from a probe of the graph (sampler_coherence_kernel: the share
// of stored neighbours that sit in the vertex's own flow tile, over every 16th vertex): merging and partitioning the
// lists (shared-memory hash per vertex) pays for itself only where the tile flow will run; for an incoherent order
// (subdivision numbering, an unordered cloud, the hashed order) the raw single-walk build is twice as fast and the
// per-lane flow does not care.
// Measured (B200): icosphere in subdivision order (6 neighbours), raw 0.97 ms vs merged 1.26 ms of sampling at 2.6 M
// vertices; kNN cloud (~9 neighbours, ~150 later paths that merge into ~40 entries), raw 2.1 ms vs merged 1.9 ms at
// 2 M points. Hence the second condition: at most 8 neighbours on average.
__device__ __forceinline__ bool use_raw_build(const SamplerArgs& a) {
    const unsigned long long same = *(volatile unsigned long long*)(a.stats + 2);
    const unsigned long long all = *(volatile unsigned long long*)(a.stats + 3);
    const bool sparse = all * 2ull <= (unsigned long long)a.n;          // the probe visits n / 16 vertices
    return (a.force_lane || same * 4ull < all) && sparse;
}

template <int MODE> __device__ __forceinline__ uint32_t rank_of(const SamplerArgs& a, int32_t v) {
    return MODE == GMG_SAMPLER_EXACT ? uint32_t(v) : order_rank(a.order, uint32_t(v));
}
template <int MODE> __device__ __forceinline__ uint32_t vertex_of(const SamplerArgs& a, uint32_t r) {
    return MODE == GMG_SAMPLER_EXACT ? r : order_unrank(a.order, r);
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

// One walk of every two-hop neighbourhood (8 lanes per vertex): counts the preceding covering paths (same flow
// tile / other tiles) and writes the LATER coverees as a merged list. A mesh vertex has ~24 later paths but only
// ~8 distinct later coverees: every later target goes straight into a small per-vertex hash table in shared
// memory (atomicCAS on the key, atomicAdd on its multiplicity -- both native on 32-bit shared words), and the
// occupied slots are written out, same-tile targets first, an entry standing for up to `cap` paths (the bits a
// rank leaves free in 32). A vertex with more distinct later coverees than the table holds (dense kNN graphs)
// walks again and writes its paths raw, flagged kUnpartitioned.
constexpr int kMergeSlots = 64;      // per vertex; a walk that fills more than 3/4 of them falls back to raw paths

template <int MODE>
__global__ void __launch_bounds__(kBlock) sampler_build_kernel(SamplerArgs a) {
    static_assert(kBlock / kLanes == 32, "one warp scans the group totals");
    static_assert(kMergeSlots % kLanes == 0, "the lanes of a group split the slots evenly");
    __shared__ int32_t s_group[kBlock / kLanes];
    __shared__ unsigned long long s_base;
    __shared__ uint32_t s_key[kBlock / kLanes][kMergeSlots];
    __shared__ uint32_t s_cnt[kBlock / kLanes][kMergeSlots];
    __shared__ int32_t s_used[kBlock / kLanes];        // occupied slots of the group's table (-1: overflowed)
    __shared__ int32_t s_same;
    __shared__ int s_skip;
    if (threadIdx.x == 0) s_skip = use_raw_build(a) ? 1 : 0;      // (one read per block: the words are shared by the grid)
    __syncthreads();
    if (s_skip) return;
    const int64_t gid = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    const int32_t i = int32_t(gid / kLanes);
    const int sub = threadIdx.x % kLanes;
    const int g = threadIdx.x / kLanes;
    const bool active = i < a.n;
    for (int s = sub; s < kMergeSlots; s += kLanes) { s_key[g][s] = 0xFFFFFFFFu; s_cnt[g][s] = 0u; }
    if (sub == 0) s_used[g] = 0;
    if (threadIdx.x == 0) s_same = 0;
    __syncthreads();
    int32_t prev_in = 0, prev_ext = 0, later = 0;
    uint32_t ri = 0;
    if (active) {
        ri = rank_of<MODE>(a, i);
        uint32_t* keys = s_key[g];
        uint32_t* cnts = s_cnt[g];
        for_each_covered(a, i, sub, [&](int32_t t) {
            const uint32_t rt = rank_of<MODE>(a, t);
            if (rt < ri) { if (!a.force_lane && (rt >> kTileLog) == (ri >> kTileLog)) ++prev_in; else ++prev_ext; return; }
            ++later;
            uint32_t h = (rt * 0x9E3779B1u) >> 16;
            for (int probe = 0; probe < kMergeSlots; ++probe, ++h) {
                const int sl = int(h & (kMergeSlots - 1));
                const uint32_t prev = atomicCAS(keys + sl, 0xFFFFFFFFu, rt);
                if (prev == 0xFFFFFFFFu) { if (atomicAdd(s_used + g, 1) >= kMergeSlots * 3 / 4) s_used[g] = -(1 << 20); }
                if (prev == 0xFFFFFFFFu || prev == rt) { atomicAdd(cnts + sl, 1u); return; }
            }
            s_used[g] = -(1 << 20);                      // table full
        });
    }
    const int lane = threadIdx.x & 31;
    __syncwarp();
    const bool group_over = active && *(volatile int32_t*)(s_used + g) < 0;
#pragma unroll
    for (int d = kLanes / 2; d > 0; d >>= 1) {
        prev_in += __shfl_xor_sync(0xffffffffu, prev_in, d, kLanes);
        prev_ext += __shfl_xor_sync(0xffffffffu, prev_ext, d, kLanes);
    }
    // entries this lane writes: its share of the table's slots, an entry per `cap` paths of a slot
    const uint32_t cap = 1u << (32u - a.order.kbits < 4u ? 32u - a.order.kbits : 4u);
    int32_t kept = 0, kept_in = 0;
    if (active && !group_over) {
        for (int sl = sub; sl < kMergeSlots; sl += kLanes) {
            const uint32_t x = s_key[g][sl];
            if (x == 0xFFFFFFFFu) continue;
            const int32_t pieces = int32_t((s_cnt[g][sl] + cap - 1u) / cap);
            kept += pieces;
            if (!a.force_lane && (x >> kTileLog) == (ri >> kTileLog)) kept_in += pieces;
        }
    }
    const int32_t own = group_over ? later : kept;
    int32_t incl = own, incl_in = kept_in;
#pragma unroll
    for (int d = 1; d < kLanes; d <<= 1) {
        const int32_t t = __shfl_up_sync(0xffffffffu, incl, d, kLanes);
        const int32_t t2 = __shfl_up_sync(0xffffffffu, incl_in, d, kLanes);
        if (sub >= d) { incl += t; incl_in += t2; }
    }
    const int32_t group_total = __shfl_sync(0xffffffffu, incl, kLanes - 1, kLanes);
    const int32_t group_in = __shfl_sync(0xffffffffu, incl_in, kLanes - 1, kLanes);
    if (sub == 0) s_group[g] = group_total;
    __syncthreads();
    if (sub == 0 && group_in > 0) atomicAdd(&s_same, group_in);
    if (threadIdx.x < 32) {
        const int32_t v = s_group[threadIdx.x];
        int32_t s = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int32_t t = __shfl_up_sync(0xffffffffu, s, d);
            if (int(threadIdx.x) >= d) s += t;
        }
        s_group[threadIdx.x] = s - v;
        if (threadIdx.x == 31) { s_base = atomicAdd(a.cursor, (unsigned long long)s); if (s) atomicAdd(a.stats + 1, (unsigned long long)s); }
    }
    __syncthreads();
    if (threadIdx.x == 0 && s_same > 0) atomicAdd(a.stats + 0, (unsigned long long)s_same);
    if (!active) return;
    const long long start = (long long)s_base + s_group[g];
    if (!group_over) {
        // same-tile targets first: [start, start + group_in), the others behind them
        uint32_t* dst_in = a.lst + start + (incl_in - kept_in);
        uint32_t* dst_ext = a.lst + start + group_in + ((incl - kept) - (incl_in - kept_in));
        for (int sl = sub; sl < kMergeSlots; sl += kLanes) {
            const uint32_t x = s_key[g][sl];
            if (x == 0xFFFFFFFFu) continue;
            const bool same = !a.force_lane && (x >> kTileLog) == (ri >> kTileLog);
            for (uint32_t left = s_cnt[g][sl]; left > 0u;) {
                const uint32_t mult = left < cap ? left : cap;
                const uint32_t packed = x | ((mult - 1u) << a.order.kbits);
                if (same) *dst_in++ = packed; else *dst_ext++ = packed;
                left -= mult;
            }
        }
    } else {                                          // more distinct later coverees than the table holds: walk again
        uint32_t* dst = a.lst + start + (incl - own);
        for_each_covered(a, i, sub, [&](int32_t t) {
            const uint32_t rt = rank_of<MODE>(a, t);
            if (!(rt < ri)) *dst++ = rt;
        });
    }
    if (sub == 0) {
        if (group_total > 0xFFFF || group_in >= int32_t(kUnpartitioned)) a.flags[0] = 1;
        a.lptr[ri] = start;
        a.lcnt[ri] = uint32_t(group_total) | ((group_over ? kUnpartitioned : uint32_t(group_in)) << 16);
        a.cint[ri] = uint32_t(prev_in);
        a.wext[ri] = uint32_t(prev_ext);
    }
}

// The probe behind use_raw_build: one thread per 16th vertex.
template <int MODE>
__global__ void __launch_bounds__(kBlock) sampler_coherence_kernel(SamplerArgs a) {
    const int64_t i = (int64_t(blockIdx.x) * kBlock + threadIdx.x) * 16;
    unsigned same = 0, all = 0;
    if (i < a.n) {
        const uint32_t ti = rank_of<MODE>(a, int32_t(i)) >> kTileLog;
        for (int32_t k = __ldg(a.colptr + i); k < __ldg(a.colptr + i + 1); ++k) {
            const int32_t j = __ldg(a.rowidx + k);
            if (j == i) continue;
            ++all;
            same += (rank_of<MODE>(a, j) >> kTileLog) == ti ? 1u : 0u;
        }
    }
    same = __reduce_add_sync(0xffffffffu, same);
    all = __reduce_add_sync(0xffffffffu, all);
    if ((threadIdx.x & 31) == 0 && all) { atomicAdd(a.stats + 2, (unsigned long long)same); atomicAdd(a.stats + 3, (unsigned long long)all); }
}

// The build for incoherent orders: ONE walk, later targets staged in a lane-private row of shared memory and written
// as they are (one entry per path, flagged kUnpartitioned; every preceding path counts into the global word). A
// vertex with a lane that outgrows its row walks again.
template <int MODE>
__global__ void __launch_bounds__(kBlock) sampler_build_raw_kernel(SamplerArgs a) {
    static_assert(kBlock / kLanes == 32, "one warp scans the group totals");
    __shared__ int32_t s_group[kBlock / kLanes];
    __shared__ unsigned long long s_base;
    __shared__ uint32_t s_stage[kBlock][kStage + 1];   // +1: conflict-free lane-private rows
    __shared__ int s_skip;
    if (threadIdx.x == 0) s_skip = use_raw_build(a) ? 0 : 1;
    __syncthreads();
    if (s_skip) return;
    const int64_t gid = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    const int32_t i = int32_t(gid / kLanes);
    const int sub = threadIdx.x % kLanes;
    const int g = threadIdx.x / kLanes;
    const bool active = i < a.n;
    int32_t prev = 0, later = 0;
    uint32_t ri = 0;
    if (active) {
        ri = rank_of<MODE>(a, i);
        uint32_t* stage = s_stage[threadIdx.x];
        for_each_covered(a, i, sub, [&](int32_t t) {
            const uint32_t rt = rank_of<MODE>(a, t);
            if (rt < ri) ++prev;
            else { if (later < kStage) stage[later] = rt; ++later; }
        });
    }
    const int lane = threadIdx.x & 31;
    const unsigned gmask = 0xFFu << (lane & ~(kLanes - 1));
    const bool group_over = (__ballot_sync(0xffffffffu, later > kStage) & gmask) != 0;
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
        int32_t sc = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int32_t t = __shfl_up_sync(0xffffffffu, sc, d);
            if (int(threadIdx.x) >= d) sc += t;
        }
        s_group[threadIdx.x] = sc - v;
        if (threadIdx.x == 31) { s_base = atomicAdd(a.cursor, (unsigned long long)sc); if (sc) atomicAdd(a.stats + 1, (unsigned long long)sc); }
    }
    __syncthreads();
    if (!active) return;
    const long long start = (long long)s_base + s_group[g];
    uint32_t* dst = a.lst + start + (incl - later);
    if (!group_over) {
        const uint32_t* stage = s_stage[threadIdx.x];
        for (int32_t e = 0; e < later; ++e) dst[e] = stage[e];
    } else {
        for_each_covered(a, i, sub, [&](int32_t t) {
            const uint32_t rt = rank_of<MODE>(a, t);
            if (!(rt < ri)) *dst++ = rt;
        });
    }
    if (sub == 0) {
        if (group_total > 0xFFFF) a.flags[0] = 1;
        a.lptr[ri] = start;
        a.lcnt[ri] = uint32_t(group_total) | (kUnpartitioned << 16);
        a.cint[ri] = 0u;
        a.wext[ri] = uint32_t(prev);
    }
}

__global__ void stamp_kernel(unsigned long long* out) { *out = global_ns(); }

// The flow: a vertex is IN once every preceding coverer is OUT (both path counters at 0), OUT once a preceding
// coverer is IN. Two kernels; which one does the work is decided on the device (use_lane_flow).
//
// (1) TILE flow, for spatially coherent orders. On a row-major 10000 x 1000 grid the dependency chain is tens of
// thousands of decisions long and only a handful of vertices per tile can be decided at any moment, so what counts
// is the latency of ONE decision step, not throughput. A block takes flow TILES (kTile consecutive ranks) by
// ticket, so every tile that precedes a running tile is running or done and the waits below always end. Inside a
// tile nothing polls: the tile is a WORK LIST in shared memory.
//   * Both halves of the tile's lists are staged in shared memory once (they are static, so this happens before
//     the tile has anything to wait for), next to one state word per rank: same-tile path counter | PENDING (the
//     other-tile counter is not 0 yet) | DECIDED.
//   * Whoever makes the atomic that leaves a word at "counter 0, not pending, not decided" queues that rank as
//     IN; whoever first ORs DECIDED into a word queues it as OUT. Atomics on a word are totally ordered, so each
//     rank is queued exactly once, by the operation that decided it, at the moment it is decided.
//   * FAST warps pop the queue: the 32 lanes apply the popped rank's same-tile list in parallel (one shared-memory
//     atomic each), queueing what that decides, and hand the rank on to the second queue.
//   * EXT warps pop the second queue and apply the other-tile half with atomics on the global words. These return
//     the old value: the lane that brings a word to 0, or sets its OUT bit first, appends the rank to the MAILBOX of
//     the tile that owns it (<= 1 notification per rank). Two L2 round trips, kept away from the fast warps.
//   * One WATCH warp reads the tile's own mailbox, 32 slots per look, and turns notifications into state-word
//     atomics. A waiting tile therefore costs one L2 load per look, and is told exactly which ranks changed.
//
// (2) LANE flow, for incoherent orders (hashed priorities, subdivision orders): nearly every target sits in
// another tile, the dependency depth is O(log n), and one rank per lane with no block-wide structure is faster.
constexpr int32_t kRemMask = 0x000FFFFF;      // same-tile preceding paths still undecided
constexpr int32_t kPending = 1 << 20;         // the global word (other-tile preceding paths) has not reached 0
constexpr int32_t kDecided = 1 << 21;         // queued as OUT, or not a vertex at all
constexpr int32_t kSlow = 1 << 22;            // lists not staged (no room, or not partitioned): read from global memory
constexpr uint32_t kEmpty = 0xFFFFu;          // queue / mailbox slot not written yet
// Shape of a flow block (31 worker warps + the watcher). Measured on C3 level 0 (B200, bit-exact in every shape):
// FAST x lanes per rank, EXT lanes per rank -> ms: 16x8,8 -> 24.8 (four ranks per fast warp in lock step: a group cannot
// look at its slot while a sibling applies a list; the trace shows a median 450 ns between a rank being queued and being
// popped); 24x32,8 -> 24.1; 26x32,8 -> 22.3 (one rank per fast warp: that wait drops to 290 ns, but 5 EXT warps are
// about as few as the other-tile halves tolerate: their queue wait doubles to 2.7 us); 27x32,8 -> 22.9; 28x32,8 -> 24.6;
// 20x32,8 -> 30.5; 26x32,4 -> 26.4; 24x16,8 -> 25.3.
constexpr int kFastWarps = 26;                // warps 1 .. kFastWarps pop the tile's queue; warp 0 watches the mailbox
constexpr int kGroup = 32;                    // lanes that apply one rank's same-tile list
constexpr int kXGroup = 8;                    // lanes that apply one rank's other-tile list
constexpr int kPtrLen = kTile + 2;            // prefix arrays (kTile + 1 used; even length keeps 4-byte alignment)

// DBG: the instantiation with the timeline / per-rank trace instrumentation (GMG_SAMPLER_DEBUG, GMG_SAMPLER_TRACE); the
// production one carries none of it -- the loops below are latency chains, every instruction in them counts (an
// experimental extra mode in the pop loop once cost 10 % just by being compiled in).
template <int MODE, bool DBG>
__global__ void __launch_bounds__(kFlowThreads, 1) sampler_flow_kernel(SamplerArgs a) {
    extern __shared__ unsigned char smem_raw[];
    int32_t* s_state = reinterpret_cast<int32_t*>(smem_raw);                     // kTile state words
    uint32_t* s_ext = reinterpret_cast<uint32_t*>(s_state + kTile);              // kExtCap staged other-tile entries
    uint16_t* s_list = reinterpret_cast<uint16_t*>(s_ext + kExtCap);             // kListCap staged same-tile entries
    uint16_t* s_lptr = s_list + kListCap;                                        // same-tile list of rank i: [s_lptr[i], s_lptr[i+1])
    uint16_t* s_xptr = s_lptr + kPtrLen;                                         // other-tile list of rank i, likewise
    uint16_t* s_queue = s_xptr + kPtrLen;                                        // decided ranks: offset | IN << kTileLog
    uint16_t* s_xqueue = s_queue + kTile;                                        // ... whose other-tile half is still to apply
    __shared__ int32_t s_warp[2][kFlowThreads / 32];
    __shared__ unsigned int s_tile;
    // [0] vertices of the tile, [1] queued, [2] claimed, [3] vertices with an other-tile half, [4] queued, [5] claimed
    __shared__ int32_t s_ctr[8];
    if (use_lane_flow(a)) return;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t rmask = a.order.kbits >= 32 ? 0xFFFFFFFFu : ((1u << a.order.kbits) - 1u);
    const uint64_t ranks = MODE == GMG_SAMPLER_EXACT ? uint64_t(a.n) : (uint64_t(1) << a.order.kbits);
    const unsigned long long trace_t0 = (DBG && a.trace) ? *a.trace_t0 : 0ull;
    auto push = [&](uint32_t t, uint32_t in) {
        if (DBG && a.trace) a.trace[4 * (size_t(s_tile) * kTile + t) + 0] = uint32_t(global_ns() - trace_t0);
        const int32_t slot = atomicAdd(&s_ctr[1], 1);
        *(volatile uint16_t*)(s_queue + slot) = uint16_t(t | (in << kTileLog));
    };
    // a decrement of `mult` paths / a knock-out arriving at rank t of the tile
    auto release = [&](uint32_t t, int32_t mult) {
        const int32_t now = atomicSub(s_state + t, mult) - mult;
        if ((now & (kRemMask | kPending | kDecided)) == 0) push(t, 1u);
    };
    auto knock_out = [&](uint32_t t) {
        if (!(atomicOr(s_state + t, kDecided) & kDecided)) push(t, 0u);
    };
    for (;;) {
        __syncthreads();                                   // everybody is done with the previous tile's shared state
        if (tid == 0) {
            s_tile = atomicAdd(a.ticket, 1u);
#pragma unroll
            for (int k = 0; k < 8; ++k) s_ctr[k] = 0;
        }
        __syncthreads();
        const unsigned int tile = s_tile;
        if (tile >= a.tiles) break;
        const uint32_t t0 = tile << kTileLog;
        if (DBG && a.dbg && tid == 0) { a.dbg[8 * tile + 0] = global_ns(); a.dbg[8 * tile + 7] = blockIdx.x; }
        // ---- static data of the tile's ranks; ranks [tid * kOwn, (tid + 1) * kOwn) of the tile belong to this thread
        //      for the staging (so that the scans below lay the lists out in rank order)
        long long ptr[kOwn];
        uint32_t tot[kOwn], nin[kOwn];
        int32_t st[kOwn];
        int32_t mine = 0, mine_x = 0, nvalid = 0, nx = 0;
        unsigned valid_bits = 0;
#pragma unroll
        for (int q = 0; q < kOwn; ++q) {
            const uint64_t r = uint64_t(t0) + uint64_t(tid * kOwn + q);
            const bool valid = r < ranks && vertex_of<MODE>(a, uint32_t(r)) < uint32_t(a.n);
            ptr[q] = 0; tot[q] = 0; nin[q] = 0; st[q] = kDecided;
            if (valid) {
                const uint32_t lc = a.lcnt[r];
                ptr[q] = a.lptr[r]; tot[q] = lc & 0xFFFFu; nin[q] = lc >> 16;
                // the CURRENT value: tiles that are running have pushed into it already. Whatever happens to the word
                // after this read arrives through the mailbox (and a notification about something this read has seen
                // is recognised as such: PENDING already clear / DECIDED already set).
                const uint32_t w = *(volatile const uint32_t*)(a.wext + r);
                st[q] = int32_t(a.cint[r]) & kRemMask;
                if (w & kOutFlag) st[q] |= kDecided;
                else if (w != 0u) st[q] |= kPending;
                const bool raw = nin[q] == kUnpartitioned;
                if (raw) st[q] |= kSlow;
                else { mine += int32_t(nin[q]); mine_x += int32_t(tot[q] - nin[q]); }
                valid_bits |= 1u << q;
                ++nvalid;
                if (raw ? tot[q] > 0u : tot[q] > nin[q]) ++nx;
            }
        }
        // shared-memory segments for the two halves of the lists: block-wide exclusive scans over the threads
        {
            int32_t s = mine, sx = mine_x, nv = nvalid, nxx = nx;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int32_t t = __shfl_up_sync(0xffffffffu, s, d);
                const int32_t tx = __shfl_up_sync(0xffffffffu, sx, d);
                if (lane >= d) { s += t; sx += tx; }
            }
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) {
                nv += __shfl_down_sync(0xffffffffu, nv, d);
                nxx += __shfl_down_sync(0xffffffffu, nxx, d);
            }
            if (lane == 0) { if (nv) atomicAdd(&s_ctr[0], nv); if (nxx) atomicAdd(&s_ctr[3], nxx); }
            if (lane == 31) { s_warp[0][warp] = s; s_warp[1][warp] = sx; }
            __syncthreads();
            if (warp < 2) {
                const int32_t v = s_warp[warp][lane];
                int32_t w = v;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    const int32_t t = __shfl_up_sync(0xffffffffu, w, d);
                    if (lane >= d) w += t;
                }
                s_warp[warp][lane] = w - v;
            }
            __syncthreads();
            int32_t off = s_warp[0][warp] + (s - mine), offx = s_warp[1][warp] + (sx - mine_x);
#pragma unroll
            for (int q = 0; q < kOwn; ++q) {
                const int i = tid * kOwn + q;
                const bool part = ((valid_bits >> q) & 1u) && !(st[q] & kSlow);
                const int32_t want = part ? int32_t(nin[q]) : 0, wantx = part ? int32_t(tot[q] - nin[q]) : 0;
                s_lptr[i] = uint16_t(off < 0xFFFF ? off : 0xFFFF);
                s_xptr[i] = uint16_t(offx < 0xFFFF ? offx : 0xFFFF);
                if (part && (off + want > kListCap || offx + wantx > kExtCap)) st[q] |= kSlow;   // and so is every rank after it
                off += want; offx += wantx;
                s_state[i] = st[q];
                s_queue[i] = uint16_t(kEmpty);
                s_xqueue[i] = uint16_t(kEmpty);
            }
            if (tid == kFlowThreads - 1) {
                s_lptr[kTile] = uint16_t(off < 0xFFFF ? off : 0xFFFF);
                s_xptr[kTile] = uint16_t(offx < 0xFFFF ? offx : 0xFFFF);
            }
        }
        // ---- stage the lists: same-tile entries as uint16 (offset in the tile | (multiplicity - 1) << 12), the
        //      others as they are
#pragma unroll
        for (int q = 0; q < kOwn; ++q) {
            if (!((valid_bits >> q) & 1u) || (st[q] & kSlow)) continue;
            const int i = tid * kOwn + q;
            const uint32_t lo = s_lptr[i], xo = s_xptr[i];
            for (uint32_t e = 0; e < nin[q]; ++e) {
                const uint32_t x = a.lst[ptr[q] + e];
                s_list[lo + e] = uint16_t(((x & rmask) - t0) | ((x >> a.order.kbits) << kTileLog));
            }
            for (uint32_t e = nin[q]; e < tot[q]; ++e) s_ext[xo + (e - nin[q])] = a.lst[ptr[q] + e];
        }
        __syncthreads();
        if (DBG && a.dbg && tid == 0) a.dbg[8 * tile + 1] = global_ns();
        // ---- what is decided from the start: knocked out already / nothing to wait for
#pragma unroll
        for (int q = 0; q < kOwn; ++q) {
            if (!((valid_bits >> q) & 1u)) continue;
            if (st[q] & kDecided) push(uint32_t(tid * kOwn + q), 0u);
            else if ((st[q] & (kRemMask | kPending)) == 0) push(uint32_t(tid * kOwn + q), 1u);
        }
        const int32_t total = *(volatile int32_t*)&s_ctr[0];
        if (warp == 0) {
            // ---- WATCH: the tile's mailbox, a window of 32 slots per look (a slot is written after it is reserved, so
            //      the filled slots need not be a prefix: the window moves past its leading dealt-with slots)
            const volatile uint16_t* box = a.mbox + (size_t(tile) << kTileLog);
            int next = 0;
            unsigned seen = 0;                 // bit l: slot next + l has been dealt with
            unsigned long long looks = 0;
            while (*(volatile int32_t*)&s_ctr[1] < total) {
                const int k = next + lane;
                const uint32_t x = k < kTile ? uint32_t(box[k]) : kEmpty;
                const unsigned fresh = __ballot_sync(0xffffffffu, x != kEmpty) & ~seen;
                if ((fresh >> lane) & 1u) {
                    const uint32_t t = x & uint32_t(kTile - 1);
                    if (x >> kTileLog) knock_out(t);
                    else {
                        const int32_t old = atomicAnd(s_state + t, ~kPending);
                        if ((old & kPending) && (old & (kRemMask | kDecided)) == 0) push(t, 1u);
                    }
                }
                seen |= fresh;
                const int cnt = seen == 0xFFFFFFFFu ? 32 : __ffs(int(~seen)) - 1;
                next += cnt;
                seen = cnt == 32 ? 0u : seen >> cnt;
                ++looks;
            }
            if (DBG && a.dbg && lane == 0) { a.dbg[8 * tile + 4] = looks; a.dbg[8 * tile + 5] = (unsigned long long)next; }
        } else if (warp <= kFastWarps) {
            // ---- FAST: 32 / kGroup ranks per warp at a time, kGroup lanes each. The groups of a warp move in lock step:
            //      look at the claimed slot, apply it if it is filled, claim the next (kGroup = 32: one rank per warp).
            constexpr int grp = kGroup;
            const int lead = lane & ~(grp - 1), sub = lane & (grp - 1);
            int32_t slot = 0;
            if (sub == 0) slot = atomicAdd(&s_ctr[2], 1);
            slot = __shfl_sync(0xffffffffu, slot, lead);
            while (__any_sync(0xffffffffu, slot < total)) {
                const uint32_t e = slot < total ? uint32_t(*(volatile uint16_t*)(s_queue + slot)) : kEmpty;
                if (e != kEmpty) {
                    const uint32_t v = e & uint32_t(kTile - 1);
                    const bool in = (e >> kTileLog) != 0u;
                    if (DBG && a.dbg && sub == 0) atomicMin(a.dbg + 8 * tile + 2, global_ns());
                    if (DBG && a.trace && sub == 0) a.trace[4 * (size_t(t0) + v) + 1] = uint32_t(global_ns() - trace_t0);
                    bool has_ext;
                    // (independent loads, issued together: every dependent shared-memory access is a step of the chain)
                    const int32_t sv = *(volatile int32_t*)(s_state + v);
                    const uint32_t b = s_lptr[v], end = s_lptr[v + 1], xb = s_xptr[v], xe = s_xptr[v + 1];
                    if (!(sv & kSlow)) {
                        for (uint32_t k = b + sub; k < end; k += grp) {
                            const uint32_t y = s_list[k];
                            // a selected vertex knocks out everything it covers (src/sampling.cpp:38,45); a rejected one
                            // releases its later coverees
                            if (in) knock_out(y & uint32_t(kTile - 1));
                            else release(y & uint32_t(kTile - 1), int32_t(y >> kTileLog) + 1);
                        }
                        has_ext = xe > xb;
                    } else {
                        const uint32_t lc = a.lcnt[t0 + v];
                        const long long p = a.lptr[t0 + v];
                        const uint32_t all = lc & 0xFFFFu, same = lc >> 16;
                        const bool raw = same == kUnpartitioned;
                        const uint32_t cnt = raw ? all : same;
                        for (uint32_t k = sub; k < cnt; k += grp) {
                            const uint32_t x = a.lst[p + k];
                            const uint32_t rt = x & rmask;
                            if ((rt >> kTileLog) != tile) continue;
                            if (in) knock_out(rt - t0);
                            else release(rt - t0, int32_t(x >> a.order.kbits) + 1);
                        }
                        has_ext = raw ? all > 0u : all > same;
                    }
                    if (sub == 0) {
                        a.res[t0 + v] = in ? 1 : 0;
                        if (has_ext) {
                            const int32_t xs = atomicAdd(&s_ctr[4], 1);
                            *(volatile uint16_t*)(s_xqueue + xs) = uint16_t(e);
                        }
                        slot = atomicAdd(&s_ctr[2], 1);
                    }
                }
                slot = __shfl_sync(0xffffffffu, slot, lead);
            }
        } else {
            // ---- EXT: the other-tile half, with atomics that return: the lane that decides something about a rank of
            //      another tile tells that tile. Same lock-step groups; two L2 round trips per pass.
            constexpr int grp = kXGroup;
            const int lead = lane & ~(grp - 1), sub = lane & (grp - 1);
            const int32_t total_x = *(volatile int32_t*)&s_ctr[3];
            int32_t slot = 0;
            if (sub == 0) slot = atomicAdd(&s_ctr[5], 1);
            slot = __shfl_sync(0xffffffffu, slot, lead);
            while (__any_sync(0xffffffffu, slot < total_x)) {
                const uint32_t e = slot < total_x ? uint32_t(*(volatile uint16_t*)(s_xqueue + slot)) : kEmpty;
                // (sleeping: a spinning warp takes issue slots from the warps that walk the dependency chain)
                if (!__any_sync(0xffffffffu, e != kEmpty)) { __nanosleep(64); continue; }
                if (e != kEmpty) {
                    const uint32_t v = e & uint32_t(kTile - 1);
                    const bool in = (e >> kTileLog) != 0u;
                    const bool slow = (*(volatile int32_t*)(s_state + v) & kSlow) != 0;
                    if (DBG && a.trace && sub == 0) a.trace[4 * (size_t(t0) + v) + 2] = uint32_t(global_ns() - trace_t0);
                    uint32_t first, count;
                    long long p = 0;
                    bool raw = false;
                    if (!slow) { first = s_xptr[v]; count = uint32_t(s_xptr[v + 1]) - first; }
                    else {
                        const uint32_t lc = a.lcnt[t0 + v];
                        p = a.lptr[t0 + v];
                        raw = (lc >> 16) == kUnpartitioned;
                        first = raw ? 0u : (lc >> 16);
                        count = (lc & 0xFFFFu) - first;
                    }
                    for (uint32_t k = sub; k < count; k += grp) {
                        const uint32_t x = slow ? a.lst[p + first + k] : s_ext[first + k];
                        const uint32_t rt = x & rmask;
                        const uint32_t tt = rt >> kTileLog;
                        if (raw && tt == tile) continue;
                        bool tell;
                        if (in) tell = !(atomicOr(a.wext + rt, kOutFlag) & kOutFlag);          // the first knock-out
                        else {
                            const uint32_t mult = (x >> a.order.kbits) + 1u;
                            tell = atomicSub(a.wext + rt, mult) == mult;                         // the one that reaches 0
                        }
                        if (tell) {
                            const unsigned int k2 = atomicAdd(a.epoch + tt, 1u);                 // the mailbox's fill count
                            *(volatile uint16_t*)(a.mbox + (size_t(tt) << kTileLog) + k2) =
                                uint16_t((rt & uint32_t(kTile - 1)) | (in ? uint32_t(kTile) : 0u));
                        }
                    }
                    if (DBG && a.trace && sub == 0) a.trace[4 * (size_t(t0) + v) + 3] = uint32_t(global_ns() - trace_t0);
                    if (sub == 0) slot = atomicAdd(&s_ctr[5], 1);
                }
                slot = __shfl_sync(0xffffffffu, slot, lead);
            }
            if (DBG && a.dbg && lane == 0) atomicMax(a.dbg + 8 * tile + 3, global_ns());
        }
        if (DBG && a.dbg) { __syncthreads(); if (tid == 0) a.dbg[8 * tile + 6] = global_ns(); }
    }
}

template <int MODE>
__global__ void __launch_bounds__(kBlock) sampler_merge_counts_kernel(SamplerArgs a) {
    __shared__ int s_skip;
    if (threadIdx.x == 0) s_skip = (!use_lane_flow(a) || use_raw_build(a)) ? 1 : 0;     // (the raw build leaves nothing to merge)
    __syncthreads();
    if (s_skip) return;
    const uint64_t ranks = MODE == GMG_SAMPLER_EXACT ? uint64_t(a.n) : (uint64_t(1) << a.order.kbits);
    const uint64_t r = uint64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (r < ranks && vertex_of<MODE>(a, uint32_t(r)) < uint32_t(a.n)) a.wext[r] += a.cint[r];
}

// One rank per lane, 32 consecutive ranks per warp ticket; the warp polls its 32 words together (one coalesced
// request per poll). A decided lane applies its whole list with reductions that return nothing.
template <int MODE>
__global__ void __launch_bounds__(kBlock) sampler_lane_flow_kernel(SamplerArgs a) {
    if (!use_lane_flow(a)) return;
    const int lane = threadIdx.x & 31;
    const uint32_t rmask = a.order.kbits >= 32 ? 0xFFFFFFFFu : ((1u << a.order.kbits) - 1u);
    const uint64_t ranks = MODE == GMG_SAMPLER_EXACT ? uint64_t(a.n) : (uint64_t(1) << a.order.kbits);
    for (;;) {
        unsigned int base = 0;
        if (lane == 0) base = atomicAdd(a.ticket + 1, 32u);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (uint64_t(base) >= ranks) break;
        const uint64_t r = uint64_t(base) + lane;
        bool pending = r < ranks && vertex_of<MODE>(a, uint32_t(r)) < uint32_t(a.n);
        long long p = 0;
        uint32_t c = 0;
        if (pending) { p = a.lptr[r]; c = a.lcnt[r] & 0xFFFFu; }
        unsigned ns = 0;
        while (__any_sync(0xffffffffu, pending)) {
            const uint32_t w = pending ? *(volatile uint32_t*)(a.wext + r) : 1u;
            if (pending && (w == 0u || (w & kOutFlag))) {
                pending = false;
                const bool in = w == 0u;
                for (uint32_t k = 0; k < c; k += 4) {
                    uint32_t x[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u) x[u] = k + u < c ? a.lst[p + k + u] : 0xFFFFFFFFu;
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        if (k + u >= c) continue;
                        // a selected vertex knocks out everything it covers (src/sampling.cpp:38,45); a rejected one
                        // releases its later coverees (the last release selects: :25-28)
                        if (in) atomicOr(a.wext + (x[u] & rmask), kOutFlag);
                        else atomicSub(a.wext + (x[u] & rmask), (x[u] >> a.order.kbits) + 1u);
                    }
                }
                a.res[r] = in ? 1 : 0;
                ns = 0;
            } else if (ns) {
                __nanosleep(ns);
            }
            ns = ns == 0 ? 32 : (ns < 512 ? ns << 1 : ns);
        }
    }
}

__global__ void __launch_bounds__(kBlock) sampler_init_kernel(uint64_t ranks, uint8_t* __restrict__ res) {
    const int64_t i = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (uint64_t(i) < ranks) res[i] = 0;
}
template <int MODE>
__global__ void __launch_bounds__(kBlock) flag_in_kernel(SamplerArgs a, int32_t* __restrict__ flag) {
    const int64_t i = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (i < a.n) flag[i] = a.res[rank_of<MODE>(a, int32_t(i))] ? 1 : 0;
}
__global__ void __launch_bounds__(kBlock) compact_kernel(int32_t n, const int32_t* __restrict__ flag,
                                                         const int32_t* __restrict__ pos, int32_t* __restrict__ out) {
    const int64_t i = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (i < n && flag[i]) out[pos[i]] = int32_t(i);
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

constexpr size_t kFlowSmem = sizeof(int32_t) * kTile + sizeof(uint32_t) * kExtCap + sizeof(uint16_t) * (kListCap + 2 * kPtrLen + 2 * kTile);

template <int MODE> static void run_sampler(gmg_ctx* c, SamplerArgs a, uint64_t ranks, int32_t* flag) {
    // probe, then both build kernels: the one the probe does not select returns at once (no host round trip)
    GMG_LAUNCH(c, sampler_coherence_kernel<MODE>, div_up(div_up(int64_t(a.n), 16), kBlock), kBlock, 0, a);
    GMG_LAUNCH(c, sampler_build_kernel<MODE>, div_up(int64_t(a.n) * kLanes, kBlock), kBlock, 0, a);
    GMG_LAUNCH(c, sampler_build_raw_kernel<MODE>, div_up(int64_t(a.n) * kLanes, kBlock), kBlock, 0, a);
    // both flow kernels are launched; the one the build's counts do not select returns at once
    const bool dbg = a.dbg != nullptr || a.trace != nullptr;
    auto flow = dbg ? sampler_flow_kernel<MODE, true> : sampler_flow_kernel<MODE, false>;
    GMG_CUDA(cudaFuncSetAttribute(flow, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kFlowSmem)));
    int per_sm = 0;
    GMG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, flow, kFlowThreads, kFlowSmem));
    GMG_REQUIRE(per_sm > 0, GMG_ERR_CUDA, "sampler flow kernel does not fit on an SM");
    int grid = per_sm * c->sm_count;
    if (grid > int(a.tiles)) grid = int(a.tiles);
    if (!a.force_lane) GMG_LAUNCH(c, flow, grid < 1 ? 1 : grid, kFlowThreads, kFlowSmem, a);
    GMG_LAUNCH(c, sampler_merge_counts_kernel<MODE>, div_up(int64_t(ranks), kBlock), kBlock, 0, a);
    GMG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, sampler_lane_flow_kernel<MODE>, kBlock, 0));
    GMG_REQUIRE(per_sm > 0, GMG_ERR_CUDA, "sampler flow kernel does not fit on an SM");
    grid = per_sm * c->sm_count;
    const int needed = div_up(int64_t(ranks), kBlock);
    if (grid > needed) grid = needed;
    GMG_LAUNCH(c, sampler_lane_flow_kernel<MODE>, grid < 1 ? 1 : grid, kBlock, 0, a);
    GMG_LAUNCH(c, flag_in_kernel<MODE>, div_up(a.n, kBlock), kBlock, 0, a, flag);
}

int32_t disc_sample(gmg_ctx* c, int32_t n, const int32_t* colptr, const int32_t* rowidx, const double* len,
                    double radius, int mode, uint64_t seed, DevBuf<int32_t>& selection, SamplerStats* stats,
                    int64_t path_bound) {
    if (stats) stats->rounds = 0;
    if (n == 0) { selection.alloc(c, 0); return 0; }
    GMG_REQUIRE(mode == GMG_SAMPLER_EXACT || mode == GMG_SAMPLER_PARALLEL || mode == GMG_SAMPLER_HASHED, GMG_ERR_INVALID,
                "unknown sampler mode");
    // bound on the later-path lists: every path has a mirror image, exactly one of the two is "later"
    // [0] list cursor, [1] sum deg^2, [2] tickets (tile flow | lane flow), [3] flags | epoch, [4] same-tile entries, [5] all entries,
    // [6] [7] the coherence probe (same-tile neighbours, neighbours)
    DevBuf<unsigned long long> cursor(c, 8);
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
    const Order order = make_order(n, seed, mode);
    const uint64_t ranks = mode == GMG_SAMPLER_EXACT ? uint64_t(n) : (uint64_t(1) << order.kbits);
    GMG_REQUIRE(ranks <= (uint64_t(1) << 31), GMG_ERR_INVALID, "fastDiscSample: too many vertices for 32-bit ranks");
    DevBuf<uint32_t> wext(c, int64_t(ranks)), cint(c, int64_t(ranks)), lcnt(c, int64_t(ranks)), lst(c, list_cap);
    DevBuf<long long> lptr(c, int64_t(ranks));
    DevBuf<uint8_t> res(c, int64_t(ranks));
    DevBuf<int32_t> flag(c, n);
    const unsigned ntiles = unsigned((ranks + kTile - 1) >> kTileLog);
    DevBuf<unsigned int> epochs(c, int64_t(ntiles));
    epochs.zero();
    DevBuf<uint16_t> mbox;
    if (mode != GMG_SAMPLER_HASHED) {
        mbox.alloc(c, int64_t(ntiles) << kTileLog);
        GMG_CUDA(cudaMemsetAsync(mbox.p, 0xFF, sizeof(uint16_t) * (size_t(ntiles) << kTileLog), c->stream));
    }
    DevBuf<uint32_t> trace;
    const char* trace_path = mode == GMG_SAMPLER_HASHED || n < 1000000 ? nullptr : getenv("GMG_SAMPLER_TRACE");
    DevBuf<unsigned long long> trace_t0;
    if (trace_path) { trace.alloc(c, 4 * int64_t(ranks)); trace.zero(); trace_t0.alloc(c, 1); }
    DevBuf<unsigned long long> dbg;
    const bool debug = getenv("GMG_SAMPLER_DEBUG") != nullptr && mode != GMG_SAMPLER_HASHED;
    if (debug) {
        dbg.alloc(c, 8 * int64_t(ntiles));
        GMG_CUDA(cudaMemsetAsync(dbg.p, 0, sizeof(unsigned long long) * 8 * size_t(ntiles), c->stream));
        // first-decision slots start at the maximum (atomicMin)
        std::vector<unsigned long long> init(size_t(8) * ntiles, 0ull);
        for (unsigned t = 0; t < ntiles; ++t) init[size_t(8) * t + 2] = ~0ull;
        GMG_CUDA(cudaMemcpyAsync(dbg.p, init.data(), sizeof(unsigned long long) * init.size(), cudaMemcpyHostToDevice, c->stream));
        sync(c);
    }
    GMG_LAUNCH(c, sampler_init_kernel, div_up(int64_t(ranks), kBlock), kBlock, 0, ranks, res.p);
    SamplerArgs a{n, colptr, rowidx, len, radius, order, wext.p, cint.p, lptr.p, lcnt.p, lst.p, cursor.p,
                  reinterpret_cast<unsigned int*>(cursor.p + 2), ntiles, res.p,
                  reinterpret_cast<int32_t*>(cursor.p + 3), cursor.p + 4, epochs.p,
                  mode == GMG_SAMPLER_HASHED ? 1 : 0, debug ? dbg.p : nullptr, trace.p, trace_t0.p, mbox.p};
    if (trace_path) GMG_LAUNCH(c, stamp_kernel, 1, 1, 0, trace_t0.p);
    if (mode == GMG_SAMPLER_EXACT) run_sampler<GMG_SAMPLER_EXACT>(c, a, ranks, flag.p);
    else run_sampler<GMG_SAMPLER_PARALLEL>(c, a, ranks, flag.p);
    // ascending compaction of the IN flags
    DevBuf<int64_t> total(c, 1);
    DevBuf<int32_t> pos(c, n);
    exclusive_scan_i32(c, flag.p, pos.p, n, total.p);
    int64_t m = 0;
    unsigned long long hflags = 0;
    total.download(&m, 1);
    GMG_CUDA(cudaMemcpyAsync(&hflags, cursor.p + 3, sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
    sync(c);
    GMG_REQUIRE((hflags & 0xFFFFFFFFull) == 0, GMG_ERR_INVALID, "fastDiscSample: a vertex covers more than 65535 later vertices");
    if (trace_path) {   // per-rank timestamps of the tile flow (diagnostic only)
        std::vector<uint32_t> h(size_t(4) * ranks);
        GMG_CUDA(cudaMemcpyAsync(h.data(), trace.p, sizeof(uint32_t) * h.size(), cudaMemcpyDeviceToHost, c->stream));
        sync(c);
        if (FILE* f = fopen(trace_path, "wb")) { fwrite(h.data(), sizeof(uint32_t), h.size() < 12000000 ? h.size() : 12000000, f); fclose(f); }
    }
    if (debug) {        // timeline of the tile flow (diagnostic only)
        std::vector<unsigned long long> h(size_t(8) * ntiles);
        GMG_CUDA(cudaMemcpyAsync(h.data(), dbg.p, sizeof(unsigned long long) * h.size(), cudaMemcpyDeviceToHost, c->stream));
        sync(c);
        if (h[6] != 0) {
            unsigned long long t_begin = ~0ull, t_end = 0, iters = 0, polls = 0;
            double active = 0.0, wait = 0.0;
            for (unsigned t = 0; t < ntiles; ++t) {
                const unsigned long long* w = h.data() + size_t(8) * t;
                if (w[0] < t_begin) t_begin = w[0];
                if (w[6] > t_end) t_end = w[6];
                iters += w[4]; polls += w[5];
                active += double(w[6] - w[2]); wait += double(w[2] - w[1]);
            }
            fprintf(stderr, "[gmg] sampler tile flow: n %d, %u tiles, %.3f ms; per tile: first decision -> done %.1f us, staged -> first decision %.1f us, "
                            "%.0f mailbox looks, %.0f notifications\n", n, ntiles, double(t_end - t_begin) * 1e-6, active / ntiles * 1e-3,
                    wait / ntiles * 1e-3, double(iters) / ntiles, double(polls) / ntiles);
            for (unsigned t = 0; t < ntiles && t < 400; t += (t < 24 ? 1 : 47)) {
                const unsigned long long* w = h.data() + size_t(8) * t;
                fprintf(stderr, "[gmg]   tile %u block %llu: ticket %.1f staged %.1f first %.1f ext-done %.1f done %.1f us; looks %llu notifications %llu\n", t, w[7],
                        double(w[0] - t_begin) * 1e-3, double(w[1] - t_begin) * 1e-3, double(w[2] - t_begin) * 1e-3, double(w[3] - t_begin) * 1e-3,
                        double(w[6] - t_begin) * 1e-3, w[4], w[5]);
            }
        }
    }
    selection.alloc(c, m);
    GMG_LAUNCH(c, compact_kernel, div_up(n, kBlock), kBlock, 0, n, flag.p, pos.p, selection.p);
    return int32_t(m);
}

enum : int32_t { kUndecided = 0, kIn = 1, kOut = 2 };   // the row-sharded variant's replicated state bytes

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
    uint32_t* dst = a.lst + start + (incl - prev);
    int32_t w = 0;
    for_each_covered(a, i, sub, [&](int32_t t) { if (rank_of<MODE>(a, t) < ri) dst[w++] = uint32_t(t); });
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
    GMG_REQUIRE(mode == GMG_SAMPLER_EXACT || mode == GMG_SAMPLER_PARALLEL || mode == GMG_SAMPLER_HASHED, GMG_ERR_INVALID,
                "unknown sampler mode");
    // the pull rounds need a shallow dependency depth: the row-sharded build hashes every vertex in both
    // non-exact modes (the super-tile sweeps of GMG_SAMPLER_PARALLEL are as deep as the index order)
    if (mode == GMG_SAMPLER_PARALLEL) mode = GMG_SAMPLER_HASHED;
    out.lo = lo; out.hi = hi;
    const int32_t own = hi - lo;
    out.lptr.alloc(c, own); out.lcnt.alloc(c, own);
    if (own <= 0) { out.lst.alloc(c, 0); return; }
    DevBuf<unsigned long long> cursor(c, 2);
    cursor.zero();
    SamplerArgs a{n, colptr, rowidx, len, radius, make_order(n, seed, mode), nullptr, nullptr, out.lptr.p,
                  reinterpret_cast<uint32_t*>(out.lcnt.p), nullptr, cursor.p, nullptr, 0, nullptr, nullptr, nullptr, nullptr, 1, nullptr};
    if (mode == GMG_SAMPLER_EXACT) GMG_LAUNCH(c, shard_path_bound_kernel<GMG_SAMPLER_EXACT>, div_up(own, kBlock), kBlock, 0, a, lo, hi, cursor.p + 1);
    else GMG_LAUNCH(c, shard_path_bound_kernel<GMG_SAMPLER_PARALLEL>, div_up(own, kBlock), kBlock, 0, a, lo, hi, cursor.p + 1);
    unsigned long long hb[2];
    cursor.download(hb, 2);
    sync(c);
    out.lst.alloc(c, int64_t(hb[1]) + 1);
    a.lst = reinterpret_cast<uint32_t*>(out.lst.p);
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
