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

// Precedence order. EXACT: rank(v) = v. Otherwise a vertex is cut into (super-tile, offset) = (v >> tshift,
// v & (2^tshift - 1)) and the SUPER-TILE number goes through a seeded invertible mixing of kt-bit integers:
// xor-seed, then (odd multiply, xorshift by ceil(kt/2)) twice and a last odd multiply, all modulo 2^kt; every
// step is a bijection on kt bits and the xorshift is its own inverse because twice the shift is >= kt.
//   GMG_SAMPLER_HASHED    tshift = 0: every vertex hashed, dependency depth O(log n)
//   GMG_SAMPLER_PARALLEL  tshift = ceil(log2(n / 3)): the reference's index-order sweep restarted in 2-3
//                         super-tiles taken in hashed order (why: DESIGN.md section 4, "sampler orders")
struct Order {
    uint32_t mask, shift, seed;
    uint32_t c1, c2, c3, i1, i2, i3;   // odd multipliers and their inverses modulo 2^32
    uint32_t tshift;                   // log2 of the super-tile size
    uint32_t kbits;                    // bits of a rank (kt + tshift); ranks live in [0, 2^kbits)
};
__host__ __device__ __forceinline__ uint32_t order_rank(const Order& o, uint32_t v) {
    uint32_t x = ((v >> o.tshift) ^ o.seed) & o.mask;
    x = (x * o.c1) & o.mask; x ^= x >> o.shift;
    x = (x * o.c2) & o.mask; x ^= x >> o.shift;
    x = (x * o.c3) & o.mask;
    return (x << o.tshift) | (v & ((1u << o.tshift) - 1u));
}
__host__ __device__ __forceinline__ uint32_t order_unrank(const Order& o, uint32_t r) {
    uint32_t x = ((r >> o.tshift) * o.i3) & o.mask;
    x ^= x >> o.shift; x = (x * o.i2) & o.mask;
    x ^= x >> o.shift; x = (x * o.i1) & o.mask;
    x = (x ^ o.seed) & o.mask;
    return (x << o.tshift) | (r & ((1u << o.tshift) - 1u));
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
    o.tshift = mode == GMG_SAMPLER_PARALLEL ? uint32_t(ceil_log2((int64_t(n) + 2) / 3)) : 0u;
    const int64_t tiles = (int64_t(n) + (int64_t(1) << o.tshift) - 1) >> o.tshift;
    const int kt = ceil_log2(tiles);
    o.mask = kt >= 32 ? 0xFFFFFFFFu : ((1u << kt) - 1u);
    o.shift = uint32_t((kt + 1) / 2);
    if (o.shift == 0) o.shift = 1;
    const uint64_t z = (seed + 0x9E3779B97F4A7C15ull) * 0xBF58476D1CE4E5B9ull;
    o.seed = uint32_t(z >> 32) & o.mask;
    o.c1 = 0x9E3779B1u; o.c2 = 0x85EBCA6Bu | 1u; o.c3 = 0xC2B2AE35u | 1u;
    o.i1 = inverse_odd(o.c1); o.i2 = inverse_odd(o.c2); o.i3 = inverse_odd(o.c3);
    o.kbits = uint32_t(kt) + o.tshift;
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
constexpr int kTileLog = 12;
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
    unsigned int* epoch;         // [tiles] word t is bumped by whoever pushed an effect INTO tile t: what a waiting tile watches
    int force_lane;              // GMG_SAMPLER_HASHED: no tile ever holds two coverers, always the per-lane flow
    unsigned long long* dbg;     // GMG_SAMPLER_DEBUG: 8 words per tile (timeline of the tile flow), else null
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

// The flow: a vertex is IN once every preceding coverer is OUT (both path counters at 0), OUT once a preceding
// coverer is IN. Two kernels; which one does the work is decided on the device (use_lane_flow).
//
// (1) TILE flow, for spatially coherent orders. A block takes flow TILES (kTile consecutive ranks) by ticket, so
// every tile that precedes a running tile is running or done and the waits below always end. Inside a tile the
// decisions travel through SHARED memory: both halves of the tile's lists are staged there once (they are
// static, so this happens BEFORE the tile has anything to wait for), the same-tile counters live there, and a
// round -- every thread looks at its undecided ranks, a newly decided rank applies its same-tile list -- costs a
// barrier instead of a round trip through L2. That is what the index order of a coherent numbering needs: on a
// row-major 10000 x 1000 grid a tile is four rows, its preceding coverers sit in the tile before it, and the
// dependency chain (24,000 steps at level 0) is walked at shared-memory speed.
// What crosses tiles goes through the global words: a selected rank ORs the OUT bit into its later coverees, a
// rejected one subtracts its multiplicity, both without a return value. A tile that can make no progress does
// not poll its 4096 words (148 waiting tiles doing that saturate the L2 and slow the one tile everybody waits
// for): whoever pushes an effect into a tile bumps that tile's epoch word, and a waiting tile watches its own word
// only -- a wake-up goes to the tile it concerns and to nobody else.
//
// (2) LANE flow, for incoherent orders (hashed priorities, subdivision orders): nearly every target sits in
// another tile, the dependency depth is O(log n), and one rank per lane with no block-wide barriers is faster.
template <int MODE>
__global__ void __launch_bounds__(kFlowThreads, 1) sampler_flow_kernel(SamplerArgs a) {
    extern __shared__ unsigned char smem_raw[];
    int32_t* s_rem = reinterpret_cast<int32_t*>(smem_raw);                       // kTile same-tile path counters
    uint32_t* s_ext = reinterpret_cast<uint32_t*>(s_rem + kTile);                // kExtCap staged other-tile entries
    uint16_t* s_list = reinterpret_cast<uint16_t*>(s_ext + kExtCap);             // kListCap staged same-tile entries
    uint8_t* s_hit = reinterpret_cast<uint8_t*>(s_list + kListCap);              // kTile: knocked OUT inside the tile
    __shared__ int32_t s_warp[2][kFlowThreads / 32];
    __shared__ unsigned int s_tile;
    __shared__ unsigned int s_epoch;      // the last value of the global epoch word this block has looked at
    __shared__ int32_t s_left;            // undecided ranks of the tile
    if (use_lane_flow(a)) return;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t rmask = a.order.kbits >= 32 ? 0xFFFFFFFFu : ((1u << a.order.kbits) - 1u);
    const uint64_t ranks = MODE == GMG_SAMPLER_EXACT ? uint64_t(a.n) : (uint64_t(1) << a.order.kbits);
    for (;;) {
        __syncthreads();                                   // everybody is done with the previous tile's shared state
        if (tid == 0) { s_tile = atomicAdd(a.ticket, 1u); s_left = 0; s_epoch = 0; }
        __syncthreads();
        const unsigned int tile = s_tile;
        if (tile >= a.tiles) break;
        const uint32_t t0 = tile << kTileLog;
        const unsigned int* my_epoch = a.epoch + tile;
        if (a.dbg && tid == 0) a.dbg[8 * tile + 0] = global_ns();
        // ---- static data of the tile's ranks; rank i of the tile belongs to thread i % kFlowThreads
        long long ptr[kOwn];
        uint32_t tot[kOwn], nin[kOwn], soff[kOwn], xoff[kOwn];
        int32_t mine = 0, mine_x = 0;  // same-tile / other-tile list entries this thread wants to stage
        unsigned undecided = 0;        // bit q: rank q of this thread is not decided yet
#pragma unroll
        for (int q = 0; q < kOwn; ++q) {
            const int i = tid + q * kFlowThreads;
            const uint64_t r = uint64_t(t0) + uint64_t(i);
            const bool valid = r < ranks && vertex_of<MODE>(a, uint32_t(r)) < uint32_t(a.n);
            ptr[q] = 0; tot[q] = 0; nin[q] = 0;
            int32_t rem = 0;
            if (valid) {
                const uint32_t lc = a.lcnt[r];
                ptr[q] = a.lptr[r]; tot[q] = lc & 0xFFFFu; nin[q] = lc >> 16;
                undecided |= 1u << q;
                rem = int32_t(a.cint[r]);
                if (nin[q] != kUnpartitioned) { mine += int32_t(nin[q]); mine_x += int32_t(tot[q] - nin[q]); }
            }
            s_rem[i] = rem;
            s_hit[i] = 0;
        }
        if (undecided) atomicAdd(&s_left, __popc(undecided));
        // shared-memory segments for the two halves of the lists: block-wide exclusive scans over the threads
        {
            int32_t s = mine, sx = mine_x;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int32_t t = __shfl_up_sync(0xffffffffu, s, d);
                const int32_t tx = __shfl_up_sync(0xffffffffu, sx, d);
                if (lane >= d) { s += t; sx += tx; }
            }
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
                const bool part = ((undecided >> q) & 1u) && nin[q] != kUnpartitioned;
                const int32_t want = part ? int32_t(nin[q]) : 0, wantx = part ? int32_t(tot[q] - nin[q]) : 0;
                soff[q] = (want > 0 && off + want <= kListCap) ? uint32_t(off) : 0xFFFFFFFFu;
                xoff[q] = (wantx > 0 && offx + wantx <= kExtCap) ? uint32_t(offx) : 0xFFFFFFFFu;
                off += want; offx += wantx;
            }
        }
        // ---- stage the lists: same-tile entries as uint16 (offset in the tile | (multiplicity - 1) << 12), the
        //      others as they are
#pragma unroll
        for (int q = 0; q < kOwn; ++q) {
            if (soff[q] != 0xFFFFFFFFu)
                for (uint32_t e = 0; e < nin[q]; ++e) {
                    const uint32_t x = a.lst[ptr[q] + e];
                    s_list[soff[q] + e] = uint16_t(((x & rmask) - t0) | ((x >> a.order.kbits) << kTileLog));
                }
            if (xoff[q] != 0xFFFFFFFFu)
                for (uint32_t e = nin[q]; e < tot[q]; ++e) s_ext[xoff[q] + (e - nin[q])] = a.lst[ptr[q] + e];
        }
        __syncthreads();
        // ---- the flow: no barrier inside. Every thread spins over its undecided ranks; a decision is applied at
        //      once (same-tile targets in shared memory, the others with reductions on the global words), and a
        //      thread that pushed into other tiles bumps their epoch words so that they look again.
        if (a.dbg && tid == 0) a.dbg[8 * tile + 1] = global_ns();
        unsigned dbg_iter = 0, dbg_poll = 0;
        bool dbg_first = true;
        uint32_t ext[kOwn];
#pragma unroll
        for (int q = 0; q < kOwn; ++q) ext[q] = 1u;                    // 1: the global word has not been read yet
        unsigned int seen = 0xFFFFFFFFu;                               // forces the first read
        unsigned idle = 0, watch = 0;
        while (*(volatile int32_t*)&s_left > 0) {
            // one thread of the block watches the epoch word (with a back-off while the block is making no progress)
            if (tid == 0 && (dbg_iter & 3u) == 0u) {
                // (every 4096th look also forces a fresh read of the words, whatever the epoch says: a lost
                // wake-up must cost microseconds, never the kernel)
                const unsigned int e = ((*(volatile const unsigned int*)my_epoch << 1) | ((++watch >> 12) & 1u)) & 0x7FFFFFFFu;
                if (e != *(volatile unsigned int*)&s_epoch) *(volatile unsigned int*)&s_epoch = e;
            }
            ++dbg_iter;
            const unsigned int now = *(volatile unsigned int*)&s_epoch;
            if (now != seen) {
                ++dbg_poll;
                // the epoch was read BEFORE these words: a push that lands after this poll bumps it again
                seen = now;
                __threadfence();
#pragma unroll
                for (int q = 0; q < kOwn; ++q)     // a word that has reached 0 has no outside coverer left and stays 0
                    if (((undecided >> q) & 1u) && ext[q] != 0u) ext[q] = *(volatile uint32_t*)(a.wext + t0 + tid + q * kFlowThreads);
            }
            bool decided_any = false;
            uint32_t wake[4];                      // tiles this thread pushed into during this pass
            int nwake = 0;
#pragma unroll
            for (int q = 0; q < kOwn; ++q) {
                if (!((undecided >> q) & 1u)) continue;
                const int i = tid + q * kFlowThreads;
                bool in;
                if ((ext[q] & kOutFlag) || *(volatile uint8_t*)(s_hit + i)) in = false;
                else if (ext[q] == 0u && *(volatile int32_t*)(s_rem + i) == 0) in = true;
                else continue;
                undecided &= ~(1u << q);
                decided_any = true;
                if (a.dbg && dbg_first) { dbg_first = false; atomicMin(a.dbg + 8 * tile + 2, global_ns()); }
                // a selected vertex knocks out everything it covers (src/sampling.cpp:38,45); a rejected one releases
                // its later coverees
                const bool raw = nin[q] == kUnpartitioned;
                if (soff[q] != 0xFFFFFFFFu) {
                    for (uint32_t e = 0; e < nin[q]; ++e) {
                        const uint32_t y = s_list[soff[q] + e];
                        if (in) *(volatile uint8_t*)(s_hit + (y & (kTile - 1))) = 1;
                        else atomicSub(s_rem + (y & (kTile - 1)), int32_t(y >> kTileLog) + 1);
                    }
                } else {
                    // list not staged (no room, or not partitioned): same-tile targets straight from the list
                    const uint32_t cnt = raw ? tot[q] : nin[q];
                    for (uint32_t e = 0; e < cnt; ++e) {
                        const uint32_t x = a.lst[ptr[q] + e];
                        const uint32_t rt = x & rmask;
                        if ((rt >> kTileLog) != tile) continue;
                        if (in) *(volatile uint8_t*)(s_hit + (rt - t0)) = 1;
                        else atomicSub(s_rem + (rt - t0), int32_t(x >> a.order.kbits) + 1);
                    }
                }
                const uint32_t first = raw ? 0u : nin[q];
                for (uint32_t e = first; e < tot[q]; ++e) {
                    const uint32_t x = xoff[q] != 0xFFFFFFFFu ? s_ext[xoff[q] + (e - first)] : a.lst[ptr[q] + e];
                    const uint32_t rt = x & rmask;
                    const uint32_t tt = rt >> kTileLog;
                    if (raw && tt == tile) continue;
                    if (in) atomicOr(a.wext + rt, kOutFlag);
                    else atomicSub(a.wext + rt, (x >> a.order.kbits) + 1u);
                    bool known = false;
#pragma unroll
                    for (int u = 0; u < 4; ++u) known = known || (u < nwake && wake[u] == tt);
                    if (!known) {
                        if (nwake == 4) {          // rare: flush what is noted so far (its pushes are fenced first)
                            __threadfence();
#pragma unroll
                            for (int u = 0; u < 4; ++u) atomicAdd(a.epoch + wake[u], 1u);
                            nwake = 0;
                        }
#pragma unroll
                        for (int u = 0; u < 4; ++u) if (u == nwake) wake[u] = tt;
                        ++nwake;
                    }
                }
                a.res[t0 + i] = in ? 1 : 0;
                atomicSub(&s_left, 1);
            }
            if (nwake) {                           // the pushes first, then the wake-ups
                __threadfence();
#pragma unroll
                for (int u = 0; u < 4; ++u) if (u < nwake) atomicAdd(a.epoch + wake[u], 1u);
            }
            // a thread that is done with the tile meets the block at the barrier; thread 0 stays: it is the watcher
            if (!undecided && tid != 0) { if (a.dbg) atomicMax(a.dbg + 8 * tile + 3, global_ns()); break; }
            // no sleeping here: a waiting rank must react within one pass (the dependency chain of a row-major grid
            // is tens of thousands of passes long); only a tile that has waited very long backs off
            if (decided_any) idle = 0;
            else if (++idle > 20000) __nanosleep(200);
        }
        if (a.dbg && tid == 0) { a.dbg[8 * tile + 4] = dbg_iter; a.dbg[8 * tile + 5] = dbg_poll; a.dbg[8 * tile + 6] = global_ns(); a.dbg[8 * tile + 7] = blockIdx.x; }
    }
}

template <int MODE>
__global__ void __launch_bounds__(kBlock) sampler_merge_counts_kernel(SamplerArgs a) {
    if (!use_lane_flow(a)) return;
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

constexpr size_t kFlowSmem = sizeof(int32_t) * kTile + sizeof(uint32_t) * kExtCap + sizeof(uint16_t) * kListCap + kTile;

template <int MODE> static void run_sampler(gmg_ctx* c, SamplerArgs a, uint64_t ranks, int32_t* flag) {
    GMG_LAUNCH(c, sampler_build_kernel<MODE>, div_up(int64_t(a.n) * kLanes, kBlock), kBlock, 0, a);
    // both flow kernels are launched; the one the build's counts do not select returns at once
    GMG_CUDA(cudaFuncSetAttribute(sampler_flow_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kFlowSmem)));
    int per_sm = 0;
    GMG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, sampler_flow_kernel<MODE>, kFlowThreads, kFlowSmem));
    GMG_REQUIRE(per_sm > 0, GMG_ERR_CUDA, "sampler flow kernel does not fit on an SM");
    int grid = per_sm * c->sm_count;
    if (grid > int(a.tiles)) grid = int(a.tiles);
    if (!a.force_lane) GMG_LAUNCH(c, sampler_flow_kernel<MODE>, grid < 1 ? 1 : grid, kFlowThreads, kFlowSmem, a);
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
    // [0] list cursor, [1] sum deg^2, [2] tickets (tile flow | lane flow), [3] flags | epoch, [4] same-tile entries, [5] all entries
    DevBuf<unsigned long long> cursor(c, 6);
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
                  mode == GMG_SAMPLER_HASHED ? 1 : 0, debug ? dbg.p : nullptr};
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
                            "%.0f spin iterations, %.0f polls\n", n, ntiles, double(t_end - t_begin) * 1e-6, active / ntiles * 1e-3,
                    wait / ntiles * 1e-3, double(iters) / ntiles, double(polls) / ntiles);
            for (unsigned t = 0; t < ntiles && t < 400; t += (t < 24 ? 1 : 47)) {
                const unsigned long long* w = h.data() + size_t(8) * t;
                fprintf(stderr, "[gmg]   tile %u block %llu: ticket %.1f staged %.1f first %.1f last-thread %.1f done %.1f us; iters %llu polls %llu\n", t, w[7],
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
