// Cluster CSR, coarse graph and coarse positions.
//   build_clusters  : children of every coarse point in ascending fine index (the std::set order
//                     of reference src/multigrid.cpp:179-181)
//   coarse_edges    : GravoMG::extractCoarseEdges, src/multigrid.cpp:135-169
//   coarse_points   : GravoMG::coarseFromMeanOfFineChildren, src/multigrid.cpp:171-207
#include "stages.h"

namespace gmg {

// ---------------------------------------------------------------------------------------------
// Cluster CSR: histogram -> scan -> scatter -> per-cluster sort (clusters hold ~8 children).
// Algorithmic bytes: 8n + 4m.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlock) cluster_count_kernel(int32_t n, const int32_t* __restrict__ parents,
                                                               int32_t* __restrict__ size) {
    const int64_t i = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (i < n) atomicAdd(size + __ldg(parents + i), 1);
}
__global__ void __launch_bounds__(kBlock) cluster_scatter_kernel(int32_t n, const int32_t* __restrict__ parents,
                                                                 const int32_t* __restrict__ ptr,
                                                                 int32_t* __restrict__ cursor,
                                                                 int32_t* __restrict__ child) {
    const int64_t i = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (i >= n) return;
    const int32_t p = __ldg(parents + i);
    child[__ldg(ptr + p) + atomicAdd(cursor + p, 1)] = int32_t(i);
}
// In-place ascending sort of every segment [ptr[s], ptr[s+1]); one thread per segment.
__global__ void __launch_bounds__(kBlock) segment_sort_kernel(int32_t segments, const int32_t* __restrict__ ptr,
                                                              int32_t* __restrict__ data) {
    const int64_t s = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (s >= segments) return;
    const int32_t b = ptr[s], e = ptr[s + 1];
    for (int32_t i = b + 1; i < e; ++i) {
        const int32_t v = data[i];
        int32_t j = i - 1;
        while (j >= b && data[j] > v) { data[j + 1] = data[j]; --j; }
        data[j + 1] = v;
    }
}

void build_clusters(gmg_ctx* c, int32_t n, int32_t m, const int32_t* parents, Clusters& out) {
    out.ptr.alloc(c, int64_t(m) + 1);
    out.child.alloc(c, n);
    DevBuf<int32_t> size(c, int64_t(m) + 1);
    size.zero();
    if (n > 0) GMG_LAUNCH(c, cluster_count_kernel, div_up(n, kBlock), kBlock, 0, n, parents, size.p);
    exclusive_scan_i32(c, size.p, out.ptr.p, int64_t(m) + 1, nullptr);
    size.zero();
    if (n > 0) {
        GMG_LAUNCH(c, cluster_scatter_kernel, div_up(n, kBlock), kBlock, 0, n, parents, out.ptr.p, size.p, out.child.p);
        if (m > 0) GMG_LAUNCH(c, segment_sort_kernel, div_up(m, kBlock), kBlock, 0, m, out.ptr.p, out.child.p);
    }
}

// ---------------------------------------------------------------------------------------------
// Coarse graph. One warp (or one block for oversized neighbourhoods) per cluster P = ROW P of the
// coarse matrix. For each child `fine` and stored neighbour nb with Q = parent[nb] != P:
//     t = coeff(fine, P) + value(nb, fine)                                   (:150-151)
// where coeff(fine, P) reads FINE column number P (the reference's index-space quirk, kept).
// Distinct Q are collected in a shared-memory hash with an ordered-uint64 atomicMin on t, then
// sorted. Row P is written to a scratch segment bounded by the cluster's degree sum; the
// column-major result is a gather: column P has the same pattern as row P (symmetric pattern) and
// takes its values from the rows of its neighbours.
// The reference's "a stored 0 means empty" rule (:156-164) only differs from a plain minimum when
// some t is not > 0; that case is detected and recomputed by a sequential in-order fold.
// Algorithmic bytes: 8n + 12nnz + 4m + 12nnz_c.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned long long ordered_bits(double v) {
    const unsigned long long b = (unsigned long long)__double_as_longlong(v);
    return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __forceinline__ double from_ordered_bits(unsigned long long o) {
    const unsigned long long b = (o >> 63) ? (o & 0x7FFFFFFFFFFFFFFFull) : ~o;
    return __longlong_as_double((long long)b);
}

// q0[fine] = fine_edge_matrix.coeff(fine, parent[fine]), for the children of the clusters [p_lo, p_hi)
__global__ void __launch_bounds__(kBlock) quirk_term_kernel(int32_t n, int32_t p_lo, int32_t p_hi,
                                                            const int32_t* __restrict__ colptr,
                                                            const int32_t* __restrict__ rowidx,
                                                            const double* __restrict__ val,
                                                            const int32_t* __restrict__ parents,
                                                            double* __restrict__ q0) {
    const int64_t f = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (f >= n) return;
    const int32_t P = __ldg(parents + f);
    if (P < p_lo || P >= p_hi) { q0[f] = 0.0; return; }
    const int32_t b = __ldg(colptr + P), e = __ldg(colptr + P + 1);
    const int32_t pos = lower_bound_i32(rowidx, b, e, int32_t(f));
    q0[f] = (pos < e && __ldg(rowidx + pos) == int32_t(f)) ? __ldg(val + pos) : 0.0;
}

// degsum[P - p_lo] for the clusters [p_lo, p_hi)
__global__ void __launch_bounds__(kBlock) cluster_degree_kernel(int32_t p_lo, int32_t p_hi, const int32_t* __restrict__ cptr,
                                                                const int32_t* __restrict__ child,
                                                                const int32_t* __restrict__ colptr,
                                                                int32_t* __restrict__ degsum) {
    const int64_t P = int64_t(p_lo) + int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (P >= p_hi) return;
    int32_t s = 0;
    for (int32_t i = cptr[P]; i < cptr[P + 1]; ++i) { const int32_t f = child[i]; s += colptr[f + 1] - colptr[f]; }
    degsum[P - p_lo] = s;
}

struct CoarseArgs {
    int32_t p_lo, p_hi;       // clusters handled by this call (all of them, or the rank's share)
    const int32_t* cptr;
    const int32_t* child;
    const int32_t* colptr;
    const int32_t* rowidx;
    const double* val;
    const int32_t* parents;
    const double* q0;
    const int32_t* boff;      // scratch offset of row P (pointer pre-shifted so that boff[P] is valid on [p_lo, p_hi])
    int32_t* sQ;              // scratch: sorted neighbour clusters of row P
    double* st;               // scratch: min t per neighbour
    int32_t* cnt;             // distinct neighbours of row P
    int32_t* flags;           // [0] some t not > 0, [1] overflow list tail, [2] hard overflow
    int32_t* overflow;        // clusters whose neighbourhood did not fit the warp-level table
};

// THREADS cooperate on one cluster with a table of SLOTS entries in shared memory.
template <int THREADS, int SLOTS>
__device__ __forceinline__ bool coarse_row(const CoarseArgs& a, int32_t P, int tid, int32_t* keys,
                                           unsigned long long* vals, int32_t* scount) {
    for (int s = tid; s < SLOTS; s += THREADS) { keys[s] = -1; vals[s] = ~0ull; }
    if (tid == 0) *scount = 0;
    if (THREADS == 32) __syncwarp(); else __syncthreads();
    bool nonpos = false;
    const int32_t cb = __ldg(a.cptr + P), ce = __ldg(a.cptr + P + 1);
    // 8 lanes per child (mesh columns hold ~7 entries), THREADS/8 children in flight
    constexpr int kSub = 8;
    for (int32_t ci = cb + tid / kSub; ci < ce; ci += THREADS / kSub) {
        const int32_t f = __ldg(a.child + ci);
        const double base = __ldg(a.q0 + f);
        const int32_t b = __ldg(a.colptr + f), e = __ldg(a.colptr + f + 1);
        for (int32_t k = b + tid % kSub; k < e; k += kSub) {
            const int32_t Q = __ldg(a.parents + __ldg(a.rowidx + k));
            if (Q == P) continue;                                                 // :148
            const double t = base + __ldg(a.val + k);                            // :150-151
            if (!(t > 0.0)) nonpos = true;
            uint32_t h = (uint32_t(Q) * 0x9E3779B1u) >> 7;
            for (int probe = 0; probe < SLOTS; ++probe, ++h) {
                const int s = int(h & (SLOTS - 1));
                const int32_t prev = atomicCAS(keys + s, -1, Q);
                if (prev == -1) atomicAdd(scount, 1);
                if (prev == -1 || prev == Q) { atomicMin(vals + s, ordered_bits(t)); break; }
            }
        }
    }
    if (THREADS == 32) __syncwarp(); else __syncthreads();
    const int32_t K = *scount;
    if (nonpos) a.flags[0] = 1;
    if (K > SLOTS * 3 / 4) return false;   // too full to trust probing: let the caller escalate
    const int32_t off = __ldg(a.boff + P);
    if constexpr (THREADS == 32) {
        // compact the occupied slots to the front (in place: a slot only ever moves down, and every
        // chunk is read before it is written), then rank-sort the K survivors: K*K/32 steps per lane
        // instead of SLOTS*SLOTS/32
        int32_t base = 0;
        for (int s0 = 0; s0 < SLOTS; s0 += 32) {
            const int32_t q = keys[s0 + tid];
            const unsigned long long v = vals[s0 + tid];
            const unsigned occ = __ballot_sync(0xffffffffu, q >= 0);
            __syncwarp();
            if (q >= 0) { const int pos = base + __popc(occ & ((1u << tid) - 1u)); keys[pos] = q; vals[pos] = v; }
            base += __popc(occ);
            __syncwarp();
        }
        for (int s = tid; s < K; s += 32) {
            const int32_t q = keys[s];
            int32_t rank = 0;
            for (int s2 = 0; s2 < K; ++s2) rank += keys[s2] < q ? 1 : 0;
            a.sQ[off + rank] = q;
            a.st[off + rank] = from_ordered_bits(vals[s]);
        }
        if (tid == 0) a.cnt[P] = K;
        return true;
    }
    // rank sort of the occupied slots by key into the scratch row
    for (int s = tid; s < SLOTS; s += THREADS) {
        const int32_t q = keys[s];
        if (q < 0) continue;
        int32_t rank = 0;
        for (int s2 = 0; s2 < SLOTS; ++s2) { const int32_t q2 = keys[s2]; rank += (q2 >= 0 && q2 < q) ? 1 : 0; }
        a.sQ[off + rank] = q;
        a.st[off + rank] = from_ordered_bits(vals[s]);
    }
    if (tid == 0) a.cnt[P] = K;
    return true;
}

constexpr int kWarpSlots = 128;
constexpr int kBlockSlots = 2048;

__global__ void __launch_bounds__(kBlock) coarse_rows_warp_kernel(CoarseArgs a) {
    __shared__ int32_t keys[kBlock / 32][kWarpSlots];
    __shared__ unsigned long long vals[kBlock / 32][kWarpSlots];
    __shared__ int32_t scount[kBlock / 32];
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t P = int64_t(a.p_lo) + int64_t(blockIdx.x) * (kBlock / 32) + w;
    if (P >= a.p_hi) return;
    if (!coarse_row<32, kWarpSlots>(a, int32_t(P), lane, keys[w], vals[w], &scount[w])) {
        if (lane == 0) { a.cnt[P] = 0; a.overflow[atomicAdd(a.flags + 1, 1)] = int32_t(P); }
    }
}

__global__ void __launch_bounds__(kBlock) coarse_rows_block_kernel(CoarseArgs a, int32_t* __restrict__ hard_overflow) {
    __shared__ int32_t keys[kBlockSlots];
    __shared__ unsigned long long vals[kBlockSlots];
    __shared__ int32_t scount;
    const int32_t n_overflow = a.flags[1];
    for (int32_t o = blockIdx.x; o < n_overflow; o += gridDim.x) {
        const int32_t P = a.overflow[o];
        if (!coarse_row<kBlock, kBlockSlots>(a, P, threadIdx.x, keys, vals, &scount)) {
            if (threadIdx.x == 0) *hard_overflow = 1;
        }
        __syncthreads();
    }
}

// Exact in-order fold of the reference (:156-164) into the already-known pattern of row P.
__global__ void __launch_bounds__(kBlock) coarse_rows_strict_kernel(CoarseArgs a) {
    const int64_t P = int64_t(a.p_lo) + int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (P >= a.p_hi || a.flags[0] == 0) return;                   // only when some candidate was not > 0
    const int32_t off = a.boff[P], K = a.cnt[P];
    for (int32_t i = 0; i < K; ++i) a.st[off + i] = 0.0;          // absent == reads as 0
    for (int32_t ci = a.cptr[P]; ci < a.cptr[P + 1]; ++ci) {      // children ascending = fine ascending
        const int32_t f = a.child[ci];
        const double base = a.q0[f];
        for (int32_t k = a.colptr[f]; k < a.colptr[f + 1]; ++k) {
            const int32_t Q = a.parents[a.rowidx[k]];
            if (Q == int32_t(P)) continue;
            const double t = base + a.val[k];
            const int32_t pos = lower_bound_i32(a.sQ, off, off + K, Q);
            const double cur = a.st[pos];
            if (cur == 0) a.st[pos] = t;
            else a.st[pos] = (t < cur) ? t : cur;                 // std::min(cur, t)
        }
    }
}

// Column P of the result: rows = neighbours of P (symmetric pattern), value (Q,P) from row Q.
__global__ void __launch_bounds__(kBlock) coarse_gather_kernel(int32_t m, const int32_t* __restrict__ boff,
                                                               const int32_t* __restrict__ cnt,
                                                               const int32_t* __restrict__ sQ,
                                                               const double* __restrict__ st,
                                                               const int32_t* __restrict__ colptr_out,
                                                               int32_t* __restrict__ rowidx_out,
                                                               double* __restrict__ val_out,
                                                               int32_t* __restrict__ asym_flag) {
    constexpr int G = 8;
    const int64_t gid = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    const int32_t P = int32_t(gid / G);
    const int sub = int(gid % G);
    if (P >= m) return;
    const int32_t off = boff[P], K = cnt[P], dst = colptr_out[P];
    for (int32_t e = sub; e < K; e += G) {
        const int32_t Q = sQ[off + e];
        const int32_t qoff = boff[Q], qK = cnt[Q];
        const int32_t pos = lower_bound_i32(sQ, qoff, qoff + qK, P);
        rowidx_out[dst + e] = Q;
        if (pos < qoff + qK && sQ[pos] == P) val_out[dst + e] = st[pos];
        else { val_out[dst + e] = 0.0; *asym_flag = 1; }
    }
}

// Rows [p_lo, p_hi) of the coarse matrix into degree-bounded scratch segments: out.cnt[P] entries at
// out.boff[P - p_lo], sorted by neighbour cluster.
void coarse_rows_range(gmg_ctx* c, int32_t n, const int32_t* colptr, const int32_t* rowidx, const double* val,
                       int32_t p_lo, int32_t p_hi, const int32_t* parents, const Clusters& cl, int32_t* cnt,
                       CoarseRows& out, int32_t* d_hard_overflow) {
    const int32_t own = p_hi - p_lo;
    out.p_lo = p_lo; out.p_hi = p_hi;
    out.boff.alloc(c, int64_t(own) + 1);
    if (own <= 0) { out.sQ.alloc(c, 0); out.st.alloc(c, 0); out.bound = 0; return; }
    DevBuf<double> q0(c, n);
    GMG_LAUNCH(c, quirk_term_kernel, div_up(n, kBlock), kBlock, 0, n, p_lo, p_hi, colptr, rowidx, val, parents, q0.p);
    DevBuf<int32_t> degsum(c, int64_t(own) + 1), flags(c, 4), overflow(c, own);
    degsum.zero(); flags.zero();
    GMG_CUDA(cudaMemsetAsync(cnt + p_lo, 0, sizeof(int32_t) * size_t(own), c->stream));
    GMG_LAUNCH(c, cluster_degree_kernel, div_up(own, kBlock), kBlock, 0, p_lo, p_hi, cl.ptr.p, cl.child.p, colptr, degsum.p);
    DevBuf<int64_t> total(c, 1);
    exclusive_scan_i32(c, degsum.p, out.boff.p, int64_t(own) + 1, total.p);
    if (out.bound < 0) {                      // caller does not know the number of candidates
        total.download(&out.bound, 1);
        sync(c);
    }
    out.sQ.alloc(c, out.bound);
    out.st.alloc(c, out.bound);
    CoarseArgs a{p_lo, p_hi, cl.ptr.p, cl.child.p, colptr, rowidx, val, parents, q0.p, out.boff.p - p_lo, out.sQ.p, out.st.p,
                 cnt, flags.p, overflow.p};
    GMG_LAUNCH(c, coarse_rows_warp_kernel, div_up(own, kBlock / 32), kBlock, 0, a);
    // The two repair passes read their trigger on the device and return at once when it is clear, so the
    // host never waits for the flags: oversized neighbourhoods (flags[1] clusters), and the in-order fold
    // for a candidate that is not > 0 (flags[0]).
    GMG_LAUNCH(c, coarse_rows_block_kernel, 2 * c->sm_count, kBlock, 0, a, d_hard_overflow);
    GMG_LAUNCH(c, coarse_rows_strict_kernel, div_up(own, kBlock), kBlock, 0, a);
}

// Column-major coarse matrix from ALL rows: rows of P at boff[P] with cnt[P] entries each.
void coarse_assemble(gmg_ctx* c, int32_t m, const int32_t* boff, const int32_t* cnt, const int32_t* sQ, const double* st,
                     DeviceCsc& out, int32_t* d_asym) {
    out.n = m;
    out.colptr.alloc(c, int64_t(m) + 1);
    DevBuf<int64_t> total(c, 1);
    exclusive_scan_i32(c, cnt, out.colptr.p, int64_t(m) + 1, total.p);
    total.download(&out.nnz, 1);
    sync(c);
    out.rowidx.alloc(c, out.nnz);
    out.val.alloc(c, out.nnz);
    GMG_LAUNCH(c, coarse_gather_kernel, div_up(int64_t(m) * 8, kBlock), kBlock, 0, m, boff, cnt, sQ, st, out.colptr.p,
               out.rowidx.p, out.val.p, d_asym);
}

void coarse_edges(gmg_ctx* c, int32_t n, int64_t nnz, const int32_t* colptr, const int32_t* rowidx, const double* val,
                  int32_t m, const int32_t* parents, const Clusters& cl, DeviceCsc& out, StageStatus* status) {
    out.n = m;
    if (m == 0 || n == 0) {
        out.colptr.alloc(c, int64_t(m) + 1);
        out.colptr.zero(); out.rowidx.alloc(c, 0); out.val.alloc(c, 0); out.nnz = 0;
        return;
    }
    GMG_REQUIRE(m <= n, GMG_ERR_INVALID, "extractCoarseEdges: more coarse points than fine points");
    StageStatus local;
    StageStatus* st = status ? status : &local;
    if (!status) local.begin(c);
    DevBuf<int32_t> cnt(c, int64_t(m) + 1);
    cnt.zero();
    CoarseRows rows;
    rows.bound = nnz;   // every fine vertex belongs to exactly one cluster: the scratch rows hold exactly nnz candidates
    coarse_rows_range(c, n, colptr, rowidx, val, 0, m, parents, cl, cnt.p, rows, st->dev.p + StageStatus::kCoarseHardOverflow);
    coarse_assemble(c, m, rows.boff.p, cnt.p, rows.sQ.p, rows.st.p, out, st->dev.p + StageStatus::kCoarseAsym);
    if (status) return;
    local.fetch(c);
    local.raise_coarse();
}

void StageStatus::raise_coarse() const {
    GMG_REQUIRE(host[kCoarseHardOverflow] == 0, GMG_ERR_INVALID,
                "extractCoarseEdges: a cluster has more than 1536 distinct coarse neighbours");
    GMG_REQUIRE(host[kCoarseAsym] == 0, GMG_ERR_ASYMMETRIC, "extractCoarseEdges: the fine pattern is not symmetric");
}

// ---------------------------------------------------------------------------------------------
// Coarse positions: one thread per cluster sums its children in ascending index order from
// (0,0,0) and divides by the count (:198-203). A cluster of one child also takes every stored
// neighbour of that child (:185-191), merged into the ascending order. Algorithmic bytes 28n+28m.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlock) coarse_points_kernel(int32_t n, const double* __restrict__ pts,
                                                               const int32_t* __restrict__ colptr,
                                                               const int32_t* __restrict__ rowidx, int32_t m,
                                                               const int32_t* __restrict__ cptr,
                                                               const int32_t* __restrict__ child,
                                                               double* __restrict__ out) {
    const int64_t P = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (P >= m) return;
    const int32_t b = cptr[P], e = cptr[P + 1];
    double sx = 0.0, sy = 0.0, sz = 0.0;
    int64_t count = 0;
    if (e - b == 1) {
        const int32_t only = child[b];
        bool self_done = false;
        const int32_t kb = __ldg(colptr + only), ke = __ldg(colptr + only + 1);
        for (int32_t k = kb; k <= ke; ++k) {
            const int32_t v = k < ke ? __ldg(rowidx + k) : 0x7fffffff;
            if (!self_done && only <= v) {
                sx += __ldg(pts + only); sy += __ldg(pts + n + only); sz += __ldg(pts + 2 * int64_t(n) + only);
                ++count; self_done = true;
                if (v == only) continue;
            }
            if (k < ke) {
                sx += __ldg(pts + v); sy += __ldg(pts + n + v); sz += __ldg(pts + 2 * int64_t(n) + v);
                ++count;
            }
        }
    } else {
        for (int32_t i = b; i < e; ++i) {
            const int32_t f = child[i];
            sx += __ldg(pts + f); sy += __ldg(pts + n + f); sz += __ldg(pts + 2 * int64_t(n) + f);
        }
        count = e - b;
    }
    const double d = double(count);
    out[P] = sx / d; out[m + P] = sy / d; out[2 * int64_t(m) + P] = sz / d;
}

void coarse_points(gmg_ctx* c, int32_t n, const double* pts, const int32_t* colptr, const int32_t* rowidx, int32_t m,
                   const Clusters& cl, double* out_pts) {
    if (m == 0) return;
    GMG_LAUNCH(c, coarse_points_kernel, div_up(m, kBlock), kBlock, 0, n, pts, colptr, rowidx, m, cl.ptr.p, cl.child.p,
               out_pts);
}

// nested mode: the coarse points are the samples' own positions
__global__ void __launch_bounds__(kBlock) sample_points_kernel(int32_t n, const double* __restrict__ pts, int32_t m,
                                                               const int32_t* __restrict__ samples, double* __restrict__ out) {
    const int64_t P = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (P >= m) return;
    const int32_t f = samples[P];
    out[P] = pts[f]; out[m + P] = pts[n + f]; out[2 * int64_t(m) + P] = pts[2 * int64_t(n) + f];
}
void sample_points(gmg_ctx* c, int32_t n, const double* pts, int32_t m, const int32_t* samples, double* out_pts) {
    if (m > 0) GMG_LAUNCH(c, sample_points_kernel, div_up(m, kBlock), kBlock, 0, n, pts, m, samples, out_pts);
}

}  // namespace gmg
