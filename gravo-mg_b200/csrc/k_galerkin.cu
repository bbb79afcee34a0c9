// Galerkin coarse operator A_c = P^T A P as two sparse products, each sized by a counting pass and then computed:
//     T   = P^T A      (m x n, by fine column k)      T(i,k)   = sum_l A(l,k) * P(l,i)
//     A_c = T P        (m x m, by coarse column j)    A_c(i,j) = sum_k T(i,k) * P(k,j)
// [not in reference: README.md:6 says only the P/R builders were kept; BASELINE.json asks for it.
//  Semantics are those of Eigen's un-pruned sparse product: structural pattern, sorted indices.]
//
// A and A_c are compressed by OUTER index (column-major, like Eigen::SparseMatrix<double>); both products only
// ever walk outer vectors of A (column k of T is a combination of the rows of P named by column k of A), so no
// symmetry of A is assumed.
//
// ONE THREAD PER OUTER VECTOR. The rows here are tiny (a column of T sums ~21 products into ~8 entries, a column
// of A_c ~130 products into ~19) and what costs is finding equal keys, not arithmetic: a warp per row spends ~400
// lane-slots per product on hashing with atomics, matching lanes and sorting (the first two versions of this file),
// a thread per row spends ~15 instructions per product on a PRIVATE open-addressed table. The tables live in shared
// memory as [slot][thread], so whatever slot each lane probes, the 32 lanes of a warp hit 32 different banks.
// No atomics anywhere, and the summation order is the plain sequential one: products in stored order of A's outer
// vector, then of P's row, for T; fine rows k ascending (R = P^T has sorted rows) for A_c. The result is
// reproducible bit for bit and the CPU checker (oracle/solver_oracle.c) is the same two loops.
// Sizing: column k of T holds at most ub1(k) = sum over column k of A of the row lengths of P, column j of A_c at
// most ub2(j) = sum over row j of R of the column counts of T. Both products write into slots of that size (scans
// of ub1 / ub2), record the count they found, and A_c is then compacted into its final arrays (T is only ever
// read through its slots). A thread whose table fills up (more than kCap1 / kCap2 distinct keys: dense rows of
// deep levels or kNN operators) redoes its vector as a linear-search list directly in its slot -- same order,
// same bits, just slow. The second product runs with a small table first (occupancy) and redoes the
// columns that outgrow it with a large one in a second kernel.
// Algorithmic bytes: symbolic 12n + 4nnzA + 8nnzP + 4m; numeric 12n + 12nnzA + 24nnzP + 4m + 12nnzAc.
#include "stages.h"

#include <cstdio>
#include <cstdlib>

namespace gmg {

// ---- R = P^T ------------------------------------------------------------------------------------
// Only the columns [c_lo, c_hi) of P become rows of R (the row-sharded build transposes the coarse rows it owns).
__global__ void __launch_bounds__(kBlock) count_columns_kernel(int64_t nnz, const int32_t* __restrict__ colidx,
                                                               int32_t c_lo, int32_t c_hi, int32_t* __restrict__ count) {
    const int64_t k = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (k >= nnz) return;
    const int32_t cidx = __ldg(colidx + k);
    if (cidx >= c_lo && cidx < c_hi) atomicAdd(count + (cidx - c_lo), 1);
}
__global__ void __launch_bounds__(kBlock) transpose_fill_kernel(int32_t rows, const int32_t* __restrict__ rowptr,
                                                                const int32_t* __restrict__ colidx,
                                                                const double* __restrict__ val, int32_t c_lo, int32_t c_hi,
                                                                const int32_t* __restrict__ tptr,
                                                                int32_t* __restrict__ cursor, int32_t* __restrict__ tcol,
                                                                double* __restrict__ tval) {
    const int64_t r = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (r >= rows) return;
    for (int32_t k = rowptr[r]; k < rowptr[r + 1]; ++k) {
        const int32_t cidx = colidx[k];
        if (cidx < c_lo || cidx >= c_hi) continue;
        const int32_t pos = tptr[cidx - c_lo] + atomicAdd(cursor + (cidx - c_lo), 1);
        tcol[pos] = int32_t(r);
        tval[pos] = val[k];
    }
}
// One warp per row: rows of up to 32 entries (the usual case: ~22 fine rows per coarse column) are rank-sorted
// in registers -- every lane holds one entry and counts the smaller keys with shuffles; keys within a row are
// distinct, so the ranks are a permutation. Longer rows fall back to an insertion sort by lane 0.
__global__ void __launch_bounds__(kBlock) sort_rows_kernel(int32_t rows, const int32_t* __restrict__ ptr,
                                                           int32_t* __restrict__ idx, double* __restrict__ val) {
    const int64_t r = (int64_t(blockIdx.x) * kBlock + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (r >= rows) return;
    const int32_t b = ptr[r], e = ptr[r + 1], k = e - b;
    if (k <= 32) {
        const bool have = lane < k;
        const int32_t key = have ? idx[b + lane] : 0x7fffffff;
        const double v = have ? val[b + lane] : 0.0;
        int32_t rank = 0;
        for (int j = 0; j < k; ++j) rank += __shfl_sync(0xffffffffu, key, j) < key ? 1 : 0;
        __syncwarp();
        if (have) { idx[b + rank] = key; val[b + rank] = v; }
        return;
    }
    if (lane != 0) return;
    for (int32_t i = b + 1; i < e; ++i) {
        const int32_t x = idx[i]; const double v = val[i];
        int32_t j = i - 1;
        while (j >= b && idx[j] > x) { idx[j + 1] = idx[j]; val[j + 1] = val[j]; --j; }
        idx[j + 1] = x; val[j + 1] = v;
    }
}

void transpose_csr(gmg_ctx* c, const DeviceCsr& P, DeviceCsr& R, int32_t c_lo, int32_t c_hi, bool sort_rows) {
    if (c_hi < 0) c_hi = P.cols;
    const bool all = c_lo == 0 && c_hi == P.cols;
    R.rows = c_hi - c_lo; R.cols = P.rows;
    R.rowptr.alloc(c, int64_t(R.rows) + 1);
    DevBuf<int32_t> count(c, int64_t(R.rows) + 1);
    DevBuf<int64_t> total(c, 1);
    Trace tr(c, "transpose");
    count.zero();
    if (P.nnz > 0) GMG_LAUNCH(c, count_columns_kernel, div_up(P.nnz, kBlock), kBlock, 0, P.nnz, P.colidx.p, c_lo, c_hi, count.p);
    exclusive_scan_i32(c, count.p, R.rowptr.p, int64_t(R.rows) + 1, all ? nullptr : total.p);
    R.nnz = P.nnz;
    if (!all) { total.download(&R.nnz, 1); sync(c); }
    tr.mark("count + scan");
    R.colidx.alloc(c, R.nnz);
    R.val.alloc(c, R.nnz);
    count.zero();
    if (P.rows > 0) GMG_LAUNCH(c, transpose_fill_kernel, div_up(P.rows, kBlock), kBlock, 0, P.rows, P.rowptr.p, P.colidx.p,
                               P.val.p, c_lo, c_hi, R.rowptr.p, count.p, R.colidx.p, R.val.p);
    tr.mark("fill");
    if (R.rows > 0 && sort_rows) GMG_LAUNCH(c, sort_rows_kernel, div_up(int64_t(R.rows) * 32, kBlock), kBlock, 0, R.rows, R.rowptr.p, R.colidx.p, R.val.p);
    tr.mark("sort rows");
}


// ---- SpGEMM -------------------------------------------------------------------------------------
constexpr int kCap1 = 32, kThreads1 = 128;     // first product: slots per thread, threads per block
constexpr int kCap2 = 32, kThreads2 = 128;     // second product, first pass
constexpr int kCap2b = 128, kThreads2b = 32;   // second product, columns that outgrow the first table

// A thread's private accumulator: open addressing, linear probing, [slot][thread] in shared memory.
template <int CAP, int THREADS>
struct ThreadTable {
    int32_t* keys;      // this thread's column of the block's key array (stride THREADS)
    double* vals;
    int32_t count;
    __device__ __forceinline__ ThreadTable(unsigned char* smem, int tid)
        : keys(reinterpret_cast<int32_t*>(smem + sizeof(double) * CAP * THREADS) + tid),
          vals(reinterpret_cast<double*>(smem) + tid), count(0) {}
    static constexpr size_t bytes() { return (sizeof(double) + sizeof(int32_t)) * size_t(CAP) * THREADS; }
    __device__ __forceinline__ void clear() {
#pragma unroll 8
        for (int s = 0; s < CAP; ++s) keys[s * THREADS] = -1;
        count = 0;
    }
    // acc[key] += v (the first product of a key is stored, not added to zero); false: the table is full
    __device__ __forceinline__ bool add(int32_t key, double v) {
        constexpr int kLog = CAP == 32 ? 5 : (CAP == 64 ? 6 : 7);
        static_assert((1 << kLog) == CAP, "table sizes: 32, 64 or 128 slots");
        uint32_t h = (uint32_t(key) * 0x9E3779B1u) >> (32 - kLog);
        for (int probe = 0; probe < CAP; ++probe, h = (h + 1u) & uint32_t(CAP - 1)) {
            const int32_t cur = keys[h * THREADS];
            if (cur == key) { vals[h * THREADS] = vals[h * THREADS] + v; return true; }
            if (cur < 0) { keys[h * THREADS] = key; vals[h * THREADS] = v; ++count; return true; }
        }
        return false;
    }
    // occupied slots to the front, in slot order
    __device__ __forceinline__ void compact() {
        int pos = 0;
        for (int s = 0; s < CAP && pos < count; ++s) {
            const int32_t key = keys[s * THREADS];
            if (key < 0) continue;
            if (pos != s) { keys[pos * THREADS] = key; vals[pos * THREADS] = vals[s * THREADS]; }
            ++pos;
        }
    }
    // the compacted entries ascending by key (insertion sort: ~20 entries that arrive nearly sorted)
    __device__ __forceinline__ void sort() {
        for (int i = 1; i < count; ++i) {
            const int32_t key = keys[i * THREADS];
            const double v = vals[i * THREADS];
            int j = i - 1;
            while (j >= 0 && keys[j * THREADS] > key) {
                keys[(j + 1) * THREADS] = keys[j * THREADS];
                vals[(j + 1) * THREADS] = vals[j * THREADS];
                --j;
            }
            keys[(j + 1) * THREADS] = key;
            vals[(j + 1) * THREADS] = v;
        }
    }
};

// The slow path of both products: the same accumulation as a linear-search list in global memory (the vector's own
// slot, which holds at least as many entries as it has distinct keys).
struct SlotList {
    int32_t* idx; double* val; int32_t count;
    __device__ __forceinline__ void add(int32_t key, double v) {
        for (int32_t s = 0; s < count; ++s)
            if (idx[s] == key) { val[s] = val[s] + v; return; }
        idx[count] = key; val[count] = v; ++count;
    }
    __device__ __forceinline__ void sort() {
        for (int32_t i = 1; i < count; ++i) {
            const int32_t key = idx[i];
            const double v = val[i];
            int32_t j = i - 1;
            while (j >= 0 && idx[j] > key) { idx[j + 1] = idx[j]; val[j + 1] = val[j]; --j; }
            idx[j + 1] = key; val[j + 1] = v;
        }
    }
};

using Table1 = ThreadTable<kCap1, kThreads1>;
using Table2 = ThreadTable<kCap2, kThreads2>;
using Table2b = ThreadTable<kCap2b, kThreads2b>;

struct ProductArgs {
    int32_t n, m;                                                         // fine vertices; outer vectors of A_c formed here
    int32_t k_lo, k_hi;                                                   // the window of fine columns of T that R reads (ub1, t_ptr are indexed from k_lo)
    const int32_t* a_ptr; const int32_t* a_idx; const double* a_val;      // A by outer vector
    const int32_t* p_ptr; const int32_t* p_col; const double* p_val;      // P rows
    const int32_t* r_ptr; const int32_t* r_col; const double* r_val;      // R = P^T (rows = the outer vectors formed here)
    const uint8_t* needed;                                                // fine columns of T that R reads (null: all)
    int32_t* ub1; const int32_t* t_ptr; int2* t_loc; int32_t* t_idx; double* t_val;        // T: slots per block, (start, count) per column
    int32_t* ub2; const int32_t* w_ptr; int32_t* w_cnt; int32_t* w_idx; double* w_val;     // A_c in slots
    const int32_t* out_ptr; int32_t* out_idx; double* out_val;                              // A_c
    int32_t* flags;                                                       // [0] vectors that took the slow path, [1] queued for the second pass
    int32_t* overflow;                                                    // columns of A_c queued for the second pass
};

// window[0] = smallest, window[1] = largest fine column R names
__global__ void __launch_bounds__(kBlock) mark_needed_kernel(int64_t nnz, const int32_t* __restrict__ r_col,
                                                             uint8_t* __restrict__ needed, int32_t* __restrict__ window) {
    const int64_t e = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    int32_t lo = 0x7fffffff, hi = -1;
    if (e < nnz) { const int32_t k = r_col[e]; needed[k] = 1; lo = k; hi = k; }
    lo = __reduce_min_sync(0xffffffffu, lo);
    hi = __reduce_max_sync(0xffffffffu, hi);
    if ((threadIdx.x & 31) == 0 && hi >= 0) { atomicMin(window, lo); atomicMax(window + 1, hi); }
}

// ub1(k): products (= an upper bound on the entries) of column k of T
__global__ void __launch_bounds__(kBlock) product1_bound_kernel(ProductArgs a) {
    const int64_t w = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    const int64_t k = a.k_lo + w;
    if (k > a.k_hi) return;
    int32_t ub = 0;
    if (k < a.k_hi && (!a.needed || a.needed[k]))
        for (int32_t q = __ldg(a.a_ptr + k); q < __ldg(a.a_ptr + k + 1); ++q) {
            const int32_t l = __ldg(a.a_idx + q);
            ub += __ldg(a.p_ptr + l + 1) - __ldg(a.p_ptr + l);
        }
    a.ub1[w] = ub;          // the entry behind the window = 0: the scan's total
}

// exclusive scan of one int per thread over a block of THREADS threads (all threads must call)
template <int THREADS>
__device__ __forceinline__ int32_t block_scan_exclusive(int32_t v, int32_t* warp_sums) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int32_t incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int32_t t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += t;
    }
    if (lane == 31) warp_sums[wid] = incl;
    __syncthreads();
    int32_t off = 0;
#pragma unroll
    for (int w = 0; w < THREADS / 32; ++w) off += w < wid ? warp_sums[w] : 0;
    return off + incl - v;
}

// T(:,k) = sum_l A(l,k) * P(l,:). The loads are issued in BATCHES before anything consumes them (eight entries of A's
// outer vector, then the eight row pointers of P, then the rows themselves): a thread that walks
// a_idx -> p_ptr -> p_col one entry at a time pays three dependent trips to memory per entry, and with a 48 KB
// table per block there are too few warps on an SM to hide that (ncu: 49 % of the stalls long-scoreboard at 23 %
// occupancy). The columns of a block are then written DENSELY into the block's share of the slots (a block scan of
// the counts), so that the second product reads full sectors.
__global__ void __launch_bounds__(kThreads1) product1_kernel(ProductArgs a) {
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ int32_t s_warp[kThreads1 / 32];
    const int64_t w = int64_t(blockIdx.x) * kThreads1 + threadIdx.x;
    const int64_t k = a.k_lo + w;
    int32_t ub = 0;
    if (k < a.k_hi) ub = __ldg(a.t_ptr + w + 1) - __ldg(a.t_ptr + w);
    Table1 tab(smem, threadIdx.x);
    bool ok = true;
    int32_t ab = 0, ae = 0;
    if (ub > 0) {
        tab.clear();
        ab = __ldg(a.a_ptr + k); ae = __ldg(a.a_ptr + k + 1);
        constexpr int kB = 8, kRow = 3;
        for (int32_t q0 = ab; q0 < ae && ok; q0 += kB) {
            int32_t l[kB], pb[kB], pe[kB];
            double w[kB];
#pragma unroll
            for (int u = 0; u < kB; ++u) {
                const bool have = q0 + u < ae;
                l[u] = have ? __ldg(a.a_idx + q0 + u) : -1;
                w[u] = have ? __ldg(a.a_val + q0 + u) : 0.0;
            }
#pragma unroll
            for (int u = 0; u < kB; ++u) {
                pb[u] = l[u] >= 0 ? __ldg(a.p_ptr + l[u]) : 0;
                pe[u] = l[u] >= 0 ? __ldg(a.p_ptr + l[u] + 1) : 0;
            }
#pragma unroll
            for (int half = 0; half < kB; half += kB / 2) {
                int32_t col[kB / 2][kRow];
                double val[kB / 2][kRow];
#pragma unroll
                for (int u = 0; u < kB / 2; ++u)
#pragma unroll
                    for (int t = 0; t < kRow; ++t) {
                        const bool have = pb[half + u] + t < pe[half + u];
                        col[u][t] = have ? __ldg(a.p_col + pb[half + u] + t) : -1;
                        val[u][t] = have ? __ldg(a.p_val + pb[half + u] + t) : 0.0;
                    }
#pragma unroll
                for (int u = 0; u < kB / 2; ++u) {
#pragma unroll
                    for (int t = 0; t < kRow; ++t)
                        if (col[u][t] >= 0) ok = tab.add(col[u][t], w[half + u] * val[u][t]) && ok;
                    for (int32_t t = pb[half + u] + kRow; t < pe[half + u]; ++t)        // rows of P beyond three entries
                        ok = tab.add(__ldg(a.p_col + t), w[half + u] * __ldg(a.p_val + t)) && ok;
                }
            }
        }
    }
    // where this column goes: fast columns take what they hold, a column on the slow path its whole bound
    const int32_t reserve = ub == 0 ? 0 : (ok ? tab.count : ub);
    const int32_t off = block_scan_exclusive<kThreads1>(reserve, s_warp);
    if (k >= a.k_hi) return;
    const int64_t w0 = int64_t(blockIdx.x) * kThreads1;
    const int32_t start = __ldg(a.t_ptr + w0) + off;
    int32_t* idx = a.t_idx + start;
    double* val = a.t_val + start;
    int32_t cnt = 0;
    if (ub > 0 && ok) {
        tab.compact();
        cnt = tab.count;
        for (int32_t s = 0; s < cnt; ++s) { idx[s] = tab.keys[s * kThreads1]; val[s] = tab.vals[s * kThreads1]; }
    } else if (ub > 0) {
        SlotList list{idx, val, 0};
        for (int32_t q = ab; q < ae; ++q) {
            const int32_t l = __ldg(a.a_idx + q);
            const double w = __ldg(a.a_val + q);
            for (int32_t t = __ldg(a.p_ptr + l); t < __ldg(a.p_ptr + l + 1); ++t) list.add(__ldg(a.p_col + t), w * __ldg(a.p_val + t));
        }
        cnt = list.count;
        atomicAdd(a.flags, 1);
    }
    a.t_loc[k] = make_int2(start, cnt);
}

// ub2(j): products of column j of A_c
__global__ void __launch_bounds__(kBlock) product2_bound_kernel(ProductArgs a) {
    const int64_t j = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (j > a.m) return;
    int32_t ub = 0;
    if (j < a.m)
        for (int32_t e = __ldg(a.r_ptr + j); e < __ldg(a.r_ptr + j + 1); ++e) ub += __ldg(a.t_loc + __ldg(a.r_col + e)).y;
    a.ub2[j] = ub;
}

// A_c(:,j) = sum_k T(:,k) * P(k,j), k ascending; written sorted into the column's slot. Same batching: the first
// eight entries of the NEXT column of T are in flight (and the location of the one after that) while the current
// one is added up. false: the table is full.
template <typename Table, int THREADS>
__device__ __forceinline__ bool product2_column(const ProductArgs& a, int64_t j, Table& tab) {
    tab.clear();
    const int32_t rb = __ldg(a.r_ptr + j), re = __ldg(a.r_ptr + j + 1);
    bool ok = true;
    constexpr int kB = 8;
    int32_t ni[kB];
    double nv[kB];
    int2 nloc = make_int2(0, 0), nnloc = make_int2(0, 0);
    double nw = 0.0, nnw = 0.0;
    if (rb < re) { nloc = __ldg(a.t_loc + __ldg(a.r_col + rb)); nw = __ldg(a.r_val + rb); }
    if (rb + 1 < re) { nnloc = __ldg(a.t_loc + __ldg(a.r_col + rb + 1)); nnw = __ldg(a.r_val + rb + 1); }
#pragma unroll
    for (int u = 0; u < kB; ++u) {
        const bool have = u < nloc.y;
        ni[u] = have ? __ldg(a.t_idx + nloc.x + u) : -1;
        nv[u] = have ? __ldg(a.t_val + nloc.x + u) : 0.0;
    }
    for (int32_t e = rb; e < re && ok; ++e) {
        int32_t ci[kB];
        double cv[kB];
#pragma unroll
        for (int u = 0; u < kB; ++u) { ci[u] = ni[u]; cv[u] = nv[u]; }
        const int2 loc = nloc;
        const double w = nw;
        nloc = nnloc; nw = nnw;
#pragma unroll
        for (int u = 0; u < kB; ++u) {
            const bool have = e + 1 < re && u < nloc.y;
            ni[u] = have ? __ldg(a.t_idx + nloc.x + u) : -1;
            nv[u] = have ? __ldg(a.t_val + nloc.x + u) : 0.0;
        }
        if (e + 2 < re) { nnloc = __ldg(a.t_loc + __ldg(a.r_col + e + 2)); nnw = __ldg(a.r_val + e + 2); }
#pragma unroll
        for (int u = 0; u < kB; ++u)
            if (ci[u] >= 0) ok = tab.add(ci[u], cv[u] * w) && ok;
        for (int32_t s = kB; s < loc.y; ++s) ok = tab.add(__ldg(a.t_idx + loc.x + s), __ldg(a.t_val + loc.x + s) * w) && ok;
    }
    if (!ok) return false;
    int32_t* idx = a.w_idx + __ldg(a.w_ptr + j);
    double* val = a.w_val + __ldg(a.w_ptr + j);
    tab.compact();
    tab.sort();
    for (int32_t s = 0; s < tab.count; ++s) { idx[s] = tab.keys[s * THREADS]; val[s] = tab.vals[s * THREADS]; }
    a.w_cnt[j] = tab.count;
    return true;
}

// first pass: a small table per thread, so that four blocks of 128 threads fit on an SM (the kernel waits on memory;
// what hides that is warps). A column that outgrows it is queued for the second pass.
__global__ void __launch_bounds__(kThreads2) product2_kernel(ProductArgs a) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int64_t j = int64_t(blockIdx.x) * kThreads2 + threadIdx.x;
    if (j >= a.m) return;
    Table2 tab(smem, threadIdx.x);
    if (!product2_column<Table2, kThreads2>(a, j, tab)) {
        a.w_cnt[j] = 0;
        a.overflow[atomicAdd(a.flags + 1, 1)] = int32_t(j);
    }
}
// second pass over the queued columns: a large table, and the list in global memory behind it
__global__ void __launch_bounds__(kThreads2b) product2_overflow_kernel(ProductArgs a) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int32_t queued = a.flags[1];
    Table2b tab(smem, threadIdx.x);
    for (int64_t o = int64_t(blockIdx.x) * kThreads2b + threadIdx.x; o < queued; o += int64_t(gridDim.x) * kThreads2b) {
        const int64_t j = a.overflow[o];
        if (product2_column<Table2b, kThreads2b>(a, j, tab)) continue;
        int32_t* idx = a.w_idx + __ldg(a.w_ptr + j);
        double* val = a.w_val + __ldg(a.w_ptr + j);
        SlotList list{idx, val, 0};
        for (int32_t e = __ldg(a.r_ptr + j); e < __ldg(a.r_ptr + j + 1); ++e) {
            const int2 loc = __ldg(a.t_loc + __ldg(a.r_col + e));
            const double w = __ldg(a.r_val + e);
            for (int32_t s = 0; s < loc.y; ++s) list.add(a.t_idx[loc.x + s], a.t_val[loc.x + s] * w);
        }
        list.sort();
        a.w_cnt[j] = list.count;
        atomicAdd(a.flags, 1);
    }
}

// the slots of A_c into its final arrays: 8 lanes per column
__global__ void __launch_bounds__(kBlock) compact_columns_kernel(ProductArgs a) {
    const int64_t j = (int64_t(blockIdx.x) * kBlock + threadIdx.x) >> 3;
    const int sub = threadIdx.x & 7;
    if (j >= a.m) return;
    const int32_t src = __ldg(a.w_ptr + j), dst = __ldg(a.out_ptr + j), cnt = __ldg(a.out_ptr + j + 1) - dst;
    for (int32_t s = sub; s < cnt; s += 8) { a.out_idx[dst + s] = a.w_idx[src + s]; a.out_val[dst + s] = a.w_val[src + s]; }
}

void galerkin(gmg_ctx* c, int32_t n, const int32_t* a_colptr, const int32_t* a_rowidx, const double* a_val,
              const DeviceCsr& P, DeviceCsc& Ac, int32_t c_lo, int32_t c_hi, DeviceCsr* keep_R, cudaEvent_t* marks) {
    GMG_REQUIRE(P.rows == n, GMG_ERR_INVALID, "galerkin: P and A sizes disagree");
    if (c_hi < 0) c_hi = P.cols;
    const int32_t m = c_hi - c_lo;     // outer vectors computed here; Ac is indexed from c_lo
    Ac.n = m;
    Ac.colptr.alloc(c, int64_t(m) + 1);
    if (m == 0) {
        Ac.colptr.zero(); Ac.rowidx.alloc(c, 0); Ac.val.alloc(c, 0); Ac.nnz = 0;
        if (marks) for (int s = 0; s < 4; ++s) cudaEventRecord(marks[s], c->stream);
        return;
    }
    DeviceCsr R_local;
    DeviceCsr& R = keep_R ? *keep_R : R_local;
    Trace tr(c, "galerkin");
    if (marks) cudaEventRecord(marks[0], c->stream);
    transpose_csr(c, P, R, c_lo, c_hi, true);       // sorted rows: k ascending is the summation order of A_c
    tr.mark("transpose");
    if (marks) cudaEventRecord(marks[1], c->stream);
    const bool all = c_lo == 0 && c_hi == P.cols;
    // The row-sharded build forms the outer vectors of one rank: T is only needed for the fine columns its rows of R
    // name, which in a coherent numbering is a window around the rank's own range -- the bound, the scan and the first
    // product run over that window, not over all n (at 8 ranks the passes over all n cost more than the product).
    DevBuf<uint8_t> needed;
    int32_t k_lo = 0, k_hi = n;
    if (!all) {
        needed.alloc(c, n);
        needed.zero();
        DevBuf<int32_t> window(c, 2);
        int32_t hw[2] = {0x7fffffff, -1};
        window.upload(hw, 2);
        if (R.nnz > 0) GMG_LAUNCH(c, mark_needed_kernel, div_up(R.nnz, kBlock), kBlock, 0, R.nnz, R.colidx.p, needed.p, window.p);
        window.download(hw, 2);
        sync(c);
        if (hw[1] >= hw[0]) { k_lo = hw[0]; k_hi = hw[1] + 1; } else { k_lo = 0; k_hi = 0; }
    }
    const int64_t kw = int64_t(k_hi) - k_lo;
    DevBuf<int32_t> ub1(c, kw + 1), t_ptr(c, kw + 1), flags(c, 2), overflow(c, m);
    DevBuf<int2> t_loc(c, n);
    DevBuf<int32_t> ub2(c, int64_t(m) + 1), w_ptr(c, int64_t(m) + 1), w_cnt(c, int64_t(m) + 1);
    DevBuf<int64_t> totals(c, 3);
    flags.zero();
    ProductArgs a{n, m, k_lo, k_hi, a_colptr, a_rowidx, a_val, P.rowptr.p, P.colidx.p, P.val.p, R.rowptr.p, R.colidx.p, R.val.p,
                  all ? nullptr : needed.p, ub1.p, t_ptr.p, t_loc.p, nullptr, nullptr, ub2.p, w_ptr.p, w_cnt.p, nullptr, nullptr,
                  nullptr, nullptr, nullptr, flags.p, overflow.p};
    // ---- first product
    int64_t h_total[3] = {0, 0, 0};
    GMG_LAUNCH(c, product1_bound_kernel, div_up(kw + 1, kBlock), kBlock, 0, a);
    exclusive_scan_i32(c, ub1.p, t_ptr.p, kw + 1, totals.p);
    totals.download(h_total, 1);
    sync(c);
    GMG_REQUIRE(h_total[0] < (int64_t(1) << 31), GMG_ERR_INVALID, "galerkin: P^T A exceeds 32-bit indexing");
    DevBuf<int32_t> t_idx(c, h_total[0]);
    DevBuf<double> t_val(c, h_total[0]);
    a.t_idx = t_idx.p; a.t_val = t_val.p;
    GMG_CUDA(cudaFuncSetAttribute(product1_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(Table1::bytes())));
    if (kw > 0) GMG_LAUNCH(c, product1_kernel, div_up(kw, kThreads1), kThreads1, Table1::bytes(), a);
    tr.mark("T = P^T A");
    if (marks) cudaEventRecord(marks[2], c->stream);
    // ---- second product
    GMG_LAUNCH(c, product2_bound_kernel, div_up(int64_t(m) + 1, kBlock), kBlock, 0, a);
    exclusive_scan_i32(c, ub2.p, w_ptr.p, int64_t(m) + 1, totals.p + 1);
    totals.download(h_total, 2);
    sync(c);
    GMG_REQUIRE(h_total[1] < (int64_t(1) << 31), GMG_ERR_INVALID, "galerkin: the slots of A_c exceed 32-bit indexing");
    DevBuf<int32_t> w_idx(c, h_total[1]);
    DevBuf<double> w_val(c, h_total[1]);
    a.w_idx = w_idx.p; a.w_val = w_val.p;
    GMG_CUDA(cudaMemsetAsync(w_cnt.p + m, 0, sizeof(int32_t), c->stream));
    GMG_CUDA(cudaFuncSetAttribute(product2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(Table2::bytes())));
    GMG_LAUNCH(c, product2_kernel, div_up(m, kThreads2), kThreads2, Table2::bytes(), a);
    GMG_CUDA(cudaFuncSetAttribute(product2_overflow_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(Table2b::bytes())));
    GMG_LAUNCH(c, product2_overflow_kernel, 4 * c->sm_count, kThreads2b, Table2b::bytes(), a);
    exclusive_scan_i32(c, w_cnt.p, Ac.colptr.p, int64_t(m) + 1, totals.p + 2);
    totals.download(h_total, 3);
    int32_t slow[2] = {0, 0};
    flags.download(slow, 2);
    sync(c);
    Ac.nnz = h_total[2];
    if (getenv("GMG_TRACE")) fprintf(stderr, "[gmg] galerkin: n %d, %d outer vectors, %lld entries of T, %lld of A_c, %d columns of A_c took the large table, %d vectors the list\n",
                                     n, m, (long long)h_total[0], (long long)Ac.nnz, slow[1], slow[0]);
    Ac.rowidx.alloc(c, Ac.nnz);
    Ac.val.alloc(c, Ac.nnz);
    a.out_ptr = Ac.colptr.p; a.out_idx = Ac.rowidx.p; a.out_val = Ac.val.p;
    GMG_LAUNCH(c, compact_columns_kernel, div_up(int64_t(m) * 8, kBlock), kBlock, 0, a);
    tr.mark("A_c = T P");
    if (marks) cudaEventRecord(marks[3], c->stream);
}

}  // namespace gmg
