// Galerkin coarse operator A_c = P^T A P as a two-pass (symbolic, then numeric) SpGEMM.
// [not in reference: README.md:6 says only the P/R builders were kept; BASELINE.json asks for it.
//  Semantics are those of Eigen's un-pruned sparse product: structural pattern, sorted indices.]
//
// A and A_c are compressed by OUTER index (column-major, like Eigen::SparseMatrix<double>); the
// row-wise Gustavson formulation below only ever walks outer vectors, so it computes
//     out(inner j, outer i) = sum_k sum_l P(k,i) * A(l,k) * P(l,j)
// which is exactly (P^T A P)(j,i) with A(l,k) read from outer vector k. No symmetry of A is assumed.
//
// One warp per coarse outer index i, in two steps that share the warp's shared memory:
//   step 1  c(l) = sum_k P(k,i) * A(l,k)   R = P^T lists the fine k with P(k,i) stored (sorted ascending in shared
//                                          memory); the products (k, q), q over A's outer vector k, are taken 32 at a
//                                          time, one per lane, in (k, q) order
//   step 2  out(j) = sum_l c(l) * P(l,j)   the distinct l sorted ascending, 32 per round, one per lane; entry t = 0, 1, 2
//                                          of every row of P in turn
// Grouping by l first means every row of P is expanded once per distinct l (~45 per coarse row) and not once per
// (k, l) pair (~160). Keys are found with a shared-memory hash (atomicCAS on the key only). VALUES are accumulated
// WITHOUT atomics: shared-memory atomicAdd(double) is a compare-and-swap loop on this architecture (ATOMS.CAST.SPIN
// in the SASS; the first version spent 86 % of its time there). Instead the lanes of one group of 32 that hold the
// same key find each other with __match_any_sync, the lowest lane adds the group's values in lane order and does a
// plain read-modify-write of the slot -- distinct keys own distinct slots. The summation order is therefore FIXED
// (products in (k, q) order, then l ascending; inside a group of 32 equal keys are summed first), the result is
// reproducible bit for bit, and the CPU checker (oracle/solver_oracle.c) follows the same association.
// Pass 1 (symbolic) runs both steps on keys only and counts distinct j, a scan sizes the result, pass 2 (numeric)
// computes the values, sorts the keys and writes the outer vector. Rows that outgrow the warp's tables (more than
// 64 rows of P per column, 192 distinct l or 96 distinct j) go to a block-wide single-table version of the plain
// triple loop with atomics (not bit-reproducible; none on the BASELINE meshes).
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
struct GalerkinArgs {
    int32_t m;
    const int32_t* r_ptr; const int32_t* r_col; const double* r_val;      // R = P^T
    const int32_t* a_ptr; const int32_t* a_idx; const double* a_val;      // A by outer vector
    const int32_t* p_ptr; const int32_t* p_col; const double* p_val;      // P rows
    int32_t* out_count;        // symbolic: distinct inner indices of outer vector i
    const int32_t* out_ptr;    // numeric: scan of out_count
    int32_t* out_idx; double* out_val;
    int32_t* flags;            // [1] overflow list tail, [2] hard overflow
    int32_t* overflow;
};

template <bool NUMERIC, int THREADS, int SLOTS>
__device__ __forceinline__ bool galerkin_row(const GalerkinArgs& a, int32_t i, int tid, int32_t* keys, double* vals,
                                             int32_t* scount) {
    for (int s = tid; s < SLOTS; s += THREADS) { keys[s] = -1; if (NUMERIC) vals[s] = 0.0; }
    if (tid == 0) *scount = 0;
    if (THREADS == 32) __syncwarp(); else __syncthreads();
    const int32_t rb = __ldg(a.r_ptr + i), re = __ldg(a.r_ptr + i + 1);
    // 8 lanes per fine row k (outer vectors of A hold ~7 entries), THREADS/8 rows in flight
    constexpr int kSub = 8;
    for (int32_t e = rb + tid / kSub; e < re; e += THREADS / kSub) {
        const int32_t k = __ldg(a.r_col + e);
        const double pki = NUMERIC ? __ldg(a.r_val + e) : 0.0;
        const int32_t ab = __ldg(a.a_ptr + k), ae = __ldg(a.a_ptr + k + 1);
        for (int32_t q = ab + tid % kSub; q < ae; q += kSub) {
            const int32_t l = __ldg(a.a_idx + q);
            const double w = NUMERIC ? pki * __ldg(a.a_val + q) : 0.0;
            const int32_t pb = __ldg(a.p_ptr + l), pe = __ldg(a.p_ptr + l + 1);
            for (int32_t t = pb; t < pe; ++t) {
                const int32_t j = __ldg(a.p_col + t);
                uint32_t h = (uint32_t(j) * 0x9E3779B1u) >> 7;
                for (int probe = 0; probe < SLOTS; ++probe, ++h) {
                    const int s = int(h & (SLOTS - 1));
                    const int32_t prev = atomicCAS(keys + s, -1, j);
                    if (prev == -1) atomicAdd(scount, 1);
                    if (prev == -1 || prev == j) {
                        if (NUMERIC) atomicAdd(vals + s, w * __ldg(a.p_val + t));
                        break;
                    }
                }
            }
        }
    }
    if (THREADS == 32) __syncwarp(); else __syncthreads();
    const int32_t K = *scount;
    if (K > SLOTS * 3 / 4) return false;
    if (!NUMERIC) { if (tid == 0) a.out_count[i] = K; return true; }
    const int32_t off = __ldg(a.out_ptr + i);
    if constexpr (THREADS == 32) {
        // compact the occupied slots to the front (in place: a slot only moves down and every chunk is
        // read before it is written), then rank-sort the K survivors
        int32_t base = 0;
        for (int s0 = 0; s0 < SLOTS; s0 += 32) {
            const int32_t key = keys[s0 + tid];
            const double v = vals[s0 + tid];
            const unsigned occ = __ballot_sync(0xffffffffu, key >= 0);
            __syncwarp();
            if (key >= 0) { const int pos = base + __popc(occ & ((1u << tid) - 1u)); keys[pos] = key; vals[pos] = v; }
            base += __popc(occ);
            __syncwarp();
        }
        for (int s = tid; s < K; s += 32) {
            const int32_t key = keys[s];
            int32_t rank = 0;
            for (int s2 = 0; s2 < K; ++s2) rank += keys[s2] < key ? 1 : 0;
            a.out_idx[off + rank] = key;
            a.out_val[off + rank] = vals[s];
        }
        return true;
    }
    for (int s = tid; s < SLOTS; s += THREADS) {
        const int32_t key = keys[s];
        if (key < 0) continue;
        int32_t rank = 0;
        for (int s2 = 0; s2 < SLOTS; ++s2) { const int32_t k2 = keys[s2]; rank += (k2 >= 0 && k2 < key) ? 1 : 0; }
        a.out_idx[off + rank] = key;
        a.out_val[off + rank] = vals[s];
    }
    return true;
}

// probe-insert `key` into an open-addressed table; returns the slot, or -1 when the table is full
template <int SLOTS>
__device__ __forceinline__ int hash_insert(int32_t* keys, int32_t key, int32_t* count) {
    uint32_t h = (uint32_t(key) * 0x9E3779B1u) >> 7;
    for (int probe = 0; probe < SLOTS; ++probe, ++h) {
        const int s = int(h & (SLOTS - 1));
        const int32_t prev = atomicCAS(keys + s, -1, key);
        if (prev == -1) atomicAdd(count, 1);
        if (prev == -1 || prev == key) return s;
    }
    return -1;
}

constexpr int kGWarpSlots = 128;
constexpr int kGFineSlots = 256;
constexpr int kGBlockSlots = 2048;
constexpr int kGMaxR = 64;            // rows of P per coarse column the warp path stages

struct GalerkinWarpSmem {
    int32_t fkeys[kGFineSlots];
    int32_t keys[kGWarpSlots];
    int32_t rk[kGMaxR];               // fine rows k of R's row i, ascending
    int32_t rbase[kGMaxR + 1];        // exclusive scan of the lengths of A's outer vectors k
    int32_t scount[2];                // distinct fine / coarse keys
};
struct GalerkinWarpVals {
    double fvals[kGFineSlots];
    double vals[kGWarpSlots];
    double rp[kGMaxR];                // P(k, i)
    double xch[32];                   // group_accumulate's exchange
};

// Add v to acc[slot] for every lane with valid == true; lanes holding the same key are summed first, in lane
// order, by the lowest of them (see the header comment). The values travel through 32 doubles of shared memory so
// that only the group leaders loop. All 32 lanes must call.
__device__ __forceinline__ void group_accumulate(bool valid, int32_t key, int slot, double v, double* acc, double* xch, int lane) {
    const unsigned grp = __match_any_sync(0xffffffffu, valid ? key : -1 - lane);
    xch[lane] = v;
    __syncwarp();
    if (valid && (__ffs(grp) - 1) == lane) {
        double partial = v;
        for (unsigned rest = grp & (grp - 1); rest; rest &= rest - 1) partial = partial + xch[__ffs(rest) - 1];
        acc[slot] = acc[slot] + partial;
    }
    __syncwarp();
}

// Two-step row for ONE WARP (see the header comment).
template <bool NUMERIC>
__device__ __forceinline__ bool galerkin_row_two_step(const GalerkinArgs& a, int32_t i, int lane, GalerkinWarpSmem& sm,
                                                      GalerkinWarpVals* sv) {
    const int32_t rb = __ldg(a.r_ptr + i), re = __ldg(a.r_ptr + i + 1);
    const int32_t nr = re - rb;
    if (nr > kGMaxR) return false;
    for (int s = lane; s < kGFineSlots; s += 32) { sm.fkeys[s] = -1; if (NUMERIC) sv->fvals[s] = 0.0; }
    for (int s = lane; s < kGWarpSlots; s += 32) { sm.keys[s] = -1; if (NUMERIC) sv->vals[s] = 0.0; }
    if (lane < 2) sm.scount[lane] = 0;
    // ---- R's row: sorted by fine row (the transpose fills it in arrival order), with the scan of the lengths of
    //      A's outer vectors, which numbers the products
    {
        int32_t k0 = 0x7fffffff, k1 = 0x7fffffff;
        double p0 = 0.0, p1 = 0.0;
        if (lane < nr) { k0 = __ldg(a.r_col + rb + lane); if (NUMERIC) p0 = __ldg(a.r_val + rb + lane); }
        if (lane + 32 < nr) { k1 = __ldg(a.r_col + rb + lane + 32); if (NUMERIC) p1 = __ldg(a.r_val + rb + lane + 32); }
        sm.rbase[lane] = k0; sm.rbase[lane + 32] = k1;        // scratch: the unsorted keys
        __syncwarp();
        int32_t rank0 = 0, rank1 = 0;
        for (int32_t e = 0; e < nr; ++e) { const int32_t o = sm.rbase[e]; rank0 += o < k0 ? 1 : 0; rank1 += o < k1 ? 1 : 0; }
        __syncwarp();
        if (lane < nr) { sm.rk[rank0] = k0; if (NUMERIC) sv->rp[rank0] = p0; }
        if (lane + 32 < nr) { sm.rk[rank1] = k1; if (NUMERIC) sv->rp[rank1] = p1; }
        __syncwarp();
    }
    // ---- step 1: fine-keyed. A sub-step takes 4 rows k of R (8 lanes each) and entries [q0, q0 + 8) of A's outer
    //      vectors k; lane = 8 * (k's place among the 4) + (q - q0).
    bool full = false;
    constexpr int kSub = 8;
    for (int32_t e0 = 0; e0 < nr; e0 += 32 / kSub) {
        const int32_t e = e0 + lane / kSub;
        int32_t ab = 0, ae = 0;
        double pki = 0.0;
        if (e < nr) {
            const int32_t k = sm.rk[e];
            ab = __ldg(a.a_ptr + k); ae = __ldg(a.a_ptr + k + 1);
            if (NUMERIC) pki = sv->rp[e];
        }
        const int32_t longest = __reduce_max_sync(0xffffffffu, ae - ab);
        for (int32_t q0 = 0; q0 < longest; q0 += kSub) {
            const int32_t q = ab + q0 + lane % kSub;
            const bool valid = q < ae;
            int32_t l = 0;
            double w = 0.0;
            int slot = -1;
            if (valid) {
                l = __ldg(a.a_idx + q);
                if (NUMERIC) w = pki * __ldg(a.a_val + q);
                slot = hash_insert<kGFineSlots>(sm.fkeys, l, sm.scount);
                if (slot < 0) full = true;
            }
            if (NUMERIC) group_accumulate(valid && slot >= 0, l, slot, w, sv->fvals, sv->xch, lane);
        }
    }
    __syncwarp();
    if (__any_sync(0xffffffffu, full) || sm.scount[0] > kGFineSlots * 3 / 4) return false;
    // compact the occupied fine slots to the front (in place, chunk read before chunk written)
    int32_t kf = 0;
    for (int s0 = 0; s0 < kGFineSlots; s0 += 32) {
        const int32_t key = sm.fkeys[s0 + lane];
        const double v = NUMERIC ? sv->fvals[s0 + lane] : 0.0;
        const unsigned occ = __ballot_sync(0xffffffffu, key >= 0);
        __syncwarp();
        if (key >= 0) { const int pos = kf + __popc(occ & ((1u << lane) - 1u)); sm.fkeys[pos] = key; if (NUMERIC) sv->fvals[pos] = v; }
        kf += __popc(occ);
        __syncwarp();
    }
    if (NUMERIC) {
        // sort the distinct fine indices ascending (through registers: rank every entry, then write it back)
        constexpr int kPer = kGFineSlots * 3 / 4 / 32;      // entries per lane
        int32_t key[kPer], rank[kPer];
        double val[kPer];
#pragma unroll
        for (int u = 0; u < kPer; ++u) {
            const int f = lane + 32 * u;
            key[u] = f < kf ? sm.fkeys[f] : 0x7fffffff;
            val[u] = f < kf ? sv->fvals[f] : 0.0;
            rank[u] = 0;
        }
        for (int32_t f = 0; f < kf; ++f) {
            const int32_t o = sm.fkeys[f];
#pragma unroll
            for (int u = 0; u < kPer; ++u) rank[u] += o < key[u] ? 1 : 0;
        }
        __syncwarp();
#pragma unroll
        for (int u = 0; u < kPer; ++u)
            if (lane + 32 * u < kf) { sm.fkeys[rank[u]] = key[u]; sv->fvals[rank[u]] = val[u]; }
        __syncwarp();
    }
    // ---- step 2: coarse-keyed, one lane per distinct fine index, entry t of every row of P in turn. The sorted
    //      fine indices are dealt out INTERLEAVED (round r takes f = lane * rounds + r): neighbouring fine rows
    //      share their coarse columns, and lanes of one sub-step that hold the same column are summed one by one.
    const int32_t rounds = (kf + 31) / 32;
    for (int32_t r = 0; r < rounds; ++r) {
        const int32_t f = lane * rounds + r;
        int32_t pb = 0, pe = 0;
        double cl = 0.0;
        if (f < kf) {
            const int32_t l = sm.fkeys[f];
            pb = __ldg(a.p_ptr + l); pe = __ldg(a.p_ptr + l + 1);
            if (NUMERIC) cl = sv->fvals[f];
        }
        const int32_t longest = __reduce_max_sync(0xffffffffu, pe - pb);
        for (int32_t t = 0; t < longest; ++t) {
            const bool valid = pb + t < pe;
            int32_t j = 0;
            int slot = -1;
            if (valid) {
                j = __ldg(a.p_col + pb + t);
                slot = hash_insert<kGWarpSlots>(sm.keys, j, sm.scount + 1);
                if (slot < 0) full = true;
            }
            if (NUMERIC) {
                const double v = valid ? cl * __ldg(a.p_val + pb + t) : 0.0;
                group_accumulate(valid && slot >= 0, j, slot, v, sv->vals, sv->xch, lane);
            }
        }
    }
    __syncwarp();
    const int32_t K = sm.scount[1];
    if (__any_sync(0xffffffffu, full) || K > kGWarpSlots * 3 / 4) return false;
    if (!NUMERIC) { if (lane == 0) a.out_count[i] = K; return true; }
    const int32_t off = __ldg(a.out_ptr + i);
    int32_t base = 0;
    for (int s0 = 0; s0 < kGWarpSlots; s0 += 32) {
        const int32_t key = sm.keys[s0 + lane];
        const double v = sv->vals[s0 + lane];
        const unsigned occ = __ballot_sync(0xffffffffu, key >= 0);
        __syncwarp();
        if (key >= 0) { const int pos = base + __popc(occ & ((1u << lane) - 1u)); sm.keys[pos] = key; sv->vals[pos] = v; }
        base += __popc(occ);
        __syncwarp();
    }
    for (int s = lane; s < K; s += 32) {
        const int32_t key = sm.keys[s];
        int32_t rank = 0;
        for (int s2 = 0; s2 < K; ++s2) rank += sm.keys[s2] < key ? 1 : 0;
        a.out_idx[off + rank] = key;
        a.out_val[off + rank] = sv->vals[s];
    }
    return true;
}

template <bool NUMERIC>
__global__ void __launch_bounds__(kBlock) galerkin_warp_kernel(GalerkinArgs a) {
    __shared__ GalerkinWarpSmem sm[kBlock / 32];
    __shared__ GalerkinWarpVals sv[NUMERIC ? kBlock / 32 : 1];
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t i = int64_t(blockIdx.x) * (kBlock / 32) + w;
    if (i >= a.m) return;
    if (!galerkin_row_two_step<NUMERIC>(a, int32_t(i), lane, sm[w], NUMERIC ? &sv[w] : nullptr)) {
        // symbolic pass records the overflow; the numeric pass skips rows already on the list
        if (!NUMERIC && lane == 0) { a.out_count[i] = 0; a.overflow[atomicAdd(a.flags + 1, 1)] = int32_t(i); }
    }
}

template <bool NUMERIC>
__global__ void __launch_bounds__(kBlock) galerkin_block_kernel(GalerkinArgs a, int32_t n_overflow) {
    __shared__ int32_t keys[kGBlockSlots];
    __shared__ double vals[NUMERIC ? kGBlockSlots : 1];
    __shared__ int32_t scount;
    for (int32_t o = blockIdx.x; o < n_overflow; o += gridDim.x) {
        const int32_t i = a.overflow[o];
        if (!galerkin_row<NUMERIC, kBlock, kGBlockSlots>(a, i, threadIdx.x, keys, vals, &scount)) {
            if (threadIdx.x == 0) a.flags[2] = 1;
        }
        __syncthreads();
    }
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
    // the warp kernel sorts R's rows itself while it stages them; a restriction that is KEPT is sorted once here
    transpose_csr(c, P, R, c_lo, c_hi, keep_R != nullptr);
    tr.mark("transpose");
    if (marks) cudaEventRecord(marks[1], c->stream);
    DevBuf<int32_t> count(c, int64_t(m) + 1), flags(c, 4), overflow(c, m);
    DevBuf<int64_t> total(c, 1);
    count.zero(); flags.zero();
    GalerkinArgs a{m, R.rowptr.p, R.colidx.p, R.val.p, a_colptr, a_rowidx, a_val, P.rowptr.p, P.colidx.p, P.val.p,
                   count.p, nullptr, nullptr, nullptr, flags.p, overflow.p};
    const int grid = div_up(m, kBlock / 32);
    // pass 1: symbolic
    GMG_LAUNCH(c, galerkin_warp_kernel<false>, grid, kBlock, 0, a);
    int32_t hf[4];
    flags.download(hf, 4);
    sync(c);
    const int32_t n_over = hf[1];
    if (getenv("GMG_TRACE")) fprintf(stderr, "[gmg] galerkin: %d outer vectors, %d to the block-wide table\n", m, n_over);
    const int ogrid = n_over < 4 * c->sm_count ? n_over : 4 * c->sm_count;
    if (n_over > 0) {
        GMG_LAUNCH(c, galerkin_block_kernel<false>, ogrid, kBlock, 0, a, n_over);
        flags.download(hf, 4);
        sync(c);
        GMG_REQUIRE(hf[2] == 0, GMG_ERR_INVALID, "galerkin: a coarse row has more than 1536 entries");
    }
    exclusive_scan_i32(c, count.p, Ac.colptr.p, int64_t(m) + 1, total.p);
    total.download(&Ac.nnz, 1);
    sync(c);
    tr.mark("symbolic");
    if (marks) cudaEventRecord(marks[2], c->stream);
    GMG_REQUIRE(Ac.nnz < (int64_t(1) << 31), GMG_ERR_INVALID, "galerkin: A_c exceeds 32-bit indexing");
    Ac.rowidx.alloc(c, Ac.nnz);
    Ac.val.alloc(c, Ac.nnz);
    // pass 2: numeric
    a.out_ptr = Ac.colptr.p; a.out_idx = Ac.rowidx.p; a.out_val = Ac.val.p;
    GMG_LAUNCH(c, galerkin_warp_kernel<true>, grid, kBlock, 0, a);
    if (n_over > 0) GMG_LAUNCH(c, galerkin_block_kernel<true>, ogrid, kBlock, 0, a, n_over);
    tr.mark("numeric");
    if (marks) cudaEventRecord(marks[3], c->stream);
}

}  // namespace gmg
