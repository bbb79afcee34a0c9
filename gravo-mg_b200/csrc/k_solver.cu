// SpMV, multicolour Gauss-Seidel and the V-cycle used to validate a hierarchy.
// [not in reference: the fork kept only the P/R builders (README.md:6); BASELINE.json's north star
//  asks for "the SpMV and multicolour Gauss-Seidel smoothing used to validate a V-cycle".]
//
// Matrices are compressed by outer index. spmv() computes y[i] = sum_k val[k] * x[inner[k]] over
// outer vector i: for row-major storage that is y = A x, for column-major storage y = A^T x
// (identical for the symmetric operators a V-cycle works on).
//   SpMV            8 lanes per outer vector, strided partial sums, fixed shuffle tree (deterministic).
//                   Algorithmic bytes 20n + 12nnz.
//   colouring       Jones-Plassmann: a vertex that beats every uncoloured neighbour on a hashed
//                   priority takes the smallest colour unused among its coloured neighbours.
//   Gauss-Seidel    rows bucketed by colour; one launch per colour; within a colour rows are
//                   independent, so the sweep equals sequential GS in (colour, row) order.
#include "stages.h"

#include <algorithm>
#include <cmath>

namespace gmg {

constexpr int kSLanes = 8;

__global__ void __launch_bounds__(kBlock) spmv_kernel(int32_t n, const int32_t* __restrict__ outer,
                                                      const int32_t* __restrict__ inner,
                                                      const double* __restrict__ val, const double* __restrict__ x,
                                                      double* y, double alpha, const double* b) {
    // y = alpha * (M x) + b   (b may be null)
    const int64_t gid = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    const int32_t i = int32_t(gid / kSLanes);
    const int sub = int(gid % kSLanes);
    if (i >= n) return;
    double s = 0.0;
    const int32_t e = __ldg(outer + i + 1);
    for (int32_t k = __ldg(outer + i) + sub; k < e; k += kSLanes) s += __ldg(val + k) * x[__ldg(inner + k)];
#pragma unroll
    for (int d = kSLanes / 2; d > 0; d >>= 1) s += __shfl_xor_sync(0xffffffffu, s, d, kSLanes);
    if (sub == 0) y[i] = alpha * s + (b ? b[i] : 0.0);
}

void spmv(gmg_ctx* c, int32_t n, const int32_t* outer, const int32_t* inner, const double* val, const double* x,
          double* y, double alpha, const double* b) {
    if (n > 0) GMG_LAUNCH(c, spmv_kernel, div_up(int64_t(n) * kSLanes, kBlock), kBlock, 0, n, outer, inner, val, x, y, alpha, b);
}

__global__ void __launch_bounds__(kBlock) axpy_kernel(int32_t n, double a, const double* __restrict__ x, double* __restrict__ y) {
    const int64_t i = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (i < n) y[i] += a * x[i];
}
void axpy(gmg_ctx* c, int32_t n, double a, const double* x, double* y) {
    if (n > 0) GMG_LAUNCH(c, axpy_kernel, div_up(n, kBlock), kBlock, 0, n, a, x, y);
}

// ---- colouring ------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t prio(uint32_t v) {
    v *= 0x9E3779B1u; v ^= v >> 15; v *= 0x85EBCA6Bu; v ^= v >> 13; v *= 0xC2B2AE35u; v ^= v >> 16;
    return v;
}

constexpr int kMaxColours = 1024;

// A vertex that beats every uncoloured neighbour takes the SMALLEST colour none of its coloured neighbours has.
// The used colours are collected 64 at a time (one bit mask per window of 64 colours; rows with more than 64
// distinct neighbour colours -- Galerkin operators of kNN-32 clouds, deep coarse levels -- scan again for the next
// window), so two adjacent rows can never share a colour. More than kMaxColours colours is reported.
__global__ void __launch_bounds__(kBlock) colour_round_kernel(int32_t n, const int32_t* __restrict__ outer,
                                                              const int32_t* __restrict__ inner,
                                                              const int32_t* __restrict__ colour_in,
                                                              int32_t* __restrict__ colour_out, int32_t* __restrict__ remaining) {
    const int64_t i = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (i >= n) return;
    if (colour_in[i] >= 0) { colour_out[i] = colour_in[i]; return; }
    const uint32_t pi = prio(uint32_t(i));
    const int32_t b = __ldg(outer + i), e = __ldg(outer + i + 1);
    unsigned long long used = 0ull;
    bool wins = true;
    for (int32_t k = b; k < e; ++k) {
        const int32_t j = __ldg(inner + k);
        if (j == i) continue;
        const int32_t cj = colour_in[j];
        if (cj >= 0) { if (cj < 64) used |= 1ull << cj; continue; }
        const uint32_t pj = prio(uint32_t(j));
        if (pj > pi || (pj == pi && j > i)) { wins = false; break; }
    }
    if (!wins) { colour_out[i] = -1; remaining[0] = 1; return; }
    int32_t cidx = __ffsll((long long)~used) - 1;       // smallest free colour below 64, or -1
    for (int32_t base = 64; cidx < 0 && base < kMaxColours; base += 64) {
        used = 0ull;
        for (int32_t k = b; k < e; ++k) {
            const int32_t j = __ldg(inner + k);
            if (j == i) continue;
            const int32_t cj = colour_in[j] - base;
            if (cj >= 0 && cj < 64) used |= 1ull << cj;
        }
        const int32_t f = __ffsll((long long)~used) - 1;
        if (f >= 0) cidx = base + f;
    }
    if (cidx < 0) { cidx = 0; remaining[1] = 1; }        // more than kMaxColours colours: reported by the caller
    colour_out[i] = cidx;
}

__global__ void __launch_bounds__(kBlock) colour_hist_kernel(int32_t n, const int32_t* __restrict__ colour, int32_t* __restrict__ hist) {
    const int64_t i = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (i < n) atomicAdd(hist + colour[i], 1);
}
// stable bucket fill: rows of one colour in ascending index (flag -> scan -> scatter per colour)
__global__ void __launch_bounds__(kBlock) colour_flag_kernel(int32_t n, const int32_t* __restrict__ colour, int32_t cidx,
                                                             int32_t* __restrict__ flag) {
    const int64_t i = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (i < n) flag[i] = colour[i] == cidx ? 1 : 0;
}
__global__ void __launch_bounds__(kBlock) colour_scatter_kernel(int32_t n, const int32_t* __restrict__ colour, int32_t cidx,
                                                                const int32_t* __restrict__ pos, int32_t base,
                                                                int32_t* __restrict__ rows) {
    const int64_t i = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (i < n && colour[i] == cidx) rows[base + pos[i]] = int32_t(i);
}

void colour_graph(gmg_ctx* c, int32_t n, const int32_t* outer, const int32_t* inner, Colouring& out) {
    out.colour.alloc(c, n);
    out.rows.alloc(c, n);
    out.num_colours = 0;
    out.offsets.assign(1, 0);
    if (n == 0) return;
    DevBuf<int32_t> a(c, n), b(c, n), remaining(c, 2);
    GMG_CUDA(cudaMemsetAsync(a.p, 0xFF, sizeof(int32_t) * size_t(n), c->stream));
    int32_t* cur = a.p; int32_t* nxt = b.p;
    for (int round = 0; round < 4096; ++round) {
        remaining.zero();
        GMG_LAUNCH(c, colour_round_kernel, div_up(n, kBlock), kBlock, 0, n, outer, inner, cur, nxt, remaining.p);
        int32_t h[2] = {0, 0};
        remaining.download(h, 2);
        sync(c);
        GMG_REQUIRE(h[1] == 0, GMG_ERR_INVALID, "multicolour Gauss-Seidel: a row needs more than 1024 colours");
        std::swap(cur, nxt);
        if (!h[0]) break;
    }
    GMG_CUDA(cudaMemcpyAsync(out.colour.p, cur, sizeof(int32_t) * size_t(n), cudaMemcpyDeviceToDevice, c->stream));
    DevBuf<int32_t> hist(c, kMaxColours);
    hist.zero();
    GMG_LAUNCH(c, colour_hist_kernel, div_up(n, kBlock), kBlock, 0, n, out.colour.p, hist.p);
    std::vector<int32_t> hh(kMaxColours, 0);
    GMG_CUDA(cudaMemcpyAsync(hh.data(), hist.p, sizeof(int32_t) * kMaxColours, cudaMemcpyDeviceToHost, c->stream));
    sync(c);
    int nc = 0;
    for (int k = 0; k < kMaxColours; ++k) if (hh[k] > 0) nc = k + 1;
    out.num_colours = nc;
    out.offsets.assign(size_t(nc) + 1, 0);
    for (int k = 0; k < nc; ++k) out.offsets[size_t(k) + 1] = out.offsets[size_t(k)] + hh[k];
    for (int k = 0; k < nc; ++k) {
        if (hh[k] == 0) continue;
        GMG_LAUNCH(c, colour_flag_kernel, div_up(n, kBlock), kBlock, 0, n, out.colour.p, k, nxt);
        exclusive_scan_i32(c, nxt, nxt, n, nullptr);
        GMG_LAUNCH(c, colour_scatter_kernel, div_up(n, kBlock), kBlock, 0, n, out.colour.p, k, nxt, out.offsets[size_t(k)], out.rows.p);
    }
}

// ---- Gauss-Seidel ---------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlock) gs_colour_kernel(int32_t count, const int32_t* __restrict__ rows,
                                                           const int32_t* __restrict__ outer,
                                                           const int32_t* __restrict__ inner,
                                                           const double* __restrict__ val, const double* __restrict__ b,
                                                           double* __restrict__ x) {
    const int64_t gid = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    const int32_t r = int32_t(gid / kSLanes);
    const int sub = int(gid % kSLanes);
    if (r >= count) return;
    const int32_t i = __ldg(rows + r);
    double s = 0.0, d = 0.0;
    const int32_t e = __ldg(outer + i + 1);
    for (int32_t k = __ldg(outer + i) + sub; k < e; k += kSLanes) {
        const int32_t j = __ldg(inner + k);
        const double a = __ldg(val + k);
        if (j == i) d += a; else s += a * x[j];
    }
#pragma unroll
    for (int w = kSLanes / 2; w > 0; w >>= 1) {
        s += __shfl_xor_sync(0xffffffffu, s, w, kSLanes);
        d += __shfl_xor_sync(0xffffffffu, d, w, kSLanes);
    }
    if (sub == 0) x[i] = (b[i] - s) / d;
}

void gauss_seidel_rows(gmg_ctx* c, const int32_t* rows, int32_t count, const int32_t* outer, const int32_t* inner,
                       const double* val, const double* b, double* x) {
    if (count > 0) GMG_LAUNCH(c, gs_colour_kernel, div_up(int64_t(count) * kSLanes, kBlock), kBlock, 0, count, rows, outer, inner, val, b, x);
}

void gauss_seidel(gmg_ctx* c, const Colouring& col, const int32_t* outer, const int32_t* inner, const double* val,
                  const double* b, double* x, int sweeps, bool reverse) {
    for (int s = 0; s < sweeps; ++s)
        for (int kk = 0; kk < col.num_colours; ++kk) {
            const int k = reverse ? col.num_colours - 1 - kk : kk;
            const int32_t count = col.offsets[size_t(k) + 1] - col.offsets[size_t(k)];
            if (count == 0) continue;
            GMG_LAUNCH(c, gs_colour_kernel, div_up(int64_t(count) * kSLanes, kBlock), kBlock, 0, count,
                       col.rows.p + col.offsets[size_t(k)], outer, inner, val, b, x);
        }
}

// ---- norms ----------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlock) sumsq_partial_kernel(int32_t n, const double* __restrict__ x, double* __restrict__ partial) {
    __shared__ double ws[kBlock / 32];
    double s = 0.0;
    for (int64_t i = int64_t(blockIdx.x) * kBlock + threadIdx.x; i < n; i += int64_t(gridDim.x) * kBlock) s += x[i] * x[i];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) s += __shfl_down_sync(0xffffffffu, s, d);
    if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) { double t = 0.0; for (int w = 0; w < kBlock / 32; ++w) t += ws[w]; partial[blockIdx.x] = t; }
}
__global__ void __launch_bounds__(kBlock) dot_partial_kernel(int32_t n, const double* __restrict__ x, const double* __restrict__ y,
                                                             double* __restrict__ partial) {
    __shared__ double ws[kBlock / 32];
    double s = 0.0;
    for (int64_t i = int64_t(blockIdx.x) * kBlock + threadIdx.x; i < n; i += int64_t(gridDim.x) * kBlock) s += x[i] * y[i];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) s += __shfl_down_sync(0xffffffffu, s, d);
    if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) { double t = 0.0; for (int w = 0; w < kBlock / 32; ++w) t += ws[w]; partial[blockIdx.x] = t; }
}
// fixed grid and fixed summation order: the same inputs give the same bits
double dot(gmg_ctx* c, int32_t n, const double* x, const double* y) {
    if (n == 0) return 0.0;
    const int blocks = std::min(div_up(n, kBlock), 1024);
    DevBuf<double> partial(c, blocks);
    GMG_LAUNCH(c, dot_partial_kernel, blocks, kBlock, 0, n, x, y, partial.p);
    std::vector<double> h(size_t(blocks), 0.0);
    partial.download(h.data(), blocks);
    sync(c);
    double t = 0.0;
    for (double v : h) t += v;
    return t;
}

__global__ void __launch_bounds__(kBlock) xpby_kernel(int32_t n, const double* __restrict__ x, double b, double* __restrict__ y) {
    const int64_t i = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (i < n) y[i] = x[i] + b * y[i];
}
void xpby(gmg_ctx* c, int32_t n, const double* x, double b, double* y) {
    if (n > 0) GMG_LAUNCH(c, xpby_kernel, div_up(n, kBlock), kBlock, 0, n, x, b, y);
}

double norm2(gmg_ctx* c, int32_t n, const double* x) {
    if (n == 0) return 0.0;
    const int blocks = std::min(div_up(n, kBlock), 1024);
    DevBuf<double> partial(c, blocks);
    GMG_LAUNCH(c, sumsq_partial_kernel, blocks, kBlock, 0, n, x, partial.p);
    std::vector<double> h(size_t(blocks), 0.0);
    partial.download(h.data(), blocks);
    sync(c);
    double t = 0.0;
    for (double v : h) t += v;
    return std::sqrt(t);
}

}  // namespace gmg
