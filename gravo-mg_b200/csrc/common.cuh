// Shared device/host plumbing for the gravomg_b200 kernels: error handling, the context, stream-
// ordered device buffers, a work-efficient exclusive scan, and the 3-vector arithmetic whose
// association order matches the oracle (oracle/eigen_shim/Eigen/Core): (x+y)+z, no FMA.
//
// The whole library is compiled with -fmad=false: every floating-point result on this path either
// feeds a comparison that decides an integer output or must agree with the reference to 1e-12.
#pragma once

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include <cuda_runtime.h>
#include <cooperative_groups.h>

#include <cfloat>
#include <cstdint>
#include <cstdio>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/gravomg_b200.h"

namespace gmg {

namespace cg = cooperative_groups;

struct Error {
    int status;
    std::string what;
};

#define GMG_CUDA(expr)                                                                              \
    do {                                                                                            \
        cudaError_t e__ = (expr);                                                                   \
        if (e__ != cudaSuccess)                                                                     \
            throw ::gmg::Error{GMG_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e__) +  \
                                                 " (" __FILE__ ":" + std::to_string(__LINE__) + ")"}; \
    } while (0)

#define GMG_REQUIRE(cond, status, msg)                                \
    do {                                                              \
        if (!(cond)) throw ::gmg::Error{status, std::string(msg)};    \
    } while (0)

constexpr int kBlock = 256;

inline int div_up(int64_t a, int64_t b) { return int((a + b - 1) / b); }

}  // namespace gmg

// The context owns the stream, timing events and the retained variable-size result.
struct gmg_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    cudaStream_t copy_stream = nullptr;   // device -> host copies that overlap the build (created on first use)
    int sm_count = 148;
    std::string last_error;
    double last_kernel_ms = 0.0;
    int64_t launches = 0;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    // retained result of the last two-call function (device memory, released on the next call)
    void* keep[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    int64_t keep_n[6] = {0, 0, 0, 0, 0, 0};
    int keep_kind = 0;  // 0 none, 1 sparse (inner,val), 2 triangles (tri,normal,fan)
    // Released device buffers, by size class, for reuse by later allocations on this context's stream.
    // cudaMallocAsync's own pool keeps the memory too, but with mixed multi-GB sizes it re-maps physical
    // chunks on most allocations (measured here: ~0.4 ms per allocation, 8.5 ms/GB when it has to grow); a
    // build repeats the same sizes level after level and call after call, so exact-class reuse costs nothing.
    std::unordered_map<size_t, std::vector<void*>> cache;
    size_t cached_bytes = 0;
    size_t cache_limit = size_t(64) << 30;   // set to 40 % of the device's memory at creation
    // Small device -> host reads (counts, flags, the per-level status block) go through a MAPPED pinned page that
    // a tiny kernel writes: no copy-engine descriptor, no implicit host block (cudaMemcpyAsync into pageable
    // memory returns only when the copy is done) and no queueing behind a bulk copy on the copy stream.
    // sync() hands the staged bytes to their destinations.
    unsigned char* stage = nullptr;
    size_t stage_used = 0;
    struct Pending { void* dst; size_t off, bytes; };
    std::vector<Pending> pending;
    void drop_pending() { pending.clear(); stage_used = 0; }
    void trim() {   // give everything cached back to the driver's pool
        for (auto& kv : cache)
            for (void* q : kv.second) cudaFreeAsync(q, stream);
        cache.clear();
        cached_bytes = 0;
    }
};

namespace gmg {

// Stream-ordered device buffer (cudaMallocAsync pool): allocation and release cost no device sync.
constexpr size_t kStageBytes = 8192;
static __global__ void stage_copy_kernel(const uint32_t* __restrict__ src, uint32_t* __restrict__ dst, int words) {
    for (int i = threadIdx.x; i < words; i += blockDim.x) dst[i] = src[i];
}

// Size class of a request: the next multiple of 1/8 of the largest power of two not above it (<= 12.5 % slack).
inline size_t size_class(size_t bytes) {
    if (bytes < 512) bytes = 512;
    size_t p2 = size_t(1) << (63 - __builtin_clzll((unsigned long long)bytes));
    const size_t step = p2 / 8;
    return (bytes + step - 1) / step * step;
}

// The buffers live on the context's single stream, so handing a released block to the next request of the
// same class is stream-ordered by construction (what cudaFreeAsync + cudaMallocAsync would give, minus the
// driver's re-mapping). A context must outlive its buffers.
template <typename T> struct DevBuf {
    T* p = nullptr;
    int64_t n = 0;
    cudaStream_t s = nullptr;
    gmg_ctx* ctx = nullptr;
    size_t cap = 0;             // bytes of the underlying block (its size class)
    DevBuf() {}
    DevBuf(gmg_ctx* c, int64_t count) { alloc(c, count); }
    DevBuf(const DevBuf&) = delete;
    DevBuf& operator=(const DevBuf&) = delete;
    DevBuf(DevBuf&& o) noexcept : p(o.p), n(o.n), s(o.s), ctx(o.ctx), cap(o.cap) { o.p = nullptr; o.n = 0; }
    DevBuf& operator=(DevBuf&& o) noexcept {
        if (this != &o) { release(); p = o.p; n = o.n; s = o.s; ctx = o.ctx; cap = o.cap; o.p = nullptr; o.n = 0; }
        return *this;
    }
    ~DevBuf() { release(); }
    void alloc(gmg_ctx* c, int64_t count) {
        release();
        s = c->stream; n = count; ctx = c;
        cap = size_class(sizeof(T) * size_t(count > 0 ? count : 1));
        auto it = c->cache.find(cap);
        if (it != c->cache.end() && !it->second.empty()) {
            p = static_cast<T*>(it->second.back());
            it->second.pop_back();
            c->cached_bytes -= cap;
            return;
        }
        cudaError_t e = cudaMallocAsync(reinterpret_cast<void**>(&p), cap, s);
        if (e == cudaErrorMemoryAllocation) {       // out of memory with blocks parked in the cache: drop them, retry
            cudaGetLastError();
            c->trim();
            cudaStreamSynchronize(s);
            e = cudaMallocAsync(reinterpret_cast<void**>(&p), cap, s);
        }
        if (e != cudaSuccess) { p = nullptr; n = 0; }
        GMG_CUDA(e);
    }
    void release() {
        if (p) {
            ctx->cache[cap].push_back(p);
            ctx->cached_bytes += cap;
            if (ctx->cached_bytes > ctx->cache_limit) ctx->trim();
        }
        p = nullptr; n = 0;
    }
    T* detach() { T* q = p; p = nullptr; n = 0; return q; }   // the caller frees it with cudaFreeAsync
    void upload(const T* host, int64_t count) {
        if (count > 0) GMG_CUDA(cudaMemcpyAsync(p, host, sizeof(T) * size_t(count), cudaMemcpyHostToDevice, s));
    }
    void download(T* host, int64_t count) const {
        if (count <= 0) return;
        const size_t bytes = sizeof(T) * size_t(count);
        if (ctx && ctx->stage && bytes <= 512 && bytes % 4 == 0 && ctx->stage_used + bytes <= kStageBytes) {
            const size_t off = ctx->stage_used;
            ctx->stage_used += (bytes + 15) / 16 * 16;
            stage_copy_kernel<<<1, 64, 0, s>>>(reinterpret_cast<const uint32_t*>(p), reinterpret_cast<uint32_t*>(ctx->stage + off),
                                               int(bytes / 4));
            GMG_CUDA(cudaGetLastError());
            ++ctx->launches;
            ctx->pending.push_back({host, off, bytes});     // delivered by sync()
            return;
        }
        GMG_CUDA(cudaMemcpyAsync(host, p, bytes, cudaMemcpyDeviceToHost, s));
    }
    void zero() { GMG_CUDA(cudaMemsetAsync(p, 0, sizeof(T) * size_t(n > 0 ? n : 1), s)); }
};

template <typename T> DevBuf<T> to_device(gmg_ctx* c, const T* host, int64_t count) {
    DevBuf<T> b(c, count);
    b.upload(host, count);
    return b;
}

// Column-major n x 3 points with leading dimension ld -> packed x|y|z on the device (ld = n).
inline DevBuf<double> points_to_device(gmg_ctx* c, const double* host, int64_t n, int64_t ld) {
    DevBuf<double> b(c, 3 * n);
    for (int k = 0; k < 3; ++k)
        if (n > 0)
            GMG_CUDA(cudaMemcpyAsync(b.p + k * n, host + k * ld, sizeof(double) * size_t(n), cudaMemcpyHostToDevice,
                                     c->stream));
    return b;
}
inline void points_to_host(gmg_ctx* c, const double* dev, int64_t n, double* host, int64_t ld) {
    for (int k = 0; k < 3; ++k)
        if (n > 0)
            GMG_CUDA(cudaMemcpyAsync(host + k * ld, dev + k * n, sizeof(double) * size_t(n), cudaMemcpyDeviceToHost,
                                     c->stream));
}

inline void sync(gmg_ctx* c) {
    const cudaError_t e = cudaStreamSynchronize(c->stream);
    if (e == cudaSuccess)
        for (const auto& q : c->pending) memcpy(q.dst, c->stage + q.off, q.bytes);
    c->drop_pending();
    GMG_CUDA(e);
}
// after an error: wait for the stream and forget staged reads whose destinations may be gone
inline void abort_sync(gmg_ctx* c) {
    cudaStreamSynchronize(c->stream);
    c->drop_pending();
}

// GMG_TRACE=1 in the environment: wall-clock milliseconds between marks, each mark synchronising the
// stream (diagnostic only; changes the timing it reports on, never the results).
struct Trace {
    gmg_ctx* c;
    const char* what;
    bool on;
    std::chrono::steady_clock::time_point t0;
    Trace(gmg_ctx* ctx, const char* name) : c(ctx), what(name), on(getenv("GMG_TRACE") != nullptr) {
        if (on) { cudaStreamSynchronize(c->stream); t0 = std::chrono::steady_clock::now(); }
    }
    void mark(const char* label) {
        if (!on) return;
        cudaStreamSynchronize(c->stream);
        const auto t1 = std::chrono::steady_clock::now();
        fprintf(stderr, "[gmg] %s: %s %.3f ms\n", what, label, std::chrono::duration<double, std::milli>(t1 - t0).count());
        t0 = t1;
    }
};

#define GMG_LAUNCH(ctx, kernel, grid, block, smem, ...)                         \
    do {                                                                        \
        kernel<<<(grid), (block), (smem), (ctx)->stream>>>(__VA_ARGS__);        \
        GMG_CUDA(cudaGetLastError());                                           \
        ++(ctx)->launches;                                                      \
    } while (0)

// Device-side exclusive scan of int32 counts (in place allowed); *total (device int64) receives the sum.
void exclusive_scan_i32(gmg_ctx* c, const int32_t* in, int32_t* out, int64_t n, int64_t* d_total);
// validation of caller-supplied index arrays that kernels use as addresses (throw GMG_ERR_INVALID)
void check_index_range(gmg_ctx* c, int64_t count, const int32_t* idx, int64_t upper, const char* what);
void check_outer_pointers(gmg_ctx* c, int64_t n, const int32_t* outer, const char* what);

// ---- 3-vector arithmetic: same association order as the oracle ---------------------------------
struct V3 { double x, y, z; };
__device__ __forceinline__ V3 ld3(const double* __restrict__ p, int64_t n, int64_t i) {
    return V3{__ldg(p + i), __ldg(p + n + i), __ldg(p + 2 * n + i)};
}
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return V3{a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ V3 operator*(V3 a, double s) { return V3{a.x * s, a.y * s, a.z * s}; }
__device__ __forceinline__ double dot(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
__device__ __forceinline__ double norm(V3 a) { return sqrt(dot(a, a)); }
__device__ __forceinline__ V3 cross(V3 a, V3 b) {
    return V3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
__device__ __forceinline__ V3 normalized(V3 a) {
    double z = dot(a, a);
    if (z > 0.0) { double s = sqrt(z); return V3{a.x / s, a.y / s, a.z / s}; }
    return a;
}

// Warp-aggregated queue append: the converged lanes that reach this call together take consecutive
// slots with ONE atomic on the shared counter (a single address serialises in the L2 atomic unit at
// under 1 op/cycle, so per-lane atomics on one counter cap a whole level at ~2 G appends/s).
__device__ __forceinline__ int32_t warp_append(int32_t* counter) {
    const cg::coalesced_group g = cg::coalesced_threads();
    int32_t base = 0;
    if (g.thread_rank() == 0) base = atomicAdd(counter, int32_t(g.size()));
    return g.shfl(base, 0) + int32_t(g.thread_rank());
}
__device__ __forceinline__ void warp_count(unsigned long long* counter) {
    const cg::coalesced_group g = cg::coalesced_threads();
    if (g.thread_rank() == 0) atomicAdd(counter, (unsigned long long)g.size());
}

// lower_bound of key in the ascending int32 range [lo, hi); returns hi when every entry is smaller
__device__ __forceinline__ int32_t lower_bound_i32(const int32_t* __restrict__ a, int32_t lo, int32_t hi, int32_t key) {
    while (lo < hi) {
        int32_t mid = lo + ((hi - lo) >> 1);
        if (__ldg(a + mid) < key) lo = mid + 1; else hi = mid;
    }
    return lo;
}

}  // namespace gmg
