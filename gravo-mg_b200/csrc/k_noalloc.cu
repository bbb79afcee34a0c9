// averageEdgeLength without any allocation [no counterpart in the reference]:
// for callers that cannot call cudaMalloc -- code under CUDA graph capture, tight serving loops, harnesses
// that forbid it. gmg_average_edge_length (capi.cu / k_basic.cu) copies its host arrays to the device and takes its
// scratch (one partial sum per block) from the context's pool; this variant takes DEVICE inputs and a DEVICE output
// from the caller and keeps the partial sums in a file-scope __device__ array of fixed size.
//
//   * the scratch is kNoAllocMaxBlocks doubles = 512 KiB, held by every GPU that loads the library until exit;
//   * that bounds the edge list at kNoAllocMaxBlocks * 2048 = 134,217,728 rows: more is rejected (GMG_ERR_INVALID);
//   * the entry point calls no cudaMalloc / cudaFree / allocating library routine (the once-per-device set-up is
//     cudaGetSymbolAddress, which allocates nothing; the array is part of the loaded module);
//   * every call on a device shares the array. Nothing in it has to start at zero: a call overwrites the partials it
//     reads (one per block of its own grid) and never reads beyond them, so the previous call's contents do not
//     matter. Calls that OVERLAP on the GPU (different streams, or different contexts on one device) race on it: the
//     caller serialises them;
//   * the work is queued on the context's stream and not waited for: *d_out_mean is valid once that stream has got
//     there (stream order, an event, or a synchronise -- the caller's choice).
//
// Same summation tree as the allocating entry point (2048 rows per block with the fixed block tree, then one block over
// the partials), so the two agree bit for bit.
#include "common.cuh"

#include <mutex>

using namespace gmg;

namespace {

constexpr int kNoAllocMaxBlocks = 65536;
constexpr int kNoAllocPerThread = 8;                 // = kSumPerThread of k_basic.cu: rows per thread of a block
constexpr int64_t kNoAllocMaxEdges = int64_t(kNoAllocMaxBlocks) * kBlock * kNoAllocPerThread;

// In HOST code this name is only a stand-in for the symbol, not a device address: host code must go through
// cudaGetSymbolAddress (scratch_address below). Device code names it directly.
static __device__ __align__(256) double g_noalloc_partial[kNoAllocMaxBlocks];

// the fixed block tree of k_basic.cu's block_sum (warp shuffle tree, then the 8 warp sums by one more tree)
__device__ __forceinline__ double noalloc_block_sum(double v) {
    __shared__ double ws[kBlock / 32];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v += __shfl_down_sync(0xffffffffu, v, d);
    if (lane == 0) ws[wid] = v;
    __syncthreads();
    double r = 0.0;
    if (wid == 0) {
        r = lane < kBlock / 32 ? ws[lane] : 0.0;
#pragma unroll
        for (int d = 4; d > 0; d >>= 1) r += __shfl_down_sync(0xffffffffu, r, d);
    }
    __syncthreads();
    return r;  // valid in thread 0
}

__global__ void __launch_bounds__(kBlock) noalloc_length_partial(int32_t n, const double* __restrict__ pts, int64_t n_edges,
                                                                 const int32_t* __restrict__ edges, double* __restrict__ partial) {
    const int64_t base = int64_t(blockIdx.x) * kBlock * kNoAllocPerThread;
    double s = 0.0;
#pragma unroll
    for (int k = 0; k < kNoAllocPerThread; ++k) {
        const int64_t e = base + int64_t(k) * kBlock + threadIdx.x;
        if (e < n_edges) s += norm(ld3(pts, n, __ldg(edges + n_edges + e)) - ld3(pts, n, __ldg(edges + e)));
    }
    s = noalloc_block_sum(s);
    if (threadIdx.x == 0) partial[blockIdx.x] = s;
}

__global__ void __launch_bounds__(kBlock) noalloc_sum_partials(const double* __restrict__ partial, int64_t count, double denom,
                                                               double* __restrict__ out) {
    double s = 0.0;
    for (int64_t i = threadIdx.x; i < count; i += kBlock) s += partial[i];
    s = noalloc_block_sum(s);
    if (threadIdx.x == 0) *out = s / denom;
}

// The scratch array's DEVICE address on `device`, looked up once per device with that device current.
double* scratch_address(int device) {
    constexpr int kMaxDevices = 64;
    static std::mutex mutex;
    static double* address[kMaxDevices] = {};
    GMG_REQUIRE(device >= 0 && device < kMaxDevices, GMG_ERR_INVALID, "no-alloc entry point: device index out of range");
    std::lock_guard<std::mutex> lock(mutex);
    if (!address[device]) {
        GMG_CUDA(cudaSetDevice(device));
        double* p = nullptr;
        GMG_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_noalloc_partial));
        GMG_REQUIRE(p != nullptr, GMG_ERR_CUDA, "no-alloc entry point: the scratch array has no device address");
        address[device] = p;
    }
    return address[device];
}

}  // namespace

extern "C" {

/* d_points: n x 3 column-major PACKED (x[0..n) y[0..n) z[0..n)), d_edges: n_edges x 2 column-major (first endpoints, then
 * second endpoints), d_out_mean: one double -- all DEVICE pointers of ctx's device. */
int gmg_average_edge_length_noalloc(gmg_ctx* c, int64_t n, const double* d_points, int64_t n_edges, const int32_t* d_edges,
                                    double* d_out_mean) {
    if (!c) return GMG_ERR_STATE;
    try {
        GMG_REQUIRE(n >= 0 && n < (int64_t(1) << 31) && n_edges >= 0, GMG_ERR_INVALID, "averageEdgeLength (no-alloc): bad sizes");
        GMG_REQUIRE(n_edges <= kNoAllocMaxEdges, GMG_ERR_INVALID,
                    "averageEdgeLength (no-alloc): more than 134,217,728 edge rows (the fixed scratch holds one partial sum per "
                    "2048 rows); use gmg_average_edge_length");
        GMG_REQUIRE(d_out_mean && (n == 0 || d_points) && (n_edges == 0 || d_edges), GMG_ERR_INVALID,
                    "averageEdgeLength (no-alloc): null device pointer");
        GMG_CUDA(cudaSetDevice(c->device));
        double* partial = scratch_address(c->device);
        // (an empty edge list sums zero partials: 0 / (0 - n), like the allocating entry point)
        const int64_t blocks = (n_edges + int64_t(kBlock) * kNoAllocPerThread - 1) / (int64_t(kBlock) * kNoAllocPerThread);
        if (blocks > 0)
            GMG_LAUNCH(c, noalloc_length_partial, int(blocks), kBlock, 0, int32_t(n), d_points, n_edges, d_edges, partial);
        GMG_LAUNCH(c, noalloc_sum_partials, 1, kBlock, 0, partial, blocks, double(n_edges - n), d_out_mean);
        c->last_error.clear();
        return GMG_OK;
    } catch (const Error& e) {
        c->last_error = e.what;
        return e.status;
    }
}

}  // extern "C"
