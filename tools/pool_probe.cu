// Diagnostic: what does a large cudaMallocAsync cost once the pool holds the memory? (nvcc -O2 tools/pool_probe.cu)
#include <chrono>
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
static double now() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
int main() {
    cudaSetDevice(0);
    cudaStream_t s; cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking);
    cudaMemPool_t pool; cudaDeviceGetDefaultMemPool(&pool, 0);
    uint64_t thr = UINT64_MAX; cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr);
    const size_t GB = size_t(1) << 30;
    for (int rep = 0; rep < 4; ++rep) {
        void* a; void* b; void* c;
        double t0 = now(); cudaMallocAsync(&a, 3 * GB, s); cudaStreamSynchronize(s); double t1 = now();
        cudaMallocAsync(&b, 1 * GB, s); cudaStreamSynchronize(s); double t2 = now();
        cudaFreeAsync(a, s); cudaStreamSynchronize(s); double t3 = now();
        cudaMallocAsync(&c, 2 * GB, s); cudaStreamSynchronize(s); double t4 = now();     // fits in the hole a left
        cudaFreeAsync(b, s); cudaFreeAsync(c, s); cudaStreamSynchronize(s); double t5 = now();
        uint64_t reserved = 0, used = 0;
        cudaMemPoolGetAttribute(pool, cudaMemPoolAttrReservedMemCurrent, &reserved);
        cudaMemPoolGetAttribute(pool, cudaMemPoolAttrUsedMemCurrent, &used);
        printf("rep %d: alloc 3GB %.3f ms, alloc 1GB %.3f ms, free 3GB %.3f ms, alloc 2GB %.3f ms, frees %.3f ms; reserved %.2f GB used %.2f GB\n",
               rep, t1 - t0, t2 - t1, t3 - t2, t4 - t3, t5 - t4, reserved / double(GB), used / double(GB));
    }
    // a long sequence of mixed sizes, repeated: does the pool settle?
    for (int rep = 0; rep < 4; ++rep) {
        void* p[16]; size_t sz[16];
        double t0 = now();
        for (int i = 0; i < 16; ++i) { sz[i] = ((i * 7) % 5 + 1) * (GB / 2); cudaMallocAsync(&p[i], sz[i], s); if (i % 3 == 2) { cudaFreeAsync(p[i - 1], s); p[i - 1] = nullptr; } }
        cudaStreamSynchronize(s);
        double t1 = now();
        for (int i = 0; i < 16; ++i) if (p[i]) cudaFreeAsync(p[i], s);
        cudaStreamSynchronize(s);
        uint64_t reserved = 0; cudaMemPoolGetAttribute(pool, cudaMemPoolAttrReservedMemCurrent, &reserved);
        printf("mixed rep %d: 16 allocs %.3f ms; reserved %.2f GB\n", rep, t1 - t0, reserved / double(GB));
    }
    return 0;
}
