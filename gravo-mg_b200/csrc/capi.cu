// extern "C" entry points of include/gravomg_b200.h (per-function drop-ins). Each call uploads its
// host arrays, runs the device stages of stages.h on the context's stream and downloads the result.
// There is no CPU fallback: every path below launches CUDA kernels or fails with a status code.
#include "stages.h"

#include <cstring>
#include <stdexcept>

using namespace gmg;

namespace gmg {

__global__ void __launch_bounds__(kBlock) narrow_kernel(const int64_t* __restrict__ in, int32_t* __restrict__ out,
                                                        int64_t n, int64_t upper, int32_t* __restrict__ flag) {
    const int64_t i = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (i >= n) return;
    const int64_t v = in[i];
    if (v < 0 || v >= upper) { *flag = 1; out[i] = 0; } else out[i] = int32_t(v);
}
__global__ void __launch_bounds__(kBlock) widen_kernel(const int32_t* __restrict__ in, int64_t* __restrict__ out, int64_t n) {
    const int64_t i = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (i < n) out[i] = in[i];
}
void narrow_i64(gmg_ctx* c, const int64_t* in, int32_t* out, int64_t n, int64_t upper, int32_t* d_flag) {
    if (n > 0) GMG_LAUNCH(c, narrow_kernel, div_up(n, kBlock), kBlock, 0, in, out, n, upper, d_flag);
}
void widen_i32(gmg_ctx* c, const int32_t* in, int64_t* out, int64_t n) {
    if (n > 0) GMG_LAUNCH(c, widen_kernel, div_up(n, kBlock), kBlock, 0, in, out, n);
}

// Upload a std::vector<Index>-style int64 array as checked int32 device indices.
static DevBuf<int32_t> indices_to_device(gmg_ctx* c, const int64_t* host, int64_t count, int64_t upper, const char* what) {
    DevBuf<int64_t> wide = to_device(c, host, count);
    DevBuf<int32_t> narrow(c, count), flag(c, 1);
    flag.zero();
    narrow_i64(c, wide.p, narrow.p, count, upper, flag.p);
    int32_t h = 0;
    flag.download(&h, 1);
    sync(c);
    GMG_REQUIRE(h == 0, GMG_ERR_INVALID, std::string(what) + ": index out of range");
    return narrow;
}

static void release_kept(gmg_ctx* c) {
    for (int i = 0; i < 6; ++i) {
        if (c->keep[i]) cudaFreeAsync(c->keep[i], c->stream);
        c->keep[i] = nullptr; c->keep_n[i] = 0;
    }
    c->keep_kind = 0;
}

struct KernelTimer {
    gmg_ctx* c;
    explicit KernelTimer(gmg_ctx* ctx) : c(ctx) { cudaEventRecord(c->ev0, c->stream); }
    void stop() {
        cudaEventRecord(c->ev1, c->stream);
        cudaEventSynchronize(c->ev1);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, c->ev0, c->ev1);
        c->last_kernel_ms = ms;
    }
};

template <typename F> static int guarded(gmg_ctx* c, F&& body) {
    if (!c) return GMG_ERR_STATE;
    try {
        GMG_CUDA(cudaSetDevice(c->device));
        body();
        c->last_error.clear();
        return GMG_OK;
    } catch (const Error& e) {
        c->last_error = e.what;
        abort_sync(c);
        return e.status;
    } catch (const std::exception& e) {
        c->last_error = e.what();
        abort_sync(c);
        return GMG_ERR_INVALID;
    }
}

static void check_sizes(int64_t n, const int32_t* colptr, int64_t& nnz) {
    GMG_REQUIRE(n >= 0 && n < (int64_t(1) << 31) - 1, GMG_ERR_INVALID, "vertex count does not fit 32-bit indices");
    GMG_REQUIRE(colptr != nullptr, GMG_ERR_INVALID, "null colptr");
    GMG_REQUIRE(colptr[0] == 0 && colptr[n] >= 0, GMG_ERR_INVALID, "colptr must start at 0");
    nnz = colptr[n];
}

}  // namespace gmg

static std::string g_create_error;

extern "C" {

int gmg_create(int device, void* cuda_stream, gmg_ctx** out) {
    if (!out) return GMG_ERR_INVALID;
    *out = nullptr;
    try {
        int count = 0;
        GMG_CUDA(cudaGetDeviceCount(&count));
        GMG_REQUIRE(device >= 0 && device < count, GMG_ERR_CUDA, "no such CUDA device");
        GMG_CUDA(cudaSetDevice(device));
        gmg_ctx* c = new gmg_ctx();
        c->device = device;
        if (cuda_stream) { c->stream = static_cast<cudaStream_t>(cuda_stream); c->own_stream = false; }
        else { GMG_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking)); c->own_stream = true; }
        GMG_CUDA(cudaEventCreate(&c->ev0));
        GMG_CUDA(cudaEventCreate(&c->ev1));
        GMG_CUDA(cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, device));
        int coop = 0;
        GMG_CUDA(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, device));
        GMG_REQUIRE(coop, GMG_ERR_CUDA, "device lacks cooperative launch");
        // keep freed blocks in the stream-ordered pool instead of returning them to the driver
        cudaMemPool_t pool;
        GMG_CUDA(cudaDeviceGetDefaultMemPool(&pool, device));
        uint64_t threshold = UINT64_MAX;
        GMG_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &threshold));
        GMG_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&c->stage), kStageBytes, cudaHostAllocMapped));
        size_t free_b = 0, total_b = 0;
        GMG_CUDA(cudaMemGetInfo(&free_b, &total_b));
        c->cache_limit = total_b / 5 * 2;
        *out = c;
        return GMG_OK;
    } catch (const Error& e) {
        g_create_error = e.what;
        return e.status;
    }
}

void gmg_destroy(gmg_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    release_kept(c);
    c->trim();
    cudaStreamSynchronize(c->stream);
    if (c->ev0) cudaEventDestroy(c->ev0);
    if (c->ev1) cudaEventDestroy(c->ev1);
    if (c->copy_stream) { cudaStreamSynchronize(c->copy_stream); cudaStreamDestroy(c->copy_stream); }
    if (c->stage) cudaFreeHost(c->stage);
    if (c->own_stream) cudaStreamDestroy(c->stream);
    delete c;
}

/* Device memory parked for reuse by this context goes back to the driver's pool; *cached_bytes (optional)
 * receives how much that was. */
int gmg_trim(gmg_ctx* c, int64_t* cached_bytes) {
    if (!c) return GMG_ERR_STATE;
    cudaSetDevice(c->device);
    if (cached_bytes) *cached_bytes = int64_t(c->cached_bytes);
    c->trim();
    return GMG_OK;
}

const char* gmg_last_error(const gmg_ctx* c) { return c ? c->last_error.c_str() : g_create_error.c_str(); }
double gmg_last_kernel_ms(const gmg_ctx* c) { return c ? c->last_kernel_ms : 0.0; }
int64_t gmg_launch_count(const gmg_ctx* c) { return c ? c->launches : 0; }

int gmg_to_edge_distance_matrix(gmg_ctx* c, int64_t n, const int32_t* colptr, const int32_t* rowidx,
                                const double* points, int64_t ld, double* out_val) {
    return guarded(c, [&] {
        int64_t nnz; check_sizes(n, colptr, nnz);
        auto d_cp = to_device(c, colptr, n + 1); auto d_ri = to_device(c, rowidx, nnz);
        auto d_pts = points_to_device(c, points, n, ld);
        DevBuf<double> d_len(c, nnz);
        check_pattern(c, int32_t(n), d_cp.p, d_ri.p, false);
        KernelTimer t(c);
        edge_lengths(c, int32_t(n), d_cp.p, d_ri.p, d_pts.p, d_len.p);
        t.stop();
        d_len.download(out_val, nnz);
        sync(c);
    });
}

int gmg_extract_edges(gmg_ctx* c, int64_t n, const int32_t* colptr, const int32_t* rowidx, const double* val,
                      int32_t* out_edges, double* out_values) {
    return guarded(c, [&] {
        int64_t nnz; check_sizes(n, colptr, nnz);
        auto d_cp = to_device(c, colptr, n + 1); auto d_ri = to_device(c, rowidx, nnz); auto d_va = to_device(c, val, nnz);
        DevBuf<int32_t> d_e(c, 2 * nnz); DevBuf<double> d_v(c, nnz);
        KernelTimer t(c);
        extract_edges_dev(c, int32_t(n), d_cp.p, d_ri.p, d_va.p, nnz, d_e.p, d_v.p);
        t.stop();
        d_e.download(out_edges, 2 * nnz); d_v.download(out_values, nnz);
        sync(c);
    });
}

int gmg_average_edge_length(gmg_ctx* c, int64_t n, const double* points, int64_t ld, int64_t n_edges,
                            const int32_t* edges, double* out_mean) {
    return guarded(c, [&] {
        GMG_REQUIRE(n >= 0 && n_edges >= 0, GMG_ERR_INVALID, "negative size");
        auto d_pts = points_to_device(c, points, n, ld);
        auto d_e = to_device(c, edges, 2 * n_edges);
        check_index_range(c, 2 * n_edges, d_e.p, n, "averageEdgeLength: edges");
        DevBuf<double> d_out(c, 1);
        KernelTimer t(c);
        average_edge_length_dev(c, int32_t(n), d_pts.p, n_edges, d_e.p, d_out.p);
        t.stop();
        d_out.download(out_mean, 1);
        sync(c);
    });
}

int gmg_fast_disc_sample(gmg_ctx* c, int64_t n, const double* points, int64_t ld, const int32_t* colptr,
                         const int32_t* rowidx, double radius, int mode, uint64_t seed, int64_t* selection_out,
                         int64_t* m_out) {
    return guarded(c, [&] {
        int64_t nnz; check_sizes(n, colptr, nnz);
        auto d_cp = to_device(c, colptr, n + 1); auto d_ri = to_device(c, rowidx, nnz);
        auto d_pts = points_to_device(c, points, n, ld);
        DevBuf<double> d_len(c, nnz);
        check_pattern(c, int32_t(n), d_cp.p, d_ri.p, true);
        KernelTimer t(c);
        edge_lengths(c, int32_t(n), d_cp.p, d_ri.p, d_pts.p, d_len.p);
        DevBuf<int32_t> sel;
        const int32_t m = disc_sample(c, int32_t(n), d_cp.p, d_ri.p, d_len.p, radius, mode, seed, sel, nullptr);
        DevBuf<int64_t> wide(c, m);
        widen_i32(c, sel.p, wide.p, m);
        t.stop();
        wide.download(selection_out, m);
        sync(c);
        *m_out = m;
    });
}

/* fastDiscSample with the selection bounds the reference marks as a todo (src/sampling.cpp:11-12) */
int gmg_fast_disc_sample_bounded(gmg_ctx* c, int64_t n, const double* points, int64_t ld, const int32_t* colptr,
                                 const int32_t* rowidx, double radius, int64_t min_selection, int64_t max_selection, int mode,
                                 uint64_t seed, int64_t* selection_out, int64_t* m_out, double* radius_used) {
    return guarded(c, [&] {
        int64_t nnz; check_sizes(n, colptr, nnz);
        GMG_REQUIRE(min_selection >= 0 && max_selection >= 0 && selection_out && m_out, GMG_ERR_INVALID, "fastDiscSample: bad bounds");
        auto d_cp = to_device(c, colptr, n + 1); auto d_ri = to_device(c, rowidx, nnz);
        auto d_pts = points_to_device(c, points, n, ld);
        DevBuf<double> d_len(c, nnz);
        check_pattern(c, int32_t(n), d_cp.p, d_ri.p, true);
        KernelTimer t(c);
        edge_lengths(c, int32_t(n), d_cp.p, d_ri.p, d_pts.p, d_len.p);
        DevBuf<int32_t> sel;
        int32_t m = 0;
        // too few samples: halve the radius and sample again (the lengths are computed once)
        for (int pass = 0; pass < 64; ++pass) {
            m = disc_sample(c, int32_t(n), d_cp.p, d_ri.p, d_len.p, radius, mode, seed, sel, nullptr);
            if (m >= min_selection || m >= n) break;
            radius = radius / 2;
        }
        // too many: the greedy loop would have stopped selecting = the head of the ascending selection
        if (m > max_selection) m = int32_t(max_selection);
        DevBuf<int64_t> wide(c, m);
        widen_i32(c, sel.p, wide.p, m);
        t.stop();
        wide.download(selection_out, m);
        sync(c);
        *m_out = m;
        if (radius_used) *radius_used = radius;
    });
}

/* nested mode: the coarse points are the samples' positions */
int gmg_coarse_from_samples(gmg_ctx* c, int64_t n, const double* points, int64_t ld, int64_t m, const int64_t* samples,
                            double* out_points, int64_t out_ld) {
    return guarded(c, [&] {
        GMG_REQUIRE(n >= 0 && n < (int64_t(1) << 31) && m >= 0 && m <= n && points && out_points && ld >= n && out_ld >= m,
                    GMG_ERR_INVALID, "coarseFromSamples: bad sizes");
        auto d_pts = points_to_device(c, points, n, ld);
        auto d_s = indices_to_device(c, samples, m, n, "coarseFromSamples: coarse_samples");
        DevBuf<double> d_out(c, 3 * m);
        KernelTimer t(c);
        sample_points(c, int32_t(n), d_pts.p, int32_t(m), d_s.p, d_out.p);
        t.stop();
        points_to_host(c, d_out.p, m, out_points, out_ld);
        sync(c);
    });
}

int gmg_assign_parents(gmg_ctx* c, int64_t n, const double* points, int64_t ld, const int32_t* colptr,
                       const int32_t* rowidx, int64_t m, const int64_t* samples, int64_t* parents_out) {
    return guarded(c, [&] {
        int64_t nnz; check_sizes(n, colptr, nnz);
        auto d_cp = to_device(c, colptr, n + 1); auto d_ri = to_device(c, rowidx, nnz);
        auto d_pts = points_to_device(c, points, n, ld);
        auto d_s = indices_to_device(c, samples, m, n, "assignParents: coarse_samples");
        DevBuf<double> d_len(c, nnz);
        DevBuf<int32_t> d_par(c, n);
        DevBuf<int64_t> wide(c, n);
        check_pattern(c, int32_t(n), d_cp.p, d_ri.p, true);
        KernelTimer t(c);
        edge_lengths(c, int32_t(n), d_cp.p, d_ri.p, d_pts.p, d_len.p);
        assign_parents_dev(c, int32_t(n), d_cp.p, d_ri.p, d_len.p, int32_t(m), d_s.p, d_par.p, nullptr);
        widen_i32(c, d_par.p, wide.p, n);
        t.stop();
        wide.download(parents_out, n);
        sync(c);
    });
}

int gmg_extract_coarse_edges(gmg_ctx* c, int64_t n, const int32_t* colptr, const int32_t* rowidx, const double* val,
                             int64_t m, const int64_t* parents, int32_t* out_colptr, int64_t* nnz_out) {
    return guarded(c, [&] {
        release_kept(c);
        int64_t nnz; check_sizes(n, colptr, nnz);
        GMG_REQUIRE(m >= 0 && m <= n, GMG_ERR_INVALID, "extractCoarseEdges: need 0 <= m <= n");
        auto d_cp = to_device(c, colptr, n + 1); auto d_ri = to_device(c, rowidx, nnz); auto d_va = to_device(c, val, nnz);
        auto d_par = indices_to_device(c, parents, n, m, "extractCoarseEdges: fine_to_nearest_coarse");
        check_pattern(c, int32_t(n), d_cp.p, d_ri.p, true);
        KernelTimer t(c);
        Clusters cl;
        build_clusters(c, int32_t(n), int32_t(m), d_par.p, cl);
        DeviceCsc out;
        coarse_edges(c, int32_t(n), nnz, d_cp.p, d_ri.p, d_va.p, int32_t(m), d_par.p, cl, out);
        t.stop();
        out.colptr.download(out_colptr, m + 1);
        sync(c);
        *nnz_out = out.nnz;
        c->keep_n[0] = c->keep_n[1] = out.nnz;
        c->keep[0] = out.rowidx.detach(); c->keep[1] = out.val.detach();
        c->keep_kind = 1;
    });
}

int gmg_fetch_sparse(gmg_ctx* c, int32_t* out_inner, double* out_val) {
    return guarded(c, [&] {
        GMG_REQUIRE(c->keep_kind == 1, GMG_ERR_STATE, "gmg_fetch_sparse: no sparse result is pending");
        const int64_t nnz = c->keep_n[0];
        if (nnz > 0) {
            GMG_CUDA(cudaMemcpyAsync(out_inner, c->keep[0], sizeof(int32_t) * size_t(nnz), cudaMemcpyDeviceToHost, c->stream));
            GMG_CUDA(cudaMemcpyAsync(out_val, c->keep[1], sizeof(double) * size_t(nnz), cudaMemcpyDeviceToHost, c->stream));
        }
        sync(c);
        release_kept(c);
    });
}

int gmg_coarse_from_mean_of_fine_children(gmg_ctx* c, int64_t n, const double* points, int64_t ld,
                                          const int32_t* colptr, const int32_t* rowidx, const int64_t* parents,
                                          int64_t m, double* out_points, int64_t out_ld) {
    return guarded(c, [&] {
        int64_t nnz; check_sizes(n, colptr, nnz);
        GMG_REQUIRE(m >= 0 && m < (int64_t(1) << 31), GMG_ERR_INVALID, "bad coarse count");
        auto d_cp = to_device(c, colptr, n + 1); auto d_ri = to_device(c, rowidx, nnz);
        auto d_pts = points_to_device(c, points, n, ld);
        auto d_par = indices_to_device(c, parents, n, m, "coarseFromMeanOfFineChildren: fine_to_nearest_coarse");
        check_pattern(c, int32_t(n), d_cp.p, d_ri.p, false);
        DevBuf<double> d_out(c, 3 * m);
        KernelTimer t(c);
        Clusters cl;
        build_clusters(c, int32_t(n), int32_t(m), d_par.p, cl);
        coarse_points(c, int32_t(n), d_pts.p, d_cp.p, d_ri.p, int32_t(m), cl, d_out.p);
        t.stop();
        points_to_host(c, d_out.p, m, out_points, out_ld);
        sync(c);
    });
}

int gmg_construct_voronoi_triangles(gmg_ctx* c, int64_t m, const double* points, int64_t ld, const int32_t* colptr,
                                    const int32_t* rowidx, const double* val, int64_t* T_out, int64_t* out_fan_ptr) {
    return guarded(c, [&] {
        release_kept(c);
        int64_t nnz; check_sizes(m, colptr, nnz);
        auto d_cp = to_device(c, colptr, m + 1); auto d_ri = to_device(c, rowidx, nnz); auto d_va = to_device(c, val, nnz);
        auto d_pts = points_to_device(c, points, m, ld);
        check_pattern(c, int32_t(m), d_cp.p, d_ri.p, false);
        KernelTimer t(c);
        Triangles tr;
        voronoi_triangles(c, int32_t(m), nnz, d_pts.p, d_cp.p, d_ri.p, d_va.p, tr);
        DevBuf<int64_t> tri64(c, 3 * tr.T), fan64(c, 3 * tr.T), ptr64(c, m + 1);
        widen_i32(c, tr.tri.p, tri64.p, 3 * tr.T);
        widen_i32(c, tr.fan_tri.p, fan64.p, 3 * tr.T);
        widen_i32(c, tr.fan_ptr.p, ptr64.p, m + 1);
        t.stop();
        ptr64.download(out_fan_ptr, m + 1);
        sync(c);
        *T_out = tr.T;
        c->keep_n[0] = c->keep_n[1] = c->keep_n[2] = 3 * tr.T;
        c->keep[0] = tri64.detach(); c->keep[1] = tr.normal.detach(); c->keep[2] = fan64.detach();
        c->keep_kind = 2;
    });
}

int gmg_fetch_triangles(gmg_ctx* c, int64_t* out_tri, double* out_normal, int64_t* out_fan) {
    return guarded(c, [&] {
        GMG_REQUIRE(c->keep_kind == 2, GMG_ERR_STATE, "gmg_fetch_triangles: no triangle result is pending");
        const int64_t k = c->keep_n[0];
        if (k > 0) {
            GMG_CUDA(cudaMemcpyAsync(out_tri, c->keep[0], sizeof(int64_t) * size_t(k), cudaMemcpyDeviceToHost, c->stream));
            GMG_CUDA(cudaMemcpyAsync(out_normal, c->keep[1], sizeof(double) * size_t(k), cudaMemcpyDeviceToHost, c->stream));
            GMG_CUDA(cudaMemcpyAsync(out_fan, c->keep[2], sizeof(int64_t) * size_t(k), cudaMemcpyDeviceToHost, c->stream));
        }
        sync(c);
        release_kept(c);
    });
}

int gmg_construct_prolongation(gmg_ctx* c, int64_t n, const double* fine_points, int64_t fine_ld, int64_t m,
                               const double* coarse_points, int64_t coarse_ld, const int32_t* c_colptr,
                               const int32_t* c_rowidx, const double* c_val, const int64_t* parents, int weighting,
                               int32_t* out_rowptr, int64_t* nnz_out, gmg_prolongation_stats* stats_out) {
    return guarded(c, [&] {
        release_kept(c);
        int64_t cnnz; check_sizes(m, c_colptr, cnnz);
        GMG_REQUIRE(n >= 0 && n < (int64_t(1) << 31) / 3, GMG_ERR_INVALID, "fine vertex count too large for 32-bit P indices");
        auto d_cp = to_device(c, c_colptr, m + 1); auto d_ri = to_device(c, c_rowidx, cnnz); auto d_va = to_device(c, c_val, cnnz);
        auto d_fp = points_to_device(c, fine_points, n, fine_ld);
        auto d_cpts = points_to_device(c, coarse_points, m, coarse_ld);
        auto d_par = indices_to_device(c, parents, n, m, "constructProlongation: fine_to_nearest_coarse");
        check_pattern(c, int32_t(m), d_cp.p, d_ri.p, false);
        KernelTimer t(c);
        Triangles tr;
        voronoi_triangles(c, int32_t(m), cnnz, d_cpts.p, d_cp.p, d_ri.p, d_va.p, tr);
        DeviceCsr P;
        gmg_prolongation_stats st{};
        Error pending{GMG_OK, ""};
        try {
            prolongation(c, int32_t(n), d_fp.p, int32_t(m), d_cpts.p, d_cp.p, d_ri.p, d_par.p, tr, weighting, P, &st);
        } catch (const Error& e) { pending = e; }
        t.stop();
        if (stats_out) *stats_out = st;
        if (pending.status != GMG_OK) throw pending;
        P.rowptr.download(out_rowptr, n + 1);
        sync(c);
        *nnz_out = P.nnz;
        c->keep_n[0] = c->keep_n[1] = P.nnz;
        c->keep[0] = P.colidx.detach(); c->keep[1] = P.val.detach();
        c->keep_kind = 1;
    });
}

int gmg_projected_points(gmg_ctx* c, int64_t n, int64_t m, const int32_t* rowptr, const int32_t* colidx,
                         const double* val, const double* coarse_points, int64_t coarse_ld, double* out_points,
                         int64_t out_ld) {
    return guarded(c, [&] {
        GMG_REQUIRE(n >= 0 && m >= 0 && rowptr && rowptr[0] == 0, GMG_ERR_INVALID, "bad P");
        const int64_t nnz = rowptr[n];
        auto d_rp = to_device(c, rowptr, n + 1); auto d_ci = to_device(c, colidx, nnz); auto d_va = to_device(c, val, nnz);
        auto d_cpts = points_to_device(c, coarse_points, m, coarse_ld);
        check_outer_pointers(c, n, d_rp.p, "projectedPoints: weights");
        check_index_range(c, nnz, d_ci.p, m, "projectedPoints: weights");
        DevBuf<double> d_out(c, 3 * n);
        KernelTimer t(c);
        projected_points_dev(c, int32_t(n), int32_t(m), d_rp.p, d_ci.p, d_va.p, d_cpts.p, d_out.p);
        t.stop();
        points_to_host(c, d_out.p, n, out_points, out_ld);
        sync(c);
    });
}

}  // extern "C"
