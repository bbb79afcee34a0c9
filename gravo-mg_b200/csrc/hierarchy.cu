// Resident hierarchy builder [not in reference]: repeats the per-level call order of the
// reference's demo (test/main.cpp:68-121) with every array kept in HBM between stages and levels.
#include "hierarchy_types.h"

#include <cmath>
#include <memory>
#include <stdexcept>

using namespace gmg;

namespace {

struct StageClock {
    gmg_ctx* c;
    std::vector<cudaEvent_t> ev;
    explicit StageClock(gmg_ctx* ctx) : c(ctx) {}
    ~StageClock() { for (auto e : ev) cudaEventDestroy(e); }
    void mark() {
        cudaEvent_t e;
        GMG_CUDA(cudaEventCreate(&e));
        GMG_CUDA(cudaEventRecord(e, c->stream));
        ev.push_back(e);
    }
    double ms(size_t i) const {
        float t = 0.f;
        cudaEventElapsedTime(&t, ev[i], ev[i + 1]);
        return t;
    }
};

}  // namespace

template <typename F> static int guarded_h(gmg_ctx* c, F&& body) {
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

extern "C" {

int gmg_input_create(gmg_ctx* c, int64_t n, const double* points, int64_t ld, const int32_t* colptr,
                     const int32_t* rowidx, const double* A_val, gmg_input** out) {
    if (!out) return GMG_ERR_INVALID;
    *out = nullptr;
    return guarded_h(c, [&] {
        GMG_REQUIRE(n >= 0 && n < (int64_t(1) << 31) / 3, GMG_ERR_INVALID, "vertex count too large for 32-bit indices");
        GMG_REQUIRE(colptr && colptr[0] == 0 && colptr[n] >= 0, GMG_ERR_INVALID, "bad colptr");
        std::unique_ptr<gmg_input> in(new gmg_input());
        in->ctx = c; in->n = int32_t(n); in->nnz = colptr[n];
        in->pts = points_to_device(c, points, n, ld);
        in->colptr = to_device(c, colptr, n + 1);
        in->rowidx = to_device(c, rowidx, in->nnz);
        if (A_val) { in->a_val = to_device(c, A_val, in->nnz); in->has_a = true; }
        check_pattern(c, in->n, in->colptr.p, in->rowidx.p, true);
        sync(c);
        *out = in.release();
    });
}

/* The values of A for an input created without them, copied on the context's COPY stream while the compute stream
 * goes on: a hierarchy build only needs them when it reaches the first Galerkin product, so a 10 M-vertex build
 * overlaps their 560 MB with sampling, parents and the rows of P. A_val (nnz doubles, pinned for a real overlap) must
 * stay valid until a call that uses the values (gmg_hierarchy_build with build_galerkin, gmg_solver_create, ...) has
 * returned, or gmg_copy_sync has. */
int gmg_input_set_values_async(gmg_ctx* c, gmg_input* in, const double* A_val) {
    return guarded_h(c, [&] {
        GMG_REQUIRE(in && A_val, GMG_ERR_STATE, "null input or values");
        GMG_REQUIRE(!in->has_a, GMG_ERR_STATE, "the input already has A values");
        if (!c->copy_stream) GMG_CUDA(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
        in->a_val.alloc(c, in->nnz);
        // the block comes from the compute stream's pool: the copy stream may touch it once that stream got this far
        if (!in->a_ready) GMG_CUDA(cudaEventCreateWithFlags(&in->a_ready, cudaEventDisableTiming));
        GMG_CUDA(cudaEventRecord(in->a_ready, c->stream));
        GMG_CUDA(cudaStreamWaitEvent(c->copy_stream, in->a_ready, 0));
        if (in->nnz > 0)
            GMG_CUDA(cudaMemcpyAsync(in->a_val.p, A_val, sizeof(double) * size_t(in->nnz), cudaMemcpyHostToDevice, c->copy_stream));
        GMG_CUDA(cudaEventRecord(in->a_ready, c->copy_stream));
        in->has_a = true;
    });
}

/* An input whose arrays the CALLER fills on the device (row-sharded upload: each rank copies its own slice from
 * the host and the slices are all-gathered over NVLink instead of every rank pulling the whole mesh over PCIe).
 * gmg_input_commit checks the pattern once the arrays are complete. */
int gmg_input_create_device(gmg_ctx* c, int64_t n, int64_t nnz, int with_a, gmg_input** out) {
    if (!out) return GMG_ERR_INVALID;
    *out = nullptr;
    return guarded_h(c, [&] {
        GMG_REQUIRE(n >= 0 && n < (int64_t(1) << 31) / 3, GMG_ERR_INVALID, "vertex count too large for 32-bit indices");
        GMG_REQUIRE(nnz >= 0 && nnz < (int64_t(1) << 31), GMG_ERR_INVALID, "entry count too large for 32-bit offsets");
        std::unique_ptr<gmg_input> in(new gmg_input());
        in->ctx = c; in->n = int32_t(n); in->nnz = nnz;
        in->pts.alloc(c, 3 * n); in->colptr.alloc(c, n + 1); in->rowidx.alloc(c, nnz);
        if (with_a) { in->a_val.alloc(c, nnz); in->has_a = true; }
        sync(c);
        *out = in.release();
    });
}

/* which: 0 points (f64[3n]: x | y | z), 1 colptr (int32[n+1]), 2 rowidx (int32[nnz]), 3 A values (f64[nnz]) */
int gmg_input_device_ptr(gmg_input* in, int which, void** ptr) {
    if (!in || !ptr) return GMG_ERR_STATE;
    switch (which) {
        case 0: *ptr = in->pts.p; break;
        case 1: *ptr = in->colptr.p; break;
        case 2: *ptr = in->rowidx.p; break;
        case 3: if (!in->has_a) return GMG_ERR_STATE; *ptr = in->a_val.p; break;
        default: return GMG_ERR_INVALID;
    }
    return GMG_OK;
}

int gmg_input_commit(gmg_ctx* c, gmg_input* in) {
    return guarded_h(c, [&] {
        GMG_REQUIRE(in, GMG_ERR_STATE, "null input");
        int32_t ends[2] = {-1, -1};
        GMG_CUDA(cudaMemcpyAsync(&ends[0], in->colptr.p, sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
        GMG_CUDA(cudaMemcpyAsync(&ends[1], in->colptr.p + in->n, sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
        sync(c);
        GMG_REQUIRE(ends[0] == 0 && ends[1] == in->nnz, GMG_ERR_INVALID, "bad colptr");
        check_pattern(c, in->n, in->colptr.p, in->rowidx.p, true);
        sync(c);
    });
}

void gmg_input_destroy(gmg_input* in) {
    if (!in) return;
    cudaSetDevice(in->ctx->device);
    delete in;
}

int gmg_hierarchy_build(gmg_ctx* c, const gmg_input* in, const gmg_hierarchy_options* opt, gmg_hierarchy** out) {
    if (!out) return GMG_ERR_INVALID;
    *out = nullptr;
    return guarded_h(c, [&] {
        GMG_REQUIRE(in && opt, GMG_ERR_INVALID, "null input or options");
        GMG_REQUIRE(opt->levels >= 1, GMG_ERR_INVALID, "levels must be >= 1");
        GMG_REQUIRE(!opt->build_galerkin || in->has_a, GMG_ERR_INVALID, "build_galerkin needs A values");
        std::unique_ptr<gmg_hierarchy> h(new gmg_hierarchy());
        h->ctx = c;
        cudaEventRecord(c->ev0, c->stream);

        // current fine level (views into the input or into the previous level)
        int32_t n = in->n;
        int64_t nnz = in->nnz;
        const double* pts = in->pts.p;
        const int32_t* colptr = in->colptr.p;
        const int32_t* rowidx = in->rowidx.p;
        const double* a_val = in->has_a ? in->a_val.p : nullptr;      // system matrix values of this level
        const int32_t* a_colptr = colptr;
        const int32_t* a_rowidx = rowidx;
        int32_t a_n = n;

        // level 0: fine_edge_matrix = toEdgeDistanceMatrix(L, points)  (test/main.cpp:69)
        h->fine_val.alloc(c, nnz);
        h->fine_nnz = nnz;
        const double* val = h->fine_val.p;
        DevBuf<double> len_buf;          // Euclidean lengths of the current level's entries (levels >= 1)
        DevBuf<double> d_radius(c, 1);
        DevBuf<unsigned long long> d_degsq(c, 1);
        StageStatus status;              // device-side flags and counters, fetched once per level

        for (int lvl = 0; lvl + 1 < opt->levels; ++lvl) {
            if (n == 0) break;
            std::unique_ptr<Level> Lown(new Level());
            Level* const L = Lown.get();       // (the level moves into the hierarchy before its Galerkin product)
            L->info.n = n; L->info.nnz = nnz;
            StageClock clk(c);
            status.begin(c);
            clk.mark();
            // stage 0: lengths (the sampler and the sweep use geometric lengths, never stored values)
            const double* len;
            if (lvl == 0) { edge_lengths(c, n, colptr, rowidx, pts, h->fine_val.p); len = h->fine_val.p; }
            else { len_buf.alloc(c, nnz); edge_lengths(c, n, colptr, rowidx, pts, len_buf.p); len = len_buf.p; }
            double radius = 0.0;
            const bool overridden = opt->radius_override && opt->radius_override[lvl] > 0.0;
            // one host round trip for the radius and the sampler's list bound
            d_degsq.zero();
            degree_square_sum(c, n, colptr, d_degsq.p);
            unsigned long long degsq = 0;
            double mean = 0.0;
            if (!overridden) {
                // averageEdgeLength over extractEdges(matrix): every stored entry, divisor nnz - n
                double denom = double(nnz - n);
                if (opt->radius_policy == 1) denom = double(lvl == 0 ? nnz - n : nnz);
                mean_length_dev(c, n, nnz, colptr, len, denom, d_radius.p);
                d_radius.download(&mean, 1);
            }
            d_degsq.download(&degsq, 1);
            sync(c);
            radius = overridden ? opt->radius_override[lvl] : std::cbrt(opt->ratio) * mean;
            L->info.radius = radius;
            clk.mark();
            // stage 1: fastDiscSample
            SamplerStats ss{};
            const int32_t m = disc_sample(c, n, colptr, rowidx, len, radius, opt->sampler_mode, opt->seed + uint64_t(lvl),
                                          L->samples, &ss, int64_t(degsq) + nnz);
            L->info.m = m; L->info.sampler_rounds = ss.rounds;
            clk.mark();
            // stage 2: assignParents
            L->parents.alloc(c, n);
            assign_parents_dev(c, n, colptr, rowidx, len, m, L->samples.p, L->parents.p, nullptr, &status);
            clk.mark();
            // stage 3: cluster CSR
            Clusters cl;
            build_clusters(c, n, m, L->parents.p, cl);
            clk.mark();
            // stage 4: extractCoarseEdges
            coarse_edges(c, n, nnz, colptr, rowidx, val, m, L->parents.p, cl, L->coarse, &status);
            L->info.nnz_c = L->coarse.nnz;
            clk.mark();
            // stage 5: coarseFromMeanOfFineChildren
            L->coarse_pts.alloc(c, 3 * int64_t(m));
            if (opt->weighting & GMG_NESTED) sample_points(c, n, pts, m, L->samples.p, L->coarse_pts.p);
            else coarse_points(c, n, pts, colptr, rowidx, m, cl, L->coarse_pts.p);
            clk.mark();
            // stage 6: constructVoronoiTriangles
            voronoi_triangles(c, m, L->coarse.nnz, L->coarse_pts.p, L->coarse.colptr.p, L->coarse.rowidx.p, L->coarse.val.p,
                              L->tri);
            L->info.T = L->tri.T;
            clk.mark();
            // stage 7: constructProlongation
            prolongation(c, n, pts, m, L->coarse_pts.p, L->coarse.colptr.p, L->coarse.rowidx.p, L->parents.p, L->tri,
                         opt->weighting, L->P, nullptr, &status);
            clk.mark();
            // the level's deferred checks and counters: one round trip
            status.fetch(c);
            L->info.parent_rounds = status.host[StageStatus::kParentRounds];
            finish_prolongation(status, L->P, &L->info.stats);
            L->info.nnz_p = L->P.nnz;
            status.raise_parents();
            status.raise_coarse();
            status.raise_prolong(n);
            // P is complete on the device (the status fetch synchronised): with a Galerkin product still to come, let the
            // caller start copying P out now, so that it travels while A_c is formed
            h->levels.push_back(std::move(Lown));
            if (opt->build_galerkin && opt->on_level) opt->on_level(opt->user, h.get(), (int(h->levels.size()) - 1) | GMG_LEVEL_P_READY);
            // stage 8: Galerkin coarse operator
            StageClock gclk(c);
            if (opt->build_galerkin) {
                if (h->levels.size() == 1) await_values(c, in);        // level 0 reads the input's values (maybe still arriving)
                for (int s = 0; s < 4; ++s) { cudaEvent_t e; GMG_CUDA(cudaEventCreate(&e)); gclk.ev.push_back(e); }
                galerkin(c, a_n, a_colptr, a_rowidx, a_val, L->P, L->Ac, 0, -1, nullptr, gclk.ev.data());
                L->info.nnz_ac = L->Ac.nnz;
            }
            clk.mark();
            sync(c);
            for (int s = 0; s < 9; ++s) L->info.stage_ms[s] = clk.ms(size_t(s));
            for (int s = 0; s < 3; ++s) L->info.galerkin_ms[s] = opt->build_galerkin ? gclk.ms(size_t(s)) : 0.0;

            // next level's fine inputs
            n = m; nnz = L->coarse.nnz;
            pts = L->coarse_pts.p; colptr = L->coarse.colptr.p; rowidx = L->coarse.rowidx.p; val = L->coarse.val.p;
            if (opt->build_galerkin) { a_n = m; a_colptr = L->Ac.colptr.p; a_rowidx = L->Ac.rowidx.p; a_val = L->Ac.val.p; }
            // the level is complete on the device (the status fetch above synchronised): let the caller start
            // copying it out on the copy stream while the next level builds
            if (opt->on_level) opt->on_level(opt->user, h.get(), int(h->levels.size()) - 1);
        }
        cudaEventRecord(c->ev1, c->stream);
        cudaEventSynchronize(c->ev1);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, c->ev0, c->ev1);
        c->last_kernel_ms = ms;
        *out = h.release();
    });
}

void gmg_hierarchy_destroy(gmg_hierarchy* h) {
    if (!h) return;
    cudaSetDevice(h->ctx->device);
    delete h;
}

int gmg_hierarchy_num_levels(const gmg_hierarchy* h) { return h ? int(h->levels.size()) : 0; }

int gmg_hierarchy_level_info(const gmg_hierarchy* h, int level, gmg_level_info* out) {
    if (!h || !out || level < 0 || level >= int(h->levels.size())) return GMG_ERR_STATE;
    *out = h->levels[size_t(level)]->info;
    return GMG_OK;
}

namespace {

// device -> host copy of one array of a level. async: on the context's copy stream, no synchronisation (the
// caller waits with gmg_copy_sync); arrays that are widened to int64 on the way out are not available there.
void fetch_level_array(gmg_ctx* c, const gmg_hierarchy* h, int level, int which, void* out, bool async) {
    GMG_REQUIRE(h && level >= 0 && level < int(h->levels.size()), GMG_ERR_STATE, "no such level");
    const Level& L = *h->levels[size_t(level)];
    const int64_t n = L.info.n, m = L.info.m, T = L.info.T;
    cudaStream_t stream = c->stream;
    if (async) {
        if (!c->copy_stream) GMG_CUDA(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
        stream = c->copy_stream;
    }
    auto copy = [&](const void* dev, size_t bytes) {
        if (bytes) GMG_CUDA(cudaMemcpyAsync(out, dev, bytes, cudaMemcpyDeviceToHost, stream));
    };
    auto copy_wide = [&](const int32_t* dev, int64_t count) {
        GMG_REQUIRE(!async, GMG_ERR_INVALID, "int64 outputs (samples, parents, triangles, fans) have no asynchronous fetch");
        DevBuf<int64_t> w(c, count);
        widen_i32(c, dev, w.p, count);
        copy(w.p, sizeof(int64_t) * size_t(count));
        sync(c);
    };
    switch (which) {
        case GMG_H_SAMPLES: copy_wide(L.samples.p, m); break;
        case GMG_H_PARENTS: copy_wide(L.parents.p, n); break;
        case GMG_H_COARSE_COLPTR: copy(L.coarse.colptr.p, sizeof(int32_t) * size_t(m + 1)); break;
        case GMG_H_COARSE_ROWIDX: copy(L.coarse.rowidx.p, sizeof(int32_t) * size_t(L.coarse.nnz)); break;
        case GMG_H_COARSE_VAL: copy(L.coarse.val.p, sizeof(double) * size_t(L.coarse.nnz)); break;
        case GMG_H_COARSE_POINTS: copy(L.coarse_pts.p, sizeof(double) * size_t(3 * m)); break;
        case GMG_H_TRIANGLES: copy_wide(L.tri.tri.p, 3 * T); break;
        case GMG_H_NORMALS: copy(L.tri.normal.p, sizeof(double) * size_t(3 * T)); break;
        case GMG_H_FAN_PTR: copy_wide(L.tri.fan_ptr.p, m + 1); break;
        case GMG_H_FAN: copy_wide(L.tri.fan_tri.p, 3 * T); break;
        case GMG_H_P_ROWPTR: copy(L.P.rowptr.p, sizeof(int32_t) * size_t(n + 1)); break;
        case GMG_H_P_COLIDX: copy(L.P.colidx.p, sizeof(int32_t) * size_t(L.P.nnz)); break;
        case GMG_H_P_VAL: copy(L.P.val.p, sizeof(double) * size_t(L.P.nnz)); break;
        case GMG_H_AC_COLPTR:
            GMG_REQUIRE(L.Ac.colptr.p, GMG_ERR_STATE, "hierarchy was built without the Galerkin operator");
            copy(L.Ac.colptr.p, sizeof(int32_t) * size_t(m + 1)); break;
        case GMG_H_AC_ROWIDX:
            GMG_REQUIRE(L.Ac.colptr.p, GMG_ERR_STATE, "hierarchy was built without the Galerkin operator");
            copy(L.Ac.rowidx.p, sizeof(int32_t) * size_t(L.Ac.nnz)); break;
        case GMG_H_AC_VAL:
            GMG_REQUIRE(L.Ac.colptr.p, GMG_ERR_STATE, "hierarchy was built without the Galerkin operator");
            copy(L.Ac.val.p, sizeof(double) * size_t(L.Ac.nnz)); break;
        case GMG_H_FINE_VAL:
            GMG_REQUIRE(level == 0, GMG_ERR_STATE, "fine values exist on level 0 only");
            copy(h->fine_val.p, sizeof(double) * size_t(h->fine_nnz)); break;
        case GMG_H_PROJECTED: {
            GMG_REQUIRE(!async, GMG_ERR_INVALID, "projected points are computed on the compute stream: no asynchronous fetch");
            DevBuf<double> proj(c, 3 * n);
            projected_points_dev(c, int32_t(n), int32_t(m), L.P.rowptr.p, L.P.colidx.p, L.P.val.p, L.coarse_pts.p, proj.p);
            copy(proj.p, sizeof(double) * size_t(3 * n));
            sync(c);
            break;
        }
        default: GMG_REQUIRE(false, GMG_ERR_INVALID, "unknown array id");
    }
    if (!async) sync(c);
}

}  // namespace

int gmg_hierarchy_fetch(gmg_ctx* c, const gmg_hierarchy* h, int level, int which, void* out) {
    return guarded_h(c, [&] { fetch_level_array(c, h, level, which, out, false); });
}

/* The same copy on the context's copy stream, not waited for: meant to be called from the on_level callback
 * of gmg_hierarchy_build (the level is complete by then), so that a finished level leaves the device while the
 * next one builds. `out` should be pinned host memory. gmg_copy_sync waits for every such copy. */
int gmg_hierarchy_fetch_async(gmg_ctx* c, const gmg_hierarchy* h, int level, int which, void* out) {
    return guarded_h(c, [&] { fetch_level_array(c, h, level, which, out, true); });
}

int gmg_copy_sync(gmg_ctx* c) {
    return guarded_h(c, [&] { if (c->copy_stream) GMG_CUDA(cudaStreamSynchronize(c->copy_stream)); });
}

}  // extern "C"
