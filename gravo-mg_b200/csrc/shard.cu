// Row-sharded hierarchy build [not in reference; SURVEY.md 8(e)]: one process per GPU, fine vertices
// split into contiguous ranges, one per rank. Design of this first version:
//
//   * every rank holds the whole level graph (colptr, rowidx, lengths, points): 112 B per vertex of a
//     triangle mesh, so a 100 M-vertex level is 11 GB of 180 GB. What is SHARDED is the work and the
//     large per-level products: the sampler's coverer lists, the relaxation, the coarse rows and the
//     rows of P are computed for the owned range only;
//   * a rank only ever writes the entries it owns of a replicated per-vertex array (state, distance,
//     predecessor, parent, coarse row counts); the caller all-gathers the owned slices after each
//     phase (NCCL all-gather over NVLink through torch.distributed, or an in-process loop-back when
//     several virtual ranks share one GPU in the tests). There is no reduction and no all-to-all;
//   * the sampler and the relaxation are therefore PULL formulations (k_sampler.cu, k_parents.cu):
//     same greedy set, same distances, same parents as the single-GPU kernels, bit for bit;
//   * triangle fans are enumerated only where the owned rows of P read them (a byte mask over the
//     smallest vertex of a triangle), so that stage is sharded without any exchange; the triangle
//     list of a sharded build is therefore rank-local;
//   * edge lengths are computed only for the columns a rank reads (owned + one hop, plus at level 0 the
//     children of its clusters); the mean length keeps the single-GPU summation tree because that
//     tree is defined on groups of 32 columns, whose partial sums are all-gathered;
//   * cheap stages (cluster CSR, coarse positions, compaction scans) are computed redundantly by
//     every rank from replicated inputs, so they need no exchange;
//   * P stays sharded by rows ("P rows shard trivially");
//   * Galerkin operators (build_galerkin): the ranks all-gather their rows of P (variable-size slices,
//     like the coarse rows), every rank transposes the coarse columns it OWNS and runs the single-GPU
//     two-pass SpGEMM on exactly those outer vectors of A_c; the owned outer vectors are all-gathered and
//     A_c is replicated, because it is the next level's A.
//
// The object below is driven phase by phase from the host (gravo-mg_b200/sharded.py): run a phase on
// every rank, all-gather the buffers that phase names, go on.
#include "shard_types.h"

#include <cmath>

namespace {

template <typename F> int guarded_s(gmg_ctx* c, F&& body) {
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

int32_t round_up(int64_t a, int64_t b) { return int32_t((a + b - 1) / b * b); }

// contiguous split of [0, count) into `world` slices of `chunk` (a multiple of 32) entries
void split(int64_t count, int world, int rank, int32_t& chunk, int32_t& lo, int32_t& hi) {
    chunk = round_up((count + world - 1) / world, 32);
    if (chunk == 0) chunk = 32;
    lo = int32_t(std::min<int64_t>(count, int64_t(rank) * chunk));
    hi = int32_t(std::min<int64_t>(count, int64_t(rank + 1) * chunk));
}

__global__ void __launch_bounds__(kBlock) pack_rows_kernel(int32_t p_lo, int32_t p_hi, const int32_t* __restrict__ boff,
                                                           const int32_t* __restrict__ cnt,
                                                           const int32_t* __restrict__ rowoff, int32_t base,
                                                           const int32_t* __restrict__ sQ, const double* __restrict__ st,
                                                           int32_t* __restrict__ outQ, double* __restrict__ outt) {
    const int64_t gid = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    const int32_t P = p_lo + int32_t(gid / 8);
    const int sub = int(gid % 8);
    if (P >= p_hi) return;
    const int32_t src = boff[P - p_lo], dst = rowoff[P] - base, k = cnt[P];
    for (int32_t i = sub; i < k; i += 8) { outQ[dst + i] = sQ[src + i]; outt[dst + i] = st[src + i]; }
}

// coarse vertices whose fans the owned rows of P read, and the smallest-vertex set that yields those fans
__global__ void __launch_bounds__(kBlock) mark_parents_kernel(int32_t lo, int32_t hi, const int32_t* __restrict__ parents,
                                                              uint8_t* __restrict__ need_c) {
    const int64_t f = int64_t(lo) + int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (f < hi) need_c[parents[f]] = 1;
}
__global__ void __launch_bounds__(kBlock) mark_closure_kernel(int32_t m, const int32_t* __restrict__ colptr,
                                                              const int32_t* __restrict__ rowidx,
                                                              const uint8_t* __restrict__ need_c, uint8_t* __restrict__ need_v0) {
    const int64_t v = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (v >= m || !need_c[v]) return;
    need_v0[v] = 1;
    for (int32_t k = colptr[v]; k < colptr[v + 1]; ++k) need_v0[rowidx[k]] = 1;
}

// columns whose lengths this rank reads: its own (relaxation, first hop) and their neighbours (second hop)
__global__ void __launch_bounds__(kBlock) mark_columns_kernel(int32_t lo, int32_t hi, const int32_t* __restrict__ colptr,
                                                              const int32_t* __restrict__ rowidx, uint8_t* __restrict__ need) {
    const int64_t i = int64_t(lo) + int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (i >= hi) return;
    need[i] = 1;
    for (int32_t k = colptr[i]; k < colptr[i + 1]; ++k) need[rowidx[k]] = 1;
}
// level 0 only (the stored values ARE the lengths): columns the owned coarse rows read -- the children of
// the owned clusters, and fine column number P for every owned cluster P (the index-space quirk, :151)
__global__ void __launch_bounds__(kBlock) mark_children_kernel(int32_t n, int32_t p_lo, int32_t p_hi,
                                                               const int32_t* __restrict__ parents, uint8_t* __restrict__ need) {
    const int64_t f = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (f >= n) return;
    const int32_t P = parents[f];
    if (P >= p_lo && P < p_hi) { need[f] = 1; need[P] = 1; }
}

// Where the ranks' slices of a row-major packed array start: offsets (device, exclusive scan over `count`
// items, count + 1 entries) read at the rank boundaries (multiples of `chunk`). off[world] is the total.
// One host round trip.
void rank_offsets(gmg_shard* s, const int32_t* offsets, int64_t count, int32_t chunk, std::vector<int64_t>& off) {
    gmg_ctx* c = s->ctx;
    std::vector<int32_t> b(size_t(s->world) + 1, 0);
    for (int r = 0; r <= s->world; ++r) {
        const int64_t at = std::min<int64_t>(count, int64_t(r) * chunk);
        GMG_CUDA(cudaMemcpyAsync(&b[size_t(r)], offsets + at, sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
    }
    sync(c);
    off.assign(b.begin(), b.end());
}

__global__ void __launch_bounds__(kBlock) row_counts_kernel(int32_t rows, const int32_t* __restrict__ ptr, int32_t* __restrict__ cnt) {
    const int64_t r = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (r < rows) cnt[r] = ptr[r + 1] - ptr[r];
}

void begin_level(gmg_shard* s) {
    gmg_ctx* c = s->ctx;
    split(s->n, s->world, s->rank, s->chunk, s->lo, s->hi);
    const int64_t n_pad = int64_t(s->chunk) * s->world;
    s->state.alloc(c, n_pad); s->changed[0].alloc(c, n_pad); s->changed[1].alloc(c, n_pad);
    s->dist.alloc(c, n_pad); s->srcid.alloc(c, n_pad); s->pred.alloc(c, n_pad); s->parents.alloc(c, n_pad);
    s->state.zero(); s->pred.zero(); s->parents.zero();
    s->cur.reset(new gmg_shard::Done());
    s->cur->info.n = s->n; s->cur->info.nnz = s->nnz;
    s->cur->lo = s->lo; s->cur->hi = s->hi;
    s->status.begin(c);
    // lengths: only the columns this rank reads (all of them when it is alone)
    const uint8_t* mask = nullptr;
    if (s->world > 1) {
        s->need.alloc(c, int64_t(s->n) + 1);
        s->need.zero();
        if (s->hi > s->lo) GMG_LAUNCH(c, mark_columns_kernel, div_up(s->hi - s->lo, kBlock), kBlock, 0, s->lo, s->hi, s->colptr, s->rowidx, s->need.p);
        mask = s->need.p;
    }
    s->len.alloc(c, s->nnz);
    edge_lengths(c, s->n, s->colptr, s->rowidx, s->pts, s->len.p, mask);
    if (s->level == 0) s->val = s->len.p;                    // toEdgeDistanceMatrix (test/main.cpp:69)
    s->val_is_len = s->level == 0;
    // mean length: the owned column groups' partial sums; the host gathers them, PH_RADIUS folds them
    s->gchunk = s->chunk / 32;
    s->len_partial.alloc(c, int64_t(s->gchunk) * s->world);
    s->len_partial.zero();
    if (s->hi > s->lo) length_group_partials(c, s->n, s->lo / 32, (s->hi + 31) / 32, s->colptr, s->len.p, s->len_partial.p);
}

}  // namespace

extern "C" {

int gmg_shard_create(gmg_ctx* c, int rank, int world, const gmg_input* in, const gmg_hierarchy_options* opt,
                     gmg_shard** out) {
    if (!out) return GMG_ERR_INVALID;
    *out = nullptr;
    return guarded_s(c, [&] {
        GMG_REQUIRE(in, GMG_ERR_INVALID, "null input");
        GMG_REQUIRE(opt && opt->levels >= 1, GMG_ERR_INVALID, "levels must be >= 1");
        GMG_REQUIRE(world >= 1 && rank >= 0 && rank < world, GMG_ERR_INVALID, "bad rank/world");
        GMG_REQUIRE(!opt->build_galerkin || in->has_a, GMG_ERR_INVALID, "build_galerkin needs A values");
        std::unique_ptr<gmg_shard> s(new gmg_shard());
        s->ctx = c; s->rank = rank; s->world = world; s->opt = *opt;
        if (opt->radius_override) s->radius_override.assign(opt->radius_override, opt->radius_override + (opt->levels - 1));
        s->opt.radius_override = nullptr;
        // level 0 is a VIEW of the input (which must outlive the shard): the whole graph, on every rank
        s->n = in->n; s->nnz = in->nnz;
        s->pts = in->pts.p; s->colptr = in->colptr.p; s->rowidx = in->rowidx.p;
        await_values(c, in);
        s->a_n = in->n; s->a_colptr = in->colptr.p; s->a_rowidx = in->rowidx.p; s->a_val = in->has_a ? in->a_val.p : nullptr;
        *out = s.release();
    });
}

void gmg_shard_destroy(gmg_shard* s) {
    if (!s) return;
    cudaSetDevice(s->ctx->device);
    delete s;
}

/* io: small host array, meaning per phase (see include/gravomg_b200.h) */
int gmg_shard_phase(gmg_ctx* c, gmg_shard* s, int phase, int64_t* io) {
    return guarded_s(c, [&] {
        GMG_REQUIRE(s && io, GMG_ERR_STATE, "null shard");
        switch (phase) {
            case GMG_PH_LEVEL_BEGIN: {
                io[0] = (s->level + 1 < s->opt.levels && s->n > 0) ? 1 : 0;
                if (!io[0]) break;
                begin_level(s);
                io[1] = s->n; io[2] = s->lo; io[3] = s->hi; io[4] = s->chunk;
                break;
            }
            case GMG_PH_RADIUS: {           // after the column-group partials have been gathered
                const bool overridden = !s->radius_override.empty() && s->radius_override[size_t(s->level)] > 0.0;
                if (overridden) s->radius = s->radius_override[size_t(s->level)];
                else {
                    double denom = double(s->nnz - s->n);
                    if (s->opt.radius_policy == 1) denom = double(s->level == 0 ? s->nnz - s->n : s->nnz);
                    DevBuf<double> d_mean(c, 1);
                    mean_from_group_partials(c, (int64_t(s->n) + 31) / 32, s->len_partial.p, denom, d_mean.p);
                    double mean = 0.0;
                    d_mean.download(&mean, 1);
                    sync(c);
                    s->radius = std::cbrt(s->opt.ratio) * mean;
                }
                s->cur->info.radius = s->radius;
                break;
            }
            case GMG_PH_SAMPLER_BUILD:
                shard_sampler_build(c, s->n, s->lo, s->hi, s->colptr, s->rowidx, s->len.p, s->radius, s->opt.sampler_mode,
                                    s->opt.seed + uint64_t(s->level), s->state.p, s->sampler);
                break;
            case GMG_PH_SAMPLER_ROUND:
                shard_sampler_round(c, s->sampler, s->state.p);
                break;
            case GMG_PH_SAMPLER_CHECK:      // on the gathered state: identical on every rank
                io[0] = count_state(c, s->n, s->state.p, 0);
                ++s->cur->info.sampler_rounds;
                break;
            case GMG_PH_SAMPLES: {
                s->sampler = ShardSampler();
                s->m = compact_selected(c, s->n, s->state.p, s->samples);
                split(s->m, s->world, s->rank, s->cchunk, s->clo, s->chi);
                s->cur->info.m = s->m;
                shard_parents_init(c, s->n, s->m, s->samples.p, s->dist.p, s->changed[0].p, s->srcid.p);
                s->parity = 0;
                io[0] = s->m;
                break;
            }
            case GMG_PH_PARENTS_PULL:       // reads changed[parity], writes changed[parity ^ 1] and dist on [lo, hi)
                shard_parents_pull(c, s->lo, s->hi, s->colptr, s->rowidx, s->len.p, s->dist.p, s->changed[s->parity].p,
                                   s->changed[s->parity ^ 1].p);
                s->parity ^= 1;
                break;
            case GMG_PH_PARENTS_CHECK:      // on the gathered bytes: identical on every rank
                io[0] = count_state(c, s->n, s->changed[s->parity].p, 1);
                ++s->cur->info.parent_rounds;
                break;
            case GMG_PH_PARENTS_PRED:
                shard_parents_pred(c, s->n, s->lo, s->hi, s->colptr, s->rowidx, s->len.p, s->dist.p, s->srcid.p, s->pred.p,
                                   s->status.dev.p + StageStatus::kZeroEdge);
                break;
            case GMG_PH_PARENTS_RESOLVE:
                shard_parents_resolve(c, s->n, s->lo, s->hi, s->pred.p, s->srcid.p, s->parents.p);
                break;
            case GMG_PH_COARSE_ROWS: {
                build_clusters(c, s->n, s->m, s->parents.p, s->clusters);           // replicated
                if (s->world > 1 && s->val_is_len) {
                    // level 0: the owned coarse rows read the stored values (= lengths) of more columns
                    s->need.zero();
                    GMG_LAUNCH(c, mark_children_kernel, div_up(s->n, kBlock), kBlock, 0, s->n, s->clo, s->chi, s->parents.p, s->need.p);
                    edge_lengths(c, s->n, s->colptr, s->rowidx, s->pts, s->len.p, s->need.p);
                }
                s->ccnt.alloc(c, int64_t(s->cchunk) * s->world + 1);
                s->ccnt.zero();
                s->rows = CoarseRows();
                coarse_rows_range(c, s->n, s->colptr, s->rowidx, s->val, s->clo, s->chi, s->parents.p, s->clusters,
                                  s->ccnt.p, s->rows, s->status.dev.p + StageStatus::kCoarseHardOverflow);
                break;
            }
            case GMG_PH_COARSE_PACK: {      // after the row counts have been gathered
                s->rowoff.alloc(c, int64_t(s->m) + 1);
                DevBuf<int64_t> total(c, 1);
                GMG_CUDA(cudaMemsetAsync(s->ccnt.p + s->m, 0, sizeof(int32_t), c->stream));   // scan sentinel
                exclusive_scan_i32(c, s->ccnt.p, s->rowoff.p, int64_t(s->m) + 1, total.p);
                rank_offsets(s, s->rowoff.p, s->m, s->cchunk, s->rows_off);
                s->rows_are_ac = false;
                s->xQ.alloc(c, s->rows_off.back());
                s->xt.alloc(c, s->rows_off.back());
                const int32_t own = s->chi - s->clo;
                if (own > 0)
                    GMG_LAUNCH(c, pack_rows_kernel, div_up(int64_t(own) * 8, kBlock), kBlock, 0, s->clo, s->chi, s->rows.boff.p,
                               s->ccnt.p, s->rowoff.p, 0, s->rows.sQ.p, s->rows.st.p, s->xQ.p, s->xt.p);
                s->rows = CoarseRows();
                io[0] = s->rows_off.back();
                break;
            }
            case GMG_PH_COARSE_ASSEMBLE: {  // after the packed rows have been gathered
                gmg_shard::Done& L = *s->cur;
                coarse_assemble(c, s->m, s->rowoff.p, s->ccnt.p, s->xQ.p, s->xt.p, L.coarse, s->status.dev.p + StageStatus::kCoarseAsym);
                L.info.nnz_c = L.coarse.nnz;
                s->xQ = DevBuf<int32_t>(); s->xt = DevBuf<double>();
                // replicated cheap stages
                L.coarse_pts.alloc(c, 3 * int64_t(s->m));
                if (s->opt.weighting & GMG_NESTED) sample_points(c, s->n, s->pts, s->m, s->samples.p, L.coarse_pts.p);
                else coarse_points(c, s->n, s->pts, s->colptr, s->rowidx, s->m, s->clusters, L.coarse_pts.p);
                // Triangle fans: only what the owned rows of P read. A fan of c lists every triangle that contains c,
                // and a triangle is enumerated at its smallest vertex, which is c or a neighbour of c.
                DevBuf<uint8_t> need_c(c, int64_t(s->m) + 1), need_v0(c, int64_t(s->m) + 1);
                need_c.zero(); need_v0.zero();
                if (s->hi > s->lo) GMG_LAUNCH(c, mark_parents_kernel, div_up(s->hi - s->lo, kBlock), kBlock, 0, s->lo, s->hi, s->parents.p, need_c.p);
                if (s->m > 0) GMG_LAUNCH(c, mark_closure_kernel, div_up(s->m, kBlock), kBlock, 0, s->m, L.coarse.colptr.p, L.coarse.rowidx.p, need_c.p, need_v0.p);
                voronoi_triangles(c, s->m, L.coarse.nnz, L.coarse_pts.p, L.coarse.colptr.p, L.coarse.rowidx.p, L.coarse.val.p, L.tri,
                                  s->world > 1 ? need_v0.p : nullptr);
                L.info.T = L.tri.T;       // with more than one rank: the triangles THIS rank enumerated
                break;
            }
            case GMG_PH_PROLONG: {          // owned rows of P
                gmg_shard::Done& L = *s->cur;
                prolongation(c, s->n, s->pts, s->m, L.coarse_pts.p, L.coarse.colptr.p, L.coarse.rowidx.p, s->parents.p, L.tri,
                             s->opt.weighting, L.P, nullptr, &s->status, s->lo, s->hi);
                break;
            }
            case GMG_PH_P_COUNT: {          // build_galerkin: entries per owned row of P
                GMG_REQUIRE(s->opt.build_galerkin, GMG_ERR_STATE, "shard was created without build_galerkin");
                gmg_shard::Done& L = *s->cur;
                s->status.fetch(c);
                finish_prolongation(s->status, L.P, &L.info.stats);
                s->pcnt.alloc(c, int64_t(s->chunk) * s->world + 1);
                s->pcnt.zero();
                if (s->hi > s->lo) GMG_LAUNCH(c, row_counts_kernel, div_up(s->hi - s->lo, kBlock), kBlock, 0, s->hi - s->lo, L.P.rowptr.p, s->pcnt.p + s->lo);
                break;
            }
            case GMG_PH_P_PACK: {           // after the row counts have been gathered
                gmg_shard::Done& L = *s->cur;
                s->Pfull = DeviceCsr();
                s->Pfull.rows = s->n; s->Pfull.cols = s->m;
                s->Pfull.rowptr.alloc(c, int64_t(s->n) + 1);
                DevBuf<int64_t> total(c, 1);
                GMG_CUDA(cudaMemsetAsync(s->pcnt.p + s->n, 0, sizeof(int32_t), c->stream));     // scan sentinel
                exclusive_scan_i32(c, s->pcnt.p, s->Pfull.rowptr.p, int64_t(s->n) + 1, total.p);
                rank_offsets(s, s->Pfull.rowptr.p, s->n, s->chunk, s->p_off);
                const int64_t mine = s->p_off[size_t(s->rank) + 1] - s->p_off[size_t(s->rank)];
                GMG_REQUIRE(mine == L.P.nnz, GMG_ERR_STATE, "gathered row counts disagree with the owned rows of P");
                s->Pfull.nnz = s->p_off.back();
                s->Pfull.colidx.alloc(c, s->Pfull.nnz);
                s->Pfull.val.alloc(c, s->Pfull.nnz);
                if (L.P.nnz > 0) {
                    GMG_CUDA(cudaMemcpyAsync(s->Pfull.colidx.p + s->p_off[size_t(s->rank)], L.P.colidx.p, sizeof(int32_t) * size_t(L.P.nnz), cudaMemcpyDeviceToDevice, c->stream));
                    GMG_CUDA(cudaMemcpyAsync(s->Pfull.val.p + s->p_off[size_t(s->rank)], L.P.val.p, sizeof(double) * size_t(L.P.nnz), cudaMemcpyDeviceToDevice, c->stream));
                }
                io[0] = s->Pfull.nnz;
                break;
            }
            case GMG_PH_GALERKIN_ROWS: {    // after the rows of P have been gathered: the owned outer vectors of A_c
                gmg_shard::Done& L = *s->cur;
                Trace tr(c, "galerkin_rows");
                L.clo = s->clo; L.chi = s->chi;
                galerkin(c, s->a_n, s->a_colptr, s->a_rowidx, s->a_val, s->Pfull, s->ac_own, s->clo, s->chi, &L.R);
                tr.mark("transpose + SpGEMM");
                if (s->level >= 1 && s->n <= kKeepWholeP) L.Pfull = std::move(s->Pfull);
                s->Pfull = DeviceCsr();
                // entries per owned outer vector -> the count exchange buffer (the coarse-graph counts are done with)
                s->ccnt.zero();
                if (s->chi > s->clo) GMG_LAUNCH(c, row_counts_kernel, div_up(s->chi - s->clo, kBlock), kBlock, 0, s->chi - s->clo, s->ac_own.colptr.p, s->ccnt.p + s->clo);
                break;
            }
            case GMG_PH_GALERKIN_PACK: {    // after the counts have been gathered
                gmg_shard::Done& L = *s->cur;
                L.Ac.n = s->m;
                L.Ac.colptr.alloc(c, int64_t(s->m) + 1);
                DevBuf<int64_t> total(c, 1);
                GMG_CUDA(cudaMemsetAsync(s->ccnt.p + s->m, 0, sizeof(int32_t), c->stream));
                exclusive_scan_i32(c, s->ccnt.p, L.Ac.colptr.p, int64_t(s->m) + 1, total.p);
                rank_offsets(s, L.Ac.colptr.p, s->m, s->cchunk, s->rows_off);
                s->rows_are_ac = true;
                const int64_t at = s->rows_off[size_t(s->rank)];
                GMG_REQUIRE(s->rows_off[size_t(s->rank) + 1] - at == s->ac_own.nnz, GMG_ERR_STATE, "gathered counts disagree with the owned part of A_c");
                L.Ac.nnz = s->rows_off.back();
                L.Ac.rowidx.alloc(c, L.Ac.nnz);
                L.Ac.val.alloc(c, L.Ac.nnz);
                if (s->ac_own.nnz > 0) {
                    GMG_CUDA(cudaMemcpyAsync(L.Ac.rowidx.p + at, s->ac_own.rowidx.p, sizeof(int32_t) * size_t(s->ac_own.nnz), cudaMemcpyDeviceToDevice, c->stream));
                    GMG_CUDA(cudaMemcpyAsync(L.Ac.val.p + at, s->ac_own.val.p, sizeof(double) * size_t(s->ac_own.nnz), cudaMemcpyDeviceToDevice, c->stream));
                }
                s->ac_own = DeviceCsc();
                io[0] = L.Ac.nnz;
                break;
            }
            case GMG_PH_GALERKIN_ASSEMBLE: {    // after the owned outer vectors have been gathered
                gmg_shard::Done& L = *s->cur;
                s->rows_are_ac = false;
                L.info.nnz_ac = L.Ac.nnz;
                s->a_n = s->m; s->a_colptr = L.Ac.colptr.p; s->a_rowidx = L.Ac.rowidx.p; s->a_val = L.Ac.val.p;
                break;
            }
            case GMG_PH_LEVEL_STATUS: {        // this rank's deferred flags and counters into its slice of STATUS
                gmg_shard::Done& L = *s->cur;
                s->status.fetch(c);
                finish_prolongation(s->status, L.P, &L.info.stats);
                L.info.nnz_p = L.P.nnz;
                int64_t* w = s->status_host;
                w[0] = s->status.host[StageStatus::kZeroEdge];
                w[1] = s->status.host[StageStatus::kNoOffer];
                w[2] = s->status.host[StageStatus::kCoarseHardOverflow];
                w[3] = s->status.host[StageStatus::kCoarseAsym];
                w[4] = s->status.host_i64(StageStatus::kProlongStats + 6);      // degenerate fan
                w[5] = s->status.host_i64(StageStatus::kProlongStats + 4);      // fallback rows
                w[6] = int64_t(s->hi) - int64_t(s->lo);                          // owned rows
                w[7] = 0;
                if (!s->xstatus.p) s->xstatus.alloc(c, 8 * int64_t(s->world));
                GMG_CUDA(cudaMemcpyAsync(s->xstatus.p + 8 * int64_t(s->rank), w, sizeof(int64_t) * 8, cudaMemcpyHostToDevice, c->stream));
                sync(c);
                break;
            }
            case GMG_PH_LEVEL_END: {           // after STATUS has been gathered: the checks on the values of all ranks
                gmg_shard::Done& L = *s->cur;
                GMG_REQUIRE(s->xstatus.p, GMG_ERR_STATE, "LEVEL_END before LEVEL_STATUS");
                std::vector<int64_t> all(size_t(8) * size_t(s->world));
                GMG_CUDA(cudaMemcpyAsync(all.data(), s->xstatus.p, sizeof(int64_t) * all.size(), cudaMemcpyDeviceToHost, c->stream));
                sync(c);
                int64_t g[8] = {0, 0, 0, 0, 0, 0, 0, 0};
                for (int r = 0; r < s->world; ++r) {
                    for (int k = 0; k < 5; ++k) g[k] |= all[size_t(8 * r + k)];
                    g[5] += all[size_t(8 * r + 5)]; g[6] += all[size_t(8 * r + 6)];
                }
                StageStatus global;
                global.host[StageStatus::kZeroEdge] = int32_t(g[0] != 0);
                global.host[StageStatus::kNoOffer] = int32_t(g[1] != 0);
                global.host[StageStatus::kCoarseHardOverflow] = int32_t(g[2] != 0);
                global.host[StageStatus::kCoarseAsym] = int32_t(g[3] != 0);
                memcpy(global.host + StageStatus::kProlongStats + 6, &g[4], sizeof(int64_t));
                memcpy(global.host + StageStatus::kProlongStats + 4, &g[5], sizeof(int64_t));
                global.raise_parents();
                global.raise_coarse();
                global.raise_prolong(g[6]);
                L.clo = s->clo; L.chi = s->chi; L.chunk = s->chunk; L.cchunk = s->cchunk;
                // keep what a consumer of the level needs; the rest goes back to the pool
                L.samples = std::move(s->samples);
                L.parents.alloc(c, s->n);
                GMG_CUDA(cudaMemcpyAsync(L.parents.p, s->parents.p, sizeof(int32_t) * size_t(s->n), cudaMemcpyDeviceToDevice, c->stream));
                s->clusters = Clusters();
                // next level's fine inputs
                s->n = s->m; s->nnz = L.coarse.nnz;
                s->pts = L.coarse_pts.p; s->colptr = L.coarse.colptr.p; s->rowidx = L.coarse.rowidx.p; s->val = L.coarse.val.p;
                io[0] = L.info.m; io[1] = L.info.nnz_c; io[2] = L.info.T; io[3] = L.info.nnz_p; io[4] = L.info.nnz_ac;
                s->done.push_back(std::move(s->cur));
                ++s->level;
                sync(c);
                break;
            }
            default: GMG_REQUIRE(false, GMG_ERR_INVALID, "unknown phase");
        }
    });
}

/* Exchange buffer `which` of the current level: device pointer, element size in bytes, and the slice
 * length in elements (the buffer holds slice * world elements; rank r owns [r*slice, (r+1)*slice)). */
int gmg_shard_buffer(gmg_shard* s, int which, void** ptr, int32_t* elem_bytes, int64_t* slice) {
    if (!s || !ptr || !elem_bytes || !slice) return GMG_ERR_STATE;
    switch (which) {
        case GMG_XB_STATE: *ptr = s->state.p; *elem_bytes = 1; *slice = s->chunk; break;
        case GMG_XB_CHANGED: *ptr = s->changed[s->parity].p; *elem_bytes = 1; *slice = s->chunk; break;
        case GMG_XB_DIST: *ptr = s->dist.p; *elem_bytes = 8; *slice = s->chunk; break;
        case GMG_XB_PRED: *ptr = s->pred.p; *elem_bytes = 4; *slice = s->chunk; break;
        case GMG_XB_PARENTS: *ptr = s->parents.p; *elem_bytes = 4; *slice = s->chunk; break;
        case GMG_XB_COARSE_COUNT: *ptr = s->ccnt.p; *elem_bytes = 4; *slice = s->cchunk; break;
        case GMG_XB_ROWS_ID: case GMG_XB_ROWS_VAL: case GMG_XB_P_COL: case GMG_XB_P_VAL:
            return GMG_ERR_INVALID;      /* variable-size slices: gmg_shard_buffer_v */
        case GMG_XB_LEN_PARTIAL: *ptr = s->len_partial.p; *elem_bytes = 8; *slice = s->gchunk; break;
        case GMG_XB_P_COUNT: *ptr = s->pcnt.p; *elem_bytes = 4; *slice = s->chunk; break;
        case GMG_XB_STATUS: *ptr = s->xstatus.p; *elem_bytes = 8; *slice = 8; break;
        default: return GMG_ERR_INVALID;
    }
    return GMG_OK;
}

/* Exchange buffers whose slices differ in size (ROWS_ID, ROWS_VAL, P_COL, P_VAL): rank r owns elements
 * [offsets[r], offsets[r + 1]) of the buffer; `offsets` receives world + 1 entries. */
int gmg_shard_buffer_v(gmg_shard* s, int which, void** ptr, int32_t* elem_bytes, int64_t* offsets) {
    if (!s || !ptr || !elem_bytes || !offsets) return GMG_ERR_STATE;
    const std::vector<int64_t>* off = nullptr;
    switch (which) {
        case GMG_XB_ROWS_ID:
            *ptr = s->rows_are_ac ? (void*)s->cur->Ac.rowidx.p : (void*)s->xQ.p; *elem_bytes = 4; off = &s->rows_off; break;
        case GMG_XB_ROWS_VAL:
            *ptr = s->rows_are_ac ? (void*)s->cur->Ac.val.p : (void*)s->xt.p; *elem_bytes = 8; off = &s->rows_off; break;
        case GMG_XB_P_COL: *ptr = s->Pfull.colidx.p; *elem_bytes = 4; off = &s->p_off; break;
        case GMG_XB_P_VAL: *ptr = s->Pfull.val.p; *elem_bytes = 8; off = &s->p_off; break;
        default: return GMG_ERR_INVALID;
    }
    if (off->size() != size_t(s->world) + 1) return GMG_ERR_STATE;
    for (int r = 0; r <= s->world; ++r) offsets[r] = (*off)[size_t(r)];
    return GMG_OK;
}

int gmg_shard_num_levels(const gmg_shard* s) { return s ? int(s->done.size()) : 0; }

int gmg_shard_level_info(const gmg_shard* s, int level, gmg_level_info* out, int64_t* row_lo, int64_t* row_hi) {
    if (!s || !out || level < 0 || level >= int(s->done.size())) return GMG_ERR_STATE;
    *out = s->done[size_t(level)]->info;
    if (row_lo) *row_lo = s->done[size_t(level)]->lo;
    if (row_hi) *row_hi = s->done[size_t(level)]->hi;
    return GMG_OK;
}

/* Same array ids as gmg_hierarchy_fetch. Replicated arrays are complete on every rank; the three P
 * arrays hold the owned rows [row_lo, row_hi) only (rowptr has row_hi - row_lo + 1 entries, from 0). */
int gmg_shard_fetch(gmg_ctx* c, const gmg_shard* s, int level, int which, void* out) {
    return guarded_s(c, [&] {
        GMG_REQUIRE(s && level >= 0 && level < int(s->done.size()), GMG_ERR_STATE, "no such level");
        const gmg_shard::Done& L = *s->done[size_t(level)];
        const int64_t n = L.info.n, m = L.info.m, T = L.info.T, own = L.hi - L.lo;
        auto copy = [&](const void* dev, size_t bytes) {
            if (bytes) GMG_CUDA(cudaMemcpyAsync(out, dev, bytes, cudaMemcpyDeviceToHost, c->stream));
        };
        auto copy_wide = [&](const int32_t* dev, int64_t count) {
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
            case GMG_H_P_ROWPTR: copy(L.P.rowptr.p, sizeof(int32_t) * size_t(own + 1)); break;
            case GMG_H_P_COLIDX: copy(L.P.colidx.p, sizeof(int32_t) * size_t(L.P.nnz)); break;
            case GMG_H_P_VAL: copy(L.P.val.p, sizeof(double) * size_t(L.P.nnz)); break;
            case GMG_H_AC_COLPTR: case GMG_H_AC_ROWIDX: case GMG_H_AC_VAL:
                GMG_REQUIRE(L.Ac.colptr.p, GMG_ERR_STATE, "shard was built without the Galerkin operator");
                if (which == GMG_H_AC_COLPTR) copy(L.Ac.colptr.p, sizeof(int32_t) * size_t(m + 1));
                else if (which == GMG_H_AC_ROWIDX) copy(L.Ac.rowidx.p, sizeof(int32_t) * size_t(L.Ac.nnz));
                else copy(L.Ac.val.p, sizeof(double) * size_t(L.Ac.nnz));
                break;
            default: GMG_REQUIRE(false, GMG_ERR_INVALID, "array not available from a sharded build");
        }
        sync(c);
    });
}

}  // extern "C"
