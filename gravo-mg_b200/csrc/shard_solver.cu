// Row-sharded smoothing and V-cycle on a row-sharded hierarchy [not in reference; SURVEY.md 8(e)(5)].
//
// Every rank keeps full-length vectors per level, but only its owned entries [lo, hi) and the GHOST
// entries its rows read are ever valid. Operators act on owned rows only:
//     Gauss-Seidel   one colour at a time over the owned rows of that colour (the colouring is computed
//                    from the replicated operator, so it is the same on every rank and the sweep equals
//                    the single-GPU sweep in (colour, row) order);
//     residual       r = b - A x on owned rows;
//     restriction    b_c = R r on the owned coarse rows (R = the owned rows of P^T, kept by the build);
//     prolongation   x += P e on the owned fine rows;
//     small levels   from the first level below a size threshold down (always at least the coarsest) every rank
//                    gathers the whole right-hand side and runs the rest of the cycle redundantly with the
//                    single-GPU kernels (the build keeps P whole there): finishing a small level on every GPU
//                    is cheaper than exchanging its halos colour by colour.
// Between them the HOST exchanges ghost entries (gravo-mg_b200/sharded.py: packed all-to-all over
// NCCL / NVLink; a loop-back between virtual ranks in the single-GPU tests). This file exposes what the
// host needs for that: the byte mask of the entries an operator's owned rows read outside the owned
// range (the halo pattern), the colour array, and the vectors.
#include "shard_types.h"

#include <cmath>

namespace {

struct SLevel {
    int32_t n = 0, lo = 0, hi = 0, chunk = 0;
    const int32_t* outer = nullptr;   // replicated operator of this level
    const int32_t* inner = nullptr;
    const double* val = nullptr;
    Colouring col;
    std::vector<int32_t> own_lo, own_hi;   // per colour: the owned part of the colour's bucket (positions in col.rows)
    const DeviceCsr* R = nullptr;     // owned rows of P^T towards the next level (local outer index)
    const DeviceCsr* P = nullptr;     // owned rows of P from the next level (local outer index)
    bool replicated = false;          // this level and every coarser one run whole on every rank
    const DeviceCsr* Pfull = nullptr; // replicated, not the coarsest: all rows of P from the next level
    DeviceCsr Rfull;                  // ... and its transpose
    DevBuf<double> x, b, r;
    DevBuf<uint8_t> mask;
};

}  // namespace

struct gmg_shard_solver {
    gmg_ctx* ctx = nullptr;
    int rank = 0, world = 1;
    std::vector<std::unique_ptr<SLevel>> levels;
};

namespace {

template <typename F> int guarded_v(gmg_ctx* c, F&& body) {
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

// rows of one colour are ascending, so the owned ones are a contiguous part of the bucket
__global__ void bucket_bounds_kernel(int nc, const int32_t* __restrict__ rows, const int32_t* __restrict__ offsets, int32_t lo,
                                     int32_t hi, int32_t* __restrict__ out) {
    const int k = threadIdx.x;
    if (k >= nc) return;
    out[2 * k] = lower_bound_i32(rows, offsets[k], offsets[k + 1], lo);
    out[2 * k + 1] = lower_bound_i32(rows, offsets[k], offsets[k + 1], hi);
}

// mask[j] = 1 for every inner index j outside [own_lo, own_hi) that outer vectors [0, count) store
__global__ void __launch_bounds__(kBlock) halo_mask_kernel(int32_t count, const int32_t* __restrict__ outer,
                                                           const int32_t* __restrict__ inner, int32_t own_lo, int32_t own_hi,
                                                           uint8_t* __restrict__ mask) {
    const int64_t i = int64_t(blockIdx.x) * kBlock + threadIdx.x;
    if (i >= count) return;
    for (int32_t k = outer[i]; k < outer[i + 1]; ++k) {
        const int32_t j = inner[k];
        if (j < own_lo || j >= own_hi) mask[j] = 1;
    }
}

// the rest of a V-cycle from a replicated level down: the single-GPU cycle (csrc/solver.cu), on all rows
void vcycle_all(gmg_ctx* c, gmg_shard_solver* s, size_t l, int pre, int post, int coarse_sweeps) {
    SLevel& L = *s->levels[l];
    if (l + 1 == s->levels.size()) {
        gauss_seidel(c, L.col, L.outer, L.inner, L.val, L.b.p, L.x.p, coarse_sweeps, false);
        return;
    }
    SLevel& C = *s->levels[l + 1];
    gauss_seidel(c, L.col, L.outer, L.inner, L.val, L.b.p, L.x.p, pre, false);
    spmv(c, L.n, L.outer, L.inner, L.val, L.x.p, L.r.p, -1.0, L.b.p);
    spmv(c, C.n, L.Rfull.rowptr.p, L.Rfull.colidx.p, L.Rfull.val.p, L.r.p, C.b.p, 1.0, nullptr);
    C.x.zero();
    vcycle_all(c, s, l + 1, pre, post, coarse_sweeps);
    spmv(c, L.n, L.Pfull->rowptr.p, L.Pfull->colidx.p, L.Pfull->val.p, C.x.p, L.x.p, 1.0, L.x.p);
    gauss_seidel(c, L.col, L.outer, L.inner, L.val, L.b.p, L.x.p, post, true);
}

}  // namespace

extern "C" {

/* replicate_below: levels >= 1 with at most this many vertices (and every coarser one) run whole on every
 * rank; the coarsest level always does. */
int gmg_shard_solver_create(gmg_ctx* c, const gmg_shard* sh, const gmg_input* in, int64_t replicate_below,
                            gmg_shard_solver** out) {
    if (!out) return GMG_ERR_INVALID;
    *out = nullptr;
    return guarded_v(c, [&] {
        GMG_REQUIRE(sh && in && in->has_a, GMG_ERR_INVALID, "solver needs a shard and an input with A values");
        await_values(c, in);
        GMG_REQUIRE(!sh->done.empty() && sh->done[0]->Ac.colptr.p, GMG_ERR_STATE, "shard was built without the Galerkin operator");
        std::unique_ptr<gmg_shard_solver> s(new gmg_shard_solver());
        s->ctx = c; s->rank = sh->rank; s->world = sh->world;
        const size_t nl = sh->done.size();
        for (size_t l = 0; l <= nl; ++l) {
            std::unique_ptr<SLevel> L(new SLevel());
            if (l == 0) { L->n = in->n; L->outer = in->colptr.p; L->inner = in->rowidx.p; L->val = in->a_val.p; }
            else {
                const auto& prev = *sh->done[l - 1];
                L->n = prev.Ac.n; L->outer = prev.Ac.colptr.p; L->inner = prev.Ac.rowidx.p; L->val = prev.Ac.val.p;
            }
            if (l < nl) { const auto& d = *sh->done[l]; L->lo = d.lo; L->hi = d.hi; L->chunk = d.chunk; L->R = &d.R; L->P = &d.P; }
            else { const auto& d = *sh->done[nl - 1]; L->lo = d.clo; L->hi = d.chi; L->chunk = d.cchunk; }
            colour_graph(c, L->n, L->outer, L->inner, L->col);
            const int nc = L->col.num_colours;
            GMG_REQUIRE(nc <= 1024, GMG_ERR_INVALID, "more than 1024 colours");
            L->own_lo.assign(size_t(nc), 0); L->own_hi.assign(size_t(nc), 0);
            if (nc > 0) {
                auto d_off = to_device(c, L->col.offsets.data(), int64_t(nc) + 1);
                DevBuf<int32_t> d_b(c, 2 * int64_t(nc));
                GMG_LAUNCH(c, bucket_bounds_kernel, 1, 1024, 0, nc, L->col.rows.p, d_off.p, L->lo, L->hi, d_b.p);
                std::vector<int32_t> hb(size_t(2 * nc));
                d_b.download(hb.data(), 2 * int64_t(nc));
                sync(c);
                for (int k = 0; k < nc; ++k) { L->own_lo[size_t(k)] = hb[size_t(2 * k)]; L->own_hi[size_t(k)] = hb[size_t(2 * k + 1)]; }
            }
            const int64_t n_pad = std::max<int64_t>(int64_t(L->chunk) * s->world, L->n);
            L->x.alloc(c, n_pad); L->b.alloc(c, n_pad); L->r.alloc(c, n_pad);
            L->x.zero(); L->b.zero(); L->r.zero();
            L->mask.alloc(c, n_pad);
            if (l == nl) L->replicated = true;
            else if (l >= 1 && L->n <= replicate_below && sh->done[l]->Pfull.rowptr.p) {
                L->replicated = true;
                L->Pfull = &sh->done[l]->Pfull;
                transpose_csr(c, *L->Pfull, L->Rfull);
            }
            s->levels.push_back(std::move(L));
        }
        sync(c);
        *out = s.release();
    });
}

void gmg_shard_solver_destroy(gmg_shard_solver* s) {
    if (!s) return;
    cudaSetDevice(s->ctx->device);
    delete s;
}

int gmg_shard_solver_num_levels(const gmg_shard_solver* s) { return s ? int(s->levels.size()) : 0; }

/* io = {n, lo, hi, chunk, num_colours} */
int gmg_shard_solver_level(const gmg_shard_solver* s, int level, int64_t* io) {
    if (!s || !io || level < 0 || level >= int(s->levels.size())) return GMG_ERR_STATE;
    const SLevel& L = *s->levels[size_t(level)];
    io[0] = L.n; io[1] = L.lo; io[2] = L.hi; io[3] = L.chunk; io[4] = L.col.num_colours; io[5] = L.replicated ? 1 : 0;
    return GMG_OK;
}

int gmg_shard_solver_vector(gmg_shard_solver* s, int level, int which, void** ptr) {
    if (!s || !ptr || level < 0 || level >= int(s->levels.size())) return GMG_ERR_STATE;
    SLevel& L = *s->levels[size_t(level)];
    switch (which) {
        case GMG_SV_X: *ptr = L.x.p; break;
        case GMG_SV_B: *ptr = L.b.p; break;
        case GMG_SV_R: *ptr = L.r.p; break;
        case GMG_SV_COLOUR: *ptr = L.col.colour.p; break;
        case GMG_SV_MASK: *ptr = L.mask.p; break;
        default: return GMG_ERR_INVALID;
    }
    return GMG_OK;
}

/* Fills the level's MASK vector (uint8, one byte per vertex of `*space` vertices) with the halo pattern:
 * GMG_HP_A  entries of x (this level) the owned rows of A read outside the owned range;
 * GMG_HP_R  entries of r (this level) the owned rows of the restriction to level+1 read;
 * GMG_HP_P  entries of x (level + 1) the owned rows of the prolongation read; the mask is level+1's. */
int gmg_shard_solver_halo_mask(gmg_ctx* c, gmg_shard_solver* s, int level, int pattern, int64_t* space) {
    return guarded_v(c, [&] {
        GMG_REQUIRE(s && space && level >= 0 && level < int(s->levels.size()), GMG_ERR_STATE, "no such level");
        SLevel& L = *s->levels[size_t(level)];
        const bool last = size_t(level) + 1 == s->levels.size();
        GMG_REQUIRE(pattern == GMG_HP_A || !last, GMG_ERR_INVALID, "the coarsest level has no transfer operators");
        if (pattern == GMG_HP_A) {
            L.mask.zero();
            if (L.hi > L.lo) GMG_LAUNCH(c, halo_mask_kernel, div_up(L.hi - L.lo, kBlock), kBlock, 0, L.hi - L.lo, L.outer + L.lo, L.inner, L.lo, L.hi, L.mask.p);
            *space = L.n;
        } else if (pattern == GMG_HP_R) {
            L.mask.zero();
            if (L.R->rows > 0) GMG_LAUNCH(c, halo_mask_kernel, div_up(L.R->rows, kBlock), kBlock, 0, L.R->rows, L.R->rowptr.p, L.R->colidx.p, L.lo, L.hi, L.mask.p);
            *space = L.n;
        } else if (pattern == GMG_HP_P) {
            SLevel& C = *s->levels[size_t(level) + 1];
            C.mask.zero();
            if (L.P->rows > 0) GMG_LAUNCH(c, halo_mask_kernel, div_up(L.P->rows, kBlock), kBlock, 0, L.P->rows, L.P->rowptr.p, L.P->colidx.p, C.lo, C.hi, C.mask.p);
            *space = C.n;
        } else GMG_REQUIRE(false, GMG_ERR_INVALID, "unknown halo pattern");
    });
}

/* One operator application on the owned rows (see the header for the list). `out` receives the sum of
 * squares for GMG_SO_SUMSQ_R. */
int gmg_shard_solver_op(gmg_ctx* c, gmg_shard_solver* s, int level, int op, int arg, double* out) {
    return guarded_v(c, [&] {
        GMG_REQUIRE(s && level >= 0 && level < int(s->levels.size()), GMG_ERR_STATE, "no such level");
        SLevel& L = *s->levels[size_t(level)];
        const bool last = size_t(level) + 1 == s->levels.size();
        const int32_t own = L.hi - L.lo;
        switch (op) {
            case GMG_SO_GS_COLOUR: {
                GMG_REQUIRE(arg >= 0 && arg < L.col.num_colours, GMG_ERR_INVALID, "no such colour");
                const int32_t b0 = L.own_lo[size_t(arg)], b1 = L.own_hi[size_t(arg)];   // the owned rows of colour `arg`
                gauss_seidel_rows(c, L.col.rows.p + b0, b1 - b0, L.outer, L.inner, L.val, L.b.p, L.x.p);
                break;
            }
            case GMG_SO_RESIDUAL:
                spmv(c, own, L.outer + L.lo, L.inner, L.val, L.x.p, L.r.p + L.lo, -1.0, L.b.p + L.lo);
                break;
            case GMG_SO_RESTRICT: {
                GMG_REQUIRE(!last, GMG_ERR_INVALID, "the coarsest level has no transfer operators");
                SLevel& C = *s->levels[size_t(level) + 1];
                spmv(c, L.R->rows, L.R->rowptr.p, L.R->colidx.p, L.R->val.p, L.r.p, C.b.p + C.lo, 1.0, nullptr);
                break;
            }
            case GMG_SO_PROLONG: {
                GMG_REQUIRE(!last, GMG_ERR_INVALID, "the coarsest level has no transfer operators");
                SLevel& C = *s->levels[size_t(level) + 1];
                spmv(c, L.P->rows, L.P->rowptr.p, L.P->colidx.p, L.P->val.p, C.x.p, L.x.p + L.lo, 1.0, L.x.p + L.lo);
                break;
            }
            case GMG_SO_ZERO_X: L.x.zero(); break;
            case GMG_SO_GS_ALL:       // `arg` sweeps over ALL rows: a small level every rank solves redundantly
                GMG_REQUIRE(arg >= 0, GMG_ERR_INVALID, "negative sweep count");
                gauss_seidel(c, L.col, L.outer, L.inner, L.val, L.b.p, L.x.p, arg, false);
                break;
            case GMG_SO_VCYCLE_ALL:   // arg = pre | post << 8 | coarse_sweeps << 16: the rest of the cycle, on all rows
                GMG_REQUIRE(L.replicated && arg >= 0, GMG_ERR_INVALID, "level is not replicated");
                vcycle_all(c, s, size_t(level), arg & 0xff, (arg >> 8) & 0xff, arg >> 16);
                break;
            case GMG_SO_SUMSQ_R: {
                GMG_REQUIRE(out, GMG_ERR_INVALID, "null output");
                const double nr = norm2(c, own, L.r.p + L.lo);
                *out = nr * nr;
                break;
            }
            default: GMG_REQUIRE(false, GMG_ERR_INVALID, "unknown solver op");
        }
    });
}

}  // extern "C"
